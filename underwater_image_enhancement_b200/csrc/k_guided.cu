// Dehaze core: dark-channel transmission, guided filter, scene recovery.
//   estimate_transmission / guided_filter / restore_image   six_stadigy.py:25-46, 167-188
//   estimate_transmission / guided_filter / recover_image   enhancement_strategies.py:16-46, 209-249
// Box means are cv2.boxFilter(CV_64F, (r, r)): normalised, anchor r//2 (window [-(r//2), r-1-r//2]),
// BORDER_REFLECT_101, scale applied as a multiplication by 1/(r*r).
//
// Three kernels:
//   gf_prep_kernel      frame bytes -> packed (R, G, B, gray) word plane and dark-channel plane (float32), per frame
//   gf_ab_kernel        gray, dark -> box sums of I, p, I*p, I*I -> a, b (float64 planes)
//   gf_recover_kernel   a, b -> mean_a * I + mean_b -> t -> (img - A)/t + A -> recovered frame
//                       (+ the first level of the percentile histograms, see k_select.cu)
//
// Box sums are O(1) per pixel.  Every WARP marches down its own strip of 160 columns with no
// block-level synchronisation: a lane keeps the vertical running sums of 5 columns in registers
// (row entering the window added, row leaving it subtracted - both re-read, so there is no ring
// buffer), publishes them to a per-warp shared-memory row, and then forms the horizontal window
// sums of 5 CONSECUTIVE outputs with a sliding window (r + 4 reads for 5 outputs).
// The first-pass sums are exact whatever the order: I = g/255 is carried as the integers (sum g,
// sum g*g), and p (a float32, multiple of 2^-27) and g*p are summed in float64 far below 2^53 ulps.
#include "k_arith.cuh"
#include "k_kernels.h"

namespace uwie {

namespace {

constexpr int GC = 5;                 // columns per lane (odd: conflict-free 64-bit shared accesses)
constexpr int GSPAN = 32 * GC;        // columns a warp carries, outputs = GSPAN - r + 1
constexpr int GWARPS = 8;
constexpr int GNT = 32 * GWARPS;

// ------------------------------------------------------------------ gray + dark channel planes
__global__ void __launch_bounds__(256) gf_prep_kernel(GuidedArgs a) {
    const int b = blockIdx.y;
    __shared__ float sdiv[3][256];      // img / (A + guard), float32            SS:170, ES:221
    __shared__ uint8_t sq[3][256];      // byte cv2 sees for the (colour-corrected) frame
    const CastTables& ct = a.cast[b];
    for (int i = threadIdx.x; i < 768; i += 256) {
        const int c = i >> 8, v = i & 255;
        const float lg = __fadd_rn(a.light[b * 3 + c], (float)a.lib.dark_guard);
        sdiv[c][v] = __fdiv_rn(ct.curf[c][v], lg);
        sq[c][v] = ct.curq[c][v];
    }
    __syncthreads();
    const size_t npix = (size_t)a.H * a.W;
    const uint8_t* f = a.frames + (size_t)b * npix * 3;
    unsigned* gp = a.pix_plane + (size_t)b * npix;
    float* dp = a.dark_plane + (size_t)b * npix;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < npix; i += (size_t)gridDim.x * 256) {
        const unsigned r = f[i * 3], g = f[i * 3 + 1], bl = f[i * 3 + 2];
        const unsigned gq = (unsigned)gray_q15(sq[0][r], sq[1][g], sq[2][bl]);
        gp[i] = r | (g << 8) | (bl << 16) | (gq << 24);      // one aligned word per pixel for the marching kernels
        dp[i] = fminf(fminf(sdiv[0][r], sdiv[1][g]), sdiv[2][bl]);
    }
}

__device__ __forceinline__ float transmission0(float omega, int clip_t0, float dark) {
    float t0 = __fsub_rn(1.0f, __fmul_rn(omega, dark));           // 1 - omega * dark, float32
    if (clip_t0) t0 = t0 < 0.1f ? 0.1f : (t0 > 1.0f ? 1.0f : t0); // SS:174
    return t0;
}

struct Strip {
    int r, lo, hi, outw, xs, ys, ye, warp, lane;
    bool valid;
};
__device__ __forceinline__ Strip make_strip(const GuidedArgs& a, int r) {
    Strip s;
    s.r = r; s.lo = r / 2; s.hi = r - 1 - s.lo;
    s.outw = GSPAN - r + 1;
    s.warp = threadIdx.x >> 5; s.lane = threadIdx.x & 31;
    s.xs = (blockIdx.x * GWARPS + s.warp) * s.outw;
    s.ys = blockIdx.y * a.rows_per_block;
    s.ye = min(s.ys + a.rows_per_block, a.H);
    s.valid = s.xs < a.W && s.ys < s.ye;
    return s;
}

// ------------------------------------------------------------------ a, b
__global__ void __launch_bounds__(GNT, 2) gf_ab_kernel(GuidedArgs a) {
    __shared__ double sP[GWARPS][GSPAN];
    __shared__ double sGP[GWARPS][GSPAN];
    __shared__ unsigned long long sG[GWARPS][GSPAN];     // (sum g) << 32 | sum g*g
    const int inst = blockIdx.z;                         // frame * n_dehaze + k
    const int b = inst / a.n_dehaze, k = inst % a.n_dehaze;
    const Recipe& rc = a.recipes[a.dehaze_idx[k]];
    const Strip s = make_strip(a, rc.radius);
    if (!s.valid) return;                                // warp-uniform
    const float omega = rc.omega;
    const int clip_t0 = a.lib.clip_t0;
    const double eps = rc.eps;
    const size_t npix = (size_t)a.H * a.W;
    const unsigned* pix = a.pix_plane + (size_t)b * npix;
    const float* dark = a.dark_plane + (size_t)b * npix;
    double* pa = a.ab + (size_t)inst * 2 * npix;
    double* pb = pa + npix;
    const double rr = (double)s.r * (double)s.r;
    const double scale = 1.0 / rr;                       // cv2: sum * (1 / (kw * kh))
    const double scale_i = 1.0 / (255.0 * rr), scale_ii = 1.0 / (65025.0 * rr);
    double* wP = sP[s.warp]; double* wGP = sGP[s.warp]; unsigned long long* wG = sG[s.warp];

    int cx[GC];
#pragma unroll
    for (int j = 0; j < GC; ++j) cx[j] = reflect101(s.xs - s.lo + s.lane + 32 * j, a.W);
    double vp[GC], vgp[GC];
    unsigned long long vg[GC];
#pragma unroll
    for (int j = 0; j < GC; ++j) { vp[j] = 0.0; vgp[j] = 0.0; vg[j] = 0ull; }

    // Row loads are software-pipelined: the values of the row that enters (leaves) the window in the
    // NEXT step are requested right after this step's were consumed, so their latency hides behind
    // the shared-memory phase instead of stalling a warp that has nothing else to do.
    unsigned gE[GC], gL[GC];
    float dE[GC], dL[GC];
    auto row_load = [&](int yy, unsigned* g, float* d) {
        const size_t ro = (size_t)reflect101(yy, a.H) * a.W;
#pragma unroll
        for (int j = 0; j < GC; ++j) { g[j] = pix[ro + cx[j]] >> 24; d[j] = dark[ro + cx[j]]; }
    };
    auto row_acc = [&](const unsigned* gq, const float* d, bool add) {
#pragma unroll
        for (int j = 0; j < GC; ++j) {
            const unsigned g = gq[j];
            const double p = (double)transmission0(omega, clip_t0, d[j]);
            const unsigned long long q = ((unsigned long long)g << 32) | (unsigned long long)(g * g);
            if (add) { vp[j] += p; vgp[j] += (double)g * p; vg[j] += q; }
            else { vp[j] -= p; vgp[j] -= (double)g * p; vg[j] -= q; }
        }
    };
    auto row_add = [&](int yy) { row_load(yy, gE, dE); row_acc(gE, dE, true); };

    for (int yy = s.ys - s.lo; yy < s.ys + s.hi; ++yy) row_add(yy);
    const int base = s.lane * GC;
    row_load(s.ys + s.hi, gE, dE);
    row_load(s.ys - s.lo, gL, dL);
    for (int yo = s.ys; yo < s.ye; ++yo) {
        row_acc(gE, dE, true);
        if (yo + 1 < s.ye) row_load(yo + 1 + s.hi, gE, dE);
#pragma unroll
        for (int j = 0; j < GC; ++j) {
            wP[s.lane + 32 * j] = vp[j]; wGP[s.lane + 32 * j] = vgp[j]; wG[s.lane + 32 * j] = vg[j];
        }
        __syncwarp();
        double ra[GC], rb[GC];
        if (base < s.outw) {
            double hp = 0.0, hgp = 0.0;
            unsigned long long hg = 0ull;
            for (int i = 0; i < s.r; ++i) { hp += wP[base + i]; hgp += wGP[base + i]; hg += wG[base + i]; }
#pragma unroll
            for (int j = 0; j < GC; ++j) {
                if (j > 0 && base + j < s.outw) {
                    hp += wP[base + j + s.r - 1] - wP[base + j - 1];
                    hgp += wGP[base + j + s.r - 1] - wGP[base + j - 1];
                    hg += wG[base + j + s.r - 1] - wG[base + j - 1];
                }
                const double mI = (double)(unsigned)(hg >> 32) * scale_i;
                const double mII = (double)(unsigned)(hg & 0xffffffffull) * scale_ii;
                const double mP = hp * scale;
                const double mIP = hgp * scale_i;
                const double cov = __dsub_rn(mIP, __dmul_rn(mI, mP));
                const double var = __dsub_rn(mII, __dmul_rn(mI, mI));
                ra[j] = __ddiv_rn(cov, __dadd_rn(var, eps));
                rb[j] = __dsub_rn(mP, __dmul_rn(ra[j], mI));
            }
        }
        __syncwarp();
        if (base < s.outw) {
#pragma unroll
            for (int j = 0; j < GC; ++j) { wP[base + j] = ra[j]; wGP[base + j] = rb[j]; }
        }
        __syncwarp();
        const size_t ro = (size_t)yo * a.W;
        for (int i = s.lane; i < s.outw; i += 32) {
            const int x = s.xs + i;
            if (x < a.W) { pa[ro + x] = wP[i]; pb[ro + x] = wGP[i]; }
        }
        __syncwarp();
        row_acc(gL, dL, false);
        if (yo + 1 < s.ye) row_load(yo + 1 - s.lo, gL, dL);
    }
}

// ------------------------------------------------------------------ recover
template <class T>
__global__ void __launch_bounds__(GNT, 2) gf_recover_kernel(GuidedArgs a) {
    __shared__ double sA[GWARPS][GSPAN];
    __shared__ double sB[GWARPS][GSPAN];
    __shared__ double sI[256];                 // g / 255.0
    __shared__ float sdiff[3][256];            // img - A, float32              SS:185, ES:247
    __shared__ unsigned shist[3 * kSelBinsA];
    const int inst = blockIdx.z;
    const int b = inst / a.n_dehaze, k = inst % a.n_dehaze;
    const Recipe& rc = a.recipes[a.dehaze_idx[k]];
    const float l0 = a.light[b * 3], l1 = a.light[b * 3 + 1], l2 = a.light[b * 3 + 2];
    {
        const CastTables& ct = a.cast[b];
        for (int i = threadIdx.x; i < 3 * kSelBinsA; i += GNT) shist[i] = 0;
        for (int i = threadIdx.x; i < 256; i += GNT) sI[i] = (double)i / 255.0;
        for (int i = threadIdx.x; i < 768; i += GNT) {
            const int c = i >> 8, v = i & 255;
            sdiff[c][v] = __fsub_rn(ct.curf[c][v], c == 0 ? l0 : (c == 1 ? l1 : l2));
        }
    }
    __syncthreads();
    const Strip s = make_strip(a, rc.radius);
    if (s.valid) {
        const size_t npix = (size_t)a.H * a.W;
        const unsigned* pix = a.pix_plane + (size_t)b * npix;
        const double* pa = a.ab + (size_t)inst * 2 * npix;
        const double* pb = pa + npix;
        T* out = reinterpret_cast<T*>(a.recovered) + (size_t)inst * npix * 3;
        const double scale = 1.0 / ((double)s.r * (double)s.r);
        double* wA = sA[s.warp]; double* wB = sB[s.warp];
        const double dl0 = (double)l0, dl1 = (double)l1, dl2 = (double)l2;

        int cx[GC];
#pragma unroll
        for (int j = 0; j < GC; ++j) cx[j] = reflect101(s.xs - s.lo + s.lane + 32 * j, a.W);
        double va[GC], vb[GC];
#pragma unroll
        for (int j = 0; j < GC; ++j) { va[j] = 0.0; vb[j] = 0.0; }
        double aE[GC], bE[GC], aL[GC], bL[GC];           // software-pipelined row loads (see gf_ab_kernel)
        auto row_load = [&](int yy, double* qa, double* qb) {
            const size_t ro = (size_t)reflect101(yy, a.H) * a.W;
#pragma unroll
            for (int j = 0; j < GC; ++j) { qa[j] = pa[ro + cx[j]]; qb[j] = pb[ro + cx[j]]; }
        };
        auto row_add = [&](int yy) {
            row_load(yy, aE, bE);
#pragma unroll
            for (int j = 0; j < GC; ++j) { va[j] += aE[j]; vb[j] += bE[j]; }
        };
        for (int yy = s.ys - s.lo; yy < s.ys + s.hi; ++yy) row_add(yy);
        const int base = s.lane * GC;
        row_load(s.ys + s.hi, aE, bE);
        row_load(s.ys - s.lo, aL, bL);
        for (int yo = s.ys; yo < s.ye; ++yo) {
#pragma unroll
            for (int j = 0; j < GC; ++j) { va[j] += aE[j]; vb[j] += bE[j]; }
            if (yo + 1 < s.ye) row_load(yo + 1 + s.hi, aE, bE);
            // this row's pixels (frame bytes + gray) for the recovery step below, requested early
            const size_t ro = (size_t)yo * a.W;
            unsigned pixw[GC];
#pragma unroll
            for (int j = 0; j < GC; ++j) {
                const int x = s.xs + s.lane + 32 * j;
                pixw[j] = (s.lane + 32 * j < s.outw && x < a.W) ? pix[ro + x] : 0u;
            }
#pragma unroll
            for (int j = 0; j < GC; ++j) { wA[s.lane + 32 * j] = va[j]; wB[s.lane + 32 * j] = vb[j]; }
            __syncwarp();
            double ma[GC], mb[GC];
            if (base < s.outw) {
                double ha = 0.0, hb = 0.0;
                for (int i = 0; i < s.r; ++i) { ha += wA[base + i]; hb += wB[base + i]; }
#pragma unroll
                for (int j = 0; j < GC; ++j) {
                    if (j > 0 && base + j < s.outw) {
                        ha += wA[base + j + s.r - 1] - wA[base + j - 1];
                        hb += wB[base + j + s.r - 1] - wB[base + j - 1];
                    }
                    ma[j] = ha * scale; mb[j] = hb * scale;
                }
            }
            __syncwarp();
            if (base < s.outw) {
#pragma unroll
                for (int j = 0; j < GC; ++j) { wA[base + j] = ma[j]; wB[base + j] = mb[j]; }
            }
            __syncwarp();
#pragma unroll
            for (int jj = 0; jj < GC; ++jj) {
                const int i = s.lane + 32 * jj;
                const int x = s.xs + i;
                if (i >= s.outw || x >= a.W) continue;
                const int r0 = pixw[jj] & 255, g0 = (pixw[jj] >> 8) & 255, b0 = (pixw[jj] >> 16) & 255;
                double t = __dadd_rn(__dmul_rn(wA[i], sI[pixw[jj] >> 24]), wB[i]);
                t = t < 0.1 ? 0.1 : (t > 1.0 ? 1.0 : t);
                // (img - A) in float32, / t and + A in float64, stored in the library's float type
                double q0, q1, q2;
                if (sizeof(T) == 8) {
                    q0 = __ddiv_rn((double)sdiff[0][r0], t);
                    q1 = __ddiv_rn((double)sdiff[1][g0], t);
                    q2 = __ddiv_rn((double)sdiff[2][b0], t);
                } else {
                    // float32 buffers: x * (1/t) differs from x / t by at most 1 ulp of float64, which
                    // survives the rounding to float32 with probability ~1e-8 per value
                    const double it = __drcp_rn(t);
                    q0 = __dmul_rn((double)sdiff[0][r0], it);
                    q1 = __dmul_rn((double)sdiff[1][g0], it);
                    q2 = __dmul_rn((double)sdiff[2][b0], it);
                }
                const T v0 = clip01<T>((T)__dadd_rn(q0, dl0));
                const T v1 = clip01<T>((T)__dadd_rn(q1, dl1));
                const T v2 = clip01<T>((T)__dadd_rn(q2, dl2));
                T* o = out + (ro + x) * 3;
                o[0] = v0; o[1] = v1; o[2] = v2;
                atomicAdd(&shist[0 * kSelBinsA + sel_bin_a<T>(v0)], 1u);
                atomicAdd(&shist[1 * kSelBinsA + sel_bin_a<T>(v1)], 1u);
                atomicAdd(&shist[2 * kSelBinsA + sel_bin_a<T>(v2)], 1u);
            }
            __syncwarp();
#pragma unroll
            for (int j = 0; j < GC; ++j) { va[j] -= aL[j]; vb[j] -= bL[j]; }
            if (yo + 1 < s.ye) row_load(yo + 1 - s.lo, aL, bL);
        }
    }
    __syncthreads();
    unsigned* gh = a.sel_hist_a + (size_t)inst * 3 * kSelBinsA;
    for (int i = threadIdx.x; i < 3 * kSelBinsA; i += GNT) {
        const unsigned v = shist[i];
        if (v) atomicAdd(&gh[i], v);
    }
}

}  // namespace

int launch_guided(const GuidedArgs& a, int max_radius, cudaStream_t st) {
    if (a.n_dehaze == 0) return 0;
    const int inst = a.B * a.n_dehaze;
    const size_t npix = (size_t)a.H * a.W;
    int pb = (int)((npix + 256 * 8 - 1) / (256 * 8));
    if (pb > 148 * 8) pb = 148 * 8;
    gf_prep_kernel<<<dim3(pb, a.B), 256, 0, st>>>(a);
    const int chunks = (a.H + a.rows_per_block - 1) / a.rows_per_block;
    // strips are sized by each instance's own radius; the grid covers the largest radius present
    // (fewest outputs per strip); for a smaller radius the surplus warps exit at once
    const int outw_min = GSPAN - max_radius + 1;
    const int strips = (a.W + outw_min - 1) / outw_min;
    dim3 grid((strips + GWARPS - 1) / GWARPS, chunks, inst);
    gf_ab_kernel<<<grid, GNT, 0, st>>>(a);
    if (a.lib.f64) gf_recover_kernel<double><<<grid, GNT, 0, st>>>(a);
    else gf_recover_kernel<float><<<grid, GNT, 0, st>>>(a);
    return 3;
}

}  // namespace uwie
