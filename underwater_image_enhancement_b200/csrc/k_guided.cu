// Dehaze core: dark-channel transmission, guided filter, scene recovery.
//   estimate_transmission / guided_filter / restore_image   six_stadigy.py:25-46, 167-188
//   estimate_transmission / guided_filter / recover_image   enhancement_strategies.py:16-46, 209-249
// Box means are cv2.boxFilter(CV_64F, (r, r)): normalised, anchor r//2 (window [-(r//2), r-1-r//2]),
// BORDER_REFLECT_101, scale applied as a multiplication by 1/(r*r).
//
// Three kernels:
//   gf_prep_kernel      frame bytes -> one 8-byte record per pixel {R | G<<8 | B<<16 | gray<<24, dark channel
//                       (float32)}, per frame: the marching kernels fetch both with one 64-bit load
//   gf_ab_kernel        gray, dark -> box sums of I, p, I*p, I*I -> {a, b} (one float64 pair per pixel)
//   gf_recover_kernel   {a, b} -> mean_a * I + mean_b -> t -> (img - A)/t + A -> recovered frame
//                       (+ the first level of the percentile histograms, see k_select.cu)
//
// Box sums are O(1) per pixel.  Every WARP marches down its own strip of 160 columns with no
// block-level synchronisation: a lane keeps the vertical running sums of 5 columns in registers
// (row entering the window added, row leaving it subtracted - both re-read, so there is no ring
// buffer), publishes them to a per-warp shared-memory row, and then forms the horizontal window
// sums of 5 CONSECUTIVE outputs with a sliding window (r + 4 reads for 5 outputs).
// The first-pass sums are exact whatever the order: I = g/255 is carried as the integers (sum g,
// sum g*g), and p (a float32, multiple of 2^-27) and g*p are summed in float64 far below 2^53 ulps.
// The kernels are bound by instruction issue and load latency, not by HBM: rows are addressed as a
// warp-uniform row pointer plus a 32-bit column, records are fetched whole (64 / 128 bits), and the
// {a, b} pairs leave straight from the registers that computed them.
#include <cstdlib>

#include "k_arith.cuh"
#include "k_kernels.h"

namespace uwie {

namespace {

// Columns per lane (odd: conflict-free 64-bit shared accesses).  Measured in round 2: 3 columns per lane (80 registers,
// three blocks = 24 warps per SM, but 96 - r + 1 outputs of 96 columns instead of 160 - r + 1 of 160 and the per-row
// shuffles / loop overhead spread over fewer pixels) is SLOWER than 5 (128 registers, 14 warps): 44.9 vs 32.1 ms per
// 256-frame step.  Register budget at GC = 5, also measured in round 2 (per-kernel __maxnreg__): 96 registers (three 7-warp
// blocks per SM, 40-104 bytes spilled) 55.8 ms; 144 registers (no spills, two blocks) 39.3 ms; 128 registers 32.1 ms - kept.
constexpr int GC = 5;
constexpr int GSPAN = 32 * GC;        // columns a warp carries, outputs = GSPAN - r + 1
constexpr int GWARPS = 8;
constexpr int GNT = 32 * GWARPS;
constexpr int GBLOCKS_PER_SM = GC <= 3 ? 3 : 2;

// ------------------------------------------------------------------ gray + dark channel records
// (Round 2 tried this kernel with TMA-staged tiles like uwie_prim_transmission_head_u8 - cp.async.bulk + mbarrier, persistent
// one-wave blocks or short-lived two-tile blocks: 2 x faster alone, 1.7 -> 1.0 ms per 256-frame step, but the four-lane
// step lost 4.5 %, 2 760 -> 2 640 frames/s, with either shape and with one polling thread: small blocks of plain loads
// share the SMs with the other lanes' kernels better than 24-41 KB blocks waiting on a barrier.  Kept as it was.)
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
    const unsigned npix = (unsigned)a.H * a.W;
    const uint8_t* f = a.frames + (size_t)b * npix * 3;
    uint2* gd = a.gd_plane + (size_t)b * npix;
    auto record = [&](unsigned r, unsigned g, unsigned bl) {
        const unsigned gq = (unsigned)gray_q15(sq[0][r], sq[1][g], sq[2][bl]);
        const float dk = fminf(fminf(sdiv[0][r], sdiv[1][g]), sdiv[2][bl]);
        return make_uint2(r | (g << 8) | (bl << 16) | (gq << 24), __float_as_uint(dk));
    };
    // four pixels per thread: three 32-bit loads in, two 128-bit stores out
    const unsigned groups = ((reinterpret_cast<uintptr_t>(f) & 3) == 0 && (reinterpret_cast<uintptr_t>(gd) & 15) == 0) ? npix / 4 : 0;
    const uint32_t* w = reinterpret_cast<const uint32_t*>(f);
    for (unsigned i = blockIdx.x * 256 + threadIdx.x; i < groups; i += gridDim.x * 256) {
        const uint32_t w0 = __ldg(w + 3 * i), w1 = __ldg(w + 3 * i + 1), w2 = __ldg(w + 3 * i + 2);
        const uint2 p0 = record(w0 & 255, (w0 >> 8) & 255, (w0 >> 16) & 255);
        const uint2 p1 = record(w0 >> 24, w1 & 255, (w1 >> 8) & 255);
        const uint2 p2 = record((w1 >> 16) & 255, w1 >> 24, w2 & 255);
        const uint2 p3 = record((w2 >> 8) & 255, (w2 >> 16) & 255, w2 >> 24);
        uint4* o = reinterpret_cast<uint4*>(gd + 4 * (size_t)i);
        o[0] = make_uint4(p0.x, p0.y, p1.x, p1.y);
        o[1] = make_uint4(p2.x, p2.y, p3.x, p3.y);
    }
    for (unsigned i = 4 * groups + blockIdx.x * 256 + threadIdx.x; i < npix; i += gridDim.x * 256)
        gd[i] = record(f[i * 3], f[i * 3 + 1], f[i * 3 + 2]);
}

template <bool CLIP>
__device__ __forceinline__ float transmission0(float omega, float dark) {
    float t0 = __fsub_rn(1.0f, __fmul_rn(omega, dark));           // 1 - omega * dark, float32
    if (CLIP) t0 = t0 < 0.1f ? 0.1f : (t0 > 1.0f ? 1.0f : t0);    // SS:174
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
    s.xs = (blockIdx.x * (blockDim.x >> 5) + s.warp) * s.outw;
    s.ys = blockIdx.y * a.rows_per_block;
    s.ye = min(s.ys + a.rows_per_block, a.H);
    s.valid = s.xs < a.W && s.ys < s.ye;
    return s;
}

// ------------------------------------------------------------------ a, b
template <bool CLIP, bool FAST, class AB>
__global__ void __launch_bounds__(GNT, GBLOCKS_PER_SM) gf_ab_kernel(GuidedArgs a) {
    __shared__ double sP[GWARPS][GSPAN];
    __shared__ double sGP[GWARPS][GSPAN];
    __shared__ unsigned long long sG[GWARPS][GSPAN];     // (sum g) << 32 | sum g*g
    const int inst = blockIdx.z;                         // frame * n_dehaze + k
    const int b = inst / a.n_dehaze, k = inst % a.n_dehaze;
    const Recipe& rc = a.recipes[a.dehaze_idx[k]];
    const Strip s = make_strip(a, rc.radius);
    if (!s.valid) return;                                // warp-uniform
    const float omega = rc.omega;
    const double eps = rc.eps;
    const unsigned npix = (unsigned)a.H * a.W;
    const uint2* gd = a.gd_plane + (size_t)b * npix;
    AB* pab = reinterpret_cast<AB*>(a.ab) + (size_t)inst * npix;
    const double rr = (double)s.r * (double)s.r;
    const double scale = 1.0 / rr;                       // cv2: sum * (1 / (kw * kh))
    const double scale_i = 1.0 / (255.0 * rr), scale_ii = 1.0 / (65025.0 * rr);
    double* wP = sP[s.warp]; double* wGP = sGP[s.warp]; unsigned long long* wG = sG[s.warp];

    int cx[GC];
#pragma unroll
    for (int j = 0; j < GC; ++j) cx[j] = reflect101(s.xs - s.lo + s.lane + 32 * j, a.W);
    double vp[GC], vgp[GC];
    unsigned vg[GC], vgg[GC];
#pragma unroll
    for (int j = 0; j < GC; ++j) { vp[j] = 0.0; vgp[j] = 0.0; vg[j] = 0u; vgg[j] = 0u; }

    // Row loads are software-pipelined: the records of the row that enters (leaves) the window in the
    // NEXT step are requested right after this step's were consumed, so their latency hides behind
    // the shared-memory phase instead of stalling a warp that has nothing else to do.
    uint2 E[GC], L[GC];
    auto row_load = [&](int yy, uint2 (&q)[GC]) {
        const uint2* rowp = gd + (unsigned)reflect101(yy, a.H) * (unsigned)a.W;     // warp-uniform
#pragma unroll
        for (int j = 0; j < GC; ++j) q[j] = __ldg(rowp + cx[j]);
    };
    auto row_in = [&](const uint2 (&q)[GC]) {
#pragma unroll
        for (int j = 0; j < GC; ++j) {
            const unsigned g = q[j].x >> 24;
            const double p = (double)transmission0<CLIP>(omega, __uint_as_float(q[j].y));
            vp[j] += p; vgp[j] += (double)g * p; vg[j] += g; vgg[j] += g * g;
        }
    };
    auto row_out = [&](const uint2 (&q)[GC]) {
#pragma unroll
        for (int j = 0; j < GC; ++j) {
            const unsigned g = q[j].x >> 24;
            const double p = (double)transmission0<CLIP>(omega, __uint_as_float(q[j].y));
            vp[j] -= p; vgp[j] -= (double)g * p; vg[j] -= g; vgg[j] -= g * g;
        }
    };

    for (int yy = s.ys - s.lo; yy < s.ys + s.hi; ++yy) { row_load(yy, E); row_in(E); }
    const int base = s.lane * GC;
    const int nblocks = s.r / GC;
    row_load(s.ys + s.hi, E);
    row_load(s.ys - s.lo, L);
    for (int yo = s.ys; yo < s.ye; ++yo) {
        row_in(E);
        if (yo + 1 < s.ye) row_load(yo + 1 + s.hi, E);
#pragma unroll
        for (int j = 0; j < GC; ++j) {
            wP[s.lane + 32 * j] = vp[j]; wGP[s.lane + 32 * j] = vgp[j];
            wG[s.lane + 32 * j] = ((unsigned long long)vg[j] << 32) | vgg[j];
        }
        __syncwarp();
        // window of the first output of a lane = r / 5 whole 5-column blocks (its own and its right-hand
        // neighbours', fetched by shuffle) + r % 5 single columns; the sums are exact in any order
        double hp = 0.0, hgp = 0.0;
        unsigned long long hg = 0ull;
        {
            // (these sums are exact, so their order is free: pairs first, and the neighbours' block sums are fetched
            // before any of them is added - the dependency chain of a row is what bounds this kernel)
            static_assert(GC == 5, "block sums are written out for five columns");
            const double bp = ((wP[base] + wP[base + 1]) + (wP[base + 2] + wP[base + 3])) + wP[base + 4];
            const double bgp = ((wGP[base] + wGP[base + 1]) + (wGP[base + 2] + wGP[base + 3])) + wGP[base + 4];
            const unsigned long long bg = ((wG[base] + wG[base + 1]) + (wG[base + 2] + wG[base + 3])) + wG[base + 4];
            if (nblocks > 0) { hp = bp; hgp = bgp; hg = bg; }
            if (nblocks >= 2 && nblocks <= 4) {                   // r = 10 .. 24: the radii of both libraries
                const double p1 = __shfl_down_sync(0xffffffffu, bp, 1), p2 = __shfl_down_sync(0xffffffffu, bp, 2), p3 = __shfl_down_sync(0xffffffffu, bp, 3);
                const double q1 = __shfl_down_sync(0xffffffffu, bgp, 1), q2 = __shfl_down_sync(0xffffffffu, bgp, 2), q3 = __shfl_down_sync(0xffffffffu, bgp, 3);
                const unsigned long long g1 = __shfl_down_sync(0xffffffffu, bg, 1), g2 = __shfl_down_sync(0xffffffffu, bg, 2), g3 = __shfl_down_sync(0xffffffffu, bg, 3);
                hp = bp + p1; hgp = bgp + q1; hg = bg + g1;
                if (nblocks == 3) { hp += p2; hgp += q2; hg += g2; }
                if (nblocks == 4) { hp += p2 + p3; hgp += q2 + q3; hg += g2 + g3; }
            } else {
                for (int t = 1; t < nblocks; ++t) {               // warp-uniform trip count
                    hp += __shfl_down_sync(0xffffffffu, bp, t);
                    hgp += __shfl_down_sync(0xffffffffu, bgp, t);
                    hg += __shfl_down_sync(0xffffffffu, bg, t);
                }
            }
        }
        // Every lane forms all 5 outputs in straight-line code (dead ones from clamped reads) so that the
        // five dependent chains of a / b interleave; only the stores are predicated.
        for (int i = nblocks * GC; i < s.r; ++i) {
            const int c = min(base + i, GSPAN - 1);
            hp += wP[c]; hgp += wGP[c]; hg += wG[c];
        }
        AB* orow = pab + (unsigned)yo * (unsigned)a.W + (unsigned)(s.xs + base);
        // windows of outputs 1 .. 4 = window 0 + what entered - what left; the differences do not depend on each other
        double dP[GC], dGP[GC];
        unsigned long long dG[GC];
        dP[0] = 0.0; dGP[0] = 0.0; dG[0] = 0ull;
#pragma unroll
        for (int j = 1; j < GC; ++j) {
            const int in = min(base + j + s.r - 1, GSPAN - 1);
            dP[j] = wP[in] - wP[base + j - 1];
            dGP[j] = wGP[in] - wGP[base + j - 1];
            dG[j] = wG[in] - wG[base + j - 1];
        }
        dP[2] += dP[1]; dGP[2] += dGP[1]; dG[2] += dG[1];
        dP[4] += dP[3]; dGP[4] += dGP[3]; dG[4] += dG[3];
        dP[3] += dP[2]; dGP[3] += dGP[2]; dG[3] += dG[2];
        dP[4] += dP[2]; dGP[4] += dGP[2]; dG[4] += dG[2];
        const double hp0 = hp, hgp0 = hgp;
        const unsigned long long hg0 = hg;
#pragma unroll
        for (int j = 0; j < GC; ++j) {
            hp = hp0 + dP[j]; hgp = hgp0 + dGP[j]; hg = hg0 + dG[j];
            const double mI = (double)(unsigned)(hg >> 32) * scale_i;
            const double mII = (double)(unsigned)(hg & 0xffffffffull) * scale_ii;
            const double mP = hp * scale;
            const double mIP = hgp * scale_i;
            const double cov = __dsub_rn(mIP, __dmul_rn(mI, mP));
            const double var = __dsub_rn(mII, __dmul_rn(mI, mI));
            const double den = __dadd_rn(var, eps);
            // FAST: eps >= 1e-9, so den is normal and |a| < 1e10 (a is 0 or far above the subnormals)
            const double ra = FAST ? div_2ulp_normal(cov, den) : __ddiv_rn(cov, den);
            const double rb = __dsub_rn(mP, __dmul_rn(ra, mI));
            if (base + j < s.outw && s.xs + base + j < a.W) {
                if (sizeof(AB) == 16) *reinterpret_cast<double2*>(&orow[j]) = make_double2(ra, rb);
                else *reinterpret_cast<float2*>(&orow[j]) = make_float2((float)ra, (float)rb);
            }
        }
        __syncwarp();
        row_out(L);
        if (yo + 1 < s.ye) row_load(yo + 1 - s.lo, L);
    }
}

// ------------------------------------------------------------------ recover
struct RecoverShared {
    double sA[GWARPS][GSPAN];           // vertical running sums of a, b
    double sB[GWARPS][GSPAN];
    double sMA[GWARPS][GSPAN];          // window means of a, b: written by the lane that owns 5 consecutive
    double sMB[GWARPS][GSPAN];          // columns, read back column-interleaved for coalesced stores
    double sI[256];                     // g / 255.0
    float sdiff[3][256];                // img - A, float32              SS:185, ES:247
    unsigned shist[3 * kSelBinsA];
};
template <class T, class AB>
__global__ void __launch_bounds__(GNT, GBLOCKS_PER_SM) gf_recover_kernel(GuidedArgs a) {
    // dynamic shared memory (57 KB, above the static limit), see RecoverShared
    extern __shared__ __align__(16) unsigned char gf_smem[];
    RecoverShared& sm = *reinterpret_cast<RecoverShared*>(gf_smem);
    double (&sI)[256] = sm.sI;
    float (&sdiff)[3][256] = sm.sdiff;
    unsigned (&shist)[3 * kSelBinsA] = sm.shist;
    const int inst = blockIdx.z;
    const int b = inst / a.n_dehaze, k = inst % a.n_dehaze;
    const Recipe& rc = a.recipes[a.dehaze_idx[k]];
    const float l0 = a.light[b * 3], l1 = a.light[b * 3 + 1], l2 = a.light[b * 3 + 2];
    {
        const CastTables& ct = a.cast[b];
        for (int i = threadIdx.x; i < 3 * kSelBinsA; i += blockDim.x) shist[i] = 0;
        for (int i = threadIdx.x; i < 256; i += blockDim.x) sI[i] = (double)i / 255.0;
        for (int i = threadIdx.x; i < 768; i += blockDim.x) {
            const int c = i >> 8, v = i & 255;
            sdiff[c][v] = __fsub_rn(ct.curf[c][v], c == 0 ? l0 : (c == 1 ? l1 : l2));
        }
    }
    __syncthreads();
    const Strip s = make_strip(a, rc.radius);
    if (s.valid) {
        const unsigned npix = (unsigned)a.H * a.W;
        const uint2* gd = a.gd_plane + (size_t)b * npix;
        const AB* pab = reinterpret_cast<const AB*>(a.ab) + (size_t)inst * npix;
        T* out = reinterpret_cast<T*>(a.recovered) + (size_t)inst * npix * 3;
        const double scale = 1.0 / ((double)s.r * (double)s.r);
        double* wA = sm.sA[s.warp]; double* wB = sm.sB[s.warp];
        double* mA = sm.sMA[s.warp]; double* mB = sm.sMB[s.warp];
        const double dl0 = (double)l0, dl1 = (double)l1, dl2 = (double)l2;

        int cx[GC];
#pragma unroll
        for (int j = 0; j < GC; ++j) cx[j] = reflect101(s.xs - s.lo + s.lane + 32 * j, a.W);
        double va[GC], vb[GC];
#pragma unroll
        for (int j = 0; j < GC; ++j) { va[j] = 0.0; vb[j] = 0.0; }
        // row loads: both are requested right before the long recovery step of a row, the entering
        // row for the NEXT step and the leaving row for this one (consumed after that step)
        AB E[GC], L[GC];
        auto row_load = [&](int yy, AB (&q)[GC]) {
            const AB* rowp = pab + (unsigned)reflect101(yy, a.H) * (unsigned)a.W;   // warp-uniform
#pragma unroll
            for (int j = 0; j < GC; ++j) q[j] = __ldg(rowp + cx[j]);
        };
        for (int yy = s.ys - s.lo; yy < s.ys + s.hi; ++yy) {
            row_load(yy, E);
#pragma unroll
            for (int j = 0; j < GC; ++j) { va[j] += (double)E[j].x; vb[j] += (double)E[j].y; }
        }
        const int base = s.lane * GC;
        const int nblocks = s.r / GC;
        row_load(s.ys + s.hi, E);
        for (int yo = s.ys; yo < s.ye; ++yo) {
#pragma unroll
            for (int j = 0; j < GC; ++j) { va[j] += (double)E[j].x; vb[j] += (double)E[j].y; }
            // this row's pixels (frame bytes + gray) for the recovery step below, requested early
            const unsigned ro = (unsigned)yo * (unsigned)a.W;
            unsigned pixw[GC];
#pragma unroll
            for (int j = 0; j < GC; ++j) {
                const int x = s.xs + s.lane + 32 * j;
                pixw[j] = (s.lane + 32 * j < s.outw && x < a.W) ? __ldg(&gd[ro + x].x) : 0u;
            }
#pragma unroll
            for (int j = 0; j < GC; ++j) { wA[s.lane + 32 * j] = va[j]; wB[s.lane + 32 * j] = vb[j]; }
            __syncwarp();
            double ha = 0.0, hb = 0.0;                     // first window by blocks, see gf_ab_kernel
            {
                double ba = 0.0, bb = 0.0;
#pragma unroll
                for (int j = 0; j < GC; ++j) { ba += wA[base + j]; bb += wB[base + j]; }
                if (nblocks > 0) { ha = ba; hb = bb; }
                for (int t = 1; t < nblocks; ++t) {
                    ha += __shfl_down_sync(0xffffffffu, ba, t);
                    hb += __shfl_down_sync(0xffffffffu, bb, t);
                }
            }
            for (int i = nblocks * GC; i < s.r; ++i) {
                const int c = min(base + i, GSPAN - 1);
                ha += wA[c]; hb += wB[c];
            }
#pragma unroll
            for (int j = 0; j < GC; ++j) {                  // all lanes, dead outputs from clamped reads
                if (j > 0) {
                    const int in = min(base + j + s.r - 1, GSPAN - 1);
                    ha += wA[in] - wA[base + j - 1];
                    hb += wB[in] - wB[base + j - 1];
                }
                mA[base + j] = ha * scale; mB[base + j] = hb * scale;
            }
            __syncwarp();
            // transmission of the five pixels first: it consumes this row's pixel words, so the row loads
            // that follow do not sit in front of them in the load queue
            double tt[GC];
#pragma unroll
            for (int jj = 0; jj < GC; ++jj) {
                const int i = s.lane + 32 * jj;
                const double t = __dadd_rn(__dmul_rn(mA[i], sI[pixw[jj] >> 24]), mB[i]);
                tt[jj] = t < 0.1 ? 0.1 : (t > 1.0 ? 1.0 : t);
            }
            if (yo + 1 < s.ye) row_load(yo + 1 + s.hi, E);
            row_load(yo - s.lo, L);
#pragma unroll
            for (int jj = 0; jj < GC; ++jj) {
                // straight-line for all lanes (the five chains interleave); stores and counts are predicated
                const int i = s.lane + 32 * jj;
                const int x = s.xs + i;
                const bool ok = i < s.outw && x < a.W;
                const int r0 = pixw[jj] & 255, g0 = (pixw[jj] >> 8) & 255, b0 = (pixw[jj] >> 16) & 255;
                const double t = tt[jj];
                // (img - A) in float32, / t and + A in float64, stored in the library's float type.
                // t is in [0.1, 1] and img - A is 0 or far above the subnormals: *_rn_normal are exact IEEE
                double q0, q1, q2;
                if (sizeof(T) == 8) {
                    q0 = div_rn_normal((double)sdiff[0][r0], t);
                    q1 = div_rn_normal((double)sdiff[1][g0], t);
                    q2 = div_rn_normal((double)sdiff[2][b0], t);
                } else {
                    // float32 buffers: x * (1/t) differs from x / t by at most 1 ulp of float64, which
                    // survives the rounding to float32 with probability ~1e-8 per value
                    const double it = rcp_rn_normal(t);
                    q0 = __dmul_rn((double)sdiff[0][r0], it);
                    q1 = __dmul_rn((double)sdiff[1][g0], it);
                    q2 = __dmul_rn((double)sdiff[2][b0], it);
                }
                const T v0 = clip01<T>((T)__dadd_rn(q0, dl0));
                const T v1 = clip01<T>((T)__dadd_rn(q1, dl1));
                const T v2 = clip01<T>((T)__dadd_rn(q2, dl2));
                if (ok) {
                    T* o = out + (size_t)(ro + x) * 3;
                    o[0] = v0; o[1] = v1; o[2] = v2;
                    atomicAdd(&shist[0 * kSelBinsA + sel_bin_a<T>(v0)], 1u);
                    atomicAdd(&shist[1 * kSelBinsA + sel_bin_a<T>(v1)], 1u);
                    atomicAdd(&shist[2 * kSelBinsA + sel_bin_a<T>(v2)], 1u);
                }
            }
            __syncwarp();
#pragma unroll
            for (int j = 0; j < GC; ++j) { va[j] -= (double)L[j].x; vb[j] -= (double)L[j].y; }
        }
    }
    __syncthreads();
    unsigned* gh = a.sel_hist_a + (size_t)inst * 3 * kSelBinsA;
    for (int i = threadIdx.x; i < 3 * kSelBinsA; i += blockDim.x) {
        const unsigned v = shist[i];
        if (v) atomicAdd(&gh[i], v);
    }
}

// ------------------------------------------------------------------ self-test of *_rn_normal
__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
    z += 0x9E3779B97F4A7C15ull; z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31);
}
// log-uniform magnitude in [2^lo, 2^hi) with a full random mantissa
__device__ __forceinline__ double rand_mag(unsigned long long h, int lo, int hi) {
    const int e = lo + (int)((h >> 52) % (unsigned)(hi - lo));
    const unsigned long long bits = ((unsigned long long)(e + 1023) << 52) | (h & 0xFFFFFFFFFFFFFull);
    return __longlong_as_double((long long)bits);
}
__global__ void selftest_div_kernel(unsigned long long count, unsigned long long seed, unsigned long long* bad) {
    unsigned long long mine = 0;
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < count;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned long long h0 = mix64(seed + 3 * i), h1 = mix64(seed + 3 * i + 1), h2 = mix64(seed + 3 * i + 2);
        const double y = rand_mag(h0, -30, 1);                       // [9.3e-10, 2)
        double x = rand_mag(h1, -60, 1);                             // [8.7e-19, 2)
        if (h2 & 1) x = -x;
        if ((h2 & 0xfe) == 0) x = 0.0;
        const double t = 0.1 + (double)(h2 >> 11) * (0.9 / 9007199254740992.0);
        const double tt = t > 1.0 ? 1.0 : t;
        if (__double_as_longlong(div_rn_normal(x, y)) != __double_as_longlong(__ddiv_rn(x, y))) ++mine;
        if (__double_as_longlong(div_rn_normal(x, tt)) != __double_as_longlong(__ddiv_rn(x, tt))) ++mine;
        if (__double_as_longlong(rcp_rn_normal(tt)) != __double_as_longlong(__drcp_rn(tt))) ++mine;
        // invariant-divisor stretch division (k_arith.cuh: InvDiv): divisor in [guard, 1 + guard], dividend 0 or 2^-30 <= |x| <= 1
        {
            const unsigned long long h3 = mix64(seed + 3 * i + 0x5bd1e995ull), h4 = mix64(h3);
            const double d64 = 1e-10 + rand_mag(h3, -33, 0);                     // [1e-10 + 1.2e-10, 1 + 1e-10)
            double x64 = rand_mag(h4, -60, 0);
            if (h4 & 1) x64 = -x64;
            if ((h4 & 0xfe) == 0) x64 = 0.0;
            if (__double_as_longlong(InvDiv<double>::div(x64, d64, InvDiv<double>::rcp(d64))) != __double_as_longlong(__ddiv_rn(x64, d64))) ++mine;
            const float d32 = __fadd_rn((float)rand_mag(h4 ^ h3, -20, 0), 1e-6f);  // [1e-6 + 9.5e-7, 1 + 1e-6)
            float x32 = (float)rand_mag(h3 ^ (h4 >> 7), -30, 0);
            if (h3 & 1) x32 = -x32;
            if ((h3 & 0xfe) == 0) x32 = 0.f;
            if (__float_as_uint(InvDiv<float>::div(x32, d32, InvDiv<float>::rcp(d32))) != __float_as_uint(__fdiv_rn(x32, d32))) ++mine;
        }
    }
    if (mine) atomicAdd(bad, mine);
}

}  // namespace

int selftest_division(unsigned long long count, unsigned long long seed, unsigned long long* mismatches_host) {
    unsigned long long* bad = nullptr;
    if (cudaMalloc(&bad, 8) != cudaSuccess) return 1;
    cudaMemset(bad, 0, 8);
    selftest_div_kernel<<<148 * 8, 256>>>(count, seed, bad);
    const cudaError_t rc = cudaMemcpy(mismatches_host, bad, 8, cudaMemcpyDeviceToHost);
    cudaFree(bad);
    return rc == cudaSuccess ? 0 : 1;
}

int launch_guided(const GuidedArgs& a0, int max_radius, cudaStream_t st) {
    if (a0.n_dehaze == 0) return 0;
    GuidedArgs a = a0;
    const int inst = a.B * a.n_dehaze;
    const size_t npix = (size_t)a.H * a.W;
    int pb = (int)((npix + 256 * 8 - 1) / (256 * 8));
    if (pb > 148 * 8) pb = 148 * 8;
    gf_prep_kernel<<<dim3(pb, a.B), 256, 0, st>>>(a);
    // strips are sized by each instance's own radius; the grid covers the largest radius present
    // (fewest outputs per strip); for a smaller radius the surplus warps exit at once.
    // Warps per block: the strips spread evenly over the fewest blocks (1920 columns, r = 20: 25 strips = 4 x 7 warps)
    const int outw_min = GSPAN - max_radius + 1;
    const int strips = (a.W + outw_min - 1) / outw_min;
    const int bx = (strips + GWARPS - 1) / GWARPS;
    const int nt = 32 * ((strips + bx - 1) / bx);
    // row chunks: every chunk re-reads r - 1 warm-up rows, so as few as fill the machine ONCE
    // (all blocks do the same work, a second partial wave would only be a tail)
    // UWIE_GF_BLOCKS_PER_SM=1 (experiment): one block per SM, enforced by padding the dynamic shared memory, so
    // that the kernels of OTHER lanes can share the SM with these latency-bound ones instead of queueing behind them
    static const int per_sm = [] { const char* v = std::getenv("UWIE_GF_BLOCKS_PER_SM"); return (v && v[0] == '1') ? 1 : GBLOCKS_PER_SM; }();
    const size_t pad = per_sm == 1 ? 120 * 1024 : 0;
    {
        const int per_chunk = bx * inst;
        int chunks = (148 * per_sm) / (per_chunk > 0 ? per_chunk : 1);
        const int most = a.H / (3 * (max_radius > 0 ? max_radius : 1));
        if (chunks > most) chunks = most;
        if (chunks < 1) chunks = 1;
        a.rows_per_block = (a.H + chunks - 1) / chunks;
    }
    const int chunks = (a.H + a.rows_per_block - 1) / a.rows_per_block;
    dim3 grid(bx, chunks, inst);
    // UWIE_GF_AB32=1 (experiment): the {a, b} pairs cross HBM as float32 instead of float64
    static const bool ab32 = [] { const char* v = std::getenv("UWIE_GF_AB32"); return v && v[0] == '1'; }();
    static const bool optin = [] {
        const int most = 128 * 1024, most_r = (int)sizeof(RecoverShared) + 120 * 1024;   // + static: below the 227 KB of an SM
        return cudaFuncSetAttribute(gf_recover_kernel<double, double2>, cudaFuncAttributeMaxDynamicSharedMemorySize, most_r) == cudaSuccess &&
               cudaFuncSetAttribute(gf_recover_kernel<float, double2>, cudaFuncAttributeMaxDynamicSharedMemorySize, most_r) == cudaSuccess &&
               cudaFuncSetAttribute(gf_recover_kernel<float, float2>, cudaFuncAttributeMaxDynamicSharedMemorySize, most_r) == cudaSuccess &&
               cudaFuncSetAttribute(gf_ab_kernel<true, true, double2>, cudaFuncAttributeMaxDynamicSharedMemorySize, most) == cudaSuccess &&
               cudaFuncSetAttribute(gf_ab_kernel<true, true, float2>, cudaFuncAttributeMaxDynamicSharedMemorySize, most) == cudaSuccess &&
               cudaFuncSetAttribute(gf_ab_kernel<true, false, double2>, cudaFuncAttributeMaxDynamicSharedMemorySize, most) == cudaSuccess &&
               cudaFuncSetAttribute(gf_ab_kernel<false, true, double2>, cudaFuncAttributeMaxDynamicSharedMemorySize, most) == cudaSuccess &&
               cudaFuncSetAttribute(gf_ab_kernel<false, false, double2>, cudaFuncAttributeMaxDynamicSharedMemorySize, most) == cudaSuccess;
    }();
    (void)optin;                                        // a failure shows up as a launch error below
    if (ab32 && a.lib.clip_t0 && a.eps_safe && !a.lib.f64) {
        gf_ab_kernel<true, true, float2><<<grid, nt, pad, st>>>(a);
        gf_recover_kernel<float, float2><<<grid, nt, sizeof(RecoverShared) + pad, st>>>(a);
        return 3;
    }
    if (a.lib.clip_t0) {
        if (a.eps_safe) gf_ab_kernel<true, true, double2><<<grid, nt, pad, st>>>(a);
        else gf_ab_kernel<true, false, double2><<<grid, nt, pad, st>>>(a);
    } else {
        if (a.eps_safe) gf_ab_kernel<false, true, double2><<<grid, nt, pad, st>>>(a);
        else gf_ab_kernel<false, false, double2><<<grid, nt, pad, st>>>(a);
    }
    if (a.lib.f64) gf_recover_kernel<double, double2><<<grid, nt, sizeof(RecoverShared) + pad, st>>>(a);
    else gf_recover_kernel<float, double2><<<grid, nt, sizeof(RecoverShared) + pad, st>>>(a);
    return 3;
}

}  // namespace uwie
