// Dehaze core in float64: dark-channel transmission, guided filter, scene recovery.
//   estimate_transmission / guided_filter / restore_image   six_stadigy.py:25-46, 167-188
//   estimate_transmission / guided_filter / recover_image   enhancement_strategies.py:16-46, 209-249
// Box means are cv2.boxFilter(CV_64F, (r, r)): normalised, anchor r//2 (window [-(r//2), r-1-r//2]),
// BORDER_REFLECT_101, scale applied as a multiplication by 1/(r*r).
//
// Two marching kernels.  A block owns a strip of columns and walks down a chunk of rows, keeping the
// last r rows of horizontal window sums in a shared-memory ring and the vertical sums in registers:
//   gf_ab_kernel       frame bytes -> I (gray/255), p (raw transmission) -> a, b   (4 box sums)
//   gf_recover_kernel  a, b -> mean_a * I + mean_b -> t -> (img - A)/t + A -> recovered frame
//                      (+ the first, coarse level of the percentile histograms, see k_select.cu)
#include "k_arith.cuh"
#include "k_kernels.h"

namespace uwie {

namespace {

constexpr int NT = 256;

__device__ __forceinline__ float raw_transmission(const CastTables& ct, const float* lg, float omega,
                                                  int clip_t0, int r, int g, int b) {
    // dark = min_c(img / (A + guard)); t0 = 1 - omega * dark, all float32
    float d = __fdiv_rn(ct.curf[0][r], lg[0]);
    d = fminf(d, __fdiv_rn(ct.curf[1][g], lg[1]));
    d = fminf(d, __fdiv_rn(ct.curf[2][b], lg[2]));
    float t0 = __fsub_rn(1.0f, __fmul_rn(omega, d));
    if (clip_t0) t0 = t0 < 0.1f ? 0.1f : (t0 > 1.0f ? 1.0f : t0);
    return t0;
}

// ------------------------------------------------------------------ a, b
// thread <-> (column = tid / 4, quantity = tid % 4) with quantities I, p, I*p, I*I ; TW = 64
__global__ void __launch_bounds__(NT) gf_ab_kernel(GuidedArgs a) {
    constexpr int TW = 64;
    extern __shared__ double smem[];
    const int inst = blockIdx.z;                      // frame * n_dehaze + k
    const int b = inst / a.n_dehaze, k = inst % a.n_dehaze;
    const Recipe& rc = a.recipes[a.dehaze_idx[k]];
    const int r = rc.radius, lo = r / 2, hi = r - 1 - lo;
    const int x0 = blockIdx.x * TW;
    const int ys = blockIdx.y * a.rows_per_block;
    const int ye = min(ys + a.rows_per_block, a.H);
    if (ys >= ye) return;
    const int span = TW + r - 1;
    double* sI = smem;                 // [span]
    double* sP = sI + span;            // [span]
    double* ring = sP + span;          // [r][NT]
    __shared__ float lg[3];
    const CastTables& ct = a.cast[b];
    if (threadIdx.x < 3)
        lg[threadIdx.x] = __fadd_rn(a.light[b * 3 + threadIdx.x], (float)a.lib.dark_guard);
    __syncthreads();
    const uint8_t* frame = a.frames + (size_t)b * a.H * a.W * 3;
    const int col = threadIdx.x >> 2, q = threadIdx.x & 3;
    const int x = x0 + col;
    const double scale = 1.0 / ((double)r * (double)r);
    double* pa = a.ab + (size_t)inst * 2 * a.H * a.W;
    double* pb = pa + (size_t)a.H * a.W;

    double vsum = 0.0;
    int filled = 0, slot = 0;
    for (int yy = ys - lo; yy < ye + hi; ++yy) {
        const int y = reflect101(yy, a.H);
        for (int i = threadIdx.x; i < span; i += NT) {
            const int xx = reflect101(x0 - lo + i, a.W);
            const uint8_t* p = frame + ((size_t)y * a.W + xx) * 3;
            const int r0 = p[0], g0 = p[1], b0 = p[2];
            const int gr = gray_q15(ct.curq[0][r0], ct.curq[1][g0], ct.curq[2][b0]);
            sI[i] = (double)gr / 255.0;
            sP[i] = (double)raw_transmission(ct, lg, rc.omega, a.lib.clip_t0, r0, g0, b0);
        }
        __syncthreads();
        double h = 0.0;
        if (q == 0) { for (int i = 0; i < r; ++i) h += sI[col + i]; }
        else if (q == 1) { for (int i = 0; i < r; ++i) h += sP[col + i]; }
        else if (q == 2) { for (int i = 0; i < r; ++i) h += __dmul_rn(sI[col + i], sP[col + i]); }
        else { for (int i = 0; i < r; ++i) h += __dmul_rn(sI[col + i], sI[col + i]); }
        vsum += h;
        if (filled >= r) vsum -= ring[slot * NT + threadIdx.x];
        ring[slot * NT + threadIdx.x] = h;
        slot = slot + 1 == r ? 0 : slot + 1;
        ++filled;
        if (filled >= r) {
            const int yo = yy - hi;                  // output row of the window that just filled
            const double m = vsum * scale;
            const int base = (threadIdx.x & 31) & ~3;
            const double mI = __shfl_sync(0xffffffffu, m, base);
            const double mP = __shfl_sync(0xffffffffu, m, base + 1);
            const double mIP = __shfl_sync(0xffffffffu, m, base + 2);
            const double mII = __shfl_sync(0xffffffffu, m, base + 3);
            const double cov = __dsub_rn(mIP, __dmul_rn(mI, mP));
            const double var = __dsub_rn(mII, __dmul_rn(mI, mI));
            const double av = __ddiv_rn(cov, __dadd_rn(var, rc.eps));
            const double bv = __dsub_rn(mP, __dmul_rn(av, mI));
            if (x < a.W) {
                if (q == 0) pa[(size_t)yo * a.W + x] = av;
                else if (q == 1) pb[(size_t)yo * a.W + x] = bv;
            }
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------ recover
// thread <-> (column = tid / 2, quantity = tid % 2) with quantities a, b ; TW = 128
template <class T>
__global__ void __launch_bounds__(NT) gf_recover_kernel(GuidedArgs a) {
    constexpr int TW = 128;
    extern __shared__ double smem[];
    const int inst = blockIdx.z;
    const int b = inst / a.n_dehaze, k = inst % a.n_dehaze;
    const Recipe& rc = a.recipes[a.dehaze_idx[k]];
    const int r = rc.radius, lo = r / 2, hi = r - 1 - lo;
    const int x0 = blockIdx.x * TW;
    const int ys = blockIdx.y * a.rows_per_block;
    const int ye = min(ys + a.rows_per_block, a.H);
    if (ys >= ye) return;
    const int span = TW + r - 1;
    double* sA = smem;                 // [span]
    double* sB = sA + span;            // [span]
    double* ring = sB + span;          // [r][NT]
    unsigned* shist = reinterpret_cast<unsigned*>(ring + (size_t)r * NT);   // [3][kSelBinsA]
    for (int i = threadIdx.x; i < 3 * kSelBinsA; i += NT) shist[i] = 0;
    const CastTables& ct = a.cast[b];
    const float l0 = a.light[b * 3], l1 = a.light[b * 3 + 1], l2 = a.light[b * 3 + 2];
    const uint8_t* frame = a.frames + (size_t)b * a.H * a.W * 3;
    const int col = threadIdx.x >> 1, q = threadIdx.x & 1;
    const int x = x0 + col;
    const double scale = 1.0 / ((double)r * (double)r);
    const double* pa = a.ab + (size_t)inst * 2 * a.H * a.W;
    const double* pb = pa + (size_t)a.H * a.W;
    T* out = reinterpret_cast<T*>(a.recovered) + (size_t)inst * a.H * a.W * 3;
    __syncthreads();

    double vsum = 0.0;
    int filled = 0, slot = 0;
    for (int yy = ys - lo; yy < ye + hi; ++yy) {
        const int y = reflect101(yy, a.H);
        for (int i = threadIdx.x; i < span; i += NT) {
            const int xx = reflect101(x0 - lo + i, a.W);
            sA[i] = pa[(size_t)y * a.W + xx];
            sB[i] = pb[(size_t)y * a.W + xx];
        }
        __syncthreads();
        double h = 0.0;
        const double* src = q == 0 ? sA : sB;
        for (int i = 0; i < r; ++i) h += src[col + i];
        vsum += h;
        if (filled >= r) vsum -= ring[slot * NT + threadIdx.x];
        ring[slot * NT + threadIdx.x] = h;
        slot = slot + 1 == r ? 0 : slot + 1;
        ++filled;
        if (filled >= r) {
            const int yo = yy - hi;
            const double m = vsum * scale;
            const int base = (threadIdx.x & 31) & ~1;
            const double mA = __shfl_sync(0xffffffffu, m, base);
            const double mB = __shfl_sync(0xffffffffu, m, base + 1);
            if (x < a.W) {
                const uint8_t* p = frame + ((size_t)yo * a.W + x) * 3;
                const int r0 = p[0], g0 = p[1], b0 = p[2];
                const int gr = gray_q15(ct.curq[0][r0], ct.curq[1][g0], ct.curq[2][b0]);
                double t = __dadd_rn(__dmul_rn(mA, (double)gr / 255.0), mB);
                t = t < 0.1 ? 0.1 : (t > 1.0 ? 1.0 : t);
                T* o = out + ((size_t)yo * a.W + x) * 3;
                // (img - A) in float32, / t and + A in float64, stored in the library's float type
                const int c_lo = q == 0 ? 0 : 2, c_hi = q == 0 ? 2 : 3;
                for (int c = c_lo; c < c_hi; ++c) {
                    const int vb = c == 0 ? r0 : (c == 1 ? g0 : b0);
                    const float lc = c == 0 ? l0 : (c == 1 ? l1 : l2);
                    const double rec = __dadd_rn(__ddiv_rn((double)__fsub_rn(ct.curf[c][vb], lc), t), (double)lc);
                    const T v = clip01<T>((T)rec);
                    o[c] = v;
                    atomicAdd(&shist[c * kSelBinsA + (sel_prefix<T>(v) >> 20)], 1u);
                }
            }
        }
        __syncthreads();
    }
    unsigned* gh = a.sel_hist_a + (size_t)inst * 3 * kSelBinsA;
    for (int i = threadIdx.x; i < 3 * kSelBinsA; i += NT) {
        const unsigned v = shist[i];
        if (v) atomicAdd(&gh[i], v);
    }
}

}  // namespace

int launch_guided(const GuidedArgs& a, int max_radius, cudaStream_t st) {
    if (a.n_dehaze == 0) return 0;
    const int inst = a.B * a.n_dehaze;
    const int chunks = (a.H + a.rows_per_block - 1) / a.rows_per_block;
    {
        const size_t sm = ((size_t)2 * (64 + max_radius - 1) + (size_t)max_radius * NT) * sizeof(double);
        cudaFuncSetAttribute(gf_ab_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        gf_ab_kernel<<<dim3((a.W + 63) / 64, chunks, inst), NT, sm, st>>>(a);
    }
    {
        const size_t sm = ((size_t)2 * (128 + max_radius - 1) + (size_t)max_radius * NT) * sizeof(double) +
                          (size_t)3 * kSelBinsA * sizeof(unsigned);
        dim3 grid((a.W + 127) / 128, chunks, inst);
        if (a.lib.f64) {
            cudaFuncSetAttribute(gf_recover_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
            gf_recover_kernel<double><<<grid, NT, sm, st>>>(a);
        } else {
            cudaFuncSetAttribute(gf_recover_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
            gf_recover_kernel<float><<<grid, NT, sm, st>>>(a);
        }
    }
    return 2;
}

}  // namespace uwie
