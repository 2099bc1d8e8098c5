// The 79 hand-crafted features of feature_extraction.py, one batch of uint8 frames
// per call.  C ABI: include/uwie.h (uwie_features79*).  Reference:
//   extract_color_features     FE:17-77    35   Lab moments, HSV mean/std, colour-cast factor, RGB stats
//   extract_texture_features   FE:80-120   22   LBP(8,1,'uniform') histogram, GLCM statistics of a 128x128 thumbnail
//   extract_frequency_features FE:123-158   5   orthonormal 2-D DCT-II energy split, |DCT| mean / std
//   extract_edge_features      FE:161-200   7   Sobel magnitude, Canny density, Laplacian
//   extract_quality_features   FE:203-246  10   gray / saturation statistics
// How the frame is crossed:
//   fe_point_kernel    one read of the frame: Lab / HSV / RGB / gray histograms (10 x 256) + gray plane
//   fe_stencil_kernel  one read of the gray plane: Sobel, Laplacian, LBP codes (3x3 neighbourhoods)
//   Canny              k_canny.cu on the gray plane (edge count)
//   GLCM               thumbnail (cv2.resize fixed point) -> 4 co-occurrence matrices by atomics -> properties
//   DCT                D_H * G * D_W^T as two SGEMMs (cuBLAS, plain library GEMMs in full float32) + one reduction
//   fe_final_kernel    everything that is a function of the histograms and sums, in float64
#include <cuda_runtime.h>

#include <cmath>
#include <cstring>
#include <mutex>
#include <string>

#include "../../include/uwie.h"
#include "k_arith.cuh"
#include "k_kernels.h"
#include "uwie_tables.h"

namespace uwie {
int dct_energy_blocks(int H, int W);
int launch_dct_energy(const float* grayf, float* tmp, int B, int H, int W, double* sums, double* part, cudaStream_t st);
int uwie_fail(int code, const std::string& msg);   // uwie_engine.cu
}
using namespace uwie;

namespace {

constexpr int NT = 256;
constexpr int kFe = 79;
constexpr int kThumb = 128;              // FE:104
enum { H_L = 0, H_A, H_B, H_H, H_S, H_V, H_R, H_G, H_BL, H_GRAY, H_COUNT };

struct FeAccum {                         // per frame, zeroed at the start of every call
    unsigned hist[H_COUNT][256];
    unsigned lbp[10];
    unsigned canny;
    unsigned sob_max_bits;               // float32 bit pattern of the largest Sobel magnitude (>= 0: integer order)
    // Sobel magnitude sums as exact integers, so that the result does not depend on the order in which the blocks arrive:
    // m (float32, 0 or in [2^-8, 2^3)) is a multiple of 2^-31; m * m (exact in float64) is a multiple of 2^-62 below 2^6,
    // kept as floor(m * m * 2^30) and the 32 bits under it
    unsigned long long sob_q, sob_sq_hi, sob_sq_lo;
    long long lap_abs, lap_sum, lap_sq;  // exact integer Laplacian sums
    double dct_total, dct_low, dct_mid, dct_high, dct_abs;
    double glcm[4][6];                   // per angle: contrast, dissimilarity, homogeneity, energy, correlation, ASM
};

// RGB -> HSV hue, 8-bit: h * hdiv[d], hdiv[i] = round((180 << 12) / (6 i))
__device__ __forceinline__ int hsv_hue(const int* hdiv, int r, int g, int b) {
    int v = max(r, max(g, b));
    const int d = v - min(r, min(g, b));
    int h;
    if (v == r) h = g - b;
    else if (v == g) h = b - r + 2 * d;
    else h = r - g + 4 * d;
    h = (h * hdiv[d] + 2048) >> 12;
    return h < 0 ? h + 180 : h;
}

__global__ void __launch_bounds__(NT) fe_point_kernel(const uint8_t* __restrict__ frames, long long npix,
                                                      const ColorTables* __restrict__ ct, uint8_t* __restrict__ gray,
                                                      float* __restrict__ grayf, FeAccum* __restrict__ acc) {
    __shared__ unsigned sh[H_COUNT][256];
    __shared__ int hdiv[256];
    const int b = blockIdx.y;
    for (int i = threadIdx.x; i < H_COUNT * 256; i += NT) (&sh[0][0])[i] = 0;
    for (int i = threadIdx.x; i < 256; i += NT) hdiv[i] = i ? __double2int_rn(737280.0 / (6.0 * (double)i)) : 0;
    __syncthreads();
    const uint8_t* f = frames + (size_t)b * npix * 3;
    uint8_t* gp = gray + (size_t)b * npix;
    float* gf = grayf + (size_t)b * npix;
    for (long long i = (long long)blockIdx.x * NT + threadIdx.x; i < npix; i += (long long)gridDim.x * NT) {
        const int r = f[i * 3], g = f[i * 3 + 1], bl = f[i * 3 + 2];
        int L, A, B;
        rgb_to_lab(ct->gamma_in, ct->cbrt_tab, r, g, bl, L, A, B);
        const int gq = gray_q15(r, g, bl);
        atomicAdd(&sh[H_L][L], 1u); atomicAdd(&sh[H_A][A], 1u); atomicAdd(&sh[H_B][B], 1u);
        atomicAdd(&sh[H_H][hsv_hue(hdiv, r, g, bl)], 1u);
        atomicAdd(&sh[H_S][hsv_saturation(ct->sdiv, r, g, bl)], 1u);
        atomicAdd(&sh[H_V][max(r, max(g, bl))], 1u);
        atomicAdd(&sh[H_R][r], 1u); atomicAdd(&sh[H_G][g], 1u); atomicAdd(&sh[H_BL][bl], 1u);
        atomicAdd(&sh[H_GRAY][gq], 1u);
        gp[i] = (uint8_t)gq;
        gf[i] = (float)gq;               // FE:132: the DCT runs on gray in 0..255
    }
    __syncthreads();
    unsigned* gh = &acc[b].hist[0][0];
    for (int i = threadIdx.x; i < H_COUNT * 256; i += NT) {
        const unsigned v = (&sh[0][0])[i];
        if (v) atomicAdd(&gh[i], v);
    }
}

// ---------------------------------------------------------------- 3x3 neighbourhood features
// 32 x 32 tile + 1 halo.  Sobel and Laplacian see BORDER_REFLECT_101 (cv2 default); the LBP taps see
// constant 0 outside the image (skimage mode 'C').
constexpr int ST = 32;

struct LbpTaps { double rp[8], cp[8]; };   // (-sin, cos)(2 pi p / 8) rounded to 5 decimals (host)

__global__ void __launch_bounds__(NT) fe_stencil_kernel(const uint8_t* __restrict__ gray, int H, int W, LbpTaps taps,
                                                        FeAccum* __restrict__ acc) {
    __shared__ int sg[ST + 2][ST + 2];
    __shared__ unsigned s_lbp[10];
    __shared__ unsigned long long s_red[NT / 32][3];
    __shared__ long long s_lred[NT / 32][3];
    __shared__ float s_max[NT / 32];
    const int b = blockIdx.z;
    const uint8_t* g = gray + (size_t)b * H * W;
    const int ty0 = blockIdx.y * ST, tx0 = blockIdx.x * ST;
    if (threadIdx.x < 10) s_lbp[threadIdx.x] = 0;
    for (int i = threadIdx.x; i < (ST + 2) * (ST + 2); i += NT) {
        const int ly = i / (ST + 2), lx = i - ly * (ST + 2);
        const int y = reflect101(min(ty0 + ly - 1, H), H), x = reflect101(min(tx0 + lx - 1, W), W);
        sg[ly][lx] = g[(size_t)y * W + x];
    }
    __syncthreads();
    unsigned long long sob = 0, sob2_hi = 0, sob2_lo = 0;
    float smax = 0.f;
    long long la = 0, ls = 0, lq = 0;
    for (int i = threadIdx.x; i < ST * ST; i += NT) {
        const int ly = i / ST, lx = i - ly * ST;
        const int y = ty0 + ly, x = tx0 + lx;
        if (y >= H || x >= W) continue;
        const int a00 = sg[ly][lx], a01 = sg[ly][lx + 1], a02 = sg[ly][lx + 2];
        const int a10 = sg[ly + 1][lx], a11 = sg[ly + 1][lx + 1], a12 = sg[ly + 1][lx + 2];
        const int a20 = sg[ly + 2][lx], a21 = sg[ly + 2][lx + 1], a22 = sg[ly + 2][lx + 2];
        // Sobel on gray / 255 in float32 (FE:175-177): the integer gradient is exact, one rounding for
        // the scale, float32 squares, sum and root like the NumPy expression
        const float dx = __fdiv_rn((float)((a02 + 2 * a12 + a22) - (a00 + 2 * a10 + a20)), 255.0f);
        const float dy = __fdiv_rn((float)((a20 + 2 * a21 + a22) - (a00 + 2 * a01 + a02)), 255.0f);
        const float m = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
        {
            const double md = (double)m, sq = md * md * 1073741824.0;            // m * m * 2^30, exact
            const double hi = floor(sq);
            sob += (unsigned long long)(md * 2147483648.0);                       // m * 2^31, an integer
            sob2_hi += (unsigned long long)hi;
            sob2_lo += (unsigned long long)((sq - hi) * 4294967296.0);
        }
        smax = fmaxf(smax, m);
        // Laplacian ksize=3: [2 0 2; 0 -8 0; 2 0 2], exact integers
        const int lap = 2 * (a00 + a02 + a20 + a22) - 8 * a11;
        la += abs(lap); ls += lap; lq += (long long)lap * lap;
        // LBP(8, 1, 'uniform'): bilinear taps in double, constant 0 outside
        const bool up = y > 0, dn = y + 1 < H, lf = x > 0, rt = x + 1 < W;
        const double n00 = (up && lf) ? a00 : 0, n01 = up ? a01 : 0, n02 = (up && rt) ? a02 : 0;
        const double n10 = lf ? a10 : 0, n12 = rt ? a12 : 0;
        const double n20 = (dn && lf) ? a20 : 0, n21 = dn ? a21 : 0, n22 = (dn && rt) ? a22 : 0;
        const double c = a11;
        unsigned bits = 0;
#pragma unroll
        for (int p = 0; p < 8; ++p) {
            double t;
            if (p == 0) t = n12; else if (p == 2) t = n01; else if (p == 4) t = n10; else if (p == 6) t = n21;
            else {
                // diagonal tap: fractional position in both directions
                const double r = __dadd_rn((double)y, taps.rp[p]), cc = __dadd_rn((double)x, taps.cp[p]);
                const double r0 = floor(r), c0 = floor(cc);
                const double dr = __dsub_rn(r, r0), dc = __dsub_rn(cc, c0);
                const bool rup = r0 < (double)y, cleft = c0 < (double)x;     // which 2x2 block of the 3x3 window
                const double tl = rup ? (cleft ? n00 : n01) : (cleft ? n10 : c);
                const double tr = rup ? (cleft ? n01 : n02) : (cleft ? c : n12);
                const double bl = rup ? (cleft ? n10 : c) : (cleft ? n20 : n21);
                const double br = rup ? (cleft ? c : n12) : (cleft ? n21 : n22);
                const double top = __dadd_rn(__dmul_rn(__dsub_rn(1.0, dc), tl), __dmul_rn(dc, tr));
                const double bot = __dadd_rn(__dmul_rn(__dsub_rn(1.0, dc), bl), __dmul_rn(dc, br));
                t = __dadd_rn(__dmul_rn(__dsub_rn(1.0, dr), top), __dmul_rn(dr, bot));
            }
            if (__dsub_rn(t, c) >= 0.0) bits |= 1u << p;
        }
        const int changes = __popc((bits ^ (bits >> 1)) & 0x7fu);             // the 7 adjacent pairs, not circular
        atomicAdd(&s_lbp[changes <= 2 ? __popc(bits) : 9], 1u);
    }
    // block reduction
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sob += __shfl_xor_sync(0xffffffffu, sob, o); sob2_hi += __shfl_xor_sync(0xffffffffu, sob2_hi, o);
        sob2_lo += __shfl_xor_sync(0xffffffffu, sob2_lo, o);
        smax = fmaxf(smax, __shfl_xor_sync(0xffffffffu, smax, o));
        la += __shfl_xor_sync(0xffffffffu, la, o); ls += __shfl_xor_sync(0xffffffffu, ls, o); lq += __shfl_xor_sync(0xffffffffu, lq, o);
    }
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) {
        s_red[w][0] = sob; s_red[w][1] = sob2_hi; s_red[w][2] = sob2_lo; s_max[w] = smax;
        s_lred[w][0] = la; s_lred[w][1] = ls; s_lred[w][2] = lq;
    }
    __syncthreads();
    FeAccum& A = acc[b];
    if (threadIdx.x == 0) {
        unsigned long long t0 = 0, t1 = 0, t2 = 0; float mx = 0.f; long long l0 = 0, l1 = 0, l2 = 0;
        for (int k = 0; k < NT / 32; ++k) {
            t0 += s_red[k][0]; t1 += s_red[k][1]; t2 += s_red[k][2]; mx = fmaxf(mx, s_max[k]);
            l0 += s_lred[k][0]; l1 += s_lred[k][1]; l2 += s_lred[k][2];
        }
        atomicAdd(&A.sob_q, t0); atomicAdd(&A.sob_sq_hi, t1); atomicAdd(&A.sob_sq_lo, t2);
        atomicMax(&A.sob_max_bits, __float_as_uint(mx));
        atomicAdd((unsigned long long*)&A.lap_abs, (unsigned long long)l0);
        atomicAdd((unsigned long long*)&A.lap_sum, (unsigned long long)l1);
        atomicAdd((unsigned long long*)&A.lap_sq, (unsigned long long)l2);
    }
    if (threadIdx.x < 10 && s_lbp[threadIdx.x]) atomicAdd(&A.lbp[threadIdx.x], s_lbp[threadIdx.x]);
}

// ---------------------------------------------------------------- GLCM of the 128 x 128 thumbnail
// cv2.resize(gray, (128, 128)) = INTER_LINEAR in 11-bit fixed point (exact 2x decimation
// becomes INTER_AREA), then graycomatrix(d = 1, 4 angles, symmetric, normed) and graycoprops.
struct ThumbTap { int s0, s1, c1; };
__device__ __forceinline__ ThumbTap thumb_tap(int i, double scale, int n, bool vertical) {
    float f = (float)(((double)i + 0.5) * scale - 0.5);
    int s = (int)floorf(f);
    f -= (float)s;
    ThumbTap t;
    if (vertical) { t.s0 = min(max(s, 0), n - 1); t.s1 = min(max(s + 1, 0), n - 1); }
    else {
        if (s < 0) { s = 0; f = 0.f; }
        if (s >= n - 1) { s = n - 1; f = 0.f; }
        t.s0 = s; t.s1 = min(s + 1, n - 1);
    }
    t.c1 = __float2int_rn(f * 2048.0f);
    return t;
}

__global__ void __launch_bounds__(NT) fe_thumb_kernel(const uint8_t* __restrict__ gray, int H, int W, uint8_t* __restrict__ thumb) {
    const int b = blockIdx.y;
    const int i = blockIdx.x * NT + threadIdx.x;
    if (i >= kThumb * kThumb) return;
    const int oy = i / kThumb, ox = i - oy * kThumb;
    const uint8_t* g = gray + (size_t)b * H * W;
    int v;
    if (H == 2 * kThumb && W == 2 * kThumb) {
        const uint8_t* q = g + (size_t)(2 * oy) * W + 2 * ox;
        v = (q[0] + q[1] + q[W] + q[W + 1] + 2) >> 2;
    } else {
        const ThumbTap ty = thumb_tap(oy, (double)H / kThumb, H, true), tx = thumb_tap(ox, (double)W / kThumb, W, false);
        const int s0 = g[(size_t)ty.s0 * W + tx.s0] * (2048 - tx.c1) + g[(size_t)ty.s0 * W + tx.s1] * tx.c1;
        const int s1 = g[(size_t)ty.s1 * W + tx.s0] * (2048 - tx.c1) + g[(size_t)ty.s1 * W + tx.s1] * tx.c1;
        v = ((((2048 - ty.c1) * (s0 >> 4)) >> 16) + ((ty.c1 * (s1 >> 4)) >> 16) + 2) >> 2;
    }
    thumb[(size_t)b * kThumb * kThumb + i] = (uint8_t)min(max(v, 0), 255);
}

// counts[b][angle][i][j]: pairs (i at (r, c), j at (r + dr, c + dc)) plus their mirror (symmetric = True)
__global__ void __launch_bounds__(NT) fe_glcm_count_kernel(const uint8_t* __restrict__ thumb, unsigned* __restrict__ counts) {
    const int b = blockIdx.y;
    const int i = blockIdx.x * NT + threadIdx.x;
    if (i >= kThumb * kThumb) return;
    const int r = i / kThumb, c = i - r * kThumb;
    const uint8_t* t = thumb + (size_t)b * kThumb * kThumb;
    const int a = t[i];
    const int dr[4] = {0, 1, 1, 1}, dc[4] = {1, 1, 0, -1};       // (round(sin), round(cos)) of 0, pi/4, pi/2, 3pi/4
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int r2 = r + dr[k], c2 = c + dc[k];
        if (r2 < 0 || r2 >= kThumb || c2 < 0 || c2 >= kThumb) continue;
        const int bb = t[r2 * kThumb + c2];
        unsigned* m = counts + ((size_t)b * 4 + k) * 65536;
        atomicAdd(&m[a * 256 + bb], 1u);
        atomicAdd(&m[bb * 256 + a], 1u);
    }
}

__device__ __forceinline__ double block_sum(double v, double* scratch) {
    v = warp_sum_f64(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0;
    for (int k = 0; k < NT / 32; ++k) t += scratch[k];
    return t;
}

// one block per (frame, angle): graycoprops of the normalised matrix (FE:115-118)
__global__ void __launch_bounds__(NT) fe_glcm_props_kernel(const unsigned* __restrict__ counts, FeAccum* __restrict__ acc) {
    __shared__ double scratch[NT / 32];
    const int k = blockIdx.x, b = blockIdx.y;
    const unsigned* m = counts + ((size_t)b * 4 + k) * 65536;
    double tot = 0;
    for (int e = threadIdx.x; e < 65536; e += NT) tot += (double)m[e];
    tot = block_sum(tot, scratch);
    double con = 0, dis = 0, hom = 0, asm_ = 0, mi = 0, mj = 0;
    for (int e = threadIdx.x; e < 65536; e += NT) {
        const unsigned cnt = m[e];
        if (!cnt) continue;
        const double p = (double)cnt / tot;
        const int i = e >> 8, j = e & 255;
        const double d = (double)(i - j);
        con += p * d * d; dis += p * fabs(d); hom += p / (1.0 + d * d); asm_ += p * p;
        mi += (double)i * p; mj += (double)j * p;
    }
    con = block_sum(con, scratch); dis = block_sum(dis, scratch); hom = block_sum(hom, scratch);
    asm_ = block_sum(asm_, scratch); mi = block_sum(mi, scratch); mj = block_sum(mj, scratch);
    double vi = 0, vj = 0, cov = 0;
    for (int e = threadIdx.x; e < 65536; e += NT) {
        const unsigned cnt = m[e];
        if (!cnt) continue;
        const double p = (double)cnt / tot;
        const double di = (double)(e >> 8) - mi, dj = (double)(e & 255) - mj;
        vi += p * di * di; vj += p * dj * dj; cov += p * di * dj;
    }
    vi = block_sum(vi, scratch); vj = block_sum(vj, scratch); cov = block_sum(cov, scratch);
    if (threadIdx.x == 0) {
        const double si = sqrt(vi), sj = sqrt(vj);
        double* o = acc[b].glcm[k];
        o[0] = con; o[1] = dis; o[2] = hom; o[3] = sqrt(asm_);
        o[4] = (si < 1e-15 || sj < 1e-15) ? 1.0 : cov / (si * sj);
        o[5] = asm_;
    }
}

// ---------------------------------------------------------------- DCT
__global__ void fe_copy_dct_kernel(const double* __restrict__ sums, FeAccum* __restrict__ acc, int B) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    FeAccum& A = acc[b];
    A.dct_total = sums[b * 5]; A.dct_low = sums[b * 5 + 1]; A.dct_mid = sums[b * 5 + 2];
    A.dct_high = sums[b * 5 + 3]; A.dct_abs = sums[b * 5 + 4];
}

// ---------------------------------------------------------------- final: 79 values per frame, float64
struct HStat { double mean, m2, m3, m4, mad; int vmin, vmax; };

// moments of the value v * scale (scale = 1 for Lab / HSV planes, float32(v / 255) otherwise) under histogram h
__device__ HStat hist_stats(const unsigned* h, const double* val, double n) {
    HStat s;
    double a = 0;
    s.vmin = 255; s.vmax = 0;
    for (int v = 0; v < 256; ++v) if (h[v]) { a += (double)h[v] * val[v]; if (v < s.vmin) s.vmin = v; if (v > s.vmax) s.vmax = v; }
    s.mean = a / n;
    double m2 = 0, m3 = 0, m4 = 0, mad = 0;
    for (int v = 0; v < 256; ++v) if (h[v]) {
        const double d = val[v] - s.mean, c = (double)h[v], d2 = d * d;
        m2 += c * d2; m3 += c * d2 * d; m4 += c * d2 * d2; mad += c * fabs(d);
    }
    s.m2 = m2 / n; s.m3 = m3 / n; s.m4 = m4 / n; s.mad = mad / n;
    return s;
}

__global__ void __launch_bounds__(32) fe_final_kernel(const FeAccum* __restrict__ acc, long long npix, RankPlan p25, RankPlan p75,
                                                      double* __restrict__ feat) {
    __shared__ double v_int[256], v_unit[256];
    __shared__ HStat st[H_COUNT + 1];                       // + saturation / 255 (FE:235-241)
    const int b = blockIdx.x, t = threadIdx.x;
    const FeAccum& A = acc[b];
    const double n = (double)npix;
    for (int v = t; v < 256; v += 32) { v_int[v] = (double)v; v_unit[v] = (double)__fdiv_rn((float)v, 255.0f); }
    __syncwarp();
    if (t < H_COUNT) st[t] = hist_stats(A.hist[t], t >= H_R ? v_unit : v_int, n);
    else if (t == H_COUNT) st[t] = hist_stats(A.hist[H_S], v_unit, n);
    __syncwarp();
    if (t != 0) return;
    double* f = feat + (size_t)b * kFe;
    int k = 0;
    // ---- colour (FE:17-77)
    for (int c = H_L; c <= H_B; ++c) {
        const HStat& s = st[c];
        f[k++] = s.mean; f[k++] = sqrt(s.m2);
        f[k++] = s.m3 / (s.m2 * sqrt(s.m2));                // scipy.stats.skew (biased)
        f[k++] = s.m4 / (s.m2 * s.m2) - 3.0;                // scipy.stats.kurtosis (Fisher, biased)
    }
    for (int c = H_H; c <= H_V; ++c) { f[k++] = st[c].mean; f[k++] = sqrt(st[c].m2); }
    {
        const double ma = st[H_A].mean, mb = st[H_B].mean;
        const double M = sqrt(ma * ma + mb * mb);
        const double D = sqrt(st[H_A].mad * st[H_A].mad + st[H_B].mad * st[H_B].mad);
        f[k++] = M / (D + 1e-10); f[k++] = M; f[k++] = D; f[k++] = ma; f[k++] = mb;
    }
    for (int c = H_R; c <= H_BL; ++c) {
        f[k++] = st[c].mean; f[k++] = sqrt(st[c].m2); f[k++] = v_unit[st[c].vmin]; f[k++] = v_unit[st[c].vmax];
    }
    // ---- texture (FE:80-120)
    {
        double tot = 0;
        for (int i = 0; i < 10; ++i) tot += (double)A.lbp[i];
        for (int i = 0; i < 10; ++i) f[k++] = (double)A.lbp[i] / tot;       // density, unit-width bins
        for (int p = 0; p < 6; ++p) {
            double m = 0, v = 0;
            for (int a = 0; a < 4; ++a) m += A.glcm[a][p];
            m /= 4.0;
            for (int a = 0; a < 4; ++a) v += (A.glcm[a][p] - m) * (A.glcm[a][p] - m);
            f[k++] = m; f[k++] = sqrt(v / 4.0);
        }
    }
    // ---- frequency (FE:123-158)
    {
        const double mean_abs = A.dct_abs / n;
        f[k++] = A.dct_low / A.dct_total; f[k++] = A.dct_mid / A.dct_total; f[k++] = A.dct_high / A.dct_total;
        f[k++] = mean_abs; f[k++] = sqrt(fmax(A.dct_total / n - mean_abs * mean_abs, 0.0));
    }
    // ---- edge (FE:161-200)
    {
        const double sob_sum = (double)A.sob_q * (1.0 / 2147483648.0);
        const double sob_sq = (double)A.sob_sq_hi * (1.0 / 1073741824.0) + (double)A.sob_sq_lo * (1.0 / 4611686018427387904.0);
        const double sm = sob_sum / n;
        f[k++] = sm; f[k++] = sqrt(fmax(sob_sq / n - sm * sm, 0.0)); f[k++] = (double)__uint_as_float(A.sob_max_bits);
        f[k++] = (double)A.canny / n;
        const double lm = (double)A.lap_sum / n, lv = (double)A.lap_sq / n - lm * lm;
        f[k++] = (double)A.lap_abs / n; f[k++] = sqrt(fmax(lv, 0.0)); f[k++] = lv;
    }
    // ---- quality (FE:203-246)
    {
        const HStat& g = st[H_GRAY];
        const unsigned* h = A.hist[H_GRAY];
        double ent = 0;
        for (int v = 0; v < 256; ++v) if (h[v]) { const double p = (double)h[v] / n; ent -= p * log(p); }
        const long long k0 = (npix - 1) / 2, k1 = npix / 2;
        const float m0 = (float)v_unit[hist_rank(h, k0)], m1 = (float)v_unit[hist_rank(h, k1)];
        f[k++] = sqrt(g.m2); f[k++] = ent / log(2.0); f[k++] = g.mean;
        f[k++] = (double)(k0 == k1 ? m0 : __fdiv_rn(__fadd_rn(m0, m1), 2.0f));        // np.median
        f[k++] = (double)lut_percentile<float>(h, v_unit, npix, p25);
        f[k++] = (double)lut_percentile<float>(h, v_unit, npix, p75);
        f[k++] = (double)__fsub_rn((float)v_unit[g.vmax], (float)v_unit[g.vmin]);
        f[k++] = st[H_COUNT].mean; f[k++] = sqrt(st[H_COUNT].m2);
        f[k++] = sqrt(g.m2);                                                       // RMS contrast == std
    }
}

__global__ void fe_copy_canny_kernel(const unsigned* __restrict__ counts, FeAccum* __restrict__ acc, int B) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < B) acc[b].canny = counts[b];
}

struct FeLayout { size_t ctab, acc, canny_counts, dct_sums, dct_part, gray, grayf, tmp, thumb, glcm, labels, tile_list, tile_n, total; };
FeLayout fe_layout(int B, int H, int W) {
    FeLayout L;
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t at = off; off += (bytes + 255) & ~(size_t)255; return at; };
    const size_t N = (size_t)H * W;
    L.ctab = take(sizeof(ColorTables));
    L.acc = take((size_t)B * sizeof(FeAccum));
    L.canny_counts = take((size_t)B * 4);
    L.dct_sums = take((size_t)B * 5 * 8);
    L.dct_part = take((size_t)B * dct_energy_blocks(H, W) * 5 * 8);
    L.gray = take(B * N);
    L.grayf = take(B * N * 4);
    L.tmp = take(B * N * 4);
    L.thumb = take((size_t)B * kThumb * kThumb);
    L.glcm = take((size_t)B * 4 * 65536 * 4);
    const size_t slots = canny_tile_slots(H, W, B);
    L.labels = take(canny_label_bytes(slots));
    L.tile_list = take(slots * kCannyListCap * sizeof(int2));
    L.tile_n = take(slots * 4);
    L.total = off;
    return L;
}

#define CUF(call)                                                                                  \
    do {                                                                                           \
        cudaError_t _e = (call);                                                                   \
        if (_e != cudaSuccess) return uwie_fail(UWIE_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(_e)); \
    } while (0)
const ColorTables& host_tables() {
    static ColorTables t;
    static std::once_flag once;
    std::call_once(once, [] { build_color_tables(&t); });
    return t;
}

}  // namespace

extern "C" {

size_t uwie_features79_workspace_bytes(int batch, int height, int width) {
    if (batch < 1 || height < 1 || width < 1) return 0;
    return fe_layout(batch, height, width).total;
}

int uwie_features79(const uint8_t* frames_dev, int batch, int height, int width, double* feat_dev,
                    void* workspace_dev, size_t workspace_bytes, void* stream) {
    if (!frames_dev || !feat_dev || batch < 1 || height < 2 || width < 2)
        return uwie_fail(UWIE_ERR_ARG, "uwie_features79: bad arguments (frames of at least 2 x 2)");
    if ((long long)height * width >= (1LL << 30)) return uwie_fail(UWIE_ERR_ARG, "uwie_features79: frame too large");
    const FeLayout L = fe_layout(batch, height, width);
    if (!workspace_dev || workspace_bytes < L.total) return uwie_fail(UWIE_ERR_ARG, "uwie_features79: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)workspace_dev;
    const int B = batch, H = height, W = width;
    const long long N = (long long)H * W;
    ColorTables* ctab = (ColorTables*)(ws + L.ctab);
    FeAccum* acc = (FeAccum*)(ws + L.acc);
    unsigned* canny_counts = (unsigned*)(ws + L.canny_counts);
    uint8_t* gray = (uint8_t*)(ws + L.gray);
    float* grayf = (float*)(ws + L.grayf);
    float* tmp = (float*)(ws + L.tmp);
    double* dct_sums = (double*)(ws + L.dct_sums);
    uint8_t* thumb = (uint8_t*)(ws + L.thumb);
    unsigned* glcm = (unsigned*)(ws + L.glcm);

    CUF(cudaMemcpyAsync(ctab, &host_tables(), sizeof(ColorTables), cudaMemcpyHostToDevice, st));
    CUF(cudaMemsetAsync(acc, 0, (size_t)B * sizeof(FeAccum), st));
    CUF(cudaMemsetAsync(canny_counts, 0, (size_t)B * 4, st));
    CUF(cudaMemsetAsync(dct_sums, 0, (size_t)B * 5 * 8, st));
    CUF(cudaMemsetAsync(glcm, 0, (size_t)B * 4 * 65536 * 4, st));

    int pb = (int)((N + NT * 8 - 1) / (NT * 8));
    if (pb > 148 * 4) pb = 148 * 4;
    fe_point_kernel<<<dim3(pb, B), NT, 0, st>>>(frames_dev, N, ctab, gray, grayf, acc);

    LbpTaps taps;
    for (int p = 0; p < 8; ++p) {
        const double a = 2.0 * M_PI * (double)p / 8.0;
        taps.rp[p] = std::round(-std::sin(a) * 1e5) / 1e5;     // np.round(x, 5)
        taps.cp[p] = std::round(std::cos(a) * 1e5) / 1e5;
    }
    fe_stencil_kernel<<<dim3((W + ST - 1) / ST, (H + ST - 1) / ST, B), NT, 0, st>>>(gray, H, W, taps, acc);

    CannyArgs ca;
    std::memset(&ca, 0, sizeof(ca));
    ca.mode = CANNY_PLANES; ca.H = H; ca.W = W; ca.low = 50; ca.high = 150;
    ca.gray = gray; ca.planes_per_img = 1;
    ca.labels = (int*)(ws + L.labels); ca.tile_list = (int2*)(ws + L.tile_list); ca.tile_n = (int*)(ws + L.tile_n);
    ca.counts = canny_counts;
    launch_canny(ca, H, W, B, st);
    fe_copy_canny_kernel<<<(B + 127) / 128, 128, 0, st>>>(canny_counts, acc, B);

    fe_thumb_kernel<<<dim3(kThumb * kThumb / NT, B), NT, 0, st>>>(gray, H, W, thumb);
    fe_glcm_count_kernel<<<dim3(kThumb * kThumb / NT, B), NT, 0, st>>>(thumb, glcm);
    fe_glcm_props_kernel<<<dim3(4, B), NT, 0, st>>>(glcm, acc);

    // DCT (k_dct.cu): D_H * G * D_W^T as hand-written float32 GEMMs with the even / odd butterfly; the coefficient plane is
    // never stored, its five sums land in dct_sums
    if (launch_dct_energy(grayf, tmp, B, H, W, dct_sums, (double*)(ws + L.dct_part), st) < 0)
        return uwie_fail(UWIE_ERR_NOMEM, "uwie_features79: the cached DCT matrices could not be allocated");
    fe_copy_dct_kernel<<<(B + 127) / 128, 128, 0, st>>>(dct_sums, acc, B);

    const PercentilePlan q25 = percentile_plan_f32(N, 25.0), q75 = percentile_plan_f32(N, 75.0);
    RankPlan r25, r75;
    r25.k = q25.k; r25.gamma = q25.gamma; r75.k = q75.k; r75.gamma = q75.gamma;
    fe_final_kernel<<<B, 32, 0, st>>>(acc, N, r25, r75, feat_dev);
    CUF(cudaGetLastError());
    return UWIE_OK;
}

}  // extern "C"
