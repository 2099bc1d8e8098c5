// Full-frame 2-D DCT-II (cv2.dct of the float32 gray plane, feature_extraction.py:132-135) reduced on the fly to the
// five frequency features' sums (feature_extraction.py:142-156): energy of the blocks [:h/4, :w/4], [h/4:h/2, w/4:w/2],
// [h/2:, w/2:], total energy and the sum of |coefficient|.  Hand-written float32 SIMT GEMMs, no library call:
//
//   Z = D_H * G * D_W^T        D_n[k][i] = c_k cos(pi (2 i + 1) k / (2 n)),  c_0 = sqrt(1/n), c_k = sqrt(2/n)
//
// The cosine matrix is mirror-symmetric, D[k][n-1-i] = (-1)^k D[k][i], so for an even n the even rows of the transform
// only see g_i + g_{n-1-i} and the odd rows only g_i - g_{n-1-i} (the first butterfly of the fast DCT): each of the two
// products becomes two GEMMs of half the inner length — half the flops of the plain matrix form (6.2 instead of 12.4
// GFLOP at 1080p).  The butterflies are formed while the operand tiles are loaded; the second product never stores Z:
// its epilogue squares, classifies and sums the coefficients.  An odd dimension takes the plain (unsplit) form.
// The half matrices are built once per (device, n) and cached.
#include <cuda_runtime.h>

#include <cstdlib>
#include <map>
#include <mutex>
#include <utility>

#include "k_common.cuh"

namespace uwie {

namespace {

constexpr int BM = 128, BN = 128, BK = 8, TM = 8, TN = 8, GNT = 256;

// parity matrices of the orthonormal DCT-II of length n, split = 2: A[p][k][i] = D[2k + p][i], k < ceil((n - p) / 2),
// i < n / 2, stored [p][n/2][n/2] row-major, and the transposed copy Bt[p][i][k];  split = 1: the full matrix.
__global__ void dct_matrix_kernel(float* __restrict__ A, float* __restrict__ Bt, int n, int split) {
    const int half = n / split;
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long long)split * half * half) return;
    const int p = (int)(e / ((long long)half * half));
    const int r = (int)(e - (long long)p * half * half);
    const int k = r / half, i = r - k * half;
    const int row = split * k + p;
    // reduce the angle exactly in integers before the cosine: (2i + 1) row mod 4n
    const long long q = ((2LL * i + 1) * row) % (4LL * n);
    const double ck = row == 0 ? sqrt(1.0 / n) : sqrt(2.0 / n);
    const float v = (float)(ck * cospi((double)q / (2.0 * n)));
    A[e] = v;
    Bt[(size_t)p * half * half + (size_t)i * half + k] = v;
}

struct DctArgs {
    const float* G;      // [B][H][W]
    float* T;            // [B][H][W]   D_H * G, natural row order
    const float* Ah;     // parity matrices of length H ([split][H/split][H/split])
    const float* Btw;    // transposed parity matrices of length W
    double* sums;        // [B][5]: total, low, mid, high, sum |z|
    double* part;        // [B][blocks of step 2][5]: per-block sums, added in block order by dct_reduce_kernel
    int H, W, split_h, split_w;
};

// C[M x N] = A[M x K] * B[K x N], 128 x 128 tile per block, 8 x 8 per thread, operands through functors.
// load_a4(m, k) / load_b4(k, n) return 4 consecutive k / n as one float4 (128-bit loads; the mirrored operand is one
// reversed float4) and are used wherever the four elements are in range and `vec` says the rows are 16-byte aligned;
// the scalar functors cover the ragged edges.
template <class LoadA, class LoadB, class LoadA4, class LoadB4, class Epilogue>
__device__ __forceinline__ void gemm_tile(int M, int N, int K, int m0, int n0, bool vec, LoadA load_a, LoadB load_b,
                                          LoadA4 load_a4, LoadB4 load_b4, Epilogue epi) {
    __shared__ __align__(16) float sA[2][BK][BM + 4];
    __shared__ __align__(16) float sB[2][BK][BN + 4];
    const int t = threadIdx.x;
    const int tx = t & 15, ty = t >> 4;                 // 16 x 16 threads, each 8 x 8 outputs
    const int a_row = t >> 1, a_k = (t & 1) * 4;        // A tile: 128 rows x 8 k, 4 consecutive k per thread
    const int b_k = t >> 5, b_n = (t & 31) * 4;         // B tile: 8 k x 128 n, 4 consecutive n per thread
    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;
    float ra[4], rb[4];
    auto fetch = [&](int k0) {
        const int m = m0 + a_row, ka = k0 + a_k;
        if (vec && m < M && ka + 3 < K) {
            const float4 v = load_a4(m, ka);
            ra[0] = v.x; ra[1] = v.y; ra[2] = v.z; ra[3] = v.w;
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) ra[j] = (m < M && ka + j < K) ? load_a(m, ka + j) : 0.f;
        }
        const int kb = k0 + b_k, n = n0 + b_n;
        if (vec && kb < K && n + 3 < N) {
            const float4 v = load_b4(kb, n);
            rb[0] = v.x; rb[1] = v.y; rb[2] = v.z; rb[3] = v.w;
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) rb[j] = (kb < K && n + j < N) ? load_b(kb, n + j) : 0.f;
        }
    };
    auto stash = [&](int buf) {
#pragma unroll
        for (int j = 0; j < 4; ++j) sA[buf][a_k + j][a_row] = ra[j];
        *reinterpret_cast<float4*>(&sB[buf][b_k][b_n]) = make_float4(rb[0], rb[1], rb[2], rb[3]);
    };
    fetch(0);
    stash(0);
    __syncthreads();
    int buf = 0;
    for (int k0 = 0; k0 < K; k0 += BK) {
        const bool more = k0 + BK < K;
        if (more) fetch(k0 + BK);
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            float a[TM], b[TN];
            // a thread's 8 rows / columns are two groups of 4, 64 apart: the 16 lanes of a row of threads then read 16
            // consecutive float4 (256 contiguous bytes) instead of every other one: no shared-memory bank conflicts
            const float4 a0 = *reinterpret_cast<const float4*>(&sA[buf][kk][ty * 4]);
            const float4 a1 = *reinterpret_cast<const float4*>(&sA[buf][kk][64 + ty * 4]);
            const float4 b0 = *reinterpret_cast<const float4*>(&sB[buf][kk][tx * 4]);
            const float4 b1 = *reinterpret_cast<const float4*>(&sB[buf][kk][64 + tx * 4]);
            a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
            b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
#pragma unroll
            for (int i = 0; i < TM; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        if (more) {
            stash(buf ^ 1);
            __syncthreads();
            buf ^= 1;
        }
    }
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            const int m = m0 + ty * 4 + (i & 3) + (i >> 2) * 64, n = n0 + tx * 4 + (j & 3) + (j >> 2) * 64;
            if (m < M && n < N) epi(m, n, acc[i][j]);
        }
}

__device__ __forceinline__ float4 ld4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float4 add4(float4 u, float4 v, float s) {
    return make_float4(__fadd_rn(u.x, s * v.x), __fadd_rn(u.y, s * v.y), __fadd_rn(u.z, s * v.z), __fadd_rn(u.w, s * v.w));
}

// step 1: T = D_H * G.  grid (ceil(W / 128), split_h * ceil(rows_p / 128), B); a row tile holds rows of ONE parity
__global__ void __launch_bounds__(GNT, 2) dct_rows_kernel(DctArgs a) {
    const int b = blockIdx.z;
    const int H = a.H, W = a.W, sp = a.split_h;
    const int half = H / sp;                            // inner length
    const int tiles_p = (half + BM - 1) / BM;
    const int p = blockIdx.y / tiles_p;
    const int m0 = (blockIdx.y - p * tiles_p) * BM;
    const float* G = a.G + (size_t)b * H * W;
    float* T = a.T + (size_t)b * H * W;
    const float* A = a.Ah + (size_t)p * half * half;
    const float sign = p ? -1.f : 1.f;
    const bool vec = (half & 3) == 0 && (W & 3) == 0 && ((reinterpret_cast<uintptr_t>(G) | reinterpret_cast<uintptr_t>(A)) & 15) == 0;
    gemm_tile(half, W, half, m0, blockIdx.x * BN, vec,
              [&](int m, int k) { return __ldg(&A[(size_t)m * half + k]); },
              [&](int k, int n) {
                  const float u = __ldg(&G[(size_t)k * W + n]);
                  return sp == 1 ? u : __fadd_rn(u, sign * __ldg(&G[(size_t)(H - 1 - k) * W + n]));
              },
              [&](int m, int k) { return ld4(&A[(size_t)m * half + k]); },
              [&](int k, int n) {
                  const float4 u = ld4(&G[(size_t)k * W + n]);
                  return sp == 1 ? u : add4(u, ld4(&G[(size_t)(H - 1 - k) * W + n]), sign);
              },
              [&](int m, int n, float v) { T[(size_t)(sp * m + p) * W + n] = v; });
}

// step 2: Z = T * D_W^T, reduced in the epilogue.  grid (split_w * ceil(cols_p / 128), ceil(H / 128), B)
__global__ void __launch_bounds__(GNT, 2) dct_cols_kernel(DctArgs a) {
    const int b = blockIdx.z;
    const int H = a.H, W = a.W, sp = a.split_w;
    const int half = W / sp;
    const int tiles_p = (half + BN - 1) / BN;
    const int p = blockIdx.x / tiles_p;
    const int n0 = (blockIdx.x - p * tiles_p) * BN;
    const float* T = a.T + (size_t)b * H * W;
    const float* Bt = a.Btw + (size_t)p * half * half;
    const float sign = p ? -1.f : 1.f;
    const int h4 = H / 4, h2 = H / 2, w4 = W / 4, w2 = W / 2;
    double tot = 0, lo = 0, mid = 0, hi = 0, ab = 0;
    const bool vec = (half & 3) == 0 && (W & 3) == 0 && ((reinterpret_cast<uintptr_t>(T) | reinterpret_cast<uintptr_t>(Bt)) & 15) == 0;
    gemm_tile(H, half, half, blockIdx.y * BM, n0, vec,
              [&](int m, int k) {
                  const float u = __ldg(&T[(size_t)m * W + k]);
                  return sp == 1 ? u : __fadd_rn(u, sign * __ldg(&T[(size_t)m * W + (W - 1 - k)]));
              },
              [&](int k, int n) { return __ldg(&Bt[(size_t)k * half + n]); },
              [&](int m, int k) {
                  const float4 u = ld4(&T[(size_t)m * W + k]);
                  if (sp == 1) return u;
                  const float4 r = ld4(&T[(size_t)m * W + (W - 4 - k)]);      // elements W-1-k-3 .. W-1-k, reversed below
                  return add4(u, make_float4(r.w, r.z, r.y, r.x), sign);
              },
              [&](int k, int n) { return ld4(&Bt[(size_t)k * half + n]); },
              [&](int y, int n, float v) {
                  const int x = sp * n + p;
                  const double s = (double)__fmul_rn(v, v);          // dct ** 2 is a float32 array (FE:142)
                  tot += s; ab += (double)fabsf(v);
                  if (y < h4 && x < w4) lo += s;
                  else if (y >= h4 && y < h2 && x >= w4 && x < w2) mid += s;
                  else if (y >= h2 && x >= w2) hi += s;
              });
    __shared__ double red[GNT / 32][5];
    double v5[5] = {tot, lo, mid, hi, ab};
#pragma unroll
    for (int i = 0; i < 5; ++i) v5[i] = warp_sum_f64(v5[i]);
    if ((threadIdx.x & 31) == 0) for (int i = 0; i < 5; ++i) red[threadIdx.x >> 5][i] = v5[i];
    __syncthreads();
    if (threadIdx.x < 5) {
        double s = 0;
        for (int w = 0; w < GNT / 32; ++w) s += red[w][threadIdx.x];
        const size_t blk = (size_t)b * gridDim.x * gridDim.y + (size_t)blockIdx.y * gridDim.x + blockIdx.x;
        a.part[blk * 5 + threadIdx.x] = s;
    }
}

// the block sums of one frame in a fixed order (lane l adds entries l, l + 32, ...; then a fixed shuffle tree): the five
// results do not depend on how the blocks were scheduled.  One warp per (frame, sum).
__global__ void dct_reduce_kernel(const double* __restrict__ part, int nblk, double* __restrict__ sums) {
    const int b = blockIdx.x, k = threadIdx.y, lane = threadIdx.x;
    double s = 0;
    for (int i = lane; i < nblk; i += 32) s += part[((size_t)b * nblk + i) * 5 + k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if (lane == 0) sums[(size_t)b * 5 + k] = s;
}

struct CacheEntry { float* A; float* Bt; int split; };
std::mutex g_mu;
std::map<std::pair<int, int>, CacheEntry> g_cache;       // (device, n) -> matrices

int matrices_for(int n, cudaStream_t st, CacheEntry* out) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 1;
    std::lock_guard<std::mutex> lock(g_mu);
    auto it = g_cache.find({dev, n});
    if (it == g_cache.end()) {
        CacheEntry e;
        e.split = (n % 2 == 0 && n >= 4) ? 2 : 1;
        const size_t half = (size_t)(n / e.split);
        const size_t bytes = (size_t)e.split * half * half * sizeof(float);
        if (cudaMalloc(&e.A, bytes) != cudaSuccess) return 1;
        if (cudaMalloc(&e.Bt, bytes) != cudaSuccess) { cudaFree(e.A); return 1; }
        const long long total = (long long)e.split * half * half;
        dct_matrix_kernel<<<(int)((total + 255) / 256), 256, 0, st>>>(e.A, e.Bt, n, e.split);
        // later calls may come on other streams: make the matrices visible before anybody else uses them
        if (cudaStreamSynchronize(st) != cudaSuccess) { cudaFree(e.A); cudaFree(e.Bt); return 1; }
        it = g_cache.emplace(std::make_pair(dev, n), e).first;
    }
    *out = it->second;
    return 0;
}

}  // namespace

// k_dct_fft.cu: the same five sums through shared-memory FFTs, for sizes that factor into 2, 3 and 5
bool dct_fft_usable(int H, int W);
int dct_fft_blocks(int W);
int launch_dct_fft(const float* grayf, float* tmp, int B, int H, int W, double* part, cudaStream_t st);

namespace {
// UWIE_DCT_GEMM=1 keeps the matrix form for every size (comparison runs, tests)
bool use_fft(int H, int W) {
    static const bool forced_gemm = [] { const char* v = std::getenv("UWIE_DCT_GEMM"); return v && v[0] == '1'; }();
    return !forced_gemm && dct_fft_usable(H, W);
}
}  // namespace

// blocks of the reducing pass per frame = entries of `part` per frame (the larger of the two forms: the choice can be
// overridden at run time)
int dct_energy_blocks(int H, int W) {
    const int split_w = (W % 2 == 0 && W >= 4) ? 2 : 1;
    const int gemm = split_w * ((W / split_w + BN - 1) / BN) * ((H + BM - 1) / BM);
    const int fft = dct_fft_usable(H, W) ? dct_fft_blocks(W) : 0;
    return gemm > fft ? gemm : fft;
}

// sums: [B][5] doubles (overwritten); part: [B][dct_energy_blocks(H, W)][5] doubles of scratch.
// Returns the number of kernels launched, or -1.
int launch_dct_energy(const float* grayf, float* tmp, int B, int H, int W, double* sums, double* part, cudaStream_t st) {
    if (use_fft(H, W)) {
        const int k = launch_dct_fft(grayf, tmp, B, H, W, part, st);
        if (k < 0) return -1;
        dct_reduce_kernel<<<B, dim3(32, 5), 0, st>>>(part, dct_fft_blocks(W), sums);
        return k + 1;
    }
    CacheEntry ch, cw;
    if (matrices_for(H, st, &ch) || matrices_for(W, st, &cw)) return -1;
    DctArgs a;
    a.G = grayf; a.T = tmp; a.Ah = ch.A; a.Btw = cw.Bt; a.sums = sums; a.part = part;
    a.H = H; a.W = W; a.split_h = ch.split; a.split_w = cw.split;
    const int half_h = H / ch.split, half_w = W / cw.split;
    dct_rows_kernel<<<dim3((W + BN - 1) / BN, ch.split * ((half_h + BM - 1) / BM), B), GNT, 0, st>>>(a);
    const dim3 g2(cw.split * ((half_w + BN - 1) / BN), (H + BM - 1) / BM, B);
    dct_cols_kernel<<<g2, GNT, 0, st>>>(a);
    dct_reduce_kernel<<<B, dim3(32, 5), 0, st>>>(part, (int)(g2.x * g2.y), sums);
    return 3;
}

}  // namespace uwie
