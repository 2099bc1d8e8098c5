// Full-frame 2-D DCT-II (cv2.dct of the float32 gray plane, feature_extraction.py:132-135) through FFTs in shared memory,
// for frame sizes whose factors are 2, 3 and 5 (every size of BASELINE.json: 1920x1080, 3840x2160, 1280x720, ...).
// The matrix form of k_dct.cu costs 6.2 GFLOP per 1080p frame; this form costs ~0.1 GFLOP and two passes over the plane.
//
//   length-N DCT-II from one N-point FFT (Makhoul):  v[n] = x[2n], v[N-1-n] = x[2n+1],  V = FFT(v),
//                                                    X[k] = c_k Re(e^{-i pi k / 2N} V[k]),  c_0 = sqrt(1/N), c_k = sqrt(2/N)
//   two real sequences per complex FFT:              z = v1 + i v2,  V1[k] = (Z[k] + conj Z[N-k]) / 2,  V2[k] = (Z[k] - conj Z[N-k]) / 2i
//   FFT: Stockham autosort, decimation in frequency, radices 4, 2, 3, 5, ping-pong between two shared-memory buffers;
//        stage (n, s, r), m = n / r:  y[q + s (r p + k)] = w_n^{pk} sum_j x[q + s (p + m j)] w_r^{jk},  p < m, q < s
//
// Pass 1: one block per pair of image rows  -> T (row transforms, natural layout).
// Pass 2: one block per pair of columns of T -> coefficients, never stored: squared, classified into the three frequency
//         blocks of feature_extraction.py:142-156 and summed per block; dct_reduce adds the block sums in block order.
// The arithmetic is float32 like cv2's (its own FFT-based DCT); the two differ by rounding only (~1e-6 relative on a
// coefficient), far inside the 1e-4 bar of the five frequency features.  tools/ has no copy of this: the NumPy mirror of the
// stage loop that was used to validate the index arithmetic is tests/test_dct_fft_mirror.py.
#include <cuda_runtime.h>

#include <map>
#include <mutex>
#include <utility>

#include "k_common.cuh"

namespace uwie {

namespace {

constexpr int FNT = 256;
constexpr int kMaxFactors = 14;
constexpr int kFftMaxN = 4096;

struct FftPlan {
    int n, nf;
    int fac[kMaxFactors];
};

bool make_plan(int n, FftPlan* p) {
    if (n < 8 || n > kFftMaxN) return false;
    p->n = n; p->nf = 0;
    int m = n;
    const int radices[4] = {4, 2, 3, 5};
    for (int r : radices)
        while (m % r == 0) {
            if (p->nf == kMaxFactors) return false;
            p->fac[p->nf++] = r;
            m /= r;
        }
    return m == 1;
}

__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cmul(float2 a, float2 b) { return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
__device__ __forceinline__ float2 cscale(float2 a, float s) { return make_float2(a.x * s, a.y * s); }
__device__ __forceinline__ float2 mul_neg_i(float2 a) { return make_float2(a.y, -a.x); }      // a * (-i)
__device__ __forceinline__ float2 mul_pos_i(float2 a) { return make_float2(-a.y, a.x); }      // a * (+i)

// W[t] = e^{-2 pi i t / N};  P[k] = c_k e^{-i pi k / (2 N)}
__global__ void fft_tables_kernel(float2* __restrict__ W, float2* __restrict__ P, int n) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    double s, c;
    sincospi(2.0 * (double)t / (double)n, &s, &c);
    W[t] = make_float2((float)c, (float)(-s));
    sincospi((double)t / (2.0 * (double)n), &s, &c);
    const double ck = t == 0 ? sqrt(1.0 / n) : sqrt(2.0 / n);
    P[t] = make_float2((float)(ck * c), (float)(-ck * s));
}

// All stages of one N-point FFT held in x (shared memory); returns the buffer that holds the result (natural order).
__device__ float2* fft_shared(float2* x, float2* y, const float2* __restrict__ W, const FftPlan& pl) {
    const int N = pl.n;
    int n = N, s = 1;
    for (int st = 0; st < pl.nf; ++st) {
        const int r = pl.fac[st];
        const int m = n / r;
        const int bf = N / r;                                     // butterflies of the stage = m * s
        for (int t = threadIdx.x; t < bf; t += FNT) {
            const int p = t / s, q = t - p * s;
            const float2* in = x + q + s * p;
            float2* out = y + q + s * r * p;
            const int ms = m * s, tw = p * s;                     // twiddle index of w_n^{p k} in the N-table: p k s < N
            if (r == 4) {
                const float2 a0 = in[0], a1 = in[ms], a2 = in[2 * ms], a3 = in[3 * ms];
                const float2 t0 = cadd(a0, a2), t1 = csub(a0, a2), t2 = cadd(a1, a3), t3 = mul_neg_i(csub(a1, a3));
                out[0] = cadd(t0, t2);
                out[s] = cmul(cadd(t1, t3), __ldg(W + tw));
                out[2 * s] = cmul(csub(t0, t2), __ldg(W + 2 * tw));
                out[3 * s] = cmul(csub(t1, t3), __ldg(W + 3 * tw));
            } else if (r == 2) {
                const float2 a0 = in[0], a1 = in[ms];
                out[0] = cadd(a0, a1);
                out[s] = cmul(csub(a0, a1), __ldg(W + tw));
            } else if (r == 3) {
                const float2 a0 = in[0], a1 = in[ms], a2 = in[2 * ms];
                const float2 t1 = cadd(a1, a2);
                const float2 t2 = make_float2(a0.x - 0.5f * t1.x, a0.y - 0.5f * t1.y);
                const float2 t3 = cscale(csub(a1, a2), 0.86602540378443864676f);
                out[0] = cadd(a0, t1);
                out[s] = cmul(cadd(t2, mul_neg_i(t3)), __ldg(W + tw));
                out[2 * s] = cmul(cadd(t2, mul_pos_i(t3)), __ldg(W + 2 * tw));
            } else {                                              // r == 5
                const float c1 = 0.30901699437494742410f, c2 = -0.80901699437494742410f;
                const float s1 = 0.95105651629515357212f, s2 = 0.58778525229247312917f;
                const float2 a0 = in[0], a1 = in[ms], a2 = in[2 * ms], a3 = in[3 * ms], a4 = in[4 * ms];
                const float2 t1 = cadd(a1, a4), t2 = cadd(a2, a3), t3 = csub(a1, a4), t4 = csub(a2, a3);
                const float2 m1 = make_float2(a0.x + c1 * t1.x + c2 * t2.x, a0.y + c1 * t1.y + c2 * t2.y);
                const float2 m2 = make_float2(a0.x + c2 * t1.x + c1 * t2.x, a0.y + c2 * t1.y + c1 * t2.y);
                const float2 n1 = make_float2(s1 * t3.x + s2 * t4.x, s1 * t3.y + s2 * t4.y);
                const float2 n2 = make_float2(s2 * t3.x - s1 * t4.x, s2 * t3.y - s1 * t4.y);
                out[0] = cadd(a0, cadd(t1, t2));
                out[s] = cmul(cadd(m1, mul_neg_i(n1)), __ldg(W + tw));
                out[2 * s] = cmul(cadd(m2, mul_neg_i(n2)), __ldg(W + 2 * tw));
                out[3 * s] = cmul(cadd(m2, mul_pos_i(n2)), __ldg(W + 3 * tw));
                out[4 * s] = cmul(cadd(m1, mul_pos_i(n1)), __ldg(W + 4 * tw));
            }
        }
        __syncthreads();
        float2* sw = x; x = y; y = sw;
        n = m; s *= r;
    }
    return x;
}

// position of sample c in Makhoul's reordered sequence
__device__ __forceinline__ int makhoul_pos(int c, int N) { return (c & 1) ? N - 1 - (c >> 1) : (c >> 1); }

// the two DCT coefficients k of the sequences packed as real / imaginary part, from the FFT of the packed sequence
__device__ __forceinline__ float2 unpack_dct(const float2* Z, const float2* __restrict__ P, int k, int N) {
    const float2 z = Z[k], zr0 = Z[k == 0 ? 0 : N - k];
    const float2 zr = make_float2(zr0.x, -zr0.y);                 // conj Z[N - k]
    const float2 v1 = make_float2(0.5f * (z.x + zr.x), 0.5f * (z.y + zr.y));
    const float2 d = csub(z, zr);
    const float2 v2 = make_float2(0.5f * d.y, -0.5f * d.x);       // (Z - conj Z') / 2i
    const float2 p = __ldg(P + k);
    return make_float2(p.x * v1.x - p.y * v1.y, p.x * v2.x - p.y * v2.y);
}

struct FftDctArgs {
    const float* G;      // [B][H][W]
    float* T;            // [B][H][W] row transforms
    const float2* Ww; const float2* Pw;      // tables of length W
    const float2* Wh; const float2* Ph;      // tables of length H
    double* part;        // [B][blocks of pass 2][5]
    int H, W;
    FftPlan pw, ph;
};

// pass 1: rows 2 * blockIdx.x and 2 * blockIdx.x + 1 of frame blockIdx.y
__global__ void __launch_bounds__(FNT) dct_fft_rows_kernel(FftDctArgs a) {
    extern __shared__ __align__(16) unsigned char fft_smem[];
    float2* x = reinterpret_cast<float2*>(fft_smem);
    float2* y = x + a.W;
    const int W = a.W, b = blockIdx.y;
    const int r0 = 2 * blockIdx.x, r1 = r0 + 1;
    const float* g0 = a.G + ((size_t)b * a.H + r0) * W;
    const float* g1 = g0 + W;
    const bool two = r1 < a.H;
    for (int c = threadIdx.x; c < W; c += FNT)
        x[makhoul_pos(c, W)] = make_float2(__ldg(g0 + c), two ? __ldg(g1 + c) : 0.f);
    __syncthreads();
    const float2* Z = fft_shared(x, y, a.Ww, a.pw);
    float* t0 = a.T + ((size_t)b * a.H + r0) * W;
    float* t1 = t0 + W;
    for (int k = threadIdx.x; k < W; k += FNT) {
        const float2 v = unpack_dct(Z, a.Pw, k, W);
        t0[k] = v.x;
        if (two) t1[k] = v.y;
    }
}

// pass 2: columns 2 * blockIdx.x and 2 * blockIdx.x + 1 of T of frame blockIdx.y; the coefficients go straight into the sums
__global__ void __launch_bounds__(FNT) dct_fft_cols_kernel(FftDctArgs a) {
    extern __shared__ __align__(16) unsigned char fft_smem[];
    float2* x = reinterpret_cast<float2*>(fft_smem);
    float2* y = x + a.H;
    const int H = a.H, W = a.W, b = blockIdx.y;
    const int c0 = 2 * blockIdx.x, c1 = c0 + 1;
    const bool two = c1 < W;
    const float* T = a.T + (size_t)b * H * W;
    for (int r = threadIdx.x; r < H; r += FNT) {
        const float* p = T + (size_t)r * W + c0;
        x[makhoul_pos(r, H)] = make_float2(__ldg(p), two ? __ldg(p + 1) : 0.f);
    }
    __syncthreads();
    const float2* Z = fft_shared(x, y, a.Wh, a.ph);
    const int h4 = H / 4, h2 = H / 2, w4 = W / 4, w2 = W / 2;
    double tot = 0, lo = 0, mid = 0, hi = 0, ab = 0;
    auto take = [&](int u, int xk, float v) {                   // coefficient (row frequency u, column frequency xk)
        const double s = (double)__fmul_rn(v, v);               // dct ** 2 is a float32 array (FE:142)
        tot += s; ab += (double)fabsf(v);
        if (u < h4 && xk < w4) lo += s;
        else if (u >= h4 && u < h2 && xk >= w4 && xk < w2) mid += s;
        else if (u >= h2 && xk >= w2) hi += s;
    };
    for (int u = threadIdx.x; u < H; u += FNT) {
        const float2 v = unpack_dct(Z, a.Ph, u, H);
        take(u, c0, v.x);
        if (two) take(u, c1, v.y);
    }
    __shared__ double red[FNT / 32][5];
    double v5[5] = {tot, lo, mid, hi, ab};
#pragma unroll
    for (int i = 0; i < 5; ++i) v5[i] = warp_sum_f64(v5[i]);
    if ((threadIdx.x & 31) == 0) for (int i = 0; i < 5; ++i) red[threadIdx.x >> 5][i] = v5[i];
    __syncthreads();
    if (threadIdx.x < 5) {
        double s = 0;
        for (int w = 0; w < FNT / 32; ++w) s += red[w][threadIdx.x];
        a.part[((size_t)b * gridDim.x + blockIdx.x) * 5 + threadIdx.x] = s;
    }
}

struct Tables { float2* W; float2* P; };
std::mutex g_fft_mu;
std::map<std::pair<int, int>, Tables> g_fft_cache;               // (device, n)

int tables_for(int n, cudaStream_t st, Tables* out) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 1;
    std::lock_guard<std::mutex> lock(g_fft_mu);
    auto it = g_fft_cache.find({dev, n});
    if (it == g_fft_cache.end()) {
        Tables t;
        if (cudaMalloc(&t.W, (size_t)n * sizeof(float2)) != cudaSuccess) return 1;
        if (cudaMalloc(&t.P, (size_t)n * sizeof(float2)) != cudaSuccess) { cudaFree(t.W); return 1; }
        fft_tables_kernel<<<(n + 255) / 256, 256, 0, st>>>(t.W, t.P, n);
        // later calls may come on other streams: make the tables visible before anybody else uses them
        if (cudaStreamSynchronize(st) != cudaSuccess) { cudaFree(t.W); cudaFree(t.P); return 1; }
        it = g_fft_cache.emplace(std::make_pair(dev, n), t).first;
    }
    *out = it->second;
    return 0;
}

}  // namespace

// true when both dimensions factor into 2, 3 and 5 and fit the shared-memory FFT
bool dct_fft_usable(int H, int W) {
    FftPlan p;
    return make_plan(H, &p) && make_plan(W, &p);
}

// blocks of pass 2 per frame = entries of `part` per frame
int dct_fft_blocks(int W) { return (W + 1) / 2; }

// part: [B][dct_fft_blocks(W)][5] doubles.  Returns the number of kernels launched (the caller adds the block sums), or -1.
int launch_dct_fft(const float* grayf, float* tmp, int B, int H, int W, double* part, cudaStream_t st) {
    FftDctArgs a;
    if (!make_plan(H, &a.ph) || !make_plan(W, &a.pw)) return -1;
    Tables th, tw;
    if (tables_for(H, st, &th) || tables_for(W, st, &tw)) return -1;
    a.G = grayf; a.T = tmp; a.Ww = tw.W; a.Pw = tw.P; a.Wh = th.W; a.Ph = th.P; a.part = part; a.H = H; a.W = W;
    const size_t smem_rows = (size_t)2 * W * sizeof(float2), smem_cols = (size_t)2 * H * sizeof(float2);
    static const bool optin =
        cudaFuncSetAttribute(dct_fft_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * kFftMaxN * (int)sizeof(float2)) == cudaSuccess &&
        cudaFuncSetAttribute(dct_fft_cols_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * kFftMaxN * (int)sizeof(float2)) == cudaSuccess;
    (void)optin;                                                   // a failure shows up as a launch error
    dct_fft_rows_kernel<<<dim3((H + 1) / 2, B), FNT, smem_rows, st>>>(a);
    dct_fft_cols_kernel<<<dim3((W + 1) / 2, B), FNT, smem_cols, st>>>(a);
    return 2;
}

}  // namespace uwie
