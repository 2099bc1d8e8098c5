// The eight no-reference sub-scores and the weighted total of quality_assessment.py:14-286, formed
// from the accumulators of the final pass (k_apply.cu) and the Canny edge count (k_canny.cu), then
// the first-maximum strategy selection of main.py:144-146.
//
// The reference mixes NumPy scalar types: contrast, saturation, brightness (and colourfulness on a
// float32 image) are np.float32, entropy / edge_density / naturalness are float64, sharpness is the
// Python float 50.0 (cv2.Laplacian(float32, CV_64F) raises, QA:49, and QA:250-251 substitutes 50.0).
// The weighted sum is accumulated with the same promotions so that totals agree to rounding.
#include "k_arith.cuh"
#include "k_kernels.h"

namespace uwie {

namespace {

constexpr int NT = 128;

__device__ __forceinline__ float clipf(float x, float lo, float hi) { return x < lo ? lo : (x > hi ? hi : x); }
__device__ __forceinline__ double clipd(double x, double lo, double hi) { return x < lo ? lo : (x > hi ? hi : x); }

// fixed-order tree sums of K per-thread values over the block (deterministic run to run)
template <int K>
__device__ __forceinline__ void block_sums(double (&v)[K], double (*red)[NT]) {
#pragma unroll
    for (int k = 0; k < K; ++k) red[k][threadIdx.x] = v[k];
    __syncthreads();
    for (int o = NT / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) {
#pragma unroll
            for (int k = 0; k < K; ++k) red[k][threadIdx.x] += red[k][threadIdx.x + o];
        }
        __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < K; ++k) v[k] = red[k][0];
    __syncthreads();
}

__global__ void __launch_bounds__(NT) score_kernel(ScoreArgs a) {
    const int s = blockIdx.x, b = blockIdx.y;
    const size_t bs = (size_t)b * a.S + s;
    const QaAccum& acc = a.qa[bs];
    const double n = (double)a.n;
    __shared__ double red[10][NT];
    // every thread takes bins tid and tid + 128 of the histograms and a slice of the per-block colourfulness
    // partials; ten sums leave in one tree reduction, the variance of the gray values in a second one
    double q[10];
#pragma unroll
    for (int k = 0; k < 10; ++k) q[k] = 0.0;
    for (int i = threadIdx.x; i < a.final_blocks; i += NT) {
#pragma unroll
        for (int k = 0; k < 4; ++k) q[k] += a.cf_part[(bs * a.final_blocks + i) * 4 + k];
    }
    unsigned hg[2], hs[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int v = threadIdx.x + h * NT;
        hg[h] = acc.gray_hist[v]; hs[h] = acc.sat_hist[v];
        const float fv = __fdiv_rn((float)v, 255.0f);
        q[4] += (double)hg[h] * (double)fv;                                      // mean of gray/255
        if (hg[h]) { const double pk = (double)hg[h] / n; q[5] += -pk * log(pk); }   // entropy terms
        q[6] += (double)hs[h] * (double)fv;                                      // mean of S/255
        if (fv > 0.9f) { q[7] += (double)hs[h]; q[9] += (double)hg[h]; }         // over-saturated, bright
        if (fv < 0.1f) q[8] += (double)hg[h];                                    // dark
    }
    block_sums<10>(q, red);
    const double cf[4] = {q[0], q[1], q[2], q[3]};
    const double gmean = q[4] / n;
    double var[1] = {0.0};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const double d = (double)__fdiv_rn((float)(threadIdx.x + h * NT), 255.0f) - gmean;
        var[0] += (double)hg[h] * d * d;
    }
    block_sums<1>(var, red);
    if (threadIdx.x != 0) return;

    double sub[8];
    int is_f32[8] = {1, 0, 0, 1, 1, 0, a.lib.f64 ? 0 : 1, 0};
    // contrast: std of gray/255 (float32 data)                                   QA:25-31
    {
        const float sd = (float)sqrt(var[0] / n);
        sub[0] = (double)clipf(__fmul_rn(__fdiv_rn(sd, 0.5f), 100.0f), 0.f, 100.f);
    }
    sub[1] = 50.0;                                                               // QA:49, 250-251
    // entropy of the gray value counts, scipy.stats.entropy(counts, base=2)       QA:68-73
    {
        const double S = q[5] / log(2.0);
        sub[2] = clipd((S - 4.0) / 4.0 * 100.0, 0.0, 100.0);
    }
    // saturation: mean of S/255 (float32)                                         QA:88-95
    {
        const float ms = (float)(q[6] / n);
        sub[3] = (double)clipf(__fmul_rn(ms, 100.0f), 0.f, 100.f);
    }
    // brightness: 100 - clip(|mean(L) - 128| / 128 * 100) (float32)               QA:111-121
    {
        const float mb = (float)((double)acc.lum_sum / n);
        const float dev = fabsf(__fsub_rn(mb, 128.0f));
        sub[4] = (double)__fsub_rn(100.0f, clipf(__fmul_rn(__fdiv_rn(dev, 128.0f), 100.0f), 0.f, 100.f));
    }
    // edge density (float64)                                                      QA:136-145
    sub[5] = clipd((double)a.edge_counts[bs] / n / 0.2 * 100.0, 0.0, 100.0);
    // colourfulness                                                               QA:161-182
    {
        const double mrg = cf[0] / n, myb = cf[2] / n;
        double vrg = cf[1] / n - mrg * mrg, vyb = cf[3] / n - myb * myb;
        vrg = vrg < 0 ? 0 : vrg; vyb = vyb < 0 ? 0 : vyb;
        if (a.lib.f64) {
            const double sp = sqrt(vrg + vyb), of = sqrt(mrg * mrg + myb * myb);
            sub[6] = clipd((sp + 0.3 * of) / 0.5 * 100.0, 0.0, 100.0);
        } else {
            const float srg = (float)sqrt(vrg), syb = (float)sqrt(vyb);
            const float fm_rg = (float)mrg, fm_yb = (float)myb;
            const float sp = sqrtf(__fadd_rn(__fmul_rn(srg, srg), __fmul_rn(syb, syb)));
            const float of = sqrtf(__fadd_rn(__fmul_rn(fm_rg, fm_rg), __fmul_rn(fm_yb, fm_yb)));
            const float cfv = __fadd_rn(sp, __fmul_rn(0.3f, of));
            sub[6] = (double)clipf(__fmul_rn(__fdiv_rn(cfv, 0.5f), 100.0f), 0.f, 100.f);
        }
    }
    // naturalness (float64 ratios of float32 comparisons)                         QA:198-210
    {
        // counts below 2^53: the float64 sums above are exact
        const double ratio = q[7] / n + q[8] / n + q[9] / n;
        sub[7] = 100.0 - clipd(ratio * 200.0, 0.0, 100.0);
    }
    // total = sum(scores[k] * weights.get(k, 0)) with NumPy's scalar promotions    QA:284
    bool wide = false;      // becomes float64 once a float64 (or Python-float-times-f64) term is added
    float tf = 0.f; double td = 0.0;
    for (int k = 0; k < 8; ++k) {
        const double w = a.lib.weights[k];
        if (k == 1) {                      // Python float * Python float
            const double term = 50.0 * w;
            if (wide) td += term; else tf = __fadd_rn(tf, (float)term);
            continue;
        }
        if (is_f32[k]) {
            const float term = __fmul_rn((float)sub[k], (float)w);
            if (wide) td += (double)term; else tf = __fadd_rn(tf, term);
        } else {
            const double term = sub[k] * w;
            if (!wide) { td = (double)tf; wide = true; }
            td += term;
        }
    }
    double* row = a.scores + bs * kScoreCols;
    row[0] = wide ? td : (double)tf;
    for (int k = 0; k < 8; ++k) row[1 + k] = sub[k];
}

__global__ void best_kernel(ScoreArgs a) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= a.B) return;
    int best = 0; double bv = a.scores[((size_t)b * a.S) * kScoreCols];
    for (int s = 1; s < a.S; ++s) {
        const double v = a.scores[((size_t)b * a.S + s) * kScoreCols];
        if (v > bv) { bv = v; best = s; }
    }
    a.best[b] = best;
}

}  // namespace

int launch_scores(const ScoreArgs& a, cudaStream_t st) {
    score_kernel<<<dim3(a.S, a.B), NT, 0, st>>>(a);
    best_kernel<<<(a.B + 127) / 128, 128, 0, st>>>(a);
    return 2;
}

}  // namespace uwie
