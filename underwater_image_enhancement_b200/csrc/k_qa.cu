// Stand-alone scoring of enhanced images: QualityAssessment.comprehensive_assessment
// (quality_assessment.py:215-286) on images that did not come out of the fused engine.
// C ABI: include/uwie.h (uwie_quality_scores).  The image is read once (quantisation, gray / saturation
// histograms, Lab L sum, colourfulness moments, gray plane), then k_canny.cu counts the edges and
// k_score.cu forms the eight sub-scores and the weighted total exactly as in the fused path.
#include <cuda_runtime.h>

#include <cstring>
#include <mutex>
#include <string>

#include "../../include/uwie.h"
#include "k_arith.cuh"
#include "k_kernels.h"
#include "uwie_tables.h"

namespace uwie {
int uwie_fail(int code, const std::string& msg);   // uwie_engine.cu
}
using namespace uwie;

namespace {

constexpr int NT = 256;

template <class In> struct PixelType { using F = float; };
template <> struct PixelType<double> { using F = double; };

template <class In>
__device__ __forceinline__ typename PixelType<In>::F load_unit(const In* p);
template <> __device__ __forceinline__ float load_unit<uint8_t>(const uint8_t* p) { return __fdiv_rn((float)*p, 255.0f); }
template <> __device__ __forceinline__ float load_unit<float>(const float* p) { return *p; }
template <> __device__ __forceinline__ double load_unit<double>(const double* p) { return *p; }

// In = uint8 (image = bytes / 255 in float32), float32 or float64 HWC
template <class In>
__global__ void __launch_bounds__(NT) qa_prepare_kernel(const In* __restrict__ img, long long npix, const ColorTables* __restrict__ ct,
                                                        uint8_t* __restrict__ gray, QaAccum* __restrict__ qa,
                                                        double* __restrict__ cf_part) {
    using T = typename PixelType<In>::F;
    __shared__ unsigned sh[2][256];
    const int b = blockIdx.y;
    for (int i = threadIdx.x; i < 512; i += NT) (&sh[0][0])[i] = 0;
    __syncthreads();
    const In* f = img + (size_t)b * npix * 3;
    uint8_t* gp = gray + (size_t)b * npix;
    double m_rg = 0, m_rg2 = 0, m_yb = 0, m_yb2 = 0;
    unsigned long long lum = 0;
    for (long long i = (long long)blockIdx.x * NT + threadIdx.x; i < npix; i += (long long)gridDim.x * NT) {
        const T r = load_unit<In>(f + i * 3), g = load_unit<In>(f + i * 3 + 1), bl = load_unit<In>(f + i * 3 + 2);
        const int q0 = quantise<T>(r) & 255, q1 = quantise<T>(g) & 255, q2 = quantise<T>(bl) & 255;   // (x * 255).astype(uint8)
        const int gq = gray_q15(q0, q1, q2);
        gp[i] = (uint8_t)gq;
        atomicAdd(&sh[0][gq], 1u);
        atomicAdd(&sh[1][hsv_saturation(ct->sdiv, q0, q1, q2)], 1u);
        lum += (unsigned)lab_L_only(ct->gamma_in, ct->cbrt_tab, q0, q1, q2);
        const double rg = (double)Ar<T>::sub(r, g);                                          // QA:161-166
        const double yb = (double)Ar<T>::sub(Ar<T>::mul((T)0.5, Ar<T>::add(r, g)), bl);
        m_rg += rg; m_rg2 += rg * rg; m_yb += yb; m_yb2 += yb * yb;
    }
    __syncthreads();
    QaAccum& acc = qa[b];
    for (int i = threadIdx.x; i < 256; i += NT) {
        if (sh[0][i]) atomicAdd(&acc.gray_hist[i], sh[0][i]);
        if (sh[1][i]) atomicAdd(&acc.sat_hist[i], sh[1][i]);
    }
    __shared__ double red[NT / 32][5];
    double v5[5] = {m_rg, m_rg2, m_yb, m_yb2, (double)lum};
#pragma unroll
    for (int i = 0; i < 5; ++i) v5[i] = warp_sum_f64(v5[i]);
    if ((threadIdx.x & 31) == 0) for (int i = 0; i < 5; ++i) red[threadIdx.x >> 5][i] = v5[i];
    __syncthreads();
    if (threadIdx.x < 5) {
        double t = 0;
        for (int w = 0; w < NT / 32; ++w) t += red[w][threadIdx.x];
        if (threadIdx.x < 4) cf_part[((size_t)b * gridDim.x + blockIdx.x) * 4 + threadIdx.x] = t;
        else atomicAdd(&acc.lum_sum, (unsigned long long)t);
    }
}

constexpr int kQaBlocks = 296;      // partial-sum blocks per image (2 per SM)

struct QaLayout { size_t ctab, qa, edges, best, cf_part, gray, labels, tile_list, tile_n, total; };
QaLayout qa_layout(int B, int H, int W) {
    QaLayout L;
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t at = off; off += (bytes + 255) & ~(size_t)255; return at; };
    const size_t N = (size_t)H * W;
    L.ctab = take(sizeof(ColorTables));
    L.qa = take((size_t)B * sizeof(QaAccum));
    L.edges = take((size_t)B * 4);
    L.best = take((size_t)B * 4);
    L.cf_part = take((size_t)B * kQaBlocks * 4 * 8);
    L.gray = take(B * N);
    const size_t slots = canny_tile_slots(H, W, B);
    L.labels = take(canny_label_bytes(slots));
    L.tile_list = take(slots * kCannyListCap * sizeof(int2));
    L.tile_n = take(slots * 4);
    L.total = off;
    return L;
}

const ColorTables& host_tables() {
    static ColorTables t;
    static std::once_flag once;
    std::call_once(once, [] { build_color_tables(&t); });
    return t;
}

#define CUQ(call)                                                                                  \
    do {                                                                                           \
        cudaError_t _e = (call);                                                                   \
        if (_e != cudaSuccess) return uwie_fail(UWIE_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(_e)); \
    } while (0)

}  // namespace

extern "C" {

size_t uwie_quality_workspace_bytes(int batch, int height, int width) {
    if (batch < 1 || height < 1 || width < 1) return 0;
    return qa_layout(batch, height, width).total;
}

int uwie_quality_scores(const void* img_dev, int dtype, int batch, int height, int width, const double* weights,
                        double* scores_dev, void* workspace_dev, size_t workspace_bytes, void* stream) {
    if (!img_dev || !scores_dev || !weights || batch < 1 || height < 1 || width < 1 || dtype < 0 || dtype > 2)
        return uwie_fail(UWIE_ERR_ARG, "uwie_quality_scores: bad arguments");
    if ((long long)height * width >= (1LL << 30)) return uwie_fail(UWIE_ERR_ARG, "uwie_quality_scores: frame too large");
    const QaLayout L = qa_layout(batch, height, width);
    if (!workspace_dev || workspace_bytes < L.total) return uwie_fail(UWIE_ERR_ARG, "uwie_quality_scores: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)workspace_dev;
    const long long N = (long long)height * width;
    ColorTables* ctab = (ColorTables*)(ws + L.ctab);
    QaAccum* qa = (QaAccum*)(ws + L.qa);
    unsigned* edges = (unsigned*)(ws + L.edges);
    uint8_t* gray = (uint8_t*)(ws + L.gray);
    double* cf_part = (double*)(ws + L.cf_part);
    CUQ(cudaMemcpyAsync(ctab, &host_tables(), sizeof(ColorTables), cudaMemcpyHostToDevice, st));
    CUQ(cudaMemsetAsync(qa, 0, (size_t)batch * sizeof(QaAccum), st));
    CUQ(cudaMemsetAsync(edges, 0, (size_t)batch * 4, st));
    const dim3 grid(kQaBlocks, batch);
    if (dtype == 0) qa_prepare_kernel<uint8_t><<<grid, NT, 0, st>>>((const uint8_t*)img_dev, N, ctab, gray, qa, cf_part);
    else if (dtype == 1) qa_prepare_kernel<float><<<grid, NT, 0, st>>>((const float*)img_dev, N, ctab, gray, qa, cf_part);
    else qa_prepare_kernel<double><<<grid, NT, 0, st>>>((const double*)img_dev, N, ctab, gray, qa, cf_part);

    CannyArgs ca;
    std::memset(&ca, 0, sizeof(ca));
    ca.mode = CANNY_PLANES; ca.H = height; ca.W = width; ca.low = 50; ca.high = 150;     // QA:139
    ca.gray = gray; ca.planes_per_img = 1;
    ca.labels = (int*)(ws + L.labels); ca.tile_list = (int2*)(ws + L.tile_list); ca.tile_n = (int*)(ws + L.tile_n);
    ca.counts = edges;
    launch_canny(ca, height, width, batch, st);

    ScoreArgs sc;
    std::memset(&sc, 0, sizeof(sc));
    sc.B = batch; sc.S = 1; sc.n = N;
    sc.lib.f64 = dtype == 2 ? 1 : 0;
    for (int i = 0; i < 8; ++i) sc.lib.weights[i] = weights[i];
    sc.qa = qa; sc.cf_part = cf_part; sc.final_blocks = kQaBlocks; sc.edge_counts = edges;
    sc.scores = scores_dev; sc.best = (int*)(ws + L.best);
    launch_scores(sc, st);
    CUQ(cudaGetLastError());
    return UWIE_OK;
}

}  // extern "C"
