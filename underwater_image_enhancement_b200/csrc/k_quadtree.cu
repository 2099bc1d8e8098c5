// Quadtree atmospheric-light search, device-side descent (no host round trips).
// Reference: estimate_atmospheric_light / compute_Q / get_brightest_pixel,
// enhancement_strategies.py:77-206 and six_stadigy.py:48-165.
//
// Per level the Canny pass (k_canny.cu, CANNY_QUADRANTS) leaves, for each of the four quadrants of
// the current rectangle, the per-channel byte histograms and the edge count.  The score
//   Q = mean(R,G,B) + (sumB + sumG - 2 sumR)/n - mean(var_R,var_G,var_B) - edges/n
// is then a function of 768 counters; it is evaluated here in float64 on the float32 pixel values
// the reference sees (CastTables::curf), the first maximum wins, the rectangle shrinks.
#include "k_common.cuh"
#include "k_kernels.h"

namespace uwie {

namespace {

__global__ void quad_init_kernel(QuadArgs a) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= a.B) return;
    QuadRect r;
    r.y0 = 0; r.y1 = a.H; r.x0 = 0; r.x1 = a.W;
    r.active = (a.H > 1 && a.W > 1) ? 1 : 0;   // while nRows > min_size and nCols > min_size
    r.level = 0;
    a.rects[b] = r;
}

// one block of 384 threads per frame: warp w <-> (quadrant w / 3, channel w % 3)
__global__ void __launch_bounds__(384) quad_descend_kernel(QuadArgs a, int level) {
    const int b = blockIdx.x;
    QuadRect r = a.rects[b];
    if (!r.active) return;
    __shared__ double s_sum[12], s_var[12];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = w / 3, c = w % 3;
    const int ym = r.y0 + (r.y1 - r.y0) / 2, xm = r.x0 + (r.x1 - r.x0) / 2;
    const int qy0 = (q & 2) ? ym : r.y0, qy1 = (q & 2) ? r.y1 : ym;
    const int qx0 = (q & 1) ? xm : r.x0, qx1 = (q & 1) ? r.x1 : xm;
    const double n = (double)(qy1 - qy0) * (double)(qx1 - qx0);
    const unsigned* h = a.quad_hist + ((size_t)(b * kQuadLevelsMax + level) * 4 + q) * 768 + c * 256;
    const float* f = a.cast[b].curf[c];
    double s = 0.0;
    for (int v = lane; v < 256; v += 32) s += (double)h[v] * (double)f[v];
    s = warp_sum_f64(s);
    const double mean = s / n;
    double var = 0.0;
    for (int v = lane; v < 256; v += 32) {
        const double d = (double)f[v] - mean;
        var += (double)h[v] * d * d;
    }
    var = warp_sum_f64(var) / n;
    if (lane == 0) { s_sum[w] = s; s_var[w] = var; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double best = 0.0; int best_q = 0;
        for (int k = 0; k < 4; ++k) {
            const int ky0 = (k & 2) ? ym : r.y0, ky1 = (k & 2) ? r.y1 : ym;
            const int kx0 = (k & 1) ? xm : r.x0, kx1 = (k & 1) ? r.x1 : xm;
            const double kn = (double)(ky1 - ky0) * (double)(kx1 - kx0);
            const double sr = s_sum[k * 3], sg = s_sum[k * 3 + 1], sb = s_sum[k * 3 + 2];
            const double bright = (sr + sg + sb) / (3.0 * kn);
            const double cast = (sb + sg - 2.0 * sr) / kn;
            const double spread = (s_var[k * 3] + s_var[k * 3 + 1] + s_var[k * 3 + 2]) / 3.0;
            const double edges = (double)a.counts[(size_t)(b * kQuadLevelsMax + level) * 4 + k] / kn;
            const double Q = bright + cast - spread - edges;
            a.q_trace[(size_t)(b * kQuadLevelsMax + level) * 4 + k] = Q;
            if (k == 0 || Q > best) { best = Q; best_q = k; }
        }
        QuadRect nr;
        nr.y0 = (best_q & 2) ? ym : r.y0; nr.y1 = (best_q & 2) ? r.y1 : ym;
        nr.x0 = (best_q & 1) ? xm : r.x0; nr.x1 = (best_q & 1) ? r.x1 : xm;
        nr.level = level + 1;
        nr.active = ((nr.y1 - nr.y0) > 1 && (nr.x1 - nr.x0) > 1) ? 1 : 0;
        a.rects[b] = nr;
    }
}

// first pixel (row-major) with the largest float32 (R + G) + B inside the final rectangle
__global__ void __launch_bounds__(256) quad_brightest_kernel(QuadArgs a) {
    const int b = blockIdx.x;
    const QuadRect r = a.rects[b];
    const int rows = r.y1 - r.y0, cols = r.x1 - r.x0;
    const CastTables& ct = a.cast[b];
    const uint8_t* f = a.frames + (size_t)b * a.H * a.W * 3;
    float best = -1.0f; int best_i = 0x7fffffff;
    for (int i = threadIdx.x; i < rows * cols; i += 256) {
        const int y = r.y0 + i / cols, x = r.x0 + i % cols;
        const uint8_t* p = f + ((size_t)y * a.W + x) * 3;
        const float s = __fadd_rn(__fadd_rn(ct.curf[0][p[0]], ct.curf[1][p[1]]), ct.curf[2][p[2]]);
        if (s > best) { best = s; best_i = i; }   // increasing i per thread: first max kept
    }
    __shared__ float sv[256];
    __shared__ int si[256];
    sv[threadIdx.x] = best; si[threadIdx.x] = best_i;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) {
            const float v2 = sv[threadIdx.x + o]; const int i2 = si[threadIdx.x + o];
            if (v2 > sv[threadIdx.x] || (v2 == sv[threadIdx.x] && i2 < si[threadIdx.x])) {
                sv[threadIdx.x] = v2; si[threadIdx.x] = i2;
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const int i = si[0];
        const int y = r.y0 + i / cols, x = r.x0 + i % cols;
        const uint8_t* p = f + ((size_t)y * a.W + x) * 3;
        a.light[b * 3 + 0] = ct.curf[0][p[0]];
        a.light[b * 3 + 1] = ct.curf[1][p[1]];
        a.light[b * 3 + 2] = ct.curf[2][p[2]];
    }
}

}  // namespace

int launch_quad_init(const QuadArgs& a, cudaStream_t st) {
    quad_init_kernel<<<(a.B + 127) / 128, 128, 0, st>>>(a);
    return 1;
}
int launch_quad_descend(const QuadArgs& a, int level, cudaStream_t st) {
    quad_descend_kernel<<<a.B, 384, 0, st>>>(a, level);
    return 1;
}
int launch_quad_brightest(const QuadArgs& a, cudaStream_t st) {
    quad_brightest_kernel<<<a.B, 256, 0, st>>>(a);
    return 1;
}

}  // namespace uwie
