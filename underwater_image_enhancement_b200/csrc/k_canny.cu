// Canny(50,150) edge COUNT on the device, exact w.r.t. cv2.Canny:
//   Sobel 3x3 (BORDER_REPLICATE) -> L1 magnitude (0 outside) -> NMS with TG22 fixed point ->
//   double threshold -> 8-connected hysteresis -> number of edge pixels.
// Reference call sites: quality_assessment.py:139 (whole enhanced frame) and
// enhancement_strategies.py:180-182 / six_stadigy.py:149-151 (each quadtree block, with the
// block's own borders).
//
// Hysteresis is a fixed number of launches (no host round trip): candidates are linked with a
// lock-free union-find over a per-plane label array (label = smallest pixel index of the
// component), every strong pixel flags its root, and candidates whose root is flagged are counted.
#include "k_common.cuh"
#include "k_kernels.h"

namespace uwie {

namespace {

constexpr int TS = 32;            // output tile side
constexpr int NT = 256;           // threads per block (32 x 8)
constexpr int LABEL_FLAG = 0x40000000;
constexpr int LABEL_MASK = 0x3fffffff;

struct Job {
    int y0, y1, x0, x1;   // rectangle inside the plane, empty when y0 >= y1
    int img;              // frame index
    int plane;            // which map/label plane
    int out;              // which counter
};

__device__ __forceinline__ Job make_job(const CannyArgs& a, int inst) {
    Job j;
    if (a.mode == CANNY_QUADRANTS) {
        const int img = inst >> 2, q = inst & 3;
        const QuadRect r = a.rects[img];
        j.img = img; j.plane = img; j.out = (img * kQuadLevelsMax + a.level) * 4 + q;
        if (!r.active) { j.y0 = j.y1 = j.x0 = j.x1 = 0; return j; }
        const int ym = r.y0 + (r.y1 - r.y0) / 2, xm = r.x0 + (r.x1 - r.x0) / 2;
        j.y0 = (q & 2) ? ym : r.y0; j.y1 = (q & 2) ? r.y1 : ym;
        j.x0 = (q & 1) ? xm : r.x0; j.x1 = (q & 1) ? r.x1 : xm;
    } else {
        j.img = inst / a.planes_per_img; j.plane = inst; j.out = inst;
        j.y0 = 0; j.y1 = a.H; j.x0 = 0; j.x1 = a.W;
    }
    return j;
}

__device__ __forceinline__ int uf_find(volatile int* L, int x) {
    int l;
    while ((l = (L[x] & LABEL_MASK)) != x) x = l;
    return x;
}

__device__ __forceinline__ void uf_union(int* L, int a, int b) {
    bool done;
    do {
        a = uf_find(L, a);
        b = uf_find(L, b);
        if (a < b) { const int old = atomicMin(&L[b], a); done = (old == b); b = old; }
        else if (b < a) { const int old = atomicMin(&L[a], b); done = (old == a); a = old; }
        else done = true;
    } while (!done);
}

// ---- pass 1: gradient, NMS, classification; optional per-channel histogram of the source bytes
template <int MODE>
__global__ void __launch_bounds__(NT) canny_nms_kernel(CannyArgs a) {
    const Job j = make_job(a, blockIdx.z);
    const int rows = j.y1 - j.y0, cols = j.x1 - j.x0;
    const int ty0 = blockIdx.y * TS, tx0 = blockIdx.x * TS;
    if (ty0 >= rows || tx0 >= cols) return;

    __shared__ uint8_t sg[TS + 4][TS + 4];
    __shared__ uint16_t sm[TS + 2][TS + 2];
    __shared__ unsigned shist[MODE == CANNY_QUADRANTS ? 768 : 1];

    const int tid = threadIdx.x;
    const size_t npix = (size_t)a.H * a.W;
    const uint8_t* frame = nullptr;
    const uint8_t* gplane = nullptr;
    const CastTables* ct = nullptr;
    if (MODE == CANNY_QUADRANTS) {
        frame = a.frames + (size_t)j.img * npix * 3;
        ct = a.cast + j.img;
        for (int i = tid; i < 768; i += NT) shist[i] = 0;
        __syncthreads();
    } else {
        gplane = a.gray + (size_t)j.plane * npix;
    }

    // gray with a 2-pixel replicated halo (replication at the RECTANGLE's border)
    for (int i = tid; i < (TS + 4) * (TS + 4); i += NT) {
        const int ly = i / (TS + 4), lx = i - ly * (TS + 4);
        const int ry = ty0 + ly - 2, rx = tx0 + lx - 2;
        const int y = j.y0 + clamp_idx(ry, rows), x = j.x0 + clamp_idx(rx, cols);
        int g;
        if (MODE == CANNY_QUADRANTS) {
            const uint8_t* p = frame + ((size_t)y * a.W + x) * 3;
            const int r0 = p[0], g0 = p[1], b0 = p[2];
            g = gray_q15(ct->curq[0][r0], ct->curq[1][g0], ct->curq[2][b0]);
            const bool inner = ly >= 2 && ly < TS + 2 && lx >= 2 && lx < TS + 2 && ry < rows && rx < cols;
            if (inner) {
                atomicAdd(&shist[r0], 1u);
                atomicAdd(&shist[256 + g0], 1u);
                atomicAdd(&shist[512 + b0], 1u);
            }
        } else {
            g = gplane[(size_t)y * a.W + x];
        }
        sg[ly][lx] = (uint8_t)g;
    }
    __syncthreads();

    // L1 magnitude on the tile + 1 halo; zero outside the rectangle
    for (int i = tid; i < (TS + 2) * (TS + 2); i += NT) {
        const int ly = i / (TS + 2), lx = i - ly * (TS + 2);
        const int ry = ty0 + ly - 1, rx = tx0 + lx - 1;
        int m = 0;
        if (ry >= 0 && ry < rows && rx >= 0 && rx < cols) {
            const int cy = ly + 1, cx = lx + 1;
            const int dx = (sg[cy - 1][cx + 1] + 2 * sg[cy][cx + 1] + sg[cy + 1][cx + 1]) -
                           (sg[cy - 1][cx - 1] + 2 * sg[cy][cx - 1] + sg[cy + 1][cx - 1]);
            const int dy = (sg[cy + 1][cx - 1] + 2 * sg[cy + 1][cx] + sg[cy + 1][cx + 1]) -
                           (sg[cy - 1][cx - 1] + 2 * sg[cy - 1][cx] + sg[cy - 1][cx + 1]);
            m = abs(dx) + abs(dy);
        }
        sm[ly][lx] = (uint16_t)m;
    }
    __syncthreads();

    uint8_t* map = a.map + (size_t)j.plane * npix;
    int* lab = a.labels + (size_t)j.plane * npix;
    for (int i = tid; i < TS * TS; i += NT) {
        const int ly = i / TS, lx = i - ly * TS;
        const int ry = ty0 + ly, rx = tx0 + lx;
        if (ry >= rows || rx >= cols) continue;
        const int cy = ly + 2, cx = lx + 2;
        const int dx = (sg[cy - 1][cx + 1] + 2 * sg[cy][cx + 1] + sg[cy + 1][cx + 1]) -
                       (sg[cy - 1][cx - 1] + 2 * sg[cy][cx - 1] + sg[cy + 1][cx - 1]);
        const int dy = (sg[cy + 1][cx - 1] + 2 * sg[cy + 1][cx] + sg[cy + 1][cx + 1]) -
                       (sg[cy - 1][cx - 1] + 2 * sg[cy - 1][cx] + sg[cy - 1][cx + 1]);
        const int my = ly + 1, mx = lx + 1;
        const int cls = canny_nms_class(dx, dy, sm[my][mx], sm[my][mx - 1], sm[my][mx + 1],
                                        sm[my - 1][mx], sm[my + 1][mx], sm[my - 1][mx - 1],
                                        sm[my - 1][mx + 1], sm[my + 1][mx - 1], sm[my + 1][mx + 1],
                                        a.low, a.high);
        const size_t p = (size_t)(j.y0 + ry) * a.W + (j.x0 + rx);
        map[p] = (uint8_t)cls;
        if (cls) lab[p] = (int)p;
    }

    if (MODE == CANNY_QUADRANTS) {
        __syncthreads();
        unsigned* gh = a.quad_hist + (size_t)j.out * 768;
        for (int i = tid; i < 768; i += NT) {
            const unsigned v = shist[i];
            if (v) atomicAdd(&gh[i], v);
        }
    }
}

// ---- pass 2: link every candidate with its W, NW, N, NE candidate neighbours (inside the rect)
__global__ void __launch_bounds__(NT) canny_link_kernel(CannyArgs a) {
    const Job j = make_job(a, blockIdx.z);
    const int rows = j.y1 - j.y0, cols = j.x1 - j.x0;
    const int ty0 = blockIdx.y * TS, tx0 = blockIdx.x * TS;
    if (ty0 >= rows || tx0 >= cols) return;
    const size_t npix = (size_t)a.H * a.W;
    const uint8_t* map = a.map + (size_t)j.plane * npix;
    int* lab = a.labels + (size_t)j.plane * npix;
    for (int i = threadIdx.x; i < TS * TS; i += NT) {
        const int ly = i / TS, lx = i - ly * TS;
        const int ry = ty0 + ly, rx = tx0 + lx;
        if (ry >= rows || rx >= cols) continue;
        const int y = j.y0 + ry, x = j.x0 + rx;
        const int p = y * a.W + x;
        if (!map[p]) continue;
        if (rx > 0 && map[p - 1]) uf_union(lab, p, p - 1);
        if (ry > 0) {
            const int up = p - a.W;
            if (rx > 0 && map[up - 1]) uf_union(lab, p, up - 1);
            if (map[up]) uf_union(lab, p, up);
            if (rx + 1 < cols && map[up + 1]) uf_union(lab, p, up + 1);
        }
    }
}

// ---- pass 3: strong pixels flag the root of their component (and shorten their own path)
__global__ void __launch_bounds__(NT) canny_flag_kernel(CannyArgs a) {
    const Job j = make_job(a, blockIdx.z);
    const int rows = j.y1 - j.y0, cols = j.x1 - j.x0;
    const int ty0 = blockIdx.y * TS, tx0 = blockIdx.x * TS;
    if (ty0 >= rows || tx0 >= cols) return;
    const size_t npix = (size_t)a.H * a.W;
    const uint8_t* map = a.map + (size_t)j.plane * npix;
    int* lab = a.labels + (size_t)j.plane * npix;
    for (int i = threadIdx.x; i < TS * TS; i += NT) {
        const int ly = i / TS, lx = i - ly * TS;
        const int ry = ty0 + ly, rx = tx0 + lx;
        if (ry >= rows || rx >= cols) continue;
        const int p = (j.y0 + ry) * a.W + (j.x0 + rx);
        if (map[p] != 2) continue;
        const int root = uf_find(lab, p);
        atomicOr(&lab[root], LABEL_FLAG);
    }
}

// ---- pass 4: count candidates whose root is flagged
__global__ void __launch_bounds__(NT) canny_count_kernel(CannyArgs a) {
    const Job j = make_job(a, blockIdx.z);
    const int rows = j.y1 - j.y0, cols = j.x1 - j.x0;
    const int ty0 = blockIdx.y * TS, tx0 = blockIdx.x * TS;
    if (ty0 >= rows || tx0 >= cols) return;
    const size_t npix = (size_t)a.H * a.W;
    const uint8_t* map = a.map + (size_t)j.plane * npix;
    int* lab = a.labels + (size_t)j.plane * npix;
    unsigned cnt = 0;
    for (int i = threadIdx.x; i < TS * TS; i += NT) {
        const int ly = i / TS, lx = i - ly * TS;
        const int ry = ty0 + ly, rx = tx0 + lx;
        if (ry >= rows || rx >= cols) continue;
        const int p = (j.y0 + ry) * a.W + (j.x0 + rx);
        if (!map[p]) continue;
        const int root = uf_find(lab, p);
        if (lab[root] & LABEL_FLAG) ++cnt;
    }
    cnt = warp_sum_u32(cnt);
    __shared__ unsigned sw[NT / 32];
    if ((threadIdx.x & 31) == 0) sw[threadIdx.x >> 5] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned t = 0;
        for (int w = 0; w < NT / 32; ++w) t += sw[w];
        if (t) atomicAdd(&a.counts[j.out], t);
    }
}

}  // namespace

int launch_canny(const CannyArgs& a, int max_rows, int max_cols, int instances, cudaStream_t st) {
    if (max_rows <= 0 || max_cols <= 0 || instances <= 0) return 0;
    dim3 grid((max_cols + TS - 1) / TS, (max_rows + TS - 1) / TS, instances);
    if (a.mode == CANNY_QUADRANTS) canny_nms_kernel<CANNY_QUADRANTS><<<grid, NT, 0, st>>>(a);
    else canny_nms_kernel<CANNY_PLANES><<<grid, NT, 0, st>>>(a);
    canny_link_kernel<<<grid, NT, 0, st>>>(a);
    canny_flag_kernel<<<grid, NT, 0, st>>>(a);
    canny_count_kernel<<<grid, NT, 0, st>>>(a);
    return 4;
}

}  // namespace uwie
