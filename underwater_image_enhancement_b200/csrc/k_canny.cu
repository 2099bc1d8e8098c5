// Canny(50,150) edge COUNT on the device, exact w.r.t. cv2.Canny:
//   Sobel 3x3 (BORDER_REPLICATE) -> L1 magnitude (0 outside) -> NMS with TG22 fixed point ->
//   double threshold -> 8-connected hysteresis -> number of edge pixels.
// Reference call sites: quality_assessment.py:139 (whole enhanced frame) and
// enhancement_strategies.py:180-182 / six_stadigy.py:149-151 (each quadtree block, with the
// block's own borders).
//
// Hysteresis is a fixed number of launches (no host round trip) built on union-find:
//   1. canny_tile_kernel   gradient, NMS and classification of a 32x64 tile in shared memory, then
//                          the tile's candidates are linked by a union-find that lives entirely in
//                          shared memory.  Written out: class byte, label = pixel index of the
//                          tile-local root, and on root pixels whether the component holds a
//                          strong pixel.
//   2. canny_merge_kernel  only pixels on tile borders link across tiles (lock-free atomicMin
//                          union-find on the global label plane; trees are at most one deep
//                          below each tile-local root).
//   3. canny_flag_kernel   tile-local roots that hold a strong pixel flag their global root.
//   4. canny_count_kernel  candidates whose global root is flagged are counted.
#include "k_common.cuh"
#include "k_kernels.h"

namespace uwie {

namespace {

constexpr int TH = 32, TW = 64;   // tile
constexpr int NT = 256;
constexpr int LABEL_FLAG = 0x40000000;
constexpr int LABEL_MASK = 0x3fffffff;
// class byte: bit 0 candidate, bit 1 strong pixel, bit 2 tile-local root of a component with a strong pixel
constexpr int CLS_CAND = 1, CLS_STRONG = 2, CLS_ROOT_STRONG = 4;

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

// ---- union-find on a label array (global or shared); label = smallest index of the component
template <class P>
__device__ __forceinline__ int uf_find(P L, int x) {
    int l;
    while ((l = (L[x] & LABEL_MASK)) != x) x = l;
    return x;
}
__device__ __forceinline__ void uf_union(int* L, int a, int b) {
    bool done;
    do {
        a = uf_find((volatile int*)L, a);
        b = uf_find((volatile int*)L, b);
        if (a < b) { const int old = atomicMin(&L[b], a); done = (old == b); b = old; }
        else if (b < a) { const int old = atomicMin(&L[a], b); done = (old == a); a = old; }
        else done = true;
    } while (!done);
}

__device__ __forceinline__ unsigned long long row_mask(const unsigned* m) {
    return (unsigned long long)m[0] | ((unsigned long long)m[1] << 32);
}
// first pixel of the run of set bits of `m` that contains bit x
__device__ __forceinline__ int run_start(unsigned long long m, int x) {
    const unsigned long long below = ~m & ((1ull << x) - 1ull);
    return below ? 64 - __clzll((long long)below) : 0;
}

// ---- pass 1
template <int MODE>
__global__ void __launch_bounds__(NT) canny_tile_kernel(CannyArgs a) {
    const Job j = make_job(a, blockIdx.z);
    const int rows = j.y1 - j.y0, cols = j.x1 - j.x0;
    const int ty0 = blockIdx.y * TH, tx0 = blockIdx.x * TW;
    if (ty0 >= rows || tx0 >= cols) return;

    __shared__ uint8_t sg[TH + 4][TW + 4];
    __shared__ uint16_t sm[TH + 2][TW + 2];
    __shared__ unsigned smask_c[TH][2], smask_s[TH][2];   // per tile row: candidate / strong bit masks
    __shared__ int slab[TH * TW];
    __shared__ int sstrong[TH * TW / 32];          // bit per tile pixel: root of a component with a strong pixel
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
    } else {
        gplane = a.gray + (size_t)j.plane * npix;
    }
    for (int i = tid; i < TH * TW / 32; i += NT) sstrong[i] = 0;
    __syncthreads();

    // gray with a 2-pixel replicated halo (replication at the RECTANGLE's border)
    for (int i = tid; i < (TH + 4) * (TW + 4); i += NT) {
        const int ly = i / (TW + 4), lx = i - ly * (TW + 4);
        const int ry = ty0 + ly - 2, rx = tx0 + lx - 2;
        const int y = j.y0 + clamp_idx(ry, rows), x = j.x0 + clamp_idx(rx, cols);
        int g;
        if (MODE == CANNY_QUADRANTS) {
            const uint8_t* p = frame + ((size_t)y * a.W + x) * 3;
            const int r0 = p[0], g0 = p[1], b0 = p[2];
            g = gray_q15(ct->curq[0][r0], ct->curq[1][g0], ct->curq[2][b0]);
            const bool inner = ly >= 2 && ly < TH + 2 && lx >= 2 && lx < TW + 2 && ry < rows && rx < cols;
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
    for (int i = tid; i < (TH + 2) * (TW + 2); i += NT) {
        const int ly = i / (TW + 2), lx = i - ly * (TW + 2);
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

    // classification.  Thread mapping: 8 passes, in each a warp owns half a tile row, so one ballot
    // gives the candidate / strong bit masks of that half row.
    for (int it = 0; it < TH * TW / NT; ++it) {
        const int i = it * NT + tid;
        const int ly = i / TW, lx = i - ly * TW;
        const int ry = ty0 + ly, rx = tx0 + lx;
        int cls = 0;
        if (ry < rows && rx < cols) {
            const int cy = ly + 2, cx = lx + 2;
            const int dx = (sg[cy - 1][cx + 1] + 2 * sg[cy][cx + 1] + sg[cy + 1][cx + 1]) -
                           (sg[cy - 1][cx - 1] + 2 * sg[cy][cx - 1] + sg[cy + 1][cx - 1]);
            const int dy = (sg[cy + 1][cx - 1] + 2 * sg[cy + 1][cx] + sg[cy + 1][cx + 1]) -
                           (sg[cy - 1][cx - 1] + 2 * sg[cy - 1][cx] + sg[cy - 1][cx + 1]);
            const int my = ly + 1, mx = lx + 1;
            cls = canny_nms_class(dx, dy, sm[my][mx], sm[my][mx - 1], sm[my][mx + 1], sm[my - 1][mx],
                                  sm[my + 1][mx], sm[my - 1][mx - 1], sm[my - 1][mx + 1], sm[my + 1][mx - 1],
                                  sm[my + 1][mx + 1], a.low, a.high);
        }
        const unsigned mc = __ballot_sync(0xffffffffu, cls != 0);
        const unsigned ms = __ballot_sync(0xffffffffu, cls == 2);
        if ((tid & 31) == 0) { smask_c[ly][lx >> 5] = mc; smask_s[ly][lx >> 5] = ms; }
        slab[i] = i;
    }
    __syncthreads();

    // Link inside the tile.  A horizontal run of candidates is one node (its first pixel); runs of
    // neighbouring rows are united once per contact: with U the row above,
    //   up set:            unite unless both W and NW are set (then the pixel to the left did it)
    //   up clear, NW set:  unite if W is clear          up clear, NE set:  unite if E is clear
    for (int it = 0; it < TH * TW / NT; ++it) {
        const int i = it * NT + tid;
        const int ly = i / TW, lx = i - ly * TW;
        if (ly == 0) continue;                                                  // warp-uniform
        const unsigned long long R = row_mask(smask_c[ly]), U = row_mask(smask_c[ly - 1]);
        if (!((R >> lx) & 1ull) || U == 0ull) continue;
        const bool up = (U >> lx) & 1ull;
        const bool nw = lx > 0 && ((U >> (lx - 1)) & 1ull), ne = lx < TW - 1 && ((U >> (lx + 1)) & 1ull);
        const bool w = lx > 0 && ((R >> (lx - 1)) & 1ull), e = lx < TW - 1 && ((R >> (lx + 1)) & 1ull);
        const int me = ly * TW + run_start(R, lx);
        if (up) { if (!w || !nw) uf_union(slab, me, (ly - 1) * TW + run_start(U, lx)); }
        else {
            if (nw && !w) uf_union(slab, me, (ly - 1) * TW + run_start(U, lx - 1));
            if (ne && !e) uf_union(slab, me, (ly - 1) * TW + lx + 1);           // NE starts its run (up is clear)
        }
    }
    __syncthreads();
    // a run that holds a strong pixel marks its root (done by the run's first pixel)
    for (int it = 0; it < TH * TW / NT; ++it) {
        const int i = it * NT + tid;
        const int ly = i / TW, lx = i - ly * TW;
        const unsigned long long R = row_mask(smask_c[ly]);
        if (!((R >> lx) & 1ull) || (lx > 0 && ((R >> (lx - 1)) & 1ull))) continue;   // not a run start
        const unsigned long long holes = ~R & (~0ull << lx);
        const int end = holes ? __ffsll((long long)holes) - 1 : TW;               // first clear bit after the run
        const unsigned long long run = (end >= 64 ? ~0ull : ((1ull << end) - 1)) & (~0ull << lx);
        if (row_mask(smask_s[ly]) & run) {
            const int root = uf_find((volatile int*)slab, i);
            atomicOr(&sstrong[root >> 5], 1 << (root & 31));
        }
    }
    __syncthreads();

    // write out: class byte (+ root-strong bit) and label = global pixel index of the tile-local root
    uint8_t* map = a.map + (size_t)j.plane * npix;
    int* lab = a.labels + (size_t)j.plane * npix;
    for (int it = 0; it < TH * TW / NT; ++it) {
        const int i = it * NT + tid;
        const int ly = i / TW, lx = i - ly * TW;
        const int ry = ty0 + ly, rx = tx0 + lx;
        if (ry >= rows || rx >= cols) continue;
        const size_t p = (size_t)(j.y0 + ry) * a.W + (j.x0 + rx);
        const unsigned long long R = row_mask(smask_c[ly]);
        int cls = 0;
        if ((R >> lx) & 1ull) {
            cls = CLS_CAND | (((row_mask(smask_s[ly]) >> lx) & 1ull) ? CLS_STRONG : 0);
            const int root = uf_find((volatile int*)slab, ly * TW + run_start(R, lx));
            const int rly = root / TW, rlx = root - rly * TW;
            lab[p] = (j.y0 + ty0 + rly) * a.W + (j.x0 + tx0 + rlx);
            if (root == i && ((sstrong[i >> 5] >> (i & 31)) & 1)) cls |= CLS_ROOT_STRONG;
        }
        map[p] = (uint8_t)cls;
    }

    if (MODE == CANNY_QUADRANTS) {
        unsigned* gh = a.quad_hist + (size_t)j.out * 768;
        for (int i = tid; i < 768; i += NT) {
            const unsigned v = shist[i];
            if (v) atomicAdd(&gh[i], v);
        }
    }
}

// ---- pass 2: links that cross a tile border.  One thread per border pixel of a tile:
// top row (TW), left column below it (TH - 1), right column below it (TH - 1).
__global__ void __launch_bounds__(128) canny_merge_kernel(CannyArgs a) {
    const Job j = make_job(a, blockIdx.z);
    const int rows = j.y1 - j.y0, cols = j.x1 - j.x0;
    const int ty0 = blockIdx.y * TH, tx0 = blockIdx.x * TW;
    if (ty0 >= rows || tx0 >= cols) return;
    const size_t npix = (size_t)a.H * a.W;
    const uint8_t* map = a.map + (size_t)j.plane * npix;
    int* lab = a.labels + (size_t)j.plane * npix;
    for (int i = threadIdx.x; i < TW + 2 * (TH - 1); i += 128) {
        int ly, lx;
        if (i < TW) { ly = 0; lx = i; }
        else if (i < TW + TH - 1) { ly = i - TW + 1; lx = 0; }
        else { ly = i - (TW + TH - 1) + 1; lx = TW - 1; }
        const int ry = ty0 + ly, rx = tx0 + lx;
        if (ry >= rows || rx >= cols) continue;
        const int p = (j.y0 + ry) * a.W + (j.x0 + rx);
        if (!map[p]) continue;
        if (lx == 0 && rx > 0 && map[p - 1]) uf_union(lab, p, p - 1);
        if (ry > 0) {
            const int up = p - a.W;
            if ((lx == 0 || ly == 0) && rx > 0 && map[up - 1]) uf_union(lab, p, up - 1);
            if (ly == 0 && map[up]) uf_union(lab, p, up);
            if ((ly == 0 || lx == TW - 1) && rx + 1 < cols && map[up + 1]) uf_union(lab, p, up + 1);
        }
    }
}

// ---- pass 3 / 4 walk every pixel of the rectangle, 4 pixels of a row per thread
template <bool COUNT>
__global__ void __launch_bounds__(NT) canny_resolve_kernel(CannyArgs a) {
    const Job j = make_job(a, blockIdx.z);
    const int rows = j.y1 - j.y0, cols = j.x1 - j.x0;
    const int ty0 = blockIdx.y * TH, tx0 = blockIdx.x * TW;
    if (ty0 >= rows || tx0 >= cols) return;
    const size_t npix = (size_t)a.H * a.W;
    const uint8_t* map = a.map + (size_t)j.plane * npix;
    int* lab = a.labels + (size_t)j.plane * npix;
    unsigned cnt = 0;
    for (int i = threadIdx.x; i < TH * TW; i += NT) {
        const int ly = i / TW, lx = i - ly * TW;
        const int ry = ty0 + ly, rx = tx0 + lx;
        if (ry >= rows || rx >= cols) continue;
        const int p = (j.y0 + ry) * a.W + (j.x0 + rx);
        const int cls = map[p];
        if (COUNT) {
            if (!cls) continue;
            const int root = uf_find((volatile int*)lab, p);
            if (lab[root] & LABEL_FLAG) ++cnt;
        } else {
            if (!(cls & CLS_ROOT_STRONG)) continue;
            const int root = uf_find((volatile int*)lab, p);
            atomicOr(&lab[root], LABEL_FLAG);
        }
    }
    if (COUNT) {
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
}

}  // namespace

int launch_canny(const CannyArgs& a, int max_rows, int max_cols, int instances, cudaStream_t st) {
    if (max_rows <= 0 || max_cols <= 0 || instances <= 0) return 0;
    dim3 grid((max_cols + TW - 1) / TW, (max_rows + TH - 1) / TH, instances);
    if (a.mode == CANNY_QUADRANTS) canny_tile_kernel<CANNY_QUADRANTS><<<grid, NT, 0, st>>>(a);
    else canny_tile_kernel<CANNY_PLANES><<<grid, NT, 0, st>>>(a);
    int launches = 3;
    if (grid.x > 1 || grid.y > 1) { canny_merge_kernel<<<grid, 128, 0, st>>>(a); ++launches; }
    canny_resolve_kernel<false><<<grid, NT, 0, st>>>(a);
    canny_resolve_kernel<true><<<grid, NT, 0, st>>>(a);
    return launches;
}

}  // namespace uwie
