// Canny(50,150) edge COUNT on the device, exact w.r.t. cv2.Canny:
//   Sobel 3x3 (BORDER_REPLICATE) -> L1 magnitude (0 outside) -> NMS with TG22 fixed point ->
//   double threshold -> 8-connected hysteresis -> number of edge pixels.
// Reference call sites: quality_assessment.py:139 (whole enhanced frame) and
// enhancement_strategies.py:180-182 / six_stadigy.py:149-151 (each quadtree block, with the
// block's own borders).
//
// Only the COUNT is wanted, so no edge map and no per-pixel label plane is ever written:
//   1. canny_tile_kernel   64x64 tile: gradient (column march over shared memory), NMS, double
//                          threshold -> per-row candidate / strong bit masks; runs of candidates are
//                          linked by a union-find in shared memory; every component gets
//                          (size, holds-a-strong-pixel, touches-a-neighbouring-tile).  Components
//                          that stay inside the tile are counted on the spot.  The others go to a
//                          short per-tile list {root pixel, size, strong}, and the pixels on the
//                          tile's rim get a label (root pixel index, -1 = not a candidate) in the
//                          sparse label plane — about 6 % of the plane is touched.
//   2. canny_merge_kernel  rim pixels link across tile borders (lock-free atomicMin union-find on
//                          the root entries of the label plane).
//   3. canny_list_kernel<false>  listed components that hold a strong pixel flag their global root.
//   4. canny_list_kernel<true>   listed components whose global root is flagged add their size.
// A rectangle that fits one tile needs pass 1 only.
#include "k_common.cuh"
#include "k_kernels.h"

namespace uwie {

namespace {

constexpr int TH = kCannyTile, TW = kCannyTile;   // 64 x 64: one 64-bit mask per tile row
constexpr int NT = 256;
constexpr int GW = TW + 4, GH = TH + 4;           // gray tile with a 2-pixel halo
constexpr int MW = TW + 2, MH = TH + 2;           // magnitude tile with a 1-pixel halo
constexpr int SG_BYTES = (GH * GW + 15) & ~15;
constexpr int LABEL_FLAG = 0x40000000;
constexpr int LABEL_MASK = 0x3fffffff;
constexpr int ATTR_STRONG = 1, ATTR_RIM = 2;      // component attribute word: size << 2 | flags
constexpr int RIM = 2 * TW + 2 * (TH - 2);

static_assert(TW == 64, "a tile row is one 64-bit mask");
static_assert(SG_BYTES + MH * MW * 2 <= TH * TW * 4, "phase-1 buffers alias the attribute array");

struct Job {
    int y0, y1, x0, x1;   // rectangle inside the plane, empty when y0 >= y1
    int img;              // frame index
    int plane;            // which label plane
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

__device__ __forceinline__ int tile_slot() {
    return (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
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

// first pixel of the run of set bits of `m` that contains bit x
__device__ __forceinline__ int run_start(unsigned long long m, int x) {
    const unsigned long long below = ~m & ((1ull << x) - 1ull);
    return below ? 64 - __clzll((long long)below) : 0;
}
// one past the last pixel of the run of set bits of `m` that starts at bit x
__device__ __forceinline__ int run_end(unsigned long long m, int x) {
    const unsigned long long holes = ~m & (~0ull << x);
    return holes ? __ffsll((long long)holes) - 1 : 64;
}

// rim pixel k of a tile (row-major order is irrelevant): top row, bottom row, left and right columns
__device__ __forceinline__ void rim_pixel(int k, int& ly, int& lx) {
    if (k < TW) { ly = 0; lx = k; }
    else if (k < 2 * TW) { ly = TH - 1; lx = k - TW; }
    else if (k < 2 * TW + TH - 2) { ly = k - 2 * TW + 1; lx = 0; }
    else { ly = k - (2 * TW + TH - 2) + 1; lx = TW - 1; }
}

// ---- pass 1
template <int MODE>
__global__ void __launch_bounds__(NT) canny_tile_kernel(CannyArgs a) {
    const Job j = make_job(a, blockIdx.z);
    const int rows = j.y1 - j.y0, cols = j.x1 - j.x0;
    const int ty0 = blockIdx.y * TH, tx0 = blockIdx.x * TW;
    if (ty0 >= rows || tx0 >= cols) return;

    __shared__ __align__(16) unsigned char sraw[TH * TW * 4];
    __shared__ int slab[TH * TW];
    __shared__ unsigned long long smc[TH], sms[TH];    // per tile row: candidate / strong bit masks
    __shared__ unsigned shist[MODE == CANNY_QUADRANTS ? 768 : 1];
    __shared__ int s_nlist;
    __shared__ unsigned s_cnt;
    uint8_t (*sg)[GW] = reinterpret_cast<uint8_t (*)[GW]>(sraw);
    uint16_t (*sm)[MW] = reinterpret_cast<uint16_t (*)[MW]>(sraw + SG_BYTES);
    int* sattr = reinterpret_cast<int*>(sraw);

    const int tid = threadIdx.x;
    const size_t npix = (size_t)a.H * a.W;
    if (tid == 0) { s_nlist = 0; s_cnt = 0; }

    // ---- gray with a 2-pixel replicated halo (replication at the RECTANGLE's border)
    if (MODE == CANNY_QUADRANTS) {
        const uint8_t* frame = a.frames + (size_t)j.img * npix * 3;
        const CastTables* ct = a.cast + j.img;
        for (int i = tid; i < 768; i += NT) shist[i] = 0;
        __syncthreads();
        for (int i = tid; i < GH * GW; i += NT) {
            const int ly = i / GW, lx = i - ly * GW;
            const int ry = ty0 + ly - 2, rx = tx0 + lx - 2;
            const int y = j.y0 + clamp_idx(ry, rows), x = j.x0 + clamp_idx(rx, cols);
            const uint8_t* p = frame + ((size_t)y * a.W + x) * 3;
            const int r0 = p[0], g0 = p[1], b0 = p[2];
            sg[ly][lx] = (uint8_t)gray_q15(ct->curq[0][r0], ct->curq[1][g0], ct->curq[2][b0]);
            const bool inner = ly >= 2 && ly < TH + 2 && lx >= 2 && lx < TW + 2 && ry < rows && rx < cols;
            if (inner) {
                atomicAdd(&shist[r0], 1u);
                atomicAdd(&shist[256 + g0], 1u);
                atomicAdd(&shist[512 + b0], 1u);
            }
        }
    } else {
        const uint8_t* gplane = a.gray + (size_t)j.plane * npix;
        const bool word_ok = (a.W & 3) == 0 && ty0 >= 2 && ty0 + TH + 2 <= rows && tx0 >= 4 && tx0 + TW + 4 <= cols;
        if (word_ok) {
            // interior tile: 18 aligned 32-bit words cover columns tx0-4 .. tx0+TW+3; the halo needs
            // bytes 2,3 of the first word and bytes 0,1 of the last
            for (int i = tid; i < GH * 18; i += NT) {
                const int ly = i / 18, wx = i - ly * 18;
                const unsigned v = *reinterpret_cast<const unsigned*>(gplane + (size_t)(ty0 + ly - 2) * a.W + tx0 - 4 + wx * 4);
                const int lx = wx * 4 - 2;     // tile column of byte 0 of this word
                if (wx == 0) { sg[ly][0] = (uint8_t)(v >> 16); sg[ly][1] = (uint8_t)(v >> 24); }
                else if (wx == 17) { sg[ly][lx] = (uint8_t)v; sg[ly][lx + 1] = (uint8_t)(v >> 8); }
                else {
                    // lx = 2 mod 4: two 16-bit stores
                    *reinterpret_cast<uint16_t*>(&sg[ly][lx]) = (uint16_t)v;
                    *reinterpret_cast<uint16_t*>(&sg[ly][lx + 2]) = (uint16_t)(v >> 16);
                }
            }
        } else {
            for (int i = tid; i < GH * GW; i += NT) {
                const int ly = i / GW, lx = i - ly * GW;
                const int y = clamp_idx(ty0 + ly - 2, rows), x = clamp_idx(tx0 + lx - 2, cols);
                sg[ly][lx] = gplane[(size_t)y * a.W + x];
            }
        }
    }
    __syncthreads();

    // ---- Sobel by column march: thread = one column of the magnitude tile x one third of its rows.
    // h1 = g[x+1] - g[x-1], h2 = g[x-1] + 2 g[x] + g[x+1] per gray row; dx = h1(y-1) + 2 h1(y) + h1(y+1),
    // dy = h2(y+1) - h2(y-1).  Stored: L1 magnitude (0 outside the rectangle) | NMS direction << 12
    // (0 horizontal, 1 vertical, 2 the / diagonal, 3 the \ diagonal; cv2's TG22 fixed-point test).
    {
        constexpr int CHUNK = (MH + 2) / 3;
        const int mx = tid % MW, chunk = tid / MW;
        if (chunk < 3) {
            const int my0 = chunk * CHUNK, my1 = min(my0 + CHUNK, MH);
            const int rx = tx0 + mx - 1;
            const bool col_in = rx >= 0 && rx < cols;
            int h1a, h1b, h2a, h2b;
            {
                const int g0 = sg[my0][mx], g1 = sg[my0][mx + 1], g2 = sg[my0][mx + 2];
                h1a = g2 - g0; h2a = g0 + 2 * g1 + g2;
                const int k0 = sg[my0 + 1][mx], k1 = sg[my0 + 1][mx + 1], k2 = sg[my0 + 1][mx + 2];
                h1b = k2 - k0; h2b = k0 + 2 * k1 + k2;
            }
            for (int my = my0; my < my1; ++my) {
                const int g0 = sg[my + 2][mx], g1 = sg[my + 2][mx + 1], g2 = sg[my + 2][mx + 2];
                const int h1c = g2 - g0, h2c = g0 + 2 * g1 + g2;
                const int dx = h1a + 2 * h1b + h1c, dy = h2c - h2a;
                h1a = h1b; h1b = h1c; h2a = h2b; h2b = h2c;
                const int ry = ty0 + my - 1;
                const int ax = abs(dx), ay = abs(dy);
                int v = 0;
                if (col_in && ry >= 0 && ry < rows) {
                    const int ys = ay << 15, tg22x = ax * 13573;
                    int dir;
                    if (ys < tg22x) dir = 0;
                    else if (ys > tg22x + (ax << 16)) dir = 1;
                    else dir = ((dx ^ dy) < 0) ? 2 : 3;
                    v = (ax + ay) | (dir << 12);
                }
                sm[my][mx] = (uint16_t)v;
            }
        }
    }
    __syncthreads();

    // ---- NMS + double threshold.  In each of the 16 passes a warp owns half a tile row, so one ballot
    // gives the candidate / strong bit masks of that half row.
    //   dir 0: m > left  && m >= right     dir 1: m > up && m >= down
    //   dir 2: m > up-right && m > down-left     dir 3: m > up-left && m > down-right
    for (int it = 0; it < TH * TW / NT; ++it) {
        const int i = it * NT + tid;
        const int ly = i >> 6, lx = i & 63;
        const uint16_t* c = &sm[ly + 1][lx + 1];
        const int v = c[0], m = v & 0xfff;
        bool cand = false;
        if (m > a.low) {
            const int dir = v >> 12;
            const int d = dir == 0 ? 1 : (dir == 1 ? MW : (dir == 2 ? MW - 1 : MW + 1));
            const int mf = c[-d] & 0xfff, ms = c[d] & 0xfff;
            cand = m > mf && m > ms - (dir < 2 ? 1 : 0);
        }
        const unsigned mc = __ballot_sync(0xffffffffu, cand);
        const unsigned mst = __ballot_sync(0xffffffffu, cand && m > a.high);
        if ((tid & 31) == 0) {
            reinterpret_cast<unsigned*>(smc)[ly * 2 + (lx >> 5)] = mc;
            reinterpret_cast<unsigned*>(sms)[ly * 2 + (lx >> 5)] = mst;
        }
    }
    __syncthreads();      // sg / sm are dead from here on: sattr takes their place

    // ---- link inside the tile.  A horizontal run of candidates is one node (its first pixel); runs of
    // neighbouring rows are united once per contact: with U the row above,
    //   up set:            unite unless both W and NW are set (then the pixel to the left did it)
    //   up clear, NW set:  unite if W is clear          up clear, NE set:  unite if E is clear
    for (int it = 0; it < TH * TW / NT; ++it) {
        const int i = it * NT + tid;
        const int ly = i >> 6, lx = i & 63;
        const unsigned long long R = smc[ly];
        sattr[i] = 0;
        if (((R >> lx) & 1ull) && (lx == 0 || !((R >> (lx - 1)) & 1ull))) slab[i] = i;   // run starts are the nodes
    }
    __syncthreads();
    for (int it = 0; it < TH * TW / NT; ++it) {
        const int i = it * NT + tid;
        const int ly = i >> 6, lx = i & 63;
        if (ly == 0) continue;                                                  // warp-uniform
        const unsigned long long R = smc[ly], U = smc[ly - 1];
        if ((unsigned)(R >> (lx & 32)) == 0u || U == 0ull) continue;            // warp-uniform
        if (!((R >> lx) & 1ull)) continue;
        const bool up = (U >> lx) & 1ull;
        const bool nw = lx > 0 && ((U >> (lx - 1)) & 1ull), ne = lx < TW - 1 && ((U >> (lx + 1)) & 1ull);
        const bool w = lx > 0 && ((R >> (lx - 1)) & 1ull), e = lx < TW - 1 && ((R >> (lx + 1)) & 1ull);
        if (up) {
            if (!w || !nw) uf_union(slab, ly * TW + run_start(R, lx), (ly - 1) * TW + run_start(U, lx));
        } else {
            if (nw && !w) uf_union(slab, ly * TW + run_start(R, lx), (ly - 1) * TW + run_start(U, lx - 1));
            if (ne && !e) uf_union(slab, ly * TW + run_start(R, lx), (ly - 1) * TW + lx + 1);   // NE starts its run (up is clear)
        }
    }
    __syncthreads();

    // ---- component attributes, gathered run by run (the thread of the run's first pixel)
    const bool has_top = ty0 > 0, has_bottom = ty0 + TH < rows, has_left = tx0 > 0, has_right = tx0 + TW < cols;
    const bool any_nb = has_top || has_bottom || has_left || has_right;
    for (int it = 0; it < TH * TW / NT; ++it) {
        const int i = it * NT + tid;
        const int ly = i >> 6, lx = i & 63;
        const unsigned long long R = smc[ly];
        if ((unsigned)(R >> (lx & 32)) == 0u) continue;                         // warp-uniform
        if (!((R >> lx) & 1ull) || (lx > 0 && ((R >> (lx - 1)) & 1ull))) continue;   // not a run start
        const int end = run_end(R, lx);
        const unsigned long long run = (end >= 64 ? ~0ull : ((1ull << end) - 1ull)) & (~0ull << lx);
        int flags = (sms[ly] & run) ? ATTR_STRONG : 0;
        if ((ly == 0 && has_top) || (ly == TH - 1 && has_bottom) || (lx == 0 && has_left) || (end == TW && has_right))
            flags |= ATTR_RIM;
        const int root = uf_find((volatile int*)slab, i);
        atomicAdd(&sattr[root], (end - lx) << 2);
        if (flags) atomicOr(&sattr[root], flags);
    }
    __syncthreads();

    // ---- roots: count what is settled, list what touches a neighbouring tile
    int* lab = a.labels + (size_t)j.plane * npix;
    const int slot = tile_slot();
    unsigned cnt = 0;
    for (int it = 0; it < TH * TW / NT; ++it) {
        const int i = it * NT + tid;
        const int ly = i >> 6, lx = i & 63;
        const unsigned long long R = smc[ly];
        if ((unsigned)(R >> (lx & 32)) == 0u) continue;
        if (!((R >> lx) & 1ull) || (lx > 0 && ((R >> (lx - 1)) & 1ull))) continue;
        if (slab[i] != i) continue;
        const int attr = sattr[i];
        if (attr & ATTR_RIM) {
            const int gp = (j.y0 + ty0 + ly) * a.W + (j.x0 + tx0 + lx);
            const int k = atomicAdd(&s_nlist, 1);
            if (k < kCannyListCap) a.tile_list[(size_t)slot * kCannyListCap + k] = make_int2(gp, attr);
            lab[gp] = gp;
        } else if (attr & ATTR_STRONG) {
            cnt += (unsigned)(attr >> 2);
        }
    }
    // ---- rim labels for the merge pass
    if (any_nb && tid < RIM) {
        int ly, lx;
        rim_pixel(tid, ly, lx);
        const int ry = ty0 + ly, rx = tx0 + lx;
        if (ry < rows && rx < cols) {
            const unsigned long long R = smc[ly];
            int v = -1;
            if ((R >> lx) & 1ull) {
                const int root = uf_find((volatile int*)slab, ly * TW + run_start(R, lx));
                v = (j.y0 + ty0 + (root >> 6)) * a.W + (j.x0 + tx0 + (root & 63));
            }
            const int gp = (j.y0 + ry) * a.W + (j.x0 + rx);
            lab[gp] = v;                   // (a root on the rim gets its own index, as above)
        }
    }
    cnt = warp_sum_u32(cnt);
    if ((tid & 31) == 0 && cnt) atomicAdd(&s_cnt, cnt);
    __syncthreads();
    if (tid == 0) {
        a.tile_n[slot] = min(s_nlist, kCannyListCap);
        if (s_cnt) atomicAdd(&a.counts[j.out], s_cnt);
    }

    if (MODE == CANNY_QUADRANTS) {
        unsigned* gh = a.quad_hist + (size_t)j.out * 768;
        for (int i = tid; i < 768; i += NT) {
            const unsigned v = shist[i];
            if (v) atomicAdd(&gh[i], v);
        }
    }
}

// ---- pass 2: links that cross a tile border.  One thread per rim pixel; every cross-tile pair of
// 8-neighbours is handled by its lower / right member (neighbours W, NW, N, NE).
__global__ void __launch_bounds__(NT) canny_merge_kernel(CannyArgs a) {
    const Job j = make_job(a, blockIdx.z);
    const int rows = j.y1 - j.y0, cols = j.x1 - j.x0;
    const int ty0 = blockIdx.y * TH, tx0 = blockIdx.x * TW;
    if (ty0 >= rows || tx0 >= cols || threadIdx.x >= RIM) return;
    int* lab = a.labels + (size_t)j.plane * ((size_t)a.H * a.W);
    int ly, lx;
    rim_pixel(threadIdx.x, ly, lx);
    if (ly == TH - 1 && lx > 0 && lx < TW - 1) return;       // inner bottom row: all its earlier neighbours are in this tile
    const int ry = ty0 + ly, rx = tx0 + lx;
    if (ry >= rows || rx >= cols) return;
    const int p = (j.y0 + ry) * a.W + (j.x0 + rx);
    if (lab[p] < 0) return;
    if (lx == 0 && rx > 0 && lab[p - 1] >= 0) uf_union(lab, p, p - 1);
    if (ry > 0) {
        const int up = p - a.W;
        if ((lx == 0 || ly == 0) && rx > 0 && lab[up - 1] >= 0) uf_union(lab, p, up - 1);
        if (ly == 0 && lab[up] >= 0) uf_union(lab, p, up);
        if ((ly == 0 || lx == TW - 1) && rx + 1 < cols && lab[up + 1] >= 0) uf_union(lab, p, up + 1);
    }
}

// ---- pass 3 / 4 walk the per-tile lists
template <bool COUNT>
__global__ void __launch_bounds__(kCannyListCap) canny_list_kernel(CannyArgs a) {
    const Job j = make_job(a, blockIdx.z);
    const int rows = j.y1 - j.y0, cols = j.x1 - j.x0;
    const int ty0 = blockIdx.y * TH, tx0 = blockIdx.x * TW;
    if (ty0 >= rows || tx0 >= cols) return;
    const int slot = tile_slot();
    const int n = a.tile_n[slot];
    int* lab = a.labels + (size_t)j.plane * ((size_t)a.H * a.W);
    unsigned cnt = 0;
    if ((int)threadIdx.x < n) {
        const int2 e = a.tile_list[(size_t)slot * kCannyListCap + threadIdx.x];
        if (COUNT) {
            const int root = uf_find((volatile int*)lab, e.x);
            if (lab[root] & LABEL_FLAG) cnt = (unsigned)(e.y >> 2);
        } else if (e.y & ATTR_STRONG) {
            const int root = uf_find((volatile int*)lab, e.x);
            if (!(lab[root] & LABEL_FLAG)) atomicOr(&lab[root], LABEL_FLAG);
        }
    }
    if (COUNT) {
        if (n == 0) return;                                   // block-uniform
        cnt = warp_sum_u32(cnt);
        __shared__ unsigned sw[kCannyListCap / 32];
        if ((threadIdx.x & 31) == 0) sw[threadIdx.x >> 5] = cnt;
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned t = 0;
            for (int w = 0; w < kCannyListCap / 32; ++w) t += sw[w];
            if (t) atomicAdd(&a.counts[j.out], t);
        }
    }
}

}  // namespace

size_t canny_tile_slots(int max_rows, int max_cols, int instances) {
    if (max_rows <= 0 || max_cols <= 0 || instances <= 0) return 0;
    return (size_t)((max_cols + TW - 1) / TW) * ((max_rows + TH - 1) / TH) * instances;
}

int launch_canny(const CannyArgs& a, int max_rows, int max_cols, int instances, cudaStream_t st) {
    if (max_rows <= 0 || max_cols <= 0 || instances <= 0) return 0;
    dim3 grid((max_cols + TW - 1) / TW, (max_rows + TH - 1) / TH, instances);
    if (a.mode == CANNY_QUADRANTS) canny_tile_kernel<CANNY_QUADRANTS><<<grid, NT, 0, st>>>(a);
    else canny_tile_kernel<CANNY_PLANES><<<grid, NT, 0, st>>>(a);
    if (grid.x == 1 && grid.y == 1) return 1;
    canny_merge_kernel<<<grid, NT, 0, st>>>(a);
    canny_list_kernel<false><<<grid, kCannyListCap, 0, st>>>(a);
    canny_list_kernel<true><<<grid, kCannyListCap, 0, st>>>(a);
    return 4;
}

}  // namespace uwie
