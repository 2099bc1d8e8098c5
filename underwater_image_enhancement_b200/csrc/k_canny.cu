// Canny(50,150) edge COUNT on the device, exact w.r.t. cv2.Canny:
//   Sobel 3x3 (BORDER_REPLICATE) -> L1 magnitude (0 outside) -> NMS with TG22 fixed point ->
//   double threshold -> 8-connected hysteresis -> number of edge pixels.
// Reference call sites: quality_assessment.py:139 (whole enhanced frame) and
// enhancement_strategies.py:180-182 / six_stadigy.py:149-151 (each quadtree block, with the
// block's own borders).
//
// Only the COUNT is wanted, so no edge map and no per-pixel label plane is ever written:
//   1. canny_tile_kernel   64x64 tile: gradient (column march over shared memory), NMS, double
//                          threshold -> per-row candidate / strong bit masks (64-bit words).  Two warps
//                          then flood the masks side by side: E = candidates connected inside the tile
//                          to one of its strong pixels (counted on the spot), P = other candidates that
//                          reach a side shared with a neighbouring tile (pending).  Runs of P are linked by a
//                          union-find in shared memory; each pending component goes to a short
//                          per-tile list (node k of the tile, its size); the 252 pixels on the tile's rim
//                          get a label (edge / node of a pending component / none) in the tile's own
//                          contiguous rim array - 1 KB per tile, no full-resolution label plane.
//   2. canny_rim_kernel<false>  pending rim pixels link across tile borders (lock-free atomicMin
//                          union-find over the nodes, one entry per listed component).
//   3. canny_rim_kernel<true>   a pending rim pixel next to a certain edge flags its global root.
//   4. canny_list_kernel   listed components whose global root is flagged add their size.
// A rectangle that fits one tile needs pass 1 only.
#include "k_common.cuh"
#include "k_kernels.h"

namespace uwie {

namespace {

constexpr int TH = kCannyTile, TW = kCannyTile;   // 64 x 64: one 64-bit mask per tile row
constexpr int NT = 256;
constexpr int GL = TW + 4, GH = TH + 4;           // gray tile with a 2-pixel halo (logical width)
constexpr int GPAD = 2, GW = GL + 2 * GPAD;       // stored with 2 more columns either side: tile column -4 is byte 0 of a row,
                                                  // so an interior tile copies aligned 32-bit words straight in
constexpr int MW = TW + 2, MH = TH + 2;           // magnitude tile with a 1-pixel halo
constexpr int SG_BYTES = (GH * GW + 15) & ~15;
constexpr int LABEL_FLAG = 0x40000000;
constexpr int LABEL_MASK = 0x3fffffff;
constexpr int LAB_NONE = -1, LAB_EDGE = -2;       // rim labels that are not a pending component
constexpr int RIM = 2 * TW + 2 * (TH - 2);

static_assert(TW == 64, "a tile row is one 64-bit mask");
static_assert(NT == 4 * TH, "pending runs: one thread per quarter of a tile row");

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
// Scratch of one launch (CannyArgs::labels): rim labels [slots][RIM], then the union-find nodes [slots][kCannyListCap].
// Node id = slot * kCannyListCap + k, k = position in the tile's list.
__device__ __forceinline__ int total_slots() { return gridDim.x * gridDim.y * gridDim.z; }

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

// ---- union-find of the pending runs of ONE tile, in shared memory.  A node is a run, named by its row and
// the pair of columns its first pixel lies in (two runs cannot start in the same pair): row * 32 + (x >> 1).
// A word holds label << 16 | size: atomicMin orders by label, and once the links are final the sizes are
// added into the low half of the root's word (a component has at most 4096 pixels).
__device__ __forceinline__ int run_node(int row, int x) { return row * 32 + (x >> 1); }
// column of the first pixel of node `n`, given the pending mask of its row
__device__ __forceinline__ int node_column(unsigned long long rowmask, int n) {
    const int xx = 2 * (n & 31);
    const unsigned long long starts = rowmask & ~(rowmask << 1);
    return ((starts >> xx) & 1ull) ? xx : xx + 1;
}
__device__ __forceinline__ int suf_find(volatile int* L, int x) {
    int l;
    while ((l = (L[x] >> 16)) != x) x = l;
    return x;
}
__device__ __forceinline__ void suf_union(int* L, int a, int b) {
    bool done;
    do {
        a = suf_find(L, a);
        b = suf_find(L, b);
        if (a < b) { const int old = atomicMin(&L[b], a << 16) >> 16; done = (old == b); b = old; }
        else if (b < a) { const int old = atomicMin(&L[a], b << 16) >> 16; done = (old == a); a = old; }
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

// all bits of the runs of `c` that hold a bit of `s` (s within c): the carry of c + s runs up through a
// run from its lowest seed; the bit-reversed sum runs down (rc = c bit-reversed, computed once per row)
__device__ __forceinline__ unsigned long long fill_runs(unsigned long long s, unsigned long long c, unsigned long long rc) {
    const unsigned long long up = (c ^ (c + s)) & c;
    const unsigned long long rs = __brevll(s);
    const unsigned long long dn = __brevll((rc ^ (rc + rs)) & rc);
    return up | dn | s;
}
__device__ __forceinline__ unsigned long long dilate3(unsigned long long m) { return m | (m << 1) | (m >> 1); }

// 8-connected flood of the seeds (e0, e1) inside the candidates (c0, c1) of a 64-row tile held by one
// warp, lane l owning rows 2l and 2l+1; seeds must lie inside the candidates.  Runs until nothing moves.
// Within a round the rows are updated in the order 2l (from the lane above and row 2l+1), 2l+1 (from the
// lane below and the NEW row 2l), 2l again (from the NEW row 2l+1): a front moves two rows per round in
// either direction.
__device__ __forceinline__ void flood_rows(unsigned long long& e0, unsigned long long& e1,
                                           unsigned long long c0, unsigned long long c1, int lane) {
    const unsigned long long rc0 = __brevll(c0), rc1 = __brevll(c1);
    e0 = fill_runs(e0, c0, rc0); e1 = fill_runs(e1, c1, rc1);
    while (true) {
        unsigned long long up = __shfl_up_sync(0xffffffffu, e1, 1);
        unsigned long long dn = __shfl_down_sync(0xffffffffu, e0, 1);
        if (lane == 0) up = 0ull;
        if (lane == 31) dn = 0ull;
        const unsigned long long n0 = dilate3(up | e1) & c0 & ~e0;
        if (n0) e0 |= fill_runs(n0, c0, rc0);
        const unsigned long long n1 = dilate3(dn | e0) & c1 & ~e1;
        if (n1) e1 |= fill_runs(n1, c1, rc1);
        const unsigned long long m0 = dilate3(e1) & c0 & ~e0;
        if (m0) e0 |= fill_runs(m0, c0, rc0);
        if (!__any_sync(0xffffffffu, (n0 | n1 | m0) != 0ull)) break;
    }
}

// rim pixel k of a tile (row-major order is irrelevant): top row, bottom row, left and right columns
__device__ __forceinline__ void rim_pixel(int k, int& ly, int& lx) {
    if (k < TW) { ly = 0; lx = k; }
    else if (k < 2 * TW) { ly = TH - 1; lx = k - TW; }
    else if (k < 2 * TW + TH - 2) { ly = k - 2 * TW + 1; lx = 0; }
    else { ly = k - (2 * TW + TH - 2) + 1; lx = TW - 1; }
}

// inverse of rim_pixel for a pixel that lies on the rim
__device__ __forceinline__ int rim_index(int ly, int lx) {
    if (ly == 0) return lx;
    if (ly == TH - 1) return TW + lx;
    return lx == 0 ? 2 * TW + ly - 1 : 2 * TW + (TH - 2) + ly - 1;
}

// ---- pass 1
template <int MODE>
__global__ void __launch_bounds__(NT) canny_tile_kernel(CannyArgs a) {
    const Job j = make_job(a, blockIdx.z);
    const int rows = j.y1 - j.y0, cols = j.x1 - j.x0;
    const int ty0 = blockIdx.y * TH, tx0 = blockIdx.x * TW;
    if (ty0 >= rows || tx0 >= cols) return;

    __shared__ __align__(16) unsigned char sraw[SG_BYTES + ((MH * MW * 2 + 15) & ~15)];
    __shared__ int slab[TH * 32];                      // pending runs: label << 16 | size, see run_node
    __shared__ unsigned long long smc[TH], sms[TH];    // per tile row: candidate / strong bit masks
    __shared__ unsigned long long sme[TH], smp[TH];    // per tile row: certain edges / pending candidates
    __shared__ int s_nlist;
    uint8_t (*sg)[GW] = reinterpret_cast<uint8_t (*)[GW]>(sraw);
    uint16_t (*sm)[MW] = reinterpret_cast<uint16_t (*)[MW]>(sraw + SG_BYTES);
    unsigned* shist = reinterpret_cast<unsigned*>(slab);   // quadrant histograms: live only while the tile is loaded

    const int tid = threadIdx.x;
    const size_t npix = (size_t)a.H * a.W;
    if (tid == 0) s_nlist = 0;

    // ---- gray with a 2-pixel replicated halo (replication at the RECTANGLE's border)
    if (MODE == CANNY_QUADRANTS) {
        const uint8_t* frame = a.frames + (size_t)j.img * npix * 3;
        const CastTables* ct = a.cast + j.img;
        // the frame's byte tables (colour-cast correction + quantisation) in shared memory: the magnitude tile is not live yet
        uint8_t* s_curq = sraw + SG_BYTES;
        for (int i = tid; i < 768; i += NT) { shist[i] = 0; s_curq[i] = (&ct->curq[0][0])[i]; }
        __syncthreads();
        for (int i = tid; i < GH * GL; i += NT) {
            const int ly = i / GL, lx = i - ly * GL;
            const int ry = ty0 + ly - 2, rx = tx0 + lx - 2;
            const int y = j.y0 + clamp_idx(ry, rows), x = j.x0 + clamp_idx(rx, cols);
            const uint8_t* p = frame + ((size_t)y * a.W + x) * 3;
            const int r0 = p[0], g0 = p[1], b0 = p[2];
            sg[ly][GPAD + lx] = (uint8_t)gray_q15(s_curq[r0], s_curq[256 + g0], s_curq[512 + b0]);
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
            // interior tile: 18 aligned 32-bit words cover columns tx0-4 .. tx0+TW+3 = bytes 0 .. GW-1 of a row
            for (int i = tid; i < GH * 18; i += NT) {
                const int ly = i / 18, wx = i - ly * 18;
                *reinterpret_cast<unsigned*>(&sg[ly][wx * 4]) =
                    *reinterpret_cast<const unsigned*>(gplane + (size_t)(ty0 + ly - 2) * a.W + tx0 - 4 + wx * 4);
            }
        } else {
            for (int i = tid; i < GH * GL; i += NT) {
                const int ly = i / GL, lx = i - ly * GL;
                const int y = clamp_idx(ty0 + ly - 2, rows), x = clamp_idx(tx0 + lx - 2, cols);
                sg[ly][GPAD + lx] = gplane[(size_t)y * a.W + x];
            }
        }
    }
    __syncthreads();
    if (MODE == CANNY_QUADRANTS) {                   // (slab is not written again before the barriers below)
        unsigned* gh = a.quad_hist + (size_t)j.out * 768;
        for (int i = tid; i < 768; i += NT) {
            const unsigned v = shist[i];
            if (v) atomicAdd(&gh[i], v);
        }
    }

    // ---- Sobel by column march: thread = one column of the magnitude tile x one quarter of its rows.
    // h1 = g[x+1] - g[x-1], h2 = g[x-1] + 2 g[x] + g[x+1] per gray row; dx = h1(y-1) + 2 h1(y) + h1(y+1),
    // dy = h2(y+1) - h2(y-1).  Stored: L1 magnitude (0 outside the rectangle) | NMS direction << 12
    // (0 horizontal, 1 vertical, 2 the / diagonal, 3 the \ diagonal; cv2's TG22 fixed-point test).
    {
        // All 256 threads march: column 1 + (tid & 63) of the magnitude tile (its 64 inner columns) x one quarter of its 66
        // rows.  The two halo columns (0 and 65) are 132 single outputs, taken afterwards by threads 0 .. 131.
        const int sob_low = a.low;
        auto encode = [&](int dx, int dy, bool inside) -> int {
            const int ax = abs(dx), ay = abs(dy);
            int v = 0;
            if (inside) {
                v = ax + ay;
                if (v > sob_low) {                       // the direction is only read for candidates (m > low): flat rows skip it
                    const int ys = ay << 15, tg22x = ax * 13573;
                    int dir;
                    if (ys < tg22x) dir = 0;
                    else if (ys > tg22x + (ax << 16)) dir = 1;
                    else dir = ((dx ^ dy) < 0) ? 2 : 3;
                    v |= dir << 12;
                }
            }
            return v;
        };
        {
            constexpr int CHUNK = (MH + 3) / 4;
            const int mx = 1 + (tid & 63), chunk = tid >> 6;
            const int my0 = chunk * CHUNK, my1 = min(my0 + CHUNK, MH);
            const int rx = tx0 + mx - 1;
            const bool col_in = rx >= 0 && rx < cols;
            // rows of this chunk that lie inside the rectangle: one unsigned compare per row
            const int vlo = max(my0, 1 - ty0);
            const unsigned vspan = col_in ? (unsigned)max(min(my1, rows - ty0 + 1) - vlo, 0) : 0u;
            int h1a, h1b, h2a, h2b;
            {
                const uint8_t* r0p = &sg[my0][GPAD + mx];
                const int g0 = r0p[0], g1 = r0p[1], g2 = r0p[2];
                h1a = g2 - g0; h2a = g0 + 2 * g1 + g2;
                const int k0 = r0p[GW], k1 = r0p[GW + 1], k2 = r0p[GW + 2];
                h1b = k2 - k0; h2b = k0 + 2 * k1 + k2;
            }
            for (int my = my0; my < my1; ++my) {
                const uint8_t* rp = &sg[my + 2][GPAD + mx];
                const int g0 = rp[0], g1 = rp[1], g2 = rp[2];
                const int h1c = g2 - g0, h2c = g0 + 2 * g1 + g2;
                const int dx = h1a + 2 * h1b + h1c, dy = h2c - h2a;
                h1a = h1b; h1b = h1c; h2a = h2b; h2b = h2c;
                sm[my][mx] = (uint16_t)encode(dx, dy, (unsigned)(my - vlo) < vspan);
            }
        }
        if (tid < 2 * MH) {
            const int my = tid < MH ? tid : tid - MH, mx = tid < MH ? 0 : MW - 1;
            const int ry = ty0 + my - 1, rx = tx0 + mx - 1;
            const uint8_t* p0 = &sg[my][GPAD + mx];
            const uint8_t* p1 = p0 + GW;
            const uint8_t* p2 = p1 + GW;
            const int dx = (p0[2] - p0[0]) + 2 * (p1[2] - p1[0]) + (p2[2] - p2[0]);
            const int dy = (p2[0] + 2 * p2[1] + p2[2]) - (p0[0] + 2 * p0[1] + p0[2]);
            sm[my][mx] = (uint16_t)encode(dx, dy, ry >= 0 && ry < rows && rx >= 0 && rx < cols);
        }
    }
    __syncthreads();

    // ---- NMS + double threshold.  In each of the 16 passes a warp owns half a tile row, so one ballot
    // gives the candidate / strong bit masks of that half row.
    //   dir 0: m > left  && m >= right     dir 1: m > up && m >= down
    //   dir 2: m > up-right && m > down-left     dir 3: m > up-left && m > down-right
    // The 16 mask words of a warp stay in registers (lane `it` keeps pass `it`) and are stored once.
    {
        const int lane = tid & 31, warp = tid >> 5;
        const uint16_t* c = &sm[(tid >> 6) + 1][(tid & 63) + 1];
        const int low = a.low, high = a.high;
        unsigned keep_c = 0u, keep_s = 0u;
#pragma unroll
        for (int it = 0; it < TH * TW / NT; ++it, c += (NT / TW) * MW) {
            const int v = c[0], m = v & 0xfff;
            bool cand = false;
            if (m > low) {
                const int dir = v >> 12;
                const int d = (int)((((unsigned)(MW + 1) << 24) | ((MW - 1) << 16) | (MW << 8) | 1u) >> (dir * 8)) & 0xff;
                const int mf = c[-d] & 0xfff, ms = c[d] & 0xfff;
                cand = m > mf && m > ms - (dir < 2 ? 1 : 0);
            }
            const unsigned mc = __ballot_sync(0xffffffffu, cand);
            const unsigned mst = __ballot_sync(0xffffffffu, cand && m > high);
            if (lane == it) { keep_c = mc; keep_s = mst; }
        }
        if (lane < TH * TW / NT) {                 // pass `lane`: tile row lane * 4 + warp / 2, half warp & 1
            reinterpret_cast<unsigned*>(smc)[lane * 8 + warp] = keep_c;
            reinterpret_cast<unsigned*>(sms)[lane * 8 + warp] = keep_s;
        }
    }
    __syncthreads();

    // ---- hysteresis inside the tile: two warps flood the row masks at the same time (lane l owns rows
    // 2l and 2l+1).
    //   E  (warp 0) = candidates connected, inside the tile, to one of its strong pixels: edges for certain
    //   P' (warp 1) = candidates that reach, inside the tile, a side shared with a neighbouring tile
    //   P = P' & ~E = pending: edges iff a chain through other tiles ends at an edge.  (A component of
    //       the candidates lies either wholly in E or wholly outside, so the two floods are independent.)
    // Everything else is dropped here.
    const bool has_top = ty0 > 0, has_bottom = ty0 + TH < rows, has_left = tx0 > 0, has_right = tx0 + TW < cols;
    const bool any_nb = has_top || has_bottom || has_left || has_right;
    if (tid < 64) {
        const int lane = tid & 31;
        const int r0 = 2 * lane, r1 = r0 + 1;
        const unsigned long long c0 = smc[r0], c1 = smc[r1];
        if (tid < 32) {
            unsigned long long e0 = sms[r0], e1 = sms[r1];
            flood_rows(e0, e1, c0, c1, lane);
            sme[r0] = e0; sme[r1] = e1;
            const unsigned cnt = warp_sum_u32((unsigned)(__popcll(e0) + __popcll(e1)));
            if (lane == 0 && cnt) atomicAdd(&a.counts[j.out], cnt);
        } else {
            unsigned long long p0 = 0, p1 = 0;
            if (any_nb) {                                                       // block-uniform
                const unsigned long long side = (has_left ? 1ull : 0ull) | (has_right ? (1ull << 63) : 0ull);
                p0 = c0 & ((r0 == 0 && has_top) ? ~0ull : side);
                p1 = c1 & ((r1 == TH - 1 && has_bottom) ? ~0ull : side);
                flood_rows(p0, p1, c0, c1, lane);
            }
            smp[r0] = p0; smp[r1] = p1;
        }
    }
    __syncthreads();
    unsigned long long mine = 0ull;
    if (tid < TH) { mine = smp[tid] & ~sme[tid]; smp[tid] = mine; }
    const int s_any = __syncthreads_or(mine != 0ull);

    // ---- pending components: runs of P are the nodes of a union-find in shared memory (slab)
    const int slot = tile_slot();
    int* rim = a.labels + (size_t)slot * RIM;
    int* nodes = a.labels + (size_t)total_slots() * RIM;
    if (s_any) {                                                                // block-uniform
        // P is sparse: a thread owns the runs that START in one 16-column quarter of a tile row, so the
        // walks are short and all eight warps share them
        const int row = tid >> 2;
        const unsigned long long R = smp[row];
        const unsigned long long U = row > 0 ? smp[row - 1] : 0ull;
        const unsigned long long mystarts = R & ~(R << 1) & (0xffffull << (16 * (tid & 3)));
        for (unsigned long long m = mystarts; m; m &= m - 1) {                  // nodes
            const int n = run_node(row, __ffsll((long long)m) - 1);
            slab[n] = n << 16;
        }
        __syncthreads();
        for (unsigned long long m = mystarts; m; m &= m - 1) {                  // link to the runs of the row above
            const int s0 = __ffsll((long long)m) - 1, e0 = run_end(R, s0);
            // 8-connectivity: columns s0-1 .. e0 of the row above
            const unsigned long long hi = e0 >= 63 ? ~0ull : ((1ull << (e0 + 1)) - 1ull);
            const unsigned long long lo = s0 == 0 ? ~0ull : (~0ull << (s0 - 1));
            for (unsigned long long t = U & hi & lo; t;) {
                const int b0 = __ffsll((long long)t) - 1;
                const int ue = run_end(U, b0);
                suf_union(slab, run_node(row, s0), run_node(row - 1, run_start(U, b0)));
                t &= ue >= 64 ? 0ull : (~0ull << ue);
            }
        }
        __syncthreads();
        for (unsigned long long m = mystarts; m; m &= m - 1) {                  // component sizes, into the roots' low halves
            const int s0 = __ffsll((long long)m) - 1, e0 = run_end(R, s0);
            atomicAdd(&slab[suf_find((volatile int*)slab, run_node(row, s0))], e0 - s0);
        }
        __syncthreads();
        for (unsigned long long m = mystarts; m; m &= m - 1) {                  // roots -> list = nodes of the cross-tile union-find
            const int s0 = __ffsll((long long)m) - 1;
            const int n = run_node(row, s0);
            const int w = slab[n];
            if ((w >> 16) != n) continue;
            // every pending component holds a rim pixel and two of them are never adjacent on the ring: k < 126
            const int k = min(atomicAdd(&s_nlist, 1), kCannyListCap - 1);       // (the bound above makes the clamp dead code)
            const int id = slot * kCannyListCap + k;
            a.tile_list[id] = make_int2(id, w & 0xffff);
            nodes[id] = id;
            slab[n] = (n << 16) | k;                                            // the size is stored: the low half now names the node
        }
        __syncthreads();
    }
    // ---- rim labels for the cross-tile passes: LAB_EDGE, node of a pending component, or LAB_NONE
    if (any_nb && tid < RIM) {
        int ly, lx;
        rim_pixel(tid, ly, lx);
        int v = LAB_NONE;
        if (ty0 + ly < rows && tx0 + lx < cols) {
            const unsigned long long R = smp[ly];
            if ((sme[ly] >> lx) & 1ull) v = LAB_EDGE;
            else if ((R >> lx) & 1ull) {
                const int root = suf_find((volatile int*)slab, run_node(ly, run_start(R, lx)));
                v = slot * kCannyListCap + (slab[root] & 0xffff);
            }
        }
        rim[tid] = v;
    }
    __syncthreads();
    if (tid == 0) a.tile_n[slot] = min(s_nlist, kCannyListCap);

}

// ---- pass 2 / 3 walk the rims.  One thread per rim pixel; every cross-tile pair of 8-neighbours is
// handled by its lower / right member (neighbours W, NW, N, NE).
//   FLAG = false: two pending pixels -> their components are united (lock-free atomicMin union-find
//                 on the root entries of the label plane)
//   FLAG = true : a pending pixel next to a certain edge -> its global root is flagged
template <bool FLAG>
__device__ __forceinline__ void rim_pair(int* nodes, int lp, int lq) {
    if (!FLAG) {
        if (lp >= 0 && lq >= 0) uf_union(nodes, lp, lq);
    } else {
        int t = -1;
        if (lp >= 0 && lq == LAB_EDGE) t = lp;
        else if (lp == LAB_EDGE && lq >= 0) t = lq;
        if (t >= 0) {
            const int root = uf_find((volatile int*)nodes, t);
            if (!(nodes[root] & LABEL_FLAG)) atomicOr(&nodes[root], LABEL_FLAG);
        }
    }
}

template <bool FLAG>
__global__ void __launch_bounds__(NT) canny_rim_kernel(CannyArgs a) {
    const Job j = make_job(a, blockIdx.z);
    const int rows = j.y1 - j.y0, cols = j.x1 - j.x0;
    const int ty0 = blockIdx.y * TH, tx0 = blockIdx.x * TW;
    if (ty0 >= rows || tx0 >= cols || threadIdx.x >= RIM) return;
    int ly, lx;
    rim_pixel(threadIdx.x, ly, lx);
    if (ly == TH - 1 && lx > 0 && lx < TW - 1) return;       // inner bottom row: all its earlier neighbours are in this tile
    const int ry = ty0 + ly, rx = tx0 + lx;
    if (ry >= rows || rx >= cols) return;
    const int slot = tile_slot();
    const int lp = a.labels[(size_t)slot * RIM + threadIdx.x];
    if (lp == LAB_NONE) return;
    int* nodes = a.labels + (size_t)total_slots() * RIM;
    // label of the pixel (ry + dy, rx + dx), which lies in another tile (and therefore on that tile's rim)
    auto other = [&](int dy, int dx) -> int {
        const int qy = ry + dy, qx = rx + dx;
        const int by = qy / TH, bx = qx / TW;
        const int qslot = (blockIdx.z * gridDim.y + by) * gridDim.x + bx;
        return a.labels[(size_t)qslot * RIM + rim_index(qy - by * TH, qx - bx * TW)];
    };
    if (lx == 0 && rx > 0) rim_pair<FLAG>(nodes, lp, other(0, -1));
    if (ry > 0) {
        if ((lx == 0 || ly == 0) && rx > 0) rim_pair<FLAG>(nodes, lp, other(-1, -1));
        if (ly == 0) rim_pair<FLAG>(nodes, lp, other(-1, 0));
        if ((ly == 0 || lx == TW - 1) && rx + 1 < cols) rim_pair<FLAG>(nodes, lp, other(-1, 1));
    }
}

// ---- pass 4: listed (pending) components whose global root is flagged add their size
__global__ void __launch_bounds__(kCannyListCap) canny_list_kernel(CannyArgs a) {
    const Job j = make_job(a, blockIdx.z);
    const int rows = j.y1 - j.y0, cols = j.x1 - j.x0;
    const int ty0 = blockIdx.y * TH, tx0 = blockIdx.x * TW;
    if (ty0 >= rows || tx0 >= cols) return;
    const int slot = tile_slot();
    const int n = a.tile_n[slot];
    if (n == 0) return;                                       // block-uniform
    int* nodes = a.labels + (size_t)total_slots() * RIM;
    unsigned cnt = 0;
    if ((int)threadIdx.x < n) {
        const int2 e = a.tile_list[(size_t)slot * kCannyListCap + threadIdx.x];
        const int root = uf_find((volatile int*)nodes, e.x);
        if (nodes[root] & LABEL_FLAG) cnt = (unsigned)e.y;
    }
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

}  // namespace

size_t canny_label_bytes(size_t slots) { return slots * (size_t)(RIM + kCannyListCap) * sizeof(int); }

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
    canny_rim_kernel<false><<<grid, NT, 0, st>>>(a);
    canny_rim_kernel<true><<<grid, NT, 0, st>>>(a);
    canny_list_kernel<<<grid, kCannyListCap, 0, st>>>(a);
    return 4;
}

}  // namespace uwie
