// Per-primitive entry points: the public static methods of the reference's strategy
// libraries as stand-alone CUDA operators on caller-owned device memory.
//   estimate_atmospheric_light / compute_Q / get_brightest_pixel   enhancement_strategies.py:77-206, six_stadigy.py:48-165
//   estimate_transmission (+ guided_filter)                        enhancement_strategies.py:16-46, 209-234; six_stadigy.py:25-46, 167-180
//   recover_image / restore_image                                  enhancement_strategies.py:237-249; six_stadigy.py:182-188
//   color_enhancement / enhance_contrast / white_balance           enhancement_strategies.py:252-273; six_stadigy.py:190-199, 210-219
//   gamma_correction                                               enhancement_strategies.py:276-285; six_stadigy.py:221-224
//   CLAHE on the 8-bit Lab L plane                                 enhancement_strategies.py:302-303; six_stadigy.py:205-206
//   histogram_equalization                                         enhancement_strategies.py:331-345
//   detect_image_type / color_correction                           six_stadigy.py:292-323
// Two of them are the isolated bandwidth-bound kernels of BASELINE.json configs[2]:
//   uwie_prim_transmission_head_u8   u8 RGB in -> float32 t plane out, 7 N bytes   (TMA-staged tiles)
//   uwie_prim_clahe_lab              u8 L plane in -> u8 L plane out, 2 N bytes    (atomic-free tile histograms)
// The others exist so that images that are NOT made from bytes (example_usage.py feeds np.random.rand) run
// on the device too: the fused engine works on byte-indexed tables, these work on the float pixels.
// Float stages here are within north_star's tolerance of the reference (1/255 per pixel), not operation-by-
// operation replicas: box sums are direct float64 sums, Q statistics are float64 sums.
#include <cuda_runtime.h>

#include <cstring>
#include <mutex>
#include <string>

#include "../../include/uwie.h"
#include "k_arith.cuh"
#include "k_kernels.h"
#include "tma1d.cuh"
#include "uwie_tables.h"

namespace uwie {
int uwie_fail(int code, const std::string& msg);

namespace {

#define PCU(call)                                                                               \
    do {                                                                                        \
        cudaError_t _e = (call);                                                                \
        if (_e != cudaSuccess)                                                                  \
            return uwie_fail(UWIE_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(_e)); \
    } while (0)

struct Bump {
    size_t off = 0;
    size_t take(size_t bytes) { const size_t at = off; off += (bytes + 255) & ~(size_t)255; return at; }
};

// =====================================================================================
// 1. transmission head, u8 frames:  t0 = 1 - omega * min_c( (v_c / 255) / (A_c + guard) )   [clipped to .1..1 for SS]
//    enhancement_strategies.py:221-225, six_stadigy.py:170-174.  7 N bytes per frame.
// =====================================================================================
constexpr int TH_PIX = 4096;                 // pixels per tile: 12 KB in, 16 KB out
constexpr int TH_STAGES = 3;
constexpr int TH_NT = 256;

struct THeadArgs {
    const uint8_t* frames; const float* light; float* out;
    unsigned npix; float omega; float guard; int clip;
};

__global__ void __launch_bounds__(TH_NT) thead_kernel(THeadArgs a) {
    __shared__ __align__(128) uint8_t buf[TH_STAGES][TH_PIX * 3];
    __shared__ __align__(8) uint64_t full[TH_STAGES];
    __shared__ float sdiv[3][256];
    const int b = blockIdx.y, t = threadIdx.x;
    const uint8_t* f = a.frames + (size_t)b * a.npix * 3;
    float* o = a.out + (size_t)b * a.npix;
    for (int i = t; i < 768; i += TH_NT) {
        const int c = i >> 8, v = i & 255;
        const float lg = __fadd_rn(a.light[b * 3 + c], a.guard);
        sdiv[c][v] = __fdiv_rn(__fdiv_rn((float)v, 255.0f), lg);
    }
    if (t == 0) {
        for (int s = 0; s < TH_STAGES; ++s) tma::mbar_init(&full[s], 1);
        tma::fence_barrier_init();
    }
    __syncthreads();
    const unsigned ntiles = (a.npix + TH_PIX - 1) / TH_PIX;
    const bool src_al = (reinterpret_cast<uintptr_t>(f) & 15) == 0;
    const bool out_al = (reinterpret_cast<uintptr_t>(o) & 15) == 0;
    auto tile_pix = [&](unsigned k) { return min((unsigned)TH_PIX, a.npix - k * TH_PIX); };
    auto bulk_ok = [&](unsigned k) { return src_al && ((tile_pix(k) * 3u) & 15u) == 0; };
    auto issue = [&](unsigned k, int s) {           // thread 0 only
        if (k < ntiles && bulk_ok(k)) {
            const unsigned bytes = tile_pix(k) * 3u;
            tma::mbar_expect_tx(&full[s], bytes);
            tma::bulk_g2s(buf[s], f + (size_t)k * TH_PIX * 3, bytes, &full[s]);
        }
    };
    if (t == 0)
        for (int s = 0; s < TH_STAGES; ++s) issue(blockIdx.x + (unsigned)s * gridDim.x, s);
    unsigned phases = 0;                            // bit s = parity the next wait on stage s uses
    int s = 0;
    for (unsigned k = blockIdx.x; k < ntiles; k += gridDim.x) {
        const unsigned np = tile_pix(k);
        if (bulk_ok(k)) {
            // one thread polls the barrier, the block waits at bar.sync: 256 polling threads would take issue slots from
            // the co-resident kernels of other streams (measured: -4.5 % on the four-lane step)
            if (t == 0) tma::mbar_wait(&full[s], (phases >> s) & 1u);
            __syncthreads();
            phases ^= 1u << s;
        } else {                                    // unaligned frame or ragged last tile: plain loads
            const uint8_t* src = f + (size_t)k * TH_PIX * 3;
            for (unsigned i = t; i < np * 3; i += TH_NT) buf[s][i] = src[i];
            __syncthreads();
        }
        const uint32_t* w = reinterpret_cast<const uint32_t*>(buf[s]);
        float* ot = o + (size_t)k * TH_PIX;
        const unsigned ngroups = np >> 2;
#pragma unroll
        for (int i = 0; i < TH_PIX / 4 / TH_NT; ++i) {
            const unsigned g = t + TH_NT * i;
            if (g < ngroups) {
                const uint32_t w0 = w[3 * g], w1 = w[3 * g + 1], w2 = w[3 * g + 2];
                float d0 = fminf(fminf(sdiv[0][w0 & 255], sdiv[1][(w0 >> 8) & 255]), sdiv[2][(w0 >> 16) & 255]);
                float d1 = fminf(fminf(sdiv[0][w0 >> 24], sdiv[1][w1 & 255]), sdiv[2][(w1 >> 8) & 255]);
                float d2 = fminf(fminf(sdiv[0][(w1 >> 16) & 255], sdiv[1][w1 >> 24]), sdiv[2][w2 & 255]);
                float d3 = fminf(fminf(sdiv[0][(w2 >> 8) & 255], sdiv[1][(w2 >> 16) & 255]), sdiv[2][w2 >> 24]);
                float4 r;
                r.x = __fsub_rn(1.0f, __fmul_rn(a.omega, d0)); r.y = __fsub_rn(1.0f, __fmul_rn(a.omega, d1));
                r.z = __fsub_rn(1.0f, __fmul_rn(a.omega, d2)); r.w = __fsub_rn(1.0f, __fmul_rn(a.omega, d3));
                if (a.clip) {
                    r.x = fminf(fmaxf(r.x, 0.1f), 1.0f); r.y = fminf(fmaxf(r.y, 0.1f), 1.0f);
                    r.z = fminf(fmaxf(r.z, 0.1f), 1.0f); r.w = fminf(fmaxf(r.w, 0.1f), 1.0f);
                }
                if (out_al) __stcs(reinterpret_cast<float4*>(ot) + g, r);
                else { ot[4 * g] = r.x; ot[4 * g + 1] = r.y; ot[4 * g + 2] = r.z; ot[4 * g + 3] = r.w; }
            }
        }
        for (unsigned p = (ngroups << 2) + t; p < np; p += TH_NT) {        // last 1-3 pixels of a frame
            const uint8_t* q = buf[s] + p * 3;
            float r = __fsub_rn(1.0f, __fmul_rn(a.omega, fminf(fminf(sdiv[0][q[0]], sdiv[1][q[1]]), sdiv[2][q[2]])));
            if (a.clip) r = fminf(fmaxf(r, 0.1f), 1.0f);
            ot[p] = r;
        }
        __syncthreads();                            // everyone is done with buf[s]
        if (t == 0) issue(k + (unsigned)TH_STAGES * gridDim.x, s);
        s = s + 1 == TH_STAGES ? 0 : s + 1;
    }
}

// =====================================================================================
// 2. CLAHE on a uint8 plane (cv2.createCLAHE(clip, (tx, ty)).apply), 2 N bytes per plane.
//    Pass 1: tile histograms WITHOUT atomics: every thread owns a private histogram of 16-bit counters in
//    shared memory (two bins per word, word (bin >> 1) * NT + thread: each lane its own bank), summed once
//    per tile; the last block of a tile clips, redistributes and writes the LUT.
//    Pass 2: one block per interpolation cell (the pixels whose four neighbouring tiles are the same): the
//    four LUTs are merged into ONE float4 table in shared memory, so a pixel costs a single 128-bit look-up
//    and nine float operations (cv2's CLAHE_Interpolation_Body arithmetic, unfused).
// =====================================================================================
constexpr int CH_NT = 128;
constexpr int CH_SMEM = 128 * CH_NT * 4;       // 64 KB private histograms

struct ClaheLArgs {
    const uint8_t* in; uint8_t* out;
    int H, W, tiles_x, tiles_y, tw, th, clip, split, rows_per_slab;
    unsigned* hist;          // [B][tiles][256]  (zeroed; only used when split > 1)
    unsigned* done;          // [B][tiles]       (zeroed)
    uint8_t* lut;            // [B][tiles][256]
};

__device__ __forceinline__ void clahe_lut_from_hist_128(const unsigned* h /* shared, 256 */, int clip, int area,
                                                        uint8_t* lut_out) {
    // 128 threads, thread t owns bins 2t and 2t + 1 (clahe_clip_redistribute of uwie_hd.h, in parallel)
    __shared__ int s_part[CH_NT / 32];
    __shared__ int s_tot;
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    int v0 = (int)h[2 * t], v1 = (int)h[2 * t + 1];
    int over = (v0 > clip ? v0 - clip : 0) + (v1 > clip ? v1 - clip : 0);
    v0 = min(v0, clip); v1 = min(v1, clip);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) over += __shfl_xor_sync(0xffffffffu, over, o);
    if (lane == 0) s_part[w] = over;
    __syncthreads();
    if (t == 0) { int tot = 0; for (int k = 0; k < CH_NT / 32; ++k) tot += s_part[k]; s_tot = tot; }
    __syncthreads();
    const int excess = s_tot;
    const int batch = excess / 256, residual = excess - batch * 256;
    v0 += batch; v1 += batch;
    if (residual != 0) {
        int step = 256 / residual;
        if (step < 1) step = 1;
        if ((2 * t) % step == 0 && (2 * t) / step < residual) v0 += 1;
        if ((2 * t + 1) % step == 0 && (2 * t + 1) / step < residual) v1 += 1;
    }
    int incl = v0 + v1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += u; }
    __syncthreads();
    if (lane == 31) s_part[w] = incl;
    __syncthreads();
    int base = 0;
    for (int k = 0; k < w; ++k) base += s_part[k];
    const float scale = clahe_lut_scale(area);
    const int c1 = base + incl, c0 = c1 - v1;
    lut_out[2 * t] = clahe_lut_entry(c0, scale);
    lut_out[2 * t + 1] = clahe_lut_entry(c1, scale);
}

__global__ void __launch_bounds__(CH_NT) clahe_l_hist_kernel(ClaheLArgs a) {
    extern __shared__ __align__(16) unsigned priv[];        // [128 words][CH_NT threads]
    __shared__ unsigned s_hist[256];
    __shared__ int s_last;
    const int tile = blockIdx.x, slab = blockIdx.y, b = blockIdx.z, t = threadIdx.x;
    const int tyi = tile / a.tiles_x, txi = tile - tyi * a.tiles_x;
    for (int i = t; i < 128 * CH_NT; i += CH_NT) priv[i] = 0;
    __syncthreads();
    const uint8_t* img = a.in + (size_t)b * a.H * a.W;
    const int y0 = tyi * a.th + slab * a.rows_per_slab;
    const int y1 = min(tyi * a.th + a.th, y0 + a.rows_per_slab);
    const int x0 = txi * a.tw;
    unsigned* mine = priv + t;
    auto count = [&](unsigned v) {
        unsigned* p = mine + (v >> 1) * CH_NT;
        *p += 1u << ((v & 1u) << 4);
    };
    // two pixels at a time: both counter words are fetched before either is written back, so two shared-memory
    // round trips overlap; a pair that hits the same word is merged (neighbouring pixels often do)
    auto count2 = [&](unsigned v0, unsigned v1) {
        unsigned* p0 = mine + (v0 >> 1) * CH_NT;
        unsigned* p1 = mine + (v1 >> 1) * CH_NT;
        const unsigned w0 = *p0, w1 = *p1;
        const unsigned i0 = 1u << ((v0 & 1u) << 4), i1 = 1u << ((v1 & 1u) << 4);
        const bool same = p0 == p1;
        *p0 = w0 + i0 + (same ? i1 : 0u);
        if (!same) *p1 = w1 + i1;
    };
    auto count_word = [&](unsigned w) {
        count2(w & 255u, (w >> 8) & 255u);
        count2((w >> 16) & 255u, w >> 24);
    };
    const bool fast = x0 + a.tw <= a.W && y1 <= a.H && (a.tw & 15) == 0 && (a.W & 15) == 0 &&
                      (reinterpret_cast<uintptr_t>(img) & 15) == 0;
    if (fast) {
        const int vpr = a.tw >> 4;                          // 16-pixel vectors per tile row
        const int nvec = vpr * (y1 - y0);
        auto load = [&](int i) {
            const int ly = i / vpr, vx = i - ly * vpr;
            return __ldg(reinterpret_cast<const uint4*>(img + (size_t)(y0 + ly) * a.W + x0) + vx);
        };
        // the vectors of the next two rounds are in flight while this round's 16 pixels are counted
        uint4 n0 = make_uint4(0, 0, 0, 0), n1 = n0;
        if (t < nvec) n0 = load(t);
        if (t + CH_NT < nvec) n1 = load(t + CH_NT);
        for (int i = t; i < nvec; i += CH_NT) {
            const uint4 q = n0;
            n0 = n1;
            if (i + 2 * CH_NT < nvec) n1 = load(i + 2 * CH_NT);
            count_word(q.x); count_word(q.y); count_word(q.z); count_word(q.w);
        }
    } else {
        const int n = a.tw * max(y1 - y0, 0);
        for (int i = t; i < n; i += CH_NT) {
            const int ly = i / a.tw, lx = i - ly * a.tw;
            int y = y0 + ly, x = x0 + lx;
            if (y >= a.H) y = reflect101(y, a.H);           // cv2 pads the plane with REFLECT_101 up to a tile multiple
            if (x >= a.W) x = reflect101(x, a.W);
            count(img[(size_t)y * a.W + x]);
        }
    }
    __syncthreads();
    // bins 2t, 2t + 1 = the two halves of word t of every thread's histogram; the rotation keeps the 32 lanes
    // of a warp on 32 different banks
    unsigned lo = 0, hi = 0;
    for (int k = 0; k < CH_NT; ++k) {
        const unsigned v = priv[t * CH_NT + ((k + t) & (CH_NT - 1))];
        lo += v & 0xffffu; hi += v >> 16;
    }
    const size_t slot = (size_t)b * (a.tiles_x * a.tiles_y) + tile;
    if (a.split == 1) {
        s_hist[2 * t] = lo; s_hist[2 * t + 1] = hi;
        __syncthreads();
    } else {
        unsigned* gh = a.hist + slot * 256;
        if (lo) atomicAdd(&gh[2 * t], lo);
        if (hi) atomicAdd(&gh[2 * t + 1], hi);
        __threadfence();
        __syncthreads();
        if (t == 0) s_last = (atomicAdd(&a.done[slot], 1u) == (unsigned)(a.split - 1));
        __syncthreads();
        if (!s_last) return;
        __threadfence();
        s_hist[2 * t] = __ldcg(&gh[2 * t]); s_hist[2 * t + 1] = __ldcg(&gh[2 * t + 1]);
        __syncthreads();
    }
    clahe_lut_from_hist_128(s_hist, a.clip, a.tw * a.th, a.lut + slot * 256);
}

constexpr int CA_NT = 256;
constexpr int CA_ROWS = 64;              // rows of a cell one block takes

__global__ void __launch_bounds__(CA_NT) clahe_l_apply_kernel(ClaheLArgs a) {
    __shared__ float4 tab[256];
    __shared__ float2 s_yw[CA_ROWS];        // (ya, ya1) of a row; ya < 0 marks a row whose own floor() lands in another cell
    const int cells_x = a.tiles_x + 1;
    const int cyi = blockIdx.x / cells_x, cxi = blockIdx.x - cyi * cells_x;
    const int b = blockIdx.z, t = threadIdx.x;
    const int tx1 = cxi - 1, ty1 = cyi - 1;                 // the floor() every pixel of this cell should get
    // nominal cell: x in [ceil((tx1 + .5) tw), ceil((tx1 + 1.5) tw)); pixels whose own float arithmetic lands in a
    // neighbouring cell (possible only on the boundary columns / rows) take the exact slow path below
    auto edge = [](int k, int sz) { return (2 * k + 1) * sz <= 0 ? 0 : ((2 * k + 1) * sz + 1) / 2; };
    const int cx0 = min(edge(tx1, a.tw), a.W), cx1 = min(edge(tx1 + 1, a.tw), a.W);
    const int cy0 = min(edge(ty1, a.th), a.H), cy1 = min(edge(ty1 + 1, a.th), a.H);
    const int ys = cy0 + blockIdx.y * CA_ROWS, ye = min(cy1, ys + CA_ROWS);
    if (cx0 >= cx1 || ys >= ye) return;
    const float inv_tw = __fdiv_rn(1.0f, (float)a.tw), inv_th = __fdiv_rn(1.0f, (float)a.th);
    const int t1x = max(tx1, 0), t2x = min(tx1 + 1, a.tiles_x - 1);
    const int t1y = max(ty1, 0), t2y = min(ty1 + 1, a.tiles_y - 1);
    const uint8_t* luts = a.lut + (size_t)b * (a.tiles_x * a.tiles_y) * 256;
    {
        const uint8_t* l11 = luts + (t1y * a.tiles_x + t1x) * 256; const uint8_t* l12 = luts + (t1y * a.tiles_x + t2x) * 256;
        const uint8_t* l21 = luts + (t2y * a.tiles_x + t1x) * 256; const uint8_t* l22 = luts + (t2y * a.tiles_x + t2x) * 256;
        for (int v = t; v < 256; v += CA_NT) tab[v] = make_float4((float)l11[v], (float)l12[v], (float)l21[v], (float)l22[v]);
        for (int r = t; r < ye - ys; r += CA_NT) {
            const float tyf = __fsub_rn(__fmul_rn((float)(ys + r), inv_th), 0.5f);
            const float fl = floorf(tyf);
            const float ya = __fsub_rn(tyf, fl);
            s_yw[r] = ((int)fl == ty1) ? make_float2(ya, __fsub_rn(1.0f, ya)) : make_float2(-1.f, 0.f);
        }
    }
    __syncthreads();
    const uint8_t* img = a.in + (size_t)b * a.H * a.W;
    uint8_t* out = a.out + (size_t)b * a.H * a.W;
    auto slow = [&](int x, int y, int v) -> int {           // cv2's formula with this pixel's own tile indices
        const float tyf = __fsub_rn(__fmul_rn((float)y, inv_th), 0.5f);
        int q1 = (int)floorf(tyf);
        const float ya = __fsub_rn(tyf, (float)q1), ya1 = __fsub_rn(1.0f, ya);
        int q2 = q1 + 1; q1 = max(q1, 0); q2 = min(q2, a.tiles_y - 1);
        const float txf = __fsub_rn(__fmul_rn((float)x, inv_tw), 0.5f);
        int p1 = (int)floorf(txf);
        const float xa = __fsub_rn(txf, (float)p1), xa1 = __fsub_rn(1.0f, xa);
        int p2 = p1 + 1; p1 = max(p1, 0); p2 = min(p2, a.tiles_x - 1);
        const float top = __fadd_rn(__fmul_rn((float)luts[(q1 * a.tiles_x + p1) * 256 + v], xa1), __fmul_rn((float)luts[(q1 * a.tiles_x + p2) * 256 + v], xa));
        const float bot = __fadd_rn(__fmul_rn((float)luts[(q2 * a.tiles_x + p1) * 256 + v], xa1), __fmul_rn((float)luts[(q2 * a.tiles_x + p2) * 256 + v], xa));
        return clampi(__float2int_rn(__fadd_rn(__fmul_rn(top, ya1), __fmul_rn(bot, ya))), 0, 255);
    };
    // a blend of values in [0, 255] with weights xa1 + xa = 1 (to an ulp) cannot leave [0, 255.0001]: cv2's saturate_cast
    // is a plain round here, and so is this
    auto blend = [&](int v, float xa, float xa1, float ya, float ya1) -> unsigned {
        const float4 q = tab[v];
        const float top = __fadd_rn(__fmul_rn(q.x, xa1), __fmul_rn(q.y, xa));
        const float bot = __fadd_rn(__fmul_rn(q.z, xa1), __fmul_rn(q.w, xa));
        return (unsigned)__float2int_rn(__fadd_rn(__fmul_rn(top, ya1), __fmul_rn(bot, ya)));
    };
    const bool words = (cx0 & 3) == 0 && (a.W & 3) == 0 && ((reinterpret_cast<uintptr_t>(img) | reinterpret_cast<uintptr_t>(out)) & 3) == 0;
    if (words) {
        const int gpr = (cx1 - cx0 + 3) >> 2;               // 4-pixel groups per row (the last may overhang the cell)
        // one column group per thread, its x weights computed once; rows r0, r0 + rstep, ...
        auto column_group = [&](int g, int r0, int rstep) {
            const int xg = cx0 + 4 * g;
            float xa[4], xa1[4];
            bool okx[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float txf = __fsub_rn(__fmul_rn((float)(xg + j), inv_tw), 0.5f);
                const float fl = floorf(txf);
                xa[j] = __fsub_rn(txf, fl); xa1[j] = __fsub_rn(1.0f, xa[j]);
                okx[j] = ((int)fl == tx1);
            }
            const bool overhang = xg + 3 >= cx1;            // the last pixels of the word belong to the next cell
            const bool clean = okx[0] && okx[1] && okx[2] && okx[3] && !overhang;
            const int nrows = ye - ys;
            const uint8_t* src = img + (size_t)ys * a.W + xg;
            uint8_t* dst = out + (size_t)ys * a.W + xg;
            uint32_t nxt = r0 < nrows ? __ldg(reinterpret_cast<const uint32_t*>(src + (size_t)r0 * a.W)) : 0u;
            for (int r = r0; r < nrows; r += rstep) {
                const uint32_t wv = nxt;
                if (r + rstep < nrows) nxt = __ldg(reinterpret_cast<const uint32_t*>(src + (size_t)(r + rstep) * a.W));
                const float2 yw = s_yw[r];
                if (clean && yw.x >= 0.f) {
                    const unsigned o0 = blend(wv & 255u, xa[0], xa1[0], yw.x, yw.y);
                    const unsigned o1 = blend((wv >> 8) & 255u, xa[1], xa1[1], yw.x, yw.y);
                    const unsigned o2 = blend((wv >> 16) & 255u, xa[2], xa1[2], yw.x, yw.y);
                    const unsigned o3 = blend(wv >> 24, xa[3], xa1[3], yw.x, yw.y);
                    *reinterpret_cast<uint32_t*>(dst + (size_t)r * a.W) = o0 | (o1 << 8) | (o2 << 16) | (o3 << 24);
                } else {
                    const int y = ys + r;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        if (xg + j >= cx1) continue;
                        const int v = (wv >> (8 * j)) & 255;
                        const unsigned res = (okx[j] && yw.x >= 0.f) ? blend(v, xa[j], xa1[j], yw.x, yw.y) : (unsigned)slow(xg + j, y, v);
                        dst[(size_t)r * a.W + j] = (uint8_t)res;
                    }
                }
            }
        };
        if (gpr >= CA_NT) {
            for (int g = t; g < gpr; g += CA_NT) column_group(g, 0, 1);
        } else {
            const int nrl = CA_NT / gpr;                    // row lanes
            if (t / gpr < nrl) column_group(t % gpr, t / gpr, nrl);
        }
    } else {
        const int cw = cx1 - cx0;
        for (int i = t; i < cw * (ye - ys); i += CA_NT) {
            const int r = i / cw, x = cx0 + (i - r * cw), y = ys + r;
            out[(size_t)y * a.W + x] = (uint8_t)slow(x, y, img[(size_t)y * a.W + x]);
        }
    }
}

// =====================================================================================
// 3. float-image primitives
// =====================================================================================
// (img * 255).astype(uint8): truncation, values outside [0, 1] wrap like NumPy's cast of a value that fits int
template <class T>
__global__ void quantise_kernel(const T* __restrict__ img, size_t n, uint8_t* __restrict__ q) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const T v = Ar<T>::mul(img[i], (T)255);
        q[i] = (uint8_t)(int)fmin(fmax((double)v, -2147483648.0), 2147483647.0);
    }
}

// per-channel sums of a rectangle (float64 accumulation) as per-block partials part[q][block][c] (sum, or sum of squares
// about mean[q][c]); stats_reduce_kernel adds them in block order, so the result does not depend on scheduling
constexpr int kStatBlocks = 64;
template <class T, bool SECOND>
__global__ void __launch_bounds__(256) rect_stats_kernel(const T* __restrict__ img, int W, const QuadRect* rect, int quadrants,
                                                         const double* __restrict__ mean, double* __restrict__ part) {
    // blockIdx.y = quadrant (or 0 when quadrants == 0: the whole rectangle)
    const QuadRect r = *rect;
    int y0 = r.y0, y1 = r.y1, x0 = r.x0, x1 = r.x1;
    const int q = blockIdx.y;
    if (quadrants) {
        if (!r.active) {
            if (threadIdx.x < 3) part[((size_t)q * kStatBlocks + blockIdx.x) * 3 + threadIdx.x] = 0.0;
            return;
        }
        const int ym = r.y0 + (r.y1 - r.y0) / 2, xm = r.x0 + (r.x1 - r.x0) / 2;
        y0 = (q & 2) ? ym : r.y0; y1 = (q & 2) ? r.y1 : ym;
        x0 = (q & 1) ? xm : r.x0; x1 = (q & 1) ? r.x1 : xm;
    }
    const int cols = x1 - x0;
    const long long n = (long long)(y1 - y0) * cols;
    double s[3] = {0, 0, 0};
    double m[3] = {0, 0, 0};
    if (SECOND) { m[0] = mean[q * 3]; m[1] = mean[q * 3 + 1]; m[2] = mean[q * 3 + 2]; }
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
        const int y = y0 + (int)(i / cols), x = x0 + (int)(i % cols);
        const T* p = img + ((size_t)y * W + x) * 3;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            if (SECOND) { const double d = (double)p[c] - m[c]; s[c] += d * d; }
            else s[c] += (double)p[c];
        }
    }
    __shared__ double red[8][3];
#pragma unroll
    for (int c = 0; c < 3; ++c) s[c] = warp_sum_f64(s[c]);
    if ((threadIdx.x & 31) == 0) for (int c = 0; c < 3; ++c) red[threadIdx.x >> 5][c] = s[c];
    __syncthreads();
    if (threadIdx.x < 3) {
        double v = 0;
        for (int w = 0; w < 8; ++w) v += red[w][threadIdx.x];
        part[((size_t)q * kStatBlocks + blockIdx.x) * 3 + threadIdx.x] = v;
    }
}
__global__ void stats_reduce_kernel(const double* __restrict__ part, int nblocks, int nq, double* __restrict__ out) {
    const int i = threadIdx.x;                       // q * 3 + c
    if (i >= nq * 3) return;
    const int q = i / 3, c = i - q * 3;
    double v = 0;
    for (int b = 0; b < nblocks; ++b) v += part[((size_t)q * kStatBlocks + b) * 3 + c];
    out[i] = v;
}

// sums -> means (per quadrant), and zero the second-pass accumulators
__global__ void quad_means_kernel(const QuadRect* rect, const double* sums, double* means, double* ssq) {
    const int q = threadIdx.x / 3;
    if (threadIdx.x >= 12) return;
    const QuadRect r = *rect;
    const int ym = r.y0 + (r.y1 - r.y0) / 2, xm = r.x0 + (r.x1 - r.x0) / 2;
    const int qy0 = (q & 2) ? ym : r.y0, qy1 = (q & 2) ? r.y1 : ym;
    const int qx0 = (q & 1) ? xm : r.x0, qx1 = (q & 1) ? r.x1 : xm;
    const double n = (double)(qy1 - qy0) * (double)(qx1 - qx0);
    means[threadIdx.x] = n > 0 ? sums[threadIdx.x] / n : 0.0;
    ssq[threadIdx.x] = 0.0;
}

// Q of the four quadrants from float statistics + Canny counts, first maximum, next rectangle (ES:122-139)
__global__ void quad_descend_float_kernel(QuadRect* rect, const double* sums, const double* ssq, const unsigned* counts,
                                          int level, double* q_trace, double* sums_next) {
    if (threadIdx.x != 0) return;
    QuadRect r = *rect;
    if (!r.active) return;
    const int ym = r.y0 + (r.y1 - r.y0) / 2, xm = r.x0 + (r.x1 - r.x0) / 2;
    double best = 0.0; int best_q = 0;
    for (int k = 0; k < 4; ++k) {
        const int ky0 = (k & 2) ? ym : r.y0, ky1 = (k & 2) ? r.y1 : ym;
        const int kx0 = (k & 1) ? xm : r.x0, kx1 = (k & 1) ? r.x1 : xm;
        const double kn = (double)(ky1 - ky0) * (double)(kx1 - kx0);
        const double sr = sums[k * 3], sg = sums[k * 3 + 1], sb = sums[k * 3 + 2];
        const double bright = (sr + sg + sb) / (3.0 * kn);
        const double cast = (sb + sg - 2.0 * sr) / kn;
        const double spread = (ssq[k * 3] / kn + ssq[k * 3 + 1] / kn + ssq[k * 3 + 2] / kn) / 3.0;
        const double edges = (double)counts[level * 4 + k] / kn;
        const double Q = bright + cast - spread - edges;
        q_trace[level * 4 + k] = Q;
        if (k == 0 || Q > best) { best = Q; best_q = k; }
    }
    QuadRect nr;
    nr.y0 = (best_q & 2) ? ym : r.y0; nr.y1 = (best_q & 2) ? r.y1 : ym;
    nr.x0 = (best_q & 1) ? xm : r.x0; nr.x1 = (best_q & 1) ? r.x1 : xm;
    nr.level = level + 1;
    nr.active = ((nr.y1 - nr.y0) > 1 && (nr.x1 - nr.x0) > 1) ? 1 : 0;
    *rect = nr;
    for (int i = 0; i < 12; ++i) sums_next[i] = 0.0;
}

__global__ void identity_cast_kernel(CastTables* ct) {
    for (int i = threadIdx.x; i < 768; i += blockDim.x) {
        const int c = i >> 8, v = i & 255;
        ct->curf[c][v] = __fdiv_rn((float)v, 255.0f);
        ct->curq[c][v] = (uint8_t)v;
    }
}
__global__ void rect_init_kernel(QuadRect* r, int H, int W, double* zero12a, double* zero12b) {
    if (threadIdx.x == 0) {
        r->y0 = 0; r->y1 = H; r->x0 = 0; r->x1 = W; r->level = 0;
        r->active = (H > 1 && W > 1) ? 1 : 0;
    }
    if (threadIdx.x < 12) { zero12a[threadIdx.x] = 0.0; zero12b[threadIdx.x] = 0.0; }
}

// first pixel (row-major) of the final rectangle with the largest (R + G) + B in the image's own float type
template <class T>
__global__ void __launch_bounds__(256) brightest_float_kernel(const T* __restrict__ img, int W, const QuadRect* rect, T* light) {
    const QuadRect r = *rect;
    const int rows = r.y1 - r.y0, cols = r.x1 - r.x0;
    T best = (T)-1e300; int best_i = 0x7fffffff;
    if (sizeof(T) == 4) best = (T)-3.0e38f;
    for (int i = threadIdx.x; i < rows * cols; i += 256) {
        const int y = r.y0 + i / cols, x = r.x0 + i % cols;
        const T* p = img + ((size_t)y * W + x) * 3;
        const T s = Ar<T>::add(Ar<T>::add(p[0], p[1]), p[2]);
        if (s > best) { best = s; best_i = i; }
    }
    __shared__ double sv[256];
    __shared__ int si[256];
    sv[threadIdx.x] = (double)best; si[threadIdx.x] = best_i;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) {
            const double v2 = sv[threadIdx.x + o]; const int i2 = si[threadIdx.x + o];
            if (v2 > sv[threadIdx.x] || (v2 == sv[threadIdx.x] && i2 < si[threadIdx.x])) { sv[threadIdx.x] = v2; si[threadIdx.x] = i2; }
        }
        __syncthreads();
    }
    if (threadIdx.x < 3) {
        const int i = si[0] == 0x7fffffff ? 0 : si[0];
        const int y = r.y0 + i / cols, x = r.x0 + i % cols;
        light[threadIdx.x] = img[((size_t)y * W + x) * 3 + threadIdx.x];
    }
}

// gray guide I = GRAY(trunc(img * 255)) / 255 (float64) and raw transmission p = 1 - omega * min_c img / (A + guard)
template <class T>
__global__ void thead_float_kernel(const T* __restrict__ img, const T* __restrict__ light, size_t npix, T omega, T guard, int clip,
                                   double* __restrict__ I, double* __restrict__ P) {
    const T l0 = Ar<T>::add(light[0], guard), l1 = Ar<T>::add(light[1], guard), l2 = Ar<T>::add(light[2], guard);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (size_t)gridDim.x * blockDim.x) {
        const T r = img[i * 3], g = img[i * 3 + 1], b = img[i * 3 + 2];
        const int qr = (uint8_t)quantise<T>(r), qg = (uint8_t)quantise<T>(g), qb = (uint8_t)quantise<T>(b);
        I[i] = (double)gray_q15(qr, qg, qb) / 255.0;
        T d = Ar<T>::div(r, l0);
        const T d1 = Ar<T>::div(g, l1), d2 = Ar<T>::div(b, l2);
        d = d1 < d ? d1 : d; d = d2 < d ? d2 : d;
        T t0 = Ar<T>::sub((T)1, Ar<T>::mul(omega, d));
        if (clip) t0 = t0 < (T)0.1 ? (T)0.1 : (t0 > (T)1 ? (T)1 : t0);
        P[i] = (double)t0;
    }
}

// guided filter, float64, direct separable box sums (window [-(r/2), r-1-r/2], REFLECT_101, scale 1/(r*r)).
// pass H: horizontal sums of up to four planes derived from (I, p); pass V: vertical sums -> means.
template <int MODE>   // 0: planes I, p, I*p, I*I of (A = I, Bp = p);  1: planes a, b (A = a, Bp = b)
__global__ void gf_box_h_kernel(const double* __restrict__ A, const double* __restrict__ Bp, int H, int W, int r,
                                double* __restrict__ out /* [MODE ? 2 : 4][H][W] */) {
    const size_t npix = (size_t)H * W;
    const int lo = r / 2;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (size_t)gridDim.x * blockDim.x) {
        const int y = (int)(i / W), x = (int)(i % W);
        const double* ra = A + (size_t)y * W; const double* rb = Bp + (size_t)y * W;
        double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
        for (int k = 0; k < r; ++k) {
            const int xx = reflect101(x - lo + k, W);
            const double u = ra[xx], v = rb[xx];
            s0 += u; s1 += v;
            if (MODE == 0) { s2 += __dmul_rn(u, v); s3 += __dmul_rn(u, u); }
        }
        out[i] = s0; out[npix + i] = s1;
        if (MODE == 0) { out[2 * npix + i] = s2; out[3 * npix + i] = s3; }
    }
}
__global__ void gf_box_v_ab_kernel(const double* __restrict__ hs, int H, int W, int r, double eps,
                                   double* __restrict__ a_out, double* __restrict__ b_out) {
    const size_t npix = (size_t)H * W;
    const int lo = r / 2;
    const double scale = 1.0 / ((double)r * (double)r);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (size_t)gridDim.x * blockDim.x) {
        const int y = (int)(i / W), x = (int)(i % W);
        double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
        for (int k = 0; k < r; ++k) {
            const size_t j = (size_t)reflect101(y - lo + k, H) * W + x;
            s0 += hs[j]; s1 += hs[npix + j]; s2 += hs[2 * npix + j]; s3 += hs[3 * npix + j];
        }
        const double mI = s0 * scale, mP = s1 * scale, mIP = s2 * scale, mII = s3 * scale;
        const double cov = __dsub_rn(mIP, __dmul_rn(mI, mP));
        const double var = __dsub_rn(mII, __dmul_rn(mI, mI));
        const double ra = __ddiv_rn(cov, __dadd_rn(var, eps));
        a_out[i] = ra;
        b_out[i] = __dsub_rn(mP, __dmul_rn(ra, mI));
    }
}
__global__ void gf_box_v_q_kernel(const double* __restrict__ hs, const double* __restrict__ I, int H, int W, int r,
                                  int clip, double* __restrict__ q) {
    const size_t npix = (size_t)H * W;
    const int lo = r / 2;
    const double scale = 1.0 / ((double)r * (double)r);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (size_t)gridDim.x * blockDim.x) {
        const int y = (int)(i / W), x = (int)(i % W);
        double s0 = 0, s1 = 0;
        for (int k = 0; k < r; ++k) {
            const size_t j = (size_t)reflect101(y - lo + k, H) * W + x;
            s0 += hs[j]; s1 += hs[npix + j];
        }
        double v = __dadd_rn(__dmul_rn(s0 * scale, I[i]), s1 * scale);
        if (clip) v = v < 0.1 ? 0.1 : (v > 1.0 ? 1.0 : v);
        q[i] = v;
    }
}

// recover: clip((img - A) / t + A, 0, 1); (img - A) in the image's type, the rest in float64, stored as TO
template <class T, class TO>
__global__ void recover_kernel(const T* __restrict__ img, const double* __restrict__ t, const T* __restrict__ light,
                               size_t npix, TO* __restrict__ out) {
    const T l0 = light[0], l1 = light[1], l2 = light[2];
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (size_t)gridDim.x * blockDim.x) {
        const double tt = t[i];
        const T l[3] = {l0, l1, l2};
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const double d = (double)Ar<T>::sub(img[i * 3 + c], l[c]);
            const TO v = (TO)__dadd_rn(__ddiv_rn(d, tt), (double)l[c]);
            out[i * 3 + c] = clip01<TO>(v);
        }
    }
}

// level-A histogram of the percentile selection + range check (the selection keys need x in [0, 1])
template <class T>
__global__ void __launch_bounds__(256) level_a_kernel(const T* __restrict__ img, size_t npix, unsigned* __restrict__ hist, int* bad) {
    __shared__ unsigned sh[3 * kSelBinsA];
    for (int i = threadIdx.x; i < 3 * kSelBinsA; i += 256) sh[i] = 0;
    __syncthreads();
    int flag = 0;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < npix * 3; i += (size_t)gridDim.x * 256) {
        const T v = img[i];
        if (!(v >= (T)0 && v <= (T)1)) { flag = 1; continue; }
        atomicAdd(&sh[(i % 3) * kSelBinsA + sel_bin_a<T>(v)], 1u);
    }
    if (flag) atomicOr(bad, 1);
    __syncthreads();
    for (int i = threadIdx.x; i < 3 * kSelBinsA; i += 256) if (sh[i]) atomicAdd(&hist[i], sh[i]);
}
template <class T>
__global__ void stretch_kernel(const T* __restrict__ img, size_t npix, const StratTables* st, T* __restrict__ out) {
    T lo[3], den[3];
    for (int c = 0; c < 3; ++c) { lo[c] = (T)st->st_lo[0][c]; den[c] = (T)st->st_den[0][c]; }
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (size_t)gridDim.x * blockDim.x) {
#pragma unroll
        for (int c = 0; c < 3; ++c) out[i * 3 + c] = stretch_apply<T>(img[i * 3 + c], lo[c], den[c]);
    }
}
template <class T>
__global__ void gamma_kernel(const T* __restrict__ img, size_t n, double e, int clip, T* __restrict__ out) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        T y = Ar<T>::power(img[i], e);
        out[i] = clip ? clip01<T>(y) : y;
    }
}
// out = clip(img * scale[c], 0, 1)  (color_correction, six_stadigy.py:305-323)
template <class T>
__global__ void scale_clip_kernel(const T* __restrict__ img, size_t npix, T s0, T s1, T s2, T* __restrict__ out) {
    const T s[3] = {s0, s1, s2};
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix * 3; i += (size_t)gridDim.x * blockDim.x)
        out[i] = clip01<T>(Ar<T>::mul(img[i], s[i % 3]));
}

// equalizeHist per channel (enhancement_strategies.py:331-345): LUTs from the frame histograms, then a byte map
__global__ void eq_lut_kernel(const unsigned* hist /* [B][3][256] */, long long npix, uint8_t* lut /* [B][3][256] */) {
    const int b = blockIdx.x, c = threadIdx.x;
    if (c < 3) equalize_lut(hist + ((size_t)b * 3 + c) * 256, npix, lut + ((size_t)b * 3 + c) * 256);
}
__global__ void __launch_bounds__(256) byte_map_kernel(const uint8_t* __restrict__ in, size_t npix, const uint8_t* __restrict__ lut,
                                                       uint8_t* __restrict__ out) {
    __shared__ uint8_t s[3][256];
    const int b = blockIdx.y;
    for (int i = threadIdx.x; i < 768; i += 256) (&s[0][0])[i] = lut[(size_t)b * 768 + i];
    __syncthreads();
    const uint8_t* f = in + (size_t)b * npix * 3;
    uint8_t* o = out + (size_t)b * npix * 3;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < npix * 3; i += (size_t)gridDim.x * 256) o[i] = s[i % 3][f[i]];
}

__global__ void gray_from_bytes_kernel(const uint8_t* __restrict__ q, size_t npix, uint8_t* __restrict__ gray) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (size_t)gridDim.x * blockDim.x)
        gray[i] = (uint8_t)gray_q15(q[i * 3], q[i * 3 + 1], q[i * 3 + 2]);
}
// Q of a whole block (compute_Q, ES:147-188): sums[0..2], ssq[0..2], Canny count -> one float64
__global__ void region_score_kernel(const double* sums, const double* ssq, const unsigned* count, double n, double* q) {
    if (threadIdx.x != 0) return;
    const double sr = sums[0], sg = sums[1], sb = sums[2];
    const double bright = (sr + sg + sb) / (3.0 * n);
    const double cast = (sb + sg - 2.0 * sr) / n;
    const double spread = (ssq[0] / n + ssq[1] / n + ssq[2] / n) / 3.0;
    *q = bright + cast - spread - (double)count[0] / n;
}
__global__ void means3_kernel(const double* sums, double n, double* means, double* ssq) {
    if (threadIdx.x < 3) { means[threadIdx.x] = sums[threadIdx.x] / n; ssq[threadIdx.x] = 0.0; }
}
__global__ void divide3_kernel(const double* s, double n, double* m) { if (threadIdx.x < 3) m[threadIdx.x] = s[threadIdx.x] / n; }

int grid_for(size_t n, int nt = 256, int cap = 148 * 8) {
    size_t g = (n + nt - 1) / nt;
    return (int)(g < 1 ? 1 : (g > (size_t)cap ? cap : g));
}

}  // namespace
}  // namespace uwie

using namespace uwie;

// =====================================================================================
// C ABI
// =====================================================================================
extern "C" {

int uwie_prim_transmission_head_u8(const uint8_t* frames_dev, int batch, int height, int width, const float* light_dev,
                                   float omega, float guard, int clip_t0, float* t_dev, void* stream) {
    if (!frames_dev || !light_dev || !t_dev || batch < 1 || height < 1 || width < 1)
        return uwie_fail(UWIE_ERR_ARG, "transmission_head: bad arguments");
    THeadArgs a;
    a.frames = frames_dev; a.light = light_dev; a.out = t_dev; a.npix = (unsigned)height * (unsigned)width;
    a.omega = omega; a.guard = guard; a.clip = clip_t0;
    const unsigned ntiles = (a.npix + TH_PIX - 1) / TH_PIX;
    int per = (148 * 4 + batch - 1) / batch;                   // four resident blocks per SM, one wave
    if ((unsigned)per > ntiles) per = (int)ntiles;
    thead_kernel<<<dim3(per, batch), TH_NT, 0, static_cast<cudaStream_t>(stream)>>>(a);
    PCU(cudaGetLastError());
    return UWIE_OK;
}

size_t uwie_prim_clahe_lab_workspace_bytes(int batch, int tiles_x, int tiles_y) {
    if (batch < 1 || tiles_x < 1 || tiles_y < 1 || tiles_x * tiles_y > kMaxTiles) return 0;
    const size_t tiles = (size_t)batch * tiles_x * tiles_y;
    Bump b;
    b.take(tiles * 256 * 4); b.take(tiles * 4); b.take(tiles * 256);
    return b.off;
}

int uwie_prim_clahe_lab(const uint8_t* plane_dev, int batch, int height, int width, double clip_limit, int tiles_x,
                        int tiles_y, uint8_t* out_dev, void* workspace_dev, size_t workspace_bytes, void* stream) {
    if (!plane_dev || !out_dev || !workspace_dev || batch < 1 || height < 1 || width < 1)
        return uwie_fail(UWIE_ERR_ARG, "clahe_lab: bad arguments");
    const size_t need = uwie_prim_clahe_lab_workspace_bytes(batch, tiles_x, tiles_y);
    if (!need) return uwie_fail(UWIE_ERR_ARG, "clahe_lab: tile grid must have between 1 and 256 tiles");
    if (workspace_bytes < need) return uwie_fail(UWIE_ERR_NOMEM, "clahe_lab: workspace too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t tiles = (size_t)batch * tiles_x * tiles_y;
    char* ws = static_cast<char*>(workspace_dev);
    Bump b;
    ClaheLArgs a;
    a.in = plane_dev; a.out = out_dev; a.H = height; a.W = width; a.tiles_x = tiles_x; a.tiles_y = tiles_y;
    if (width % tiles_x == 0 && height % tiles_y == 0) { a.tw = width / tiles_x; a.th = height / tiles_y; }
    else { a.tw = (width + (tiles_x - width % tiles_x)) / tiles_x; a.th = (height + (tiles_y - height % tiles_y)) / tiles_y; }
    a.clip = clahe_clip_count(clip_limit, a.tw * a.th);
    a.hist = reinterpret_cast<unsigned*>(ws + b.take(tiles * 256 * 4));
    a.done = reinterpret_cast<unsigned*>(ws + b.take(tiles * 4));
    a.lut = reinterpret_cast<uint8_t*>(ws + b.take(tiles * 256));
    // slabs: enough blocks for two per SM, and at most 65535 pixels per thread (16-bit private counters)
    int split = (int)((148 * 2 + tiles - 1) / tiles);
    if (split > a.th / 8) split = a.th / 8;
    if (split < 1) split = 1;
    while ((long long)((a.th + split - 1) / split) * a.tw > 60000LL * CH_NT) ++split;
    a.split = split;
    a.rows_per_slab = (a.th + split - 1) / split;
    a.split = (a.th + a.rows_per_slab - 1) / a.rows_per_slab;
    if (a.split > 1) PCU(cudaMemsetAsync(ws, 0, (char*)a.lut - ws, st));
    static const bool optin = cudaFuncSetAttribute(clahe_l_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CH_SMEM) == cudaSuccess;
    (void)optin;
    clahe_l_hist_kernel<<<dim3(tiles_x * tiles_y, a.split, batch), CH_NT, CH_SMEM, st>>>(a);
    const int cells = (tiles_x + 1) * (tiles_y + 1);
    const int slabs = (a.th + CA_ROWS - 1) / CA_ROWS;
    clahe_l_apply_kernel<<<dim3(cells, slabs, batch), CA_NT, 0, st>>>(a);
    PCU(cudaGetLastError());
    return UWIE_OK;
}

int uwie_prim_equalize_hist(const uint8_t* frames_dev, int batch, int height, int width, uint8_t* out_dev,
                            void* workspace_dev, size_t workspace_bytes, void* stream) {
    if (!frames_dev || !out_dev || !workspace_dev || batch < 1 || height < 1 || width < 1)
        return uwie_fail(UWIE_ERR_ARG, "equalize_hist: bad arguments");
    if (workspace_bytes < (size_t)batch * 768 * 5) return uwie_fail(UWIE_ERR_NOMEM, "equalize_hist: workspace too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    unsigned* hist = static_cast<unsigned*>(workspace_dev);
    uint8_t* lut = reinterpret_cast<uint8_t*>(hist + (size_t)batch * 768);
    PCU(cudaMemsetAsync(hist, 0, (size_t)batch * 768 * 4, st));
    launch_frame_hist(frames_dev, batch, height, width, hist, st);
    eq_lut_kernel<<<batch, 32, 0, st>>>(hist, (long long)height * width, lut);
    const size_t npix = (size_t)height * width;
    byte_map_kernel<<<dim3(grid_for(npix * 3, 256, 148 * 4), batch), 256, 0, st>>>(frames_dev, npix, lut, out_dev);
    PCU(cudaGetLastError());
    return UWIE_OK;
}

// ---------------------------------------------------------------- float images (dtype 1 = float32, 2 = float64)
static size_t esz(int dtype) { return dtype == 2 ? 8 : 4; }

int uwie_prim_quantise(const void* img_dev, int dtype, size_t count, uint8_t* out_dev, void* stream) {
    if (!img_dev || !out_dev || (dtype != 1 && dtype != 2)) return uwie_fail(UWIE_ERR_ARG, "quantise: bad arguments");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == 1) quantise_kernel<float><<<grid_for(count), 256, 0, st>>>(static_cast<const float*>(img_dev), count, out_dev);
    else quantise_kernel<double><<<grid_for(count), 256, 0, st>>>(static_cast<const double*>(img_dev), count, out_dev);
    PCU(cudaGetLastError());
    return UWIE_OK;
}

namespace {
struct LightLayout { size_t q, cast, rect, sums, ssq, means, part, counts, quad_hist, trace, labels, tile_list, tile_n, total; };
LightLayout light_layout(int H, int W) {
    Bump b; LightLayout L;
    const size_t N = (size_t)H * W;
    L.q = b.take(N * 3); L.cast = b.take(sizeof(CastTables)); L.rect = b.take(sizeof(QuadRect));
    L.sums = b.take(12 * 8); L.ssq = b.take(12 * 8); L.means = b.take(12 * 8);
    L.part = b.take((size_t)4 * kStatBlocks * 3 * 8);
    L.counts = b.take(kQuadLevelsMax * 4 * 4); L.quad_hist = b.take((size_t)kQuadLevelsMax * 4 * 768 * 4);
    L.trace = b.take(kQuadLevelsMax * 4 * 8);
    const size_t slots = canny_tile_slots(H - H / 2, W - W / 2, 4);
    L.labels = b.take(canny_label_bytes(slots));
    L.tile_list = b.take(slots * kCannyListCap * sizeof(int2)); L.tile_n = b.take(slots * 4);
    L.total = b.off;
    return L;
}
}  // namespace

size_t uwie_prim_atmospheric_light_workspace_bytes(int height, int width) {
    if (height < 1 || width < 1) return 0;
    return light_layout(height, width).total;
}

// estimate_atmospheric_light on a float image [H][W][3]: light_dev receives the (3,) pixel in the image's dtype.
int uwie_prim_atmospheric_light(const void* img_dev, int dtype, int height, int width, void* light_dev,
                                void* workspace_dev, size_t workspace_bytes, void* stream) {
    if (!img_dev || !light_dev || !workspace_dev || (dtype != 1 && dtype != 2) || height < 1 || width < 1)
        return uwie_fail(UWIE_ERR_ARG, "atmospheric_light: bad arguments");
    const LightLayout L = light_layout(height, width);
    if (workspace_bytes < L.total) return uwie_fail(UWIE_ERR_NOMEM, "atmospheric_light: workspace too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char* ws = static_cast<char*>(workspace_dev);
    const size_t N = (size_t)height * width;
    uint8_t* q = reinterpret_cast<uint8_t*>(ws + L.q);
    QuadRect* rect = reinterpret_cast<QuadRect*>(ws + L.rect);
    double* sums = reinterpret_cast<double*>(ws + L.sums);
    double* ssq = reinterpret_cast<double*>(ws + L.ssq);
    double* means = reinterpret_cast<double*>(ws + L.means);
    unsigned* counts = reinterpret_cast<unsigned*>(ws + L.counts);
    PCU(cudaMemsetAsync(ws + L.counts, 0, L.trace + kQuadLevelsMax * 4 * 8 - L.counts, st));
    int rc = uwie_prim_quantise(img_dev, dtype, N * 3, q, stream);
    if (rc) return rc;
    identity_cast_kernel<<<1, 256, 0, st>>>(reinterpret_cast<CastTables*>(ws + L.cast));
    rect_init_kernel<<<1, 32, 0, st>>>(rect, height, width, sums, ssq);
    CannyArgs ca;
    std::memset(&ca, 0, sizeof(ca));
    ca.mode = CANNY_QUADRANTS; ca.H = height; ca.W = width; ca.low = 50; ca.high = 150;
    ca.frames = q; ca.cast = reinterpret_cast<CastTables*>(ws + L.cast); ca.rects = rect;
    ca.quad_hist = reinterpret_cast<unsigned*>(ws + L.quad_hist);
    ca.labels = reinterpret_cast<int*>(ws + L.labels);
    ca.tile_list = reinterpret_cast<int2*>(ws + L.tile_list); ca.tile_n = reinterpret_cast<int*>(ws + L.tile_n);
    ca.counts = counts;
    int rows = height, cols = width, level = 0;
    while (rows > 1 && cols > 1 && level < kQuadLevelsMax) {
        rows = rows - rows / 2; cols = cols - cols / 2;
        ca.level = level;
        launch_canny(ca, rows, cols, 4, st);
        const int gx = grid_for((size_t)rows * cols, 256, kStatBlocks);
        double* part = reinterpret_cast<double*>(ws + L.part);
        if (dtype == 1) {
            rect_stats_kernel<float, false><<<dim3(gx, 4), 256, 0, st>>>(static_cast<const float*>(img_dev), width, rect, 1, nullptr, part);
            stats_reduce_kernel<<<1, 32, 0, st>>>(part, gx, 4, sums);
            quad_means_kernel<<<1, 32, 0, st>>>(rect, sums, means, ssq);
            rect_stats_kernel<float, true><<<dim3(gx, 4), 256, 0, st>>>(static_cast<const float*>(img_dev), width, rect, 1, means, part);
        } else {
            rect_stats_kernel<double, false><<<dim3(gx, 4), 256, 0, st>>>(static_cast<const double*>(img_dev), width, rect, 1, nullptr, part);
            stats_reduce_kernel<<<1, 32, 0, st>>>(part, gx, 4, sums);
            quad_means_kernel<<<1, 32, 0, st>>>(rect, sums, means, ssq);
            rect_stats_kernel<double, true><<<dim3(gx, 4), 256, 0, st>>>(static_cast<const double*>(img_dev), width, rect, 1, means, part);
        }
        stats_reduce_kernel<<<1, 32, 0, st>>>(part, gx, 4, ssq);
        quad_descend_float_kernel<<<1, 32, 0, st>>>(rect, sums, ssq, counts, level, reinterpret_cast<double*>(ws + L.trace), sums);
        ++level;
    }
    if (dtype == 1) brightest_float_kernel<float><<<1, 256, 0, st>>>(static_cast<const float*>(img_dev), width, rect, static_cast<float*>(light_dev));
    else brightest_float_kernel<double><<<1, 256, 0, st>>>(static_cast<const double*>(img_dev), width, rect, static_cast<double*>(light_dev));
    PCU(cudaGetLastError());
    return UWIE_OK;
}

namespace {
struct ScoreLayout { size_t q, gray, rect, sums, ssq, means, part, count, labels, tile_list, tile_n, total; };
ScoreLayout score_layout(int H, int W) {
    Bump b; ScoreLayout L;
    const size_t N = (size_t)H * W;
    L.q = b.take(N * 3); L.gray = b.take(N); L.rect = b.take(sizeof(QuadRect));
    L.sums = b.take(12 * 8); L.ssq = b.take(12 * 8); L.means = b.take(12 * 8);
    L.part = b.take((size_t)kStatBlocks * 3 * 8); L.count = b.take(256);
    const size_t slots = canny_tile_slots(H, W, 1);
    L.labels = b.take(canny_label_bytes(slots));
    L.tile_list = b.take(slots * kCannyListCap * sizeof(int2)); L.tile_n = b.take(slots * 4);
    L.total = b.off;
    return L;
}
}  // namespace

size_t uwie_prim_region_score_workspace_bytes(int height, int width) {
    if (height < 1 || width < 1) return 0;
    return score_layout(height, width).total;
}

// compute_Q(block) (ES:147-188, SS:115-157) of a whole float image [H][W][3]: q_dev receives one float64.
int uwie_prim_region_score(const void* img_dev, int dtype, int height, int width, double* q_dev, void* workspace_dev,
                           size_t workspace_bytes, void* stream) {
    if (!img_dev || !q_dev || !workspace_dev || (dtype != 1 && dtype != 2) || height < 1 || width < 1)
        return uwie_fail(UWIE_ERR_ARG, "region_score: bad arguments");
    const ScoreLayout L = score_layout(height, width);
    if (workspace_bytes < L.total) return uwie_fail(UWIE_ERR_NOMEM, "region_score: workspace too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char* ws = static_cast<char*>(workspace_dev);
    const size_t N = (size_t)height * width;
    uint8_t* q = reinterpret_cast<uint8_t*>(ws + L.q);
    uint8_t* gray = reinterpret_cast<uint8_t*>(ws + L.gray);
    QuadRect r; r.y0 = 0; r.y1 = height; r.x0 = 0; r.x1 = width; r.active = 1; r.level = 0;
    PCU(cudaMemcpyAsync(ws + L.rect, &r, sizeof(r), cudaMemcpyHostToDevice, st));
    PCU(cudaMemsetAsync(ws + L.sums, 0, L.count + 256 - L.sums, st));
    double* sums = reinterpret_cast<double*>(ws + L.sums);
    double* ssq = reinterpret_cast<double*>(ws + L.ssq);
    double* means = reinterpret_cast<double*>(ws + L.means);
    unsigned* count = reinterpret_cast<unsigned*>(ws + L.count);
    int rc = uwie_prim_quantise(img_dev, dtype, N * 3, q, stream);
    if (rc) return rc;
    gray_from_bytes_kernel<<<grid_for(N), 256, 0, st>>>(q, N, gray);
    const QuadRect* rect = reinterpret_cast<QuadRect*>(ws + L.rect);
    const int gx = grid_for(N, 256, kStatBlocks);
    double* part = reinterpret_cast<double*>(ws + L.part);
    if (dtype == 1) {
        rect_stats_kernel<float, false><<<dim3(gx, 1), 256, 0, st>>>((const float*)img_dev, width, rect, 0, nullptr, part);
        stats_reduce_kernel<<<1, 32, 0, st>>>(part, gx, 1, sums);
        means3_kernel<<<1, 32, 0, st>>>(sums, (double)N, means, ssq);
        rect_stats_kernel<float, true><<<dim3(gx, 1), 256, 0, st>>>((const float*)img_dev, width, rect, 0, means, part);
    } else {
        rect_stats_kernel<double, false><<<dim3(gx, 1), 256, 0, st>>>((const double*)img_dev, width, rect, 0, nullptr, part);
        stats_reduce_kernel<<<1, 32, 0, st>>>(part, gx, 1, sums);
        means3_kernel<<<1, 32, 0, st>>>(sums, (double)N, means, ssq);
        rect_stats_kernel<double, true><<<dim3(gx, 1), 256, 0, st>>>((const double*)img_dev, width, rect, 0, means, part);
    }
    stats_reduce_kernel<<<1, 32, 0, st>>>(part, gx, 1, ssq);
    CannyArgs ca;
    std::memset(&ca, 0, sizeof(ca));
    ca.mode = CANNY_PLANES; ca.H = height; ca.W = width; ca.low = 50; ca.high = 150;
    ca.gray = gray; ca.planes_per_img = 1;
    ca.labels = reinterpret_cast<int*>(ws + L.labels);
    ca.tile_list = reinterpret_cast<int2*>(ws + L.tile_list); ca.tile_n = reinterpret_cast<int*>(ws + L.tile_n);
    ca.counts = count;
    launch_canny(ca, height, width, 1, st);
    region_score_kernel<<<1, 32, 0, st>>>(sums, ssq, count, (double)N, q_dev);
    PCU(cudaGetLastError());
    return UWIE_OK;
}

// get_brightest_pixel(block) (ES:191-206): first pixel with the largest R + G + B; light_dev: 3 values of the image's dtype.
// Workspace: 256 bytes.
int uwie_prim_brightest_pixel(const void* img_dev, int dtype, int height, int width, void* light_dev, void* workspace_dev,
                              size_t workspace_bytes, void* stream) {
    if (!img_dev || !light_dev || !workspace_dev || workspace_bytes < 256 || (dtype != 1 && dtype != 2) || height < 1 || width < 1)
        return uwie_fail(UWIE_ERR_ARG, "brightest_pixel: bad arguments (workspace of 256 bytes)");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    QuadRect r; r.y0 = 0; r.y1 = height; r.x0 = 0; r.x1 = width; r.active = 0; r.level = 0;
    PCU(cudaMemcpyAsync(workspace_dev, &r, sizeof(r), cudaMemcpyHostToDevice, st));
    const QuadRect* rect = static_cast<const QuadRect*>(workspace_dev);
    if (dtype == 1) brightest_float_kernel<float><<<1, 256, 0, st>>>((const float*)img_dev, width, rect, (float*)light_dev);
    else brightest_float_kernel<double><<<1, 256, 0, st>>>((const double*)img_dev, width, rect, (double*)light_dev);
    PCU(cudaGetLastError());
    return UWIE_OK;
}

size_t uwie_prim_guided_filter_workspace_bytes(int height, int width) {
    if (height < 1 || width < 1) return 0;
    return ((size_t)height * width * 8 * 6 + 4096);
}

// guided_filter(I, p, r, eps) on float64 planes [H][W] (ES:16-46 / SS:25-46); clip_01_1: clip the result to [.1, 1]
int uwie_prim_guided_filter(const double* guide_dev, const double* src_dev, int height, int width, int radius, double eps,
                            int clip_01_1, double* out_dev, void* workspace_dev, size_t workspace_bytes, void* stream) {
    if (!guide_dev || !src_dev || !out_dev || !workspace_dev || height < 1 || width < 1 || radius < 1 || radius > 256)
        return uwie_fail(UWIE_ERR_ARG, "guided_filter: bad arguments (radius must be in [1, 256])");
    if (workspace_bytes < uwie_prim_guided_filter_workspace_bytes(height, width))
        return uwie_fail(UWIE_ERR_NOMEM, "guided_filter: workspace too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t N = (size_t)height * width;
    double* hs = static_cast<double*>(workspace_dev);          // 4 planes
    double* pa = hs + 4 * N; double* pb = pa + N;
    const int g = grid_for(N);
    gf_box_h_kernel<0><<<g, 256, 0, st>>>(guide_dev, src_dev, height, width, radius, hs);
    gf_box_v_ab_kernel<<<g, 256, 0, st>>>(hs, height, width, radius, eps, pa, pb);
    gf_box_h_kernel<1><<<g, 256, 0, st>>>(pa, pb, height, width, radius, hs);
    gf_box_v_q_kernel<<<g, 256, 0, st>>>(hs, guide_dev, height, width, radius, clip_01_1, out_dev);
    PCU(cudaGetLastError());
    return UWIE_OK;
}

size_t uwie_prim_transmission_workspace_bytes(int height, int width) {
    if (height < 1 || width < 1) return 0;
    return uwie_prim_guided_filter_workspace_bytes(height, width) + (size_t)height * width * 16 + 512;
}

// estimate_transmission(img, A, omega, r, eps): float image [H][W][3] and A (3 values of the image's dtype, device)
// -> float64 t [H][W] clipped to [.1, 1].  clip_t0 = 1 and guard = 1e-6 give six_stadigy.py:167-180, clip_t0 = 0 and
// guard = 1e-10 enhancement_strategies.py:209-234.
int uwie_prim_transmission(const void* img_dev, int dtype, int height, int width, const void* light_dev, double omega,
                           int radius, double eps, double guard, int clip_t0, double* t_dev, void* workspace_dev,
                           size_t workspace_bytes, void* stream) {
    if (!img_dev || !light_dev || !t_dev || !workspace_dev || (dtype != 1 && dtype != 2) || height < 1 || width < 1)
        return uwie_fail(UWIE_ERR_ARG, "transmission: bad arguments");
    if (workspace_bytes < uwie_prim_transmission_workspace_bytes(height, width))
        return uwie_fail(UWIE_ERR_NOMEM, "transmission: workspace too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t N = (size_t)height * width;
    double* I = static_cast<double*>(workspace_dev);
    double* P = I + N;
    void* gws = P + N;
    if (dtype == 1)
        thead_float_kernel<float><<<grid_for(N), 256, 0, st>>>(static_cast<const float*>(img_dev), static_cast<const float*>(light_dev), N,
                                                               (float)omega, (float)guard, clip_t0, I, P);
    else
        thead_float_kernel<double><<<grid_for(N), 256, 0, st>>>(static_cast<const double*>(img_dev), static_cast<const double*>(light_dev), N,
                                                                omega, guard, clip_t0, I, P);
    return uwie_prim_guided_filter(I, P, height, width, radius, eps, 1, t_dev, gws, workspace_bytes - 16 * N, stream);
}

// recover_image / restore_image: out = clip((img - A) / t + A, 0, 1); out_dtype 1 float32 (SS:182-188), 2 float64 (ES:237-249)
int uwie_prim_recover(const void* img_dev, int dtype, int height, int width, const double* t_dev, const void* light_dev,
                      void* out_dev, int out_dtype, void* stream) {
    if (!img_dev || !t_dev || !light_dev || !out_dev || (dtype != 1 && dtype != 2) || (out_dtype != 1 && out_dtype != 2))
        return uwie_fail(UWIE_ERR_ARG, "recover: bad arguments");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t N = (size_t)height * width;
    const int g = grid_for(N);
    if (dtype == 1 && out_dtype == 1) recover_kernel<float, float><<<g, 256, 0, st>>>((const float*)img_dev, t_dev, (const float*)light_dev, N, (float*)out_dev);
    else if (dtype == 1) recover_kernel<float, double><<<g, 256, 0, st>>>((const float*)img_dev, t_dev, (const float*)light_dev, N, (double*)out_dev);
    else if (out_dtype == 1) recover_kernel<double, float><<<g, 256, 0, st>>>((const double*)img_dev, t_dev, (const double*)light_dev, N, (float*)out_dev);
    else recover_kernel<double, double><<<g, 256, 0, st>>>((const double*)img_dev, t_dev, (const double*)light_dev, N, (double*)out_dev);
    PCU(cudaGetLastError());
    return UWIE_OK;
}

namespace {
struct StretchLayout { size_t zero_begin, hist_a, hist_b, hist_c, tab_keys, tab_counts, cand_n, stats, status, bad, zero_end,
                       recipe, targets, tables, cand, total; };
StretchLayout stretch_layout(int dtype) {
    Bump b; StretchLayout L;
    L.zero_begin = b.off;
    L.hist_a = b.take(3 * kSelBinsA * 4); L.hist_b = b.take((size_t)kSelTargets * kSelBinsB * 4);
    L.hist_c = b.take((size_t)kSelTargets * kSelBinsC * 4); L.tab_keys = b.take((size_t)kSelTargets * kSelTable * 8);
    L.tab_counts = b.take((size_t)kSelTargets * kSelTable * 4); L.cand_n = b.take(kSelTargets * 4);
    L.stats = b.take(256); L.status = b.take(256); L.bad = b.take(256);
    L.zero_end = b.off;
    L.recipe = b.take(sizeof(Recipe)); L.targets = b.take(kSelTargets * sizeof(SelTarget));
    L.tables = b.take(sizeof(StratTables));
    L.cand = b.take((size_t)kSelTargets * kSelListCap * esz(dtype));
    L.total = b.off;
    return L;
}
}  // namespace

size_t uwie_prim_percentile_stretch_workspace_bytes(int dtype) {
    if (dtype != 1 && dtype != 2) return 0;
    return stretch_layout(dtype).total;
}

// color_enhancement / enhance_contrast / white_balance: per channel clip((x - p_lo) / (p_hi - p_lo + guard), 0, 1) with
// np.percentile(channel, lo / hi) (linear method, exact order statistics).  The image must lie in [0, 1] (it does wherever
// the reference calls this: recovered frames are clipped, CLAHE / equalizeHist outputs are bytes / 255); otherwise
// UWIE_ERR_ARG.  Synchronises `stream` once (the range check and the selection status come back to the host).
int uwie_prim_percentile_stretch(const void* img_dev, int dtype, int height, int width, double lo_pct, double hi_pct,
                                 double guard, void* out_dev, void* workspace_dev, size_t workspace_bytes, void* stream) {
    if (!img_dev || !out_dev || !workspace_dev || (dtype != 1 && dtype != 2) || height < 1 || width < 1)
        return uwie_fail(UWIE_ERR_ARG, "percentile_stretch: bad arguments");
    if (!(lo_pct >= 0 && lo_pct <= 100 && hi_pct >= 0 && hi_pct <= 100))
        return uwie_fail(UWIE_ERR_ARG, "percentile_stretch: percentiles must be in [0, 100]");
    const StretchLayout L = stretch_layout(dtype);
    if (workspace_bytes < L.total) return uwie_fail(UWIE_ERR_NOMEM, "percentile_stretch: workspace too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char* ws = static_cast<char*>(workspace_dev);
    const long long npix = (long long)height * width;
    Recipe rc;
    std::memset(&rc, 0, sizeof(rc));
    rc.dehaze = 1; rc.n_pre = 1;
    auto plan = [&](double pct) {
        PercentilePlan p = dtype == 2 ? percentile_plan_f64(npix, pct) : percentile_plan_f32(npix, pct);
        RankPlan r; r.k = p.k; r.gamma = p.gamma; return r;
    };
    rc.pre[0].lo = plan(lo_pct); rc.pre[0].hi = plan(hi_pct);
    PCU(cudaMemsetAsync(ws + L.zero_begin, 0, L.zero_end - L.zero_begin, st));
    PCU(cudaMemcpyAsync(ws + L.recipe, &rc, sizeof(rc), cudaMemcpyHostToDevice, st));
    unsigned* hist_a = reinterpret_cast<unsigned*>(ws + L.hist_a);
    int* bad = reinterpret_cast<int*>(ws + L.bad);
    if (dtype == 1) level_a_kernel<float><<<grid_for((size_t)npix * 3, 256, 148 * 4), 256, 0, st>>>((const float*)img_dev, (size_t)npix, hist_a, bad);
    else level_a_kernel<double><<<grid_for((size_t)npix * 3, 256, 148 * 4), 256, 0, st>>>((const double*)img_dev, (size_t)npix, hist_a, bad);
    int h_bad = 0;
    PCU(cudaMemcpyAsync(&h_bad, bad, sizeof(int), cudaMemcpyDeviceToHost, st));
    PCU(cudaStreamSynchronize(st));
    if (h_bad) return uwie_fail(UWIE_ERR_ARG, "percentile_stretch: image values outside [0, 1] (or NaN)");
    SelectArgs se;
    std::memset(&se, 0, sizeof(se));
    se.B = 1; se.S = 1; se.n = npix;
    se.lib.f64 = dtype == 2; se.lib.stretch_guard = guard; se.lib.n_strategies = 1;
    se.recipes = reinterpret_cast<Recipe*>(ws + L.recipe);
    se.n_dehaze = 1; se.dehaze_idx[0] = 0;
    se.recovered = img_dev; se.sel_hist_a = hist_a;
    se.sel_hist_b = reinterpret_cast<unsigned*>(ws + L.hist_b); se.sel_hist_c = reinterpret_cast<unsigned*>(ws + L.hist_c);
    se.targets = reinterpret_cast<SelTarget*>(ws + L.targets);
    se.tab_keys = reinterpret_cast<unsigned long long*>(ws + L.tab_keys); se.tab_counts = reinterpret_cast<unsigned*>(ws + L.tab_counts);
    se.cand = ws + L.cand; se.cand_n = reinterpret_cast<unsigned*>(ws + L.cand_n);
    se.stats = reinterpret_cast<unsigned*>(ws + L.stats);
    se.tables = reinterpret_cast<StratTables*>(ws + L.tables);
    se.status = reinterpret_cast<int*>(ws + L.status);
    launch_select(se, st);
    const int g = grid_for((size_t)npix);
    if (dtype == 1) stretch_kernel<float><<<g, 256, 0, st>>>((const float*)img_dev, (size_t)npix, se.tables, (float*)out_dev);
    else stretch_kernel<double><<<g, 256, 0, st>>>((const double*)img_dev, (size_t)npix, se.tables, (double*)out_dev);
    int status = 0;
    PCU(cudaMemcpyAsync(&status, se.status, sizeof(int), cudaMemcpyDeviceToHost, st));
    PCU(cudaStreamSynchronize(st));
    PCU(cudaGetLastError());
    if (status) return uwie_fail(UWIE_ERR_SELECT, "percentile_stretch: selection failed on the device");
    return UWIE_OK;
}

// gamma_correction: out = img ** exponent, clipped to [0, 1] when `clip` (ES:276-285 passes 1 / gamma and clips;
// SS:221-224 passes gamma and does not)
int uwie_prim_gamma(const void* img_dev, int dtype, size_t count, double exponent, int clip, void* out_dev, void* stream) {
    if (!img_dev || !out_dev || (dtype != 1 && dtype != 2)) return uwie_fail(UWIE_ERR_ARG, "gamma: bad arguments");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == 1) gamma_kernel<float><<<grid_for(count), 256, 0, st>>>((const float*)img_dev, count, (double)(float)exponent, clip, (float*)out_dev);
    else gamma_kernel<double><<<grid_for(count), 256, 0, st>>>((const double*)img_dev, count, exponent, clip, (double*)out_dev);
    PCU(cudaGetLastError());
    return UWIE_OK;
}

// color_correction: out = clip(img * scale[c], 0, 1) (six_stadigy.py:305-323: G or B x 0.85)
int uwie_prim_scale_clip(const void* img_dev, int dtype, int height, int width, const double* scale3, void* out_dev, void* stream) {
    if (!img_dev || !out_dev || !scale3 || (dtype != 1 && dtype != 2)) return uwie_fail(UWIE_ERR_ARG, "scale_clip: bad arguments");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t N = (size_t)height * width;
    if (dtype == 1) scale_clip_kernel<float><<<grid_for(N * 3), 256, 0, st>>>((const float*)img_dev, N, (float)scale3[0], (float)scale3[1], (float)scale3[2], (float*)out_dev);
    else scale_clip_kernel<double><<<grid_for(N * 3), 256, 0, st>>>((const double*)img_dev, N, scale3[0], scale3[1], scale3[2], (double*)out_dev);
    PCU(cudaGetLastError());
    return UWIE_OK;
}

// channel means of a float image (detect_image_type, six_stadigy.py:292-302): means_dev receives 3 float64
int uwie_prim_channel_means(const void* img_dev, int dtype, int height, int width, double* means_dev, void* workspace_dev,
                            size_t workspace_bytes, void* stream) {
    if (!img_dev || !means_dev || !workspace_dev || (dtype != 1 && dtype != 2) || workspace_bytes < 4096)
        return uwie_fail(UWIE_ERR_ARG, "channel_means: bad arguments (workspace of 4096 bytes)");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    QuadRect r; r.y0 = 0; r.y1 = height; r.x0 = 0; r.x1 = width; r.active = 1; r.level = 0;
    char* ws = static_cast<char*>(workspace_dev);
    double* sums = reinterpret_cast<double*>(ws + 256);
    double* part = reinterpret_cast<double*>(ws + 512);                     // kStatBlocks x 3 doubles = 1536 bytes
    PCU(cudaMemcpyAsync(ws, &r, sizeof(r), cudaMemcpyHostToDevice, st));
    const int gx = grid_for((size_t)height * width, 256, kStatBlocks);
    if (dtype == 1) rect_stats_kernel<float, false><<<dim3(gx, 1), 256, 0, st>>>((const float*)img_dev, width, reinterpret_cast<QuadRect*>(ws), 0, nullptr, part);
    else rect_stats_kernel<double, false><<<dim3(gx, 1), 256, 0, st>>>((const double*)img_dev, width, reinterpret_cast<QuadRect*>(ws), 0, nullptr, part);
    stats_reduce_kernel<<<1, 32, 0, st>>>(part, gx, 1, sums);
    divide3_kernel<<<1, 32, 0, st>>>(sums, (double)height * (double)width, means_dev);
    PCU(cudaGetLastError());
    return UWIE_OK;
}

}  // extern "C"

// =====================================================================================
// Allocation-free variant of uwie_prim_equalize_hist (no counterpart in the reference; added for callers that
// cannot allocate: graph capture, serving loops).  The channel histograms and the byte tables live in a
// file-scope device array instead of a caller workspace: UWIE_EQ_STATIC_MAX_BATCH frames x 3840 bytes = 240 KB,
// held by every GPU that loads the library until the process exits.
// In host code `g_eq_scratch` is only a handle of the symbol, not a device address: its address on the CURRENT
// device comes from cudaGetSymbolAddress, looked up once per device and cached; the kernels get that pointer.
// All calls on one device share the array: the histogram part is cleared on `stream` before every use, and two
// calls whose work overlaps on the GPU (different streams, no ordering) race on it.
// =====================================================================================
static __device__ __align__(256) unsigned char g_eq_scratch[(size_t)UWIE_EQ_STATIC_MAX_BATCH * 3840];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_eq_mutex;
void* g_eq_addr[kMaxDevices] = {};          // device address of g_eq_scratch, per device ordinal

int eq_scratch_address(void** out) {
    int dev = -1;
    PCU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return uwie_fail(UWIE_ERR_ARG, "equalize_hist_static: device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_eq_mutex);
    if (!g_eq_addr[dev]) {
        void* p = nullptr;
        PCU(cudaGetSymbolAddress(&p, g_eq_scratch));
        if (!p) return uwie_fail(UWIE_ERR_CUDA, "equalize_hist_static: cudaGetSymbolAddress returned a null address");
        g_eq_addr[dev] = p;
    }
    *out = g_eq_addr[dev];
    return UWIE_OK;
}
}  // namespace

extern "C" int uwie_prim_equalize_hist_static(const uint8_t* frames_dev, int batch, int height, int width, uint8_t* out_dev,
                                              void* stream) {
    if (!frames_dev || !out_dev || batch < 1 || height < 1 || width < 1)
        return uwie_fail(UWIE_ERR_ARG, "equalize_hist_static: bad arguments");
    if (batch > UWIE_EQ_STATIC_MAX_BATCH)
        return uwie_fail(UWIE_ERR_ARG, "equalize_hist_static: batch " + std::to_string(batch) + " exceeds the static scratch's maximum of " +
                                           std::to_string(UWIE_EQ_STATIC_MAX_BATCH) + " frames per call");
    void* base = nullptr;
    const int rc = eq_scratch_address(&base);
    if (rc != UWIE_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    unsigned* hist = static_cast<unsigned*>(base);                              // [batch][3][256] counts
    uint8_t* lut = reinterpret_cast<uint8_t*>(hist + (size_t)batch * 768);      // [batch][3][256] byte tables
    PCU(cudaMemsetAsync(hist, 0, (size_t)batch * 768 * 4, st));                 // the array keeps the last call's counts
    launch_frame_hist(frames_dev, batch, height, width, hist, st);
    eq_lut_kernel<<<batch, 32, 0, st>>>(hist, (long long)height * width, lut);
    const size_t npix = (size_t)height * width;
    byte_map_kernel<<<dim3(grid_for(npix * 3, 256, 148 * 4), batch), 256, 0, st>>>(frames_dev, npix, lut, out_dev);
    PCU(cudaGetLastError());
    return UWIE_OK;
}
