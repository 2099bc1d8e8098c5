// CLAHE on the 8-bit Lab L plane and the final per-pixel pass of every strategy.
//   clahe_enhancement / apply_clahe   enhancement_strategies.py:288-307, six_stadigy.py:201-208
//   per-strategy tails                six_stadigy.py:230-285, enhancement_strategies.py:349-474
//   scoring inputs                    quality_assessment.py:15-212 (histograms and moments only;
//                                     the scores themselves are formed in k_score.cu)
// CLAHE follows cv2: tiles on the REFLECT_101-padded plane, clip + redistribute,
// LUT = round(cumsum * 255/area), bilinear blend of the four neighbouring tile LUTs in float32.
#include "k_arith.cuh"
#include "k_kernels.h"

namespace uwie {

namespace {

constexpr int NT = 256;

struct TileGeo { int tw, th; };
__device__ __forceinline__ TileGeo tile_geo(int H, int W, int tx, int ty) {
    TileGeo g;
    if (W % tx == 0 && H % ty == 0) { g.tw = W / tx; g.th = H / ty; }
    else { g.tw = (W + (tx - W % tx)) / tx; g.th = (H + (ty - H % ty)) / ty; }   // cv2 pads BOTH sides' remainders
    return g;
}

// RGB byte triple entering the histogram stage (CLAHE) of strategy s at pixel p
template <class T>
__device__ __forceinline__ void stage_bytes(const ApplyArgs& a, const Recipe& rc, const StratTables& st,
                                            int b, int s, size_t p, int& r, int& g, int& bl) {
    if (!rc.dehaze) {
        const uint8_t* f = a.frames + ((size_t)b * a.H * a.W + p) * 3;
        r = st.preq[0][f[0]]; g = st.preq[1][f[1]]; bl = st.preq[2][f[2]];
    } else {
        const int inst = b * a.n_dehaze + a.dehaze_slot[s];
        const T* x = reinterpret_cast<const T*>(a.recovered) + ((size_t)inst * a.H * a.W + p) * 3;
        int q[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            T v = x[c];
            for (int i = 0; i < rc.n_pre; ++i) v = stretch_apply<T>(v, (T)st.st_lo[i][c], (T)st.st_den[i][c]);
            q[c] = quantise<T>(v);
        }
        r = q[0]; g = q[1]; bl = q[2];
    }
}

// ---------------------------------------------------------------- tile histograms -> tile LUTs
// One block per (tile, strategy, frame).  Tables in shared memory, 4-pixel groups moved as words when the
// tile lies inside the frame on 4-pixel boundaries (always at the BASELINE sizes); the clip / redistribute /
// cumulative-sum tail runs on all 256 threads.
// RGB -> Lab runs ONCE per pixel and strategy: this pass leaves the (L, a, b) bytes of every pixel in the strategy's
// slot of `out`, and apply_kernel continues from them in place (no second conversion, and for the dehaze strategies
// no second read of the float planes and no second pass through the pre-stretches).
template <class T>
__global__ void __launch_bounds__(NT, 6) clahe_lut_kernel(ApplyArgs a) {
    const int s = blockIdx.y, b = blockIdx.z;
    const Recipe& rc = a.recipes[s];
    if (!rc.clahe || (int)blockIdx.x >= rc.tiles_x * rc.tiles_y) return;
    const StratTables& st = a.tables[(size_t)b * a.S + s];
    const TileGeo geo = tile_geo(a.H, a.W, rc.tiles_x, rc.tiles_y);
    const int tyi = blockIdx.x / rc.tiles_x, txi = blockIdx.x % rc.tiles_x;
    __shared__ unsigned sh[8][256];
    __shared__ uint16_t s_gamma[256];
    __shared__ uint16_t s_cbrt[2048 + 8];
    __shared__ uint8_t s_preq[3][256];
    for (int i = threadIdx.x; i < 8 * 256; i += NT) (&sh[0][0])[i] = 0;
    for (int i = threadIdx.x; i < 256; i += NT) s_gamma[i] = a.ctab->gamma_in[i];
    for (int i = threadIdx.x; i < 2048 + 8; i += NT) s_cbrt[i] = a.ctab->cbrt_tab[i];
    for (int i = threadIdx.x; i < 768; i += NT) (&s_preq[0][0])[i] = (&st.preq[0][0])[i];
    const bool dehaze = rc.dehaze != 0;
    const int n_pre = rc.n_pre;
    T lo[kMaxStretch][3], den[kMaxStretch][3], rden[kMaxStretch][3];
#pragma unroll
    for (int i = 0; i < kMaxStretch; ++i)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            lo[i][c] = (T)st.st_lo[i][c]; den[i][c] = (T)st.st_den[i][c];
            rden[i][c] = (i < n_pre && dehaze) ? InvDiv<T>::rcp(den[i][c]) : (T)0;   // unused slots hold garbage tables
        }
    __syncthreads();
    unsigned* h = sh[threadIdx.x >> 5];
    const size_t npix = (size_t)a.H * a.W;
    const uint8_t* frame = a.frames + (size_t)b * npix * 3;
    const T* rec = dehaze ? reinterpret_cast<const T*>(a.recovered) + (size_t)(b * a.n_dehaze + a.dehaze_slot[s]) * npix * 3 : nullptr;
    uint8_t* lab_out = a.out + ((size_t)b * a.S + s) * npix * 3;
    // histogram entry + the pixel's Lab bytes packed as L | a << 8 | b << 16
    auto count_bytes = [&](int r8, int g8, int b8) -> uint32_t {
        int L, A, B;
        rgb_to_lab(s_gamma, s_cbrt, s_preq[0][r8], s_preq[1][g8], s_preq[2][b8], L, A, B);
        atomicAdd(&h[L], 1u);
        return (uint32_t)L | ((uint32_t)A << 8) | ((uint32_t)B << 16);
    };
    auto count_floats = [&](T v0, T v1, T v2) -> uint32_t {
        if (n_pre > 0) {
            v0 = stretch_apply_inv<T>(v0, lo[0][0], den[0][0], rden[0][0]); v1 = stretch_apply_inv<T>(v1, lo[0][1], den[0][1], rden[0][1]);
            v2 = stretch_apply_inv<T>(v2, lo[0][2], den[0][2], rden[0][2]);
        }
        if (n_pre > 1) {
            v0 = stretch_apply_inv<T>(v0, lo[1][0], den[1][0], rden[1][0]); v1 = stretch_apply_inv<T>(v1, lo[1][1], den[1][1], rden[1][1]);
            v2 = stretch_apply_inv<T>(v2, lo[1][2], den[1][2], rden[1][2]);
        }
        int L, A, B;
        rgb_to_lab(s_gamma, s_cbrt, quantise<T>(v0), quantise<T>(v1), quantise<T>(v2), L, A, B);
        atomicAdd(&h[L], 1u);
        return (uint32_t)L | ((uint32_t)A << 8) | ((uint32_t)B << 16);
    };
    const int area = geo.tw * geo.th;
    const int y0 = tyi * geo.th, x0 = txi * geo.tw;
    const bool inside = y0 + geo.th <= a.H && x0 + geo.tw <= a.W && (geo.tw & 3) == 0 && (a.W & 3) == 0 &&
                        ((reinterpret_cast<uintptr_t>(frame) | reinterpret_cast<uintptr_t>(lab_out)) & 3) == 0 &&
                        (!dehaze || (sizeof(T) == 4 && (reinterpret_cast<uintptr_t>(rec) & 15) == 0 && (npix & 3) == 0));
    if (inside) {
        const int gpr = geo.tw >> 2;                       // 4-pixel groups per tile row
        const int groups = gpr * geo.th;
        // (row, group) of a thread's next item without a division per trip: NT groups further on
        const int step_y = NT / gpr, step_x = NT - step_y * gpr;
        int ly = threadIdx.x / gpr, gx = threadIdx.x - ly * gpr;
        for (int i = threadIdx.x; i < groups; i += NT, ly += step_y, gx += step_x) {
            if (gx >= gpr) { gx -= gpr; ++ly; }
            const size_t p = (size_t)(y0 + ly) * a.W + x0 + 4 * gx;
            uint32_t l0, l1, l2, l3;
            if (!dehaze) {
                const uint32_t* w = reinterpret_cast<const uint32_t*>(frame + p * 3);
                const uint32_t w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2);
                l0 = count_bytes(w0 & 255, (w0 >> 8) & 255, (w0 >> 16) & 255);
                l1 = count_bytes(w0 >> 24, w1 & 255, (w1 >> 8) & 255);
                l2 = count_bytes((w1 >> 16) & 255, w1 >> 24, w2 & 255);
                l3 = count_bytes((w2 >> 8) & 255, (w2 >> 16) & 255, w2 >> 24);
            } else {
                const float4* q = reinterpret_cast<const float4*>(rec + p * 3);
                const float4 f0 = q[0], f1 = q[1], f2 = q[2];
                l0 = count_floats((T)f0.x, (T)f0.y, (T)f0.z); l1 = count_floats((T)f0.w, (T)f1.x, (T)f1.y);
                l2 = count_floats((T)f1.z, (T)f1.w, (T)f2.x); l3 = count_floats((T)f2.y, (T)f2.z, (T)f2.w);
            }
            uint32_t* o = reinterpret_cast<uint32_t*>(lab_out + p * 3);
            o[0] = l0 | (l1 << 24); o[1] = (l1 >> 8) | (l2 << 16); o[2] = (l2 >> 16) | (l3 << 8);
        }
    } else {
        for (int i = threadIdx.x; i < area; i += NT) {
            const int ly = i / geo.tw, lx = i - ly * geo.tw;
            int y = y0 + ly, x = x0 + lx;
            if (y >= a.H) y = reflect101(y, a.H);          // cv2 pads the plane with REFLECT_101 up to a tile multiple
            if (x >= a.W) x = reflect101(x, a.W);
            const bool own = y0 + ly < a.H && x0 + lx < a.W;      // a padded (reflected) position only counts
            const size_t p = (size_t)y * a.W + x;
            const uint32_t l = !dehaze ? count_bytes(frame[p * 3], frame[p * 3 + 1], frame[p * 3 + 2])
                                       : count_floats(rec[p * 3], rec[p * 3 + 1], rec[p * 3 + 2]);
            if (own) { lab_out[p * 3] = (uint8_t)l; lab_out[p * 3 + 1] = (uint8_t)(l >> 8); lab_out[p * 3 + 2] = (uint8_t)(l >> 16); }
        }
    }
    __syncthreads();
    // clip, redistribute (clahe_clip_redistribute, in parallel), cumulative sum, LUT
    __shared__ int s_scan[NT / 32];
    __shared__ int s_excess;
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    int v = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) v += (int)sh[k][t];
    const int clip = rc.clahe_clip;
    int over = v > clip ? v - clip : 0;
    if (v > clip) v = clip;
    int e = over;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) e += __shfl_xor_sync(0xffffffffu, e, o);
    if (lane == 0) s_scan[w] = e;
    __syncthreads();
    if (t == 0) { int tot = 0; for (int k = 0; k < NT / 32; ++k) tot += s_scan[k]; s_excess = tot; }
    __syncthreads();
    const int excess = s_excess;
    const int batch = excess / 256, residual = excess - batch * 256;
    v += batch;
    if (residual != 0) {
        int step = 256 / residual;
        if (step < 1) step = 1;
        if (t % step == 0 && t / step < residual) v += 1;      // h[0], h[step], ... while residual lasts
    }
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += u; }
    __syncthreads();
    if (lane == 31) s_scan[w] = incl;
    __syncthreads();
    int base = 0;
    for (int k = 0; k < w; ++k) base += s_scan[k];
    a.tile_lut[(((size_t)b * a.S + s) * kMaxTiles + blockIdx.x) * 256 + t] = clahe_lut_entry(base + incl, clahe_lut_scale(area));
}

// cv2 CLAHE interpolation (modules/imgproc/src/clahe.cpp, CLAHE_Interpolation_Body): the row part (tile rows
// ty1 / ty2 and their weights) is shared by the pixels of a row, the column part is per pixel.
struct ClaheRow {
    int off1, off2;          // byte offsets of the LUT rows ty1, ty2
    float ya, ya1;
};
__device__ __forceinline__ ClaheRow clahe_row(int tiles_x, int tiles_y, float inv_th, int y) {
    const float tyf = __fsub_rn(__fmul_rn((float)y, inv_th), 0.5f);
    int ty1 = (int)floorf(tyf);
    ClaheRow r;
    r.ya = __fsub_rn(tyf, (float)ty1);
    r.ya1 = __fsub_rn(1.0f, r.ya);
    int ty2 = ty1 + 1;
    ty1 = max(ty1, 0);
    ty2 = min(ty2, tiles_y - 1);
    r.off1 = ty1 * tiles_x * 256; r.off2 = ty2 * tiles_x * 256;
    return r;
}
__device__ __forceinline__ int clahe_interp(const uint8_t* luts, int tiles_x, float inv_tw, const ClaheRow& row, int x, int v) {
    const float txf = __fsub_rn(__fmul_rn((float)x, inv_tw), 0.5f);
    int tx1 = (int)floorf(txf);
    const float xa = __fsub_rn(txf, (float)tx1);
    const float xa1 = __fsub_rn(1.0f, xa);
    int tx2 = tx1 + 1;
    tx1 = max(tx1, 0);
    tx2 = min(tx2, tiles_x - 1);
    const int c1 = tx1 * 256 + v, c2 = tx2 * 256 + v;
    const float l11 = (float)luts[row.off1 + c1];
    const float l12 = (float)luts[row.off1 + c2];
    const float l21 = (float)luts[row.off2 + c1];
    const float l22 = (float)luts[row.off2 + c2];
    const float top = __fadd_rn(__fmul_rn(l11, xa1), __fmul_rn(l12, xa));
    const float bot = __fadd_rn(__fmul_rn(l21, xa1), __fmul_rn(l22, xa));
    const float res = __fadd_rn(__fmul_rn(top, row.ya1), __fmul_rn(bot, row.ya));
    return clampi(__float2int_rn(res), 0, 255);
}

// ---------------------------------------------------------------- per-pixel tail of a strategy
// FINAL = false: the CLAHE output of strategies with post-stretches: its per-channel histograms (the post-stretch
//                 percentiles come from them) and its bytes, left in the strategy's slot of `out`
// FINAL = true : output bytes, gray plane, scoring accumulators.  A strategy that went through the FINAL = false pass
//                continues from the bytes in `out` (in place) as a plain table pass: the Lab round trip and the tile
//                interpolation run once per pixel, not twice.
// A block owns pix_per_block consecutive pixels; a thread takes 4 consecutive pixels at a time so that
// frame bytes, recovered floats, output bytes and gray bytes move as 32 / 128-bit words.  The colour
// tables and the strategy's byte tables are staged in shared memory once per block.
struct ApplyShared {
    ColorTables ct;                 // 10.5 KB
    uint8_t preq[3][256];
    uint8_t postq[3][256];
};

template <class T> struct Vec4 { T v[4][3]; };

// CLAHE / DEHAZE: the strategy's path is fixed at compile time (one launch per variant present in the
// library), so a pixel runs straight-line code and a kernel carries only its own path.  A CLAHE strategy starts
// from the Lab bytes clahe_lut_kernel left in its slot of `out` (whether or not a dehaze step came before: that
// is all folded into those bytes), so CLAHE = true is only instantiated with DEHAZE = false.
template <class T, bool FINAL, bool CLAHE, bool DEHAZE>
__global__ void __launch_bounds__(NT, 4) apply_kernel(ApplyArgs a) {
    constexpr int variant = (CLAHE ? 2 : 0) + (DEHAZE ? 1 : 0);
    const int s = FINAL ? a.fin_s[variant][blockIdx.y] : a.var_s[variant][blockIdx.y], b = blockIdx.z;
    const Recipe& rc = a.recipes[s];
    if (!FINAL && !(rc.clahe && rc.n_post > 0)) return;
    const StratTables& st = a.tables[(size_t)b * a.S + s];
    const unsigned npix = (unsigned)a.H * (unsigned)a.W;
    const size_t bs = (size_t)b * a.S + s;
    __shared__ unsigned sh_a[3][256];     // HIST: channel histograms; FINAL: [0] gray, [1] saturation
    __shared__ __align__(16) ApplyShared sm;
    __shared__ T s_postf[3][256];
    for (int i = threadIdx.x; i < 768; i += NT) (&sh_a[0][0])[i] = 0;
    {
        const uint32_t* src = reinterpret_cast<const uint32_t*>(a.ctab);
        uint32_t* dst = reinterpret_cast<uint32_t*>(&sm.ct);
        for (int i = threadIdx.x; i < (int)(sizeof(ColorTables) / 4); i += NT) dst[i] = src[i];
        for (int i = threadIdx.x; i < 768; i += NT) {
            (&sm.preq[0][0])[i] = (&st.preq[0][0])[i];
            (&sm.postq[0][0])[i] = (&st.postq[0][0])[i];
            (&s_postf[0][0])[i] = (T)(&st.postf[0][0])[i];
        }
    }
    constexpr bool clahe = CLAHE, dehaze = DEHAZE;
    const int n_pre = rc.n_pre;
    T lo[kMaxStretch][3], den[kMaxStretch][3], rden[kMaxStretch][3];
#pragma unroll
    for (int i = 0; i < kMaxStretch; ++i)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            lo[i][c] = (T)st.st_lo[i][c]; den[i][c] = (T)st.st_den[i][c];
            rden[i][c] = (i < n_pre && dehaze) ? InvDiv<T>::rcp(den[i][c]) : (T)0;   // unused slots hold garbage tables
        }
    __syncthreads();

    TileGeo geo = {1, 1};
    float inv_tw = 0.f, inv_th = 0.f;
    const uint8_t* luts = a.tile_lut + bs * kMaxTiles * 256;
    const int tiles_x = rc.tiles_x, tiles_y = rc.tiles_y;
    if (clahe) {
        geo = tile_geo(a.H, a.W, tiles_x, tiles_y);
        inv_tw = __fdiv_rn(1.0f, (float)geo.tw);
        inv_th = __fdiv_rn(1.0f, (float)geo.th);
    }
    uint8_t* out = a.out + bs * (size_t)npix * 3;
    uint8_t* gray = FINAL ? a.gray + bs * (size_t)npix : nullptr;
    const bool in_place = CLAHE || (FINAL && !DEHAZE && a.from_out[s] != 0);     // block-uniform
    const uint8_t* frame = in_place ? out : a.frames + (size_t)b * npix * 3;
    const T* rec = dehaze ? reinterpret_cast<const T*>(a.recovered) + (size_t)(b * a.n_dehaze + a.dehaze_slot[s]) * npix * 3 : nullptr;
    const bool words_ok = ((reinterpret_cast<uintptr_t>(frame) | reinterpret_cast<uintptr_t>(out)) & 3) == 0 &&
                          (!FINAL || (reinterpret_cast<uintptr_t>(gray) & 3) == 0);
    double m_rg = 0, m_rg2 = 0, m_yb = 0, m_yb2 = 0;
    unsigned long long lum = 0;

    static_assert(kMaxStretch == 2, "pre_stretch is written out for two stretches");
    auto pre_stretch = [&](T& v0, T& v1, T& v2) {          // static indices keep lo / den in registers
        if (n_pre > 0) {
            v0 = stretch_apply_inv<T>(v0, lo[0][0], den[0][0], rden[0][0]); v1 = stretch_apply_inv<T>(v1, lo[0][1], den[0][1], rden[0][1]);
            v2 = stretch_apply_inv<T>(v2, lo[0][2], den[0][2], rden[0][2]);
        }
        if (n_pre > 1) {
            v0 = stretch_apply_inv<T>(v0, lo[1][0], den[1][0], rden[1][0]); v1 = stretch_apply_inv<T>(v1, lo[1][1], den[1][1], rden[1][1]);
            v2 = stretch_apply_inv<T>(v2, lo[1][2], den[1][2], rden[1][2]);
        }
    };
    // one pixel: bytes (no-dehaze) or floats (dehaze) in -> float value f and byte q per channel out.
    // In the HIST pass q receives the CLAHE output bytes and f is not written.
    auto shade = [&](int r8, int g8, int b8, T fr, T fg, T fb, const ClaheRow& row, int x, T* f, int* q) -> bool {
        if (clahe) {
            const int A = g8, Bc = b8;                   // the pixel's (L, a, b) bytes, see clahe_lut_kernel
            const int L = clahe_interp(luts, tiles_x, inv_tw, row, x, r8);
            int k0, k1, k2;
            lab_to_rgb(sm.ct.lab_yf, sm.ct.gamma_out, L, A, Bc, k0, k1, k2);
            if (!FINAL) {
                atomicAdd(&sh_a[0][k0], 1u); atomicAdd(&sh_a[1][k1], 1u); atomicAdd(&sh_a[2][k2], 1u);
                q[0] = k0; q[1] = k1; q[2] = k2;
                return true;
            }
            f[0] = s_postf[0][k0]; f[1] = s_postf[1][k1]; f[2] = s_postf[2][k2];
            q[0] = sm.postq[0][k0]; q[1] = sm.postq[1][k1]; q[2] = sm.postq[2][k2];
        } else if (!dehaze) {
            f[0] = s_postf[0][r8]; f[1] = s_postf[1][g8]; f[2] = s_postf[2][b8];
            q[0] = sm.postq[0][r8]; q[1] = sm.postq[1][g8]; q[2] = sm.postq[2][b8];
        } else {
            T v0 = fr, v1 = fg, v2 = fb;
            pre_stretch(v0, v1, v2);
            f[0] = apply_gamma_bytes<T>(v0, rc); f[1] = apply_gamma_bytes<T>(v1, rc); f[2] = apply_gamma_bytes<T>(v2, rc);
            q[0] = quantise<T>(f[0]); q[1] = quantise<T>(f[1]); q[2] = quantise<T>(f[2]);
        }
        return true;
    };
    // scoring accumulators of one output pixel; returns its gray byte
    auto account = [&](const T* f, const int* q) -> int {
        const int gq = gray_q15(q[0], q[1], q[2]);
        atomicAdd(&sh_a[0][gq], 1u);
        atomicAdd(&sh_a[1][hsv_saturation(sm.ct.sdiv, q[0], q[1], q[2])], 1u);
        lum += (unsigned)lab_L_only(sm.ct.gamma_in, sm.ct.cbrt_tab, q[0], q[1], q[2]);
        // Hasler-Suesstrunk opponent channels on the FLOAT image, in its own float type
        const double rg = (double)Ar<T>::sub(f[0], f[1]);
        const double yb = (double)Ar<T>::sub(Ar<T>::mul((T)0.5, Ar<T>::add(f[0], f[1])), f[2]);
        m_rg += rg; m_rg2 += rg * rg; m_yb += yb; m_yb2 += yb * yb;
        return gq;
    };

    const unsigned p0 = blockIdx.x * (unsigned)a.pix_per_block;
    const unsigned p1 = min(p0 + (unsigned)a.pix_per_block, npix);
    const unsigned p1v = words_ok ? p0 + ((p1 - p0) & ~3u) : p0;      // full 4-pixel groups, word I/O
    for (unsigned p = p0 + 4 * threadIdx.x; p < p1v; p += 4 * NT) {
        uint32_t w0 = 0, w1 = 0, w2 = 0;
        float4 f0 = make_float4(0, 0, 0, 0), f1 = f0, f2 = f0;
        T d[12];
        if (!dehaze) {
            const uint32_t* w = reinterpret_cast<const uint32_t*>(frame + (size_t)p * 3);
            if (in_place) { w0 = w[0]; w1 = w[1]; w2 = w[2]; }      // these bytes are overwritten below: no read-only path
            else { w0 = __ldg(w); w1 = __ldg(w + 1); w2 = __ldg(w + 2); }
        } else if (sizeof(T) == 4 && (reinterpret_cast<uintptr_t>(rec + (size_t)p * 3) & 15) == 0) {
            const float4* qv = reinterpret_cast<const float4*>(rec + (size_t)p * 3);
            f0 = qv[0]; f1 = qv[1]; f2 = qv[2];
            d[0] = (T)f0.x; d[1] = (T)f0.y; d[2] = (T)f0.z; d[3] = (T)f0.w; d[4] = (T)f1.x; d[5] = (T)f1.y;
            d[6] = (T)f1.z; d[7] = (T)f1.w; d[8] = (T)f2.x; d[9] = (T)f2.y; d[10] = (T)f2.z; d[11] = (T)f2.w;
        } else {
#pragma unroll
            for (int i = 0; i < 12; ++i) d[i] = rec[(size_t)p * 3 + i];
        }
        int y = 0, x = 0;
        ClaheRow row = {0, 0, 0.f, 0.f};
        if (clahe) {
            y = (int)(p / (unsigned)a.W); x = (int)(p - (unsigned)y * (unsigned)a.W);
            row = clahe_row(tiles_x, tiles_y, inv_th, y);
        }
        uint32_t o0 = 0, o1 = 0, o2 = 0, gw = 0;
        T f[3];
        int q[3];
        // pixel 0: bytes 0..2
        if (shade(w0 & 255, (w0 >> 8) & 255, (w0 >> 16) & 255, d[0], d[1], d[2], row, x, f, q)) {
            o0 = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16);
            if (FINAL) gw = (uint32_t)account(f, q);
        }
        if (++x == a.W) { x = 0; ++y; if (clahe) row = clahe_row(tiles_x, tiles_y, inv_th, y); }
        // pixel 1: bytes 3..5
        if (shade(w0 >> 24, w1 & 255, (w1 >> 8) & 255, d[3], d[4], d[5], row, x, f, q)) {
            o0 |= (uint32_t)q[0] << 24; o1 = (uint32_t)q[1] | ((uint32_t)q[2] << 8);
            if (FINAL) gw |= (uint32_t)account(f, q) << 8;
        }
        if (++x == a.W) { x = 0; ++y; if (clahe) row = clahe_row(tiles_x, tiles_y, inv_th, y); }
        // pixel 2: bytes 6..8
        if (shade((w1 >> 16) & 255, w1 >> 24, w2 & 255, d[6], d[7], d[8], row, x, f, q)) {
            o1 |= ((uint32_t)q[0] << 16) | ((uint32_t)q[1] << 24); o2 = (uint32_t)q[2];
            if (FINAL) gw |= (uint32_t)account(f, q) << 16;
        }
        if (++x == a.W) { x = 0; ++y; if (clahe) row = clahe_row(tiles_x, tiles_y, inv_th, y); }
        // pixel 3: bytes 9..11
        if (shade((w2 >> 8) & 255, (w2 >> 16) & 255, w2 >> 24, d[9], d[10], d[11], row, x, f, q)) {
            o2 |= ((uint32_t)q[0] << 8) | ((uint32_t)q[1] << 16) | ((uint32_t)q[2] << 24);
            if (FINAL) gw |= (uint32_t)account(f, q) << 24;
        }
        {
            uint32_t* o = reinterpret_cast<uint32_t*>(out + (size_t)p * 3);
            o[0] = o0; o[1] = o1; o[2] = o2;
            if (FINAL) *reinterpret_cast<uint32_t*>(gray + p) = gw;
        }
    }
    // what is left: the last 1-3 pixels of the frame, or everything when the buffers are not word-aligned
    for (unsigned p = p1v + threadIdx.x; p < p1; p += NT) {
        const uint8_t* fp = frame + (size_t)p * 3;
        T f[3], d0 = 0, d1 = 0, d2 = 0;
        int q[3];
        if (dehaze) { d0 = rec[(size_t)p * 3]; d1 = rec[(size_t)p * 3 + 1]; d2 = rec[(size_t)p * 3 + 2]; }
        const int y = (int)(p / (unsigned)a.W), x = (int)(p - (unsigned)y * (unsigned)a.W);
        const ClaheRow row = clahe ? clahe_row(tiles_x, tiles_y, inv_th, y) : ClaheRow{0, 0, 0.f, 0.f};
        if (shade(fp[0], fp[1], fp[2], d0, d1, d2, row, x, f, q)) {
            out[(size_t)p * 3] = (uint8_t)q[0]; out[(size_t)p * 3 + 1] = (uint8_t)q[1]; out[(size_t)p * 3 + 2] = (uint8_t)q[2];
            if (FINAL) gray[p] = (uint8_t)account(f, q);
        }
    }
    __syncthreads();
    if (!FINAL) {
        unsigned* gh = a.out_hist + bs * 768;
        for (int i = threadIdx.x; i < 768; i += NT) { const unsigned v = (&sh_a[0][0])[i]; if (v) atomicAdd(&gh[i], v); }
        return;
    }
    QaAccum& acc = a.qa[bs];
    for (int i = threadIdx.x; i < 256; i += NT) {
        if (sh_a[0][i]) atomicAdd(&acc.gray_hist[i], sh_a[0][i]);
        if (sh_a[1][i]) atomicAdd(&acc.sat_hist[i], sh_a[1][i]);
    }
    // block reduction of the five running sums, fixed order -> bit-reproducible partials
    __shared__ double red[NT / 32][5];
    double v5[5] = {m_rg, m_rg2, m_yb, m_yb2, (double)lum};
#pragma unroll
    for (int i = 0; i < 5; ++i) v5[i] = warp_sum_f64(v5[i]);
    if ((threadIdx.x & 31) == 0) for (int i = 0; i < 5; ++i) red[threadIdx.x >> 5][i] = v5[i];
    __syncthreads();
    if (threadIdx.x < 5) {
        double t = 0;
        for (int w = 0; w < NT / 32; ++w) t += red[w][threadIdx.x];
        if (threadIdx.x < 4) a.cf_part[(bs * a.final_blocks + blockIdx.x) * 4 + threadIdx.x] = t;
        else atomicAdd(&acc.lum_sum, (unsigned long long)t);
    }
}

}  // namespace

int launch_clahe_luts(const ApplyArgs& a, int max_tiles, cudaStream_t st) {
    if (max_tiles <= 0) return 0;
    dim3 grid(max_tiles, a.S, a.B);
    if (a.lib.f64) clahe_lut_kernel<double><<<grid, NT, 0, st>>>(a);
    else clahe_lut_kernel<float><<<grid, NT, 0, st>>>(a);
    return 1;
}

template <class T, bool FINAL>
int launch_apply_t(const ApplyArgs& a, cudaStream_t st) {
    int launches = 0;
    for (int v = FINAL ? 0 : 2; v < 4; ++v) {              // the histogram pass exists for CLAHE strategies only
        const int n = FINAL ? a.fin_n[v] : a.var_n[v];
        if (n == 0 || (!FINAL && a.var_hist[v] == 0)) continue;
        dim3 grid(a.final_blocks, n, a.B);
        switch (v) {
            case 0: apply_kernel<T, FINAL, false, false><<<grid, NT, 0, st>>>(a); break;
            case 1: apply_kernel<T, FINAL, false, true><<<grid, NT, 0, st>>>(a); break;
            case 2: apply_kernel<T, FINAL, true, false><<<grid, NT, 0, st>>>(a); break;
            default: break;                                // variant 3 is never filled (see apply_kernel)
        }
        ++launches;
    }
    return launches;
}

int launch_apply(const ApplyArgs& a, bool final_pass, cudaStream_t st) {
    if (a.lib.f64) return final_pass ? launch_apply_t<double, true>(a, st) : launch_apply_t<double, false>(a, st);
    return final_pass ? launch_apply_t<float, true>(a, st) : launch_apply_t<float, false>(a, st);
}

}  // namespace uwie
