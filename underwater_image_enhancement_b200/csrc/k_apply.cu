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
template <class T>
__global__ void __launch_bounds__(NT) clahe_lut_kernel(ApplyArgs a) {
    const int s = blockIdx.y, b = blockIdx.z;
    const Recipe& rc = a.recipes[s];
    if (!rc.clahe || (int)blockIdx.x >= rc.tiles_x * rc.tiles_y) return;
    const StratTables& st = a.tables[(size_t)b * a.S + s];
    const TileGeo geo = tile_geo(a.H, a.W, rc.tiles_x, rc.tiles_y);
    const int tyi = blockIdx.x / rc.tiles_x, txi = blockIdx.x % rc.tiles_x;
    __shared__ unsigned sh[8][256];
    for (int i = threadIdx.x; i < 8 * 256; i += NT) (&sh[0][0])[i] = 0;
    __syncthreads();
    unsigned* h = sh[threadIdx.x >> 5];
    const int area = geo.tw * geo.th;
    for (int i = threadIdx.x; i < area; i += NT) {
        const int ly = i / geo.tw, lx = i - ly * geo.tw;
        int y = tyi * geo.th + ly, x = txi * geo.tw + lx;
        if (y >= a.H) y = reflect101(y, a.H);
        if (x >= a.W) x = reflect101(x, a.W);
        int r, g, bl;
        stage_bytes<T>(a, rc, st, b, s, (size_t)y * a.W + x, r, g, bl);
        atomicAdd(&h[lab_L_only(a.ctab->gamma_in, a.ctab->cbrt_tab, r, g, bl)], 1u);
    }
    __syncthreads();
    __shared__ int hist[256];
    {
        unsigned v = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) v += sh[w][threadIdx.x];
        hist[threadIdx.x] = (int)v;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        clahe_clip_redistribute(hist, rc.clahe_clip);
        const float scale = clahe_lut_scale(area);
        uint8_t* lut = a.tile_lut + (((size_t)b * a.S + s) * kMaxTiles + blockIdx.x) * 256;
        int acc = 0;
        for (int i = 0; i < 256; ++i) { acc += hist[i]; lut[i] = clahe_lut_entry(acc, scale); }
    }
}

__device__ __forceinline__ int clahe_interp(const uint8_t* luts, int tiles_x, int tiles_y, float inv_tw,
                                            float inv_th, int y, int x, int v) {
    const float txf = __fsub_rn(__fmul_rn((float)x, inv_tw), 0.5f);
    const float tyf = __fsub_rn(__fmul_rn((float)y, inv_th), 0.5f);
    int tx1 = (int)floorf(txf), ty1 = (int)floorf(tyf);
    const float xa = __fsub_rn(txf, (float)tx1), ya = __fsub_rn(tyf, (float)ty1);
    const float xa1 = __fsub_rn(1.0f, xa), ya1 = __fsub_rn(1.0f, ya);
    int tx2 = tx1 + 1, ty2 = ty1 + 1;
    tx1 = max(tx1, 0); ty1 = max(ty1, 0);
    tx2 = min(tx2, tiles_x - 1); ty2 = min(ty2, tiles_y - 1);
    const float l11 = (float)luts[((size_t)ty1 * tiles_x + tx1) * 256 + v];
    const float l12 = (float)luts[((size_t)ty1 * tiles_x + tx2) * 256 + v];
    const float l21 = (float)luts[((size_t)ty2 * tiles_x + tx1) * 256 + v];
    const float l22 = (float)luts[((size_t)ty2 * tiles_x + tx2) * 256 + v];
    const float top = __fadd_rn(__fmul_rn(l11, xa1), __fmul_rn(l12, xa));
    const float bot = __fadd_rn(__fmul_rn(l21, xa1), __fmul_rn(l22, xa));
    const float res = __fadd_rn(__fmul_rn(top, ya1), __fmul_rn(bot, ya));
    return clampi(__float2int_rn(res), 0, 255);
}

// ---------------------------------------------------------------- per-pixel tail of a strategy
// FINAL = false: only the per-channel histograms of the CLAHE output (strategies with post-stretches)
// FINAL = true : output bytes, gray plane, scoring accumulators
template <class T, bool FINAL>
__global__ void __launch_bounds__(NT) apply_kernel(ApplyArgs a) {
    const int s = blockIdx.y, b = blockIdx.z;
    const Recipe& rc = a.recipes[s];
    if (!FINAL && !(rc.clahe && rc.n_post > 0)) return;
    const StratTables& st = a.tables[(size_t)b * a.S + s];
    const size_t npix = (size_t)a.H * a.W;
    const size_t bs = (size_t)b * a.S + s;
    __shared__ unsigned sh_a[3][256];     // HIST: channel histograms; FINAL: [0] gray, [1] saturation
    for (int i = threadIdx.x; i < 768; i += NT) (&sh_a[0][0])[i] = 0;
    __syncthreads();

    TileGeo geo = {1, 1};
    float inv_tw = 0.f, inv_th = 0.f;
    const uint8_t* luts = a.tile_lut + bs * kMaxTiles * 256;
    if (rc.clahe) {
        geo = tile_geo(a.H, a.W, rc.tiles_x, rc.tiles_y);
        inv_tw = __fdiv_rn(1.0f, (float)geo.tw);
        inv_th = __fdiv_rn(1.0f, (float)geo.th);
    }
    const ColorTables* ct = a.ctab;
    uint8_t* out = FINAL ? a.out + bs * npix * 3 : nullptr;
    uint8_t* gray = FINAL ? a.gray + bs * npix : nullptr;
    double m_rg = 0, m_rg2 = 0, m_yb = 0, m_yb2 = 0;
    unsigned long long lum = 0;

    const size_t p0 = (size_t)blockIdx.x * a.pix_per_block;
    const size_t p1 = min(p0 + (size_t)a.pix_per_block, npix);
    for (size_t p = p0 + threadIdx.x; p < p1; p += NT) {
        T f[3];
        int q[3];
        if (rc.clahe) {
            int r, g, bl, L, A, Bc;
            stage_bytes<T>(a, rc, st, b, s, p, r, g, bl);
            rgb_to_lab(ct->gamma_in, ct->cbrt_tab, r, g, bl, L, A, Bc);
            const int y = (int)(p / a.W), x = (int)(p - (size_t)y * a.W);
            L = clahe_interp(luts, rc.tiles_x, rc.tiles_y, inv_tw, inv_th, y, x, L);
            int k[3];
            lab_to_rgb(ct->lab_yf, ct->gamma_out, L, A, Bc, k[0], k[1], k[2]);
            if (!FINAL) {
                atomicAdd(&sh_a[0][k[0]], 1u); atomicAdd(&sh_a[1][k[1]], 1u); atomicAdd(&sh_a[2][k[2]], 1u);
                continue;
            }
#pragma unroll
            for (int c = 0; c < 3; ++c) { f[c] = (T)st.postf[c][k[c]]; q[c] = st.postq[c][k[c]]; }
        } else if (!rc.dehaze) {
            const uint8_t* fr = a.frames + ((size_t)b * npix + p) * 3;
#pragma unroll
            for (int c = 0; c < 3; ++c) { f[c] = (T)st.postf[c][fr[c]]; q[c] = st.postq[c][fr[c]]; }
        } else {
            const int inst = b * a.n_dehaze + a.dehaze_slot[s];
            const T* x = reinterpret_cast<const T*>(a.recovered) + ((size_t)inst * npix + p) * 3;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                T v = x[c];
                for (int i = 0; i < rc.n_pre; ++i) v = stretch_apply<T>(v, (T)st.st_lo[i][c], (T)st.st_den[i][c]);
                v = apply_gamma<T>(v, rc);
                f[c] = v; q[c] = quantise<T>(v);
            }
        }
        if (FINAL) {
            uint8_t* o = out + p * 3;
            o[0] = (uint8_t)q[0]; o[1] = (uint8_t)q[1]; o[2] = (uint8_t)q[2];
            const int gq = gray_q15(q[0], q[1], q[2]);
            gray[p] = (uint8_t)gq;
            atomicAdd(&sh_a[0][gq], 1u);
            atomicAdd(&sh_a[1][hsv_saturation(ct->sdiv, q[0], q[1], q[2])], 1u);
            lum += (unsigned)lab_L_only(ct->gamma_in, ct->cbrt_tab, q[0], q[1], q[2]);
            // Hasler-Suesstrunk opponent channels on the FLOAT image, in its own float type
            const double rg = (double)Ar<T>::sub(f[0], f[1]);
            const double yb = (double)Ar<T>::sub(Ar<T>::mul((T)0.5, Ar<T>::add(f[0], f[1])), f[2]);
            m_rg += rg; m_rg2 += rg * rg; m_yb += yb; m_yb2 += yb * yb;
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

int launch_apply(const ApplyArgs& a, bool final_pass, cudaStream_t st) {
    dim3 grid(a.final_blocks, a.S, a.B);
    if (a.lib.f64) {
        if (final_pass) apply_kernel<double, true><<<grid, NT, 0, st>>>(a);
        else apply_kernel<double, false><<<grid, NT, 0, st>>>(a);
    } else {
        if (final_pass) apply_kernel<float, true><<<grid, NT, 0, st>>>(a);
        else apply_kernel<float, false><<<grid, NT, 0, st>>>(a);
    }
    return 1;
}

}  // namespace uwie
