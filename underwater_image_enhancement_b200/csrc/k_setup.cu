// Table-building kernels: everything in a strategy that is a per-channel function of a byte
// (frame byte, or CLAHE / equalizeHist output byte) is folded into 256-entry tables here, in the
// float type of the library being reproduced.
//   colour cast           six_stadigy.py:292-323
//   percentile stretches  six_stadigy.py:190-199, 210-219; enhancement_strategies.py:252-273
//   equalizeHist          enhancement_strategies.py:331-345
//   gamma                 six_stadigy.py:221-224; enhancement_strategies.py:276-285
#include "k_arith.cuh"
#include "k_kernels.h"

namespace uwie {

namespace {

// ---------------------------------------------------------------- colour cast + byte tables
__global__ void __launch_bounds__(256) setup_cast_kernel(const unsigned* __restrict__ hist, long long n,
                                                        int colour_cast, int* __restrict__ types,
                                                        CastTables* __restrict__ cast) {
    const int b = blockIdx.x, v = threadIdx.x;
    const unsigned* h = hist + (size_t)b * 768;
    __shared__ double s_part[3][8];
    __shared__ int s_type;
    const float fv = __fdiv_rn((float)v, 255.0f);            // img.astype(np.float32) / 255.0
    for (int c = 0; c < 3; ++c) {
        double s = warp_sum_f64((double)h[c * 256 + v] * (double)fv);
        if ((v & 31) == 0) s_part[c][v >> 5] = s;
    }
    __syncthreads();
    if (v == 0) {
        float m[3];
        for (int c = 0; c < 3; ++c) {
            double s = 0.0;
            for (int w = 0; w < 8; ++w) s += s_part[c][w];
            m[c] = (float)(s / (double)n);
        }
        int t = 0;   // 0 normal, 1 greenish, 2 bluish
        if (colour_cast) {
            const float r = m[0], g = m[1], bl = m[2];
            if (g > r && g > bl && __fsub_rn(g, r) > 0.05f) t = 1;
            else if (bl > r && bl > g && __fsub_rn(bl, r) > 0.05f) t = 2;
        }
        s_type = t;
        types[b] = t;
    }
    __syncthreads();
    const int t = s_type;
    for (int c = 0; c < 3; ++c) {
        float f = fv;
        if ((t == 1 && c == 1) || (t == 2 && c == 2)) {
            f = __fmul_rn(f, 0.85f);                          // out[:, :, ch] * 0.85 in float32
            f = f < 0.f ? 0.f : (f > 1.f ? 1.f : f);
        }
        cast[b].curf[c][v] = f;
        cast[b].curq[c][v] = (uint8_t)(int)__fmul_rn(f, 255.0f);
    }
}

// ---------------------------------------------------------------- LUT chains
// block = 768 threads: thread <-> (channel, byte)
template <class T>
__device__ void chain_stretch(const unsigned* h, double (*val)[256], long long n, const StretchPlan& sp,
                              double guard, double* s_lo, double* s_den) {
    const int c = threadIdx.x >> 8, v = threadIdx.x & 255;
    if (v == 0) {
        const T lo = lut_percentile<T>(h + c * 256, val[c], n, sp.lo);
        const T hi = lut_percentile<T>(h + c * 256, val[c], n, sp.hi);
        s_lo[c] = (double)lo;
        s_den[c] = (double)stretch_den<T>(lo, hi, guard);
    }
    __syncthreads();
    val[c][v] = (double)stretch_apply<T>((T)val[c][v], (T)s_lo[c], (T)s_den[c]);
    __syncthreads();
}

// strategies whose input to the histogram stage is a function of the FRAME byte (no dehaze)
template <class T>
__global__ void __launch_bounds__(768) setup_frame_luts_kernel(SetupArgs a) {
    const int b = blockIdx.x, s = blockIdx.y;
    const Recipe& rc = a.recipes[s];
    const int c = threadIdx.x >> 8, v = threadIdx.x & 255;
    if (rc.dehaze) {
        // dehaze -> stretch -> CLAHE -> gamma (SS:231-249): the tail is a table over the CLAHE
        // output byte; the stretches in front are float-domain and come from k_select.cu
        if (rc.clahe) {
            StratTables& sd = a.tables[(size_t)blockIdx.x * a.S + s];
            const T y = apply_gamma<T>(Ar<T>::div((T)v, (T)255), rc);
            sd.postf[c][v] = (double)y;
            sd.postq[c][v] = (uint8_t)quantise<T>(y);
        }
        return;
    }
    __shared__ double val[3][256];
    __shared__ unsigned hq[3][256];
    __shared__ uint8_t eq[3][256];
    __shared__ double s_lo[3], s_den[3];
    const unsigned* h = a.frame_hist + (size_t)b * 768;
    const CastTables& ct = a.cast[b];
    StratTables& st = a.tables[(size_t)b * a.S + s];
    val[c][v] = (double)ct.curf[c][v];
    hq[c][v] = 0;
    __syncthreads();
    if (rc.equalize) {
        atomicAdd(&hq[c][ct.curq[c][v]], h[c * 256 + v]);
        __syncthreads();
        if (v == 0) equalize_lut(hq[c], a.n, eq[c]);
        __syncthreads();
        val[c][v] = (double)Ar<T>::div((T)eq[c][ct.curq[c][v]], (T)255);
        __syncthreads();
    }
    for (int i = 0; i < rc.n_pre; ++i) chain_stretch<T>(h, val, a.n, rc.pre[i], a.lib.stretch_guard, s_lo, s_den);
    if (rc.clahe) {
        // with no stretch in front, the byte is the float32 frame's own truncation (ES:298, SS:204)
        st.preq[c][v] = rc.n_pre == 0 ? ct.curq[c][v] : (uint8_t)quantise<T>((T)val[c][v]);
        if (rc.n_post == 0) {   // tables over the CLAHE output byte need no histogram
            const T y = apply_gamma<T>(Ar<T>::div((T)v, (T)255), rc);
            st.postf[c][v] = (double)y;
            st.postq[c][v] = (uint8_t)quantise<T>(y);
        }
    } else {
        for (int i = 0; i < rc.n_post; ++i) chain_stretch<T>(h, val, a.n, rc.post[i], a.lib.stretch_guard, s_lo, s_den);
        const T y = apply_gamma<T>((T)val[c][v], rc);
        st.postf[c][v] = (double)y;
        st.postq[c][v] = (uint8_t)quantise<T>(y);
    }
}

// CLAHE strategies with post-stretches: tables over the CLAHE OUTPUT byte, from its histograms
template <class T>
__global__ void __launch_bounds__(768) setup_post_luts_kernel(SetupArgs a) {
    const int b = blockIdx.x, s = blockIdx.y;
    const Recipe& rc = a.recipes[s];
    if (!rc.clahe || rc.n_post == 0) return;
    const int c = threadIdx.x >> 8, v = threadIdx.x & 255;
    __shared__ double val[3][256];
    __shared__ double s_lo[3], s_den[3];
    const unsigned* h = a.out_hist + ((size_t)b * a.S + s) * 768;
    StratTables& st = a.tables[(size_t)b * a.S + s];
    val[c][v] = (double)Ar<T>::div((T)v, (T)255);
    __syncthreads();
    for (int i = 0; i < rc.n_post; ++i) chain_stretch<T>(h, val, a.n, rc.post[i], a.lib.stretch_guard, s_lo, s_den);
    const T y = apply_gamma<T>((T)val[c][v], rc);
    st.postf[c][v] = (double)y;
    st.postq[c][v] = (uint8_t)quantise<T>(y);
}

}  // namespace

int launch_setup_cast(const SetupArgs& a, cudaStream_t st) {
    setup_cast_kernel<<<a.B, 256, 0, st>>>(a.frame_hist, a.n, a.lib.colour_cast, a.types, a.cast);
    return 1;
}
int launch_setup_frame_luts(const SetupArgs& a, cudaStream_t st) {
    if (a.lib.f64) setup_frame_luts_kernel<double><<<dim3(a.B, a.S), 768, 0, st>>>(a);
    else setup_frame_luts_kernel<float><<<dim3(a.B, a.S), 768, 0, st>>>(a);
    return 1;
}
int launch_setup_post_luts(const SetupArgs& a, cudaStream_t st) {
    if (a.lib.f64) setup_post_luts_kernel<double><<<dim3(a.B, a.S), 768, 0, st>>>(a);
    else setup_post_luts_kernel<float><<<dim3(a.B, a.S), 768, 0, st>>>(a);
    return 1;
}

}  // namespace uwie
