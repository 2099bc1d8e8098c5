// The pieces of the VGG-guided path that are per-pixel work; the network itself
// stays on PyTorch.
//   DifferentiableEnhancement.forward / color_stretch_batch / dehaze_batch / gamma_correction
//                                      vgg_16_UIE.py:32-128
//   extract_all_features (18 statistics zero-padded to 79)         vgg_16_UIE.py:435-466
//   EnhancementPredictor._preprocess_for_vgg (8-bit bilinear resize + ImageNet normalisation)
//                                      use_trained_model.py:39-46
// C ABI: include/uwie.h (uwie_vgg_*).  Stateless: the caller owns the workspace.
//
// color_stretch_batch takes sorted[int(L/100 * n)] of every (image, channel) plane: an exact order
// statistic, found here by a three-pass most-significant-digit radix selection on the sortable
// bit pattern of the float32 values (11 / 11 / 10 bits, shared-memory histograms) instead of a sort.
#include <cuda_runtime.h>

#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/uwie.h"
#include "k_arith.cuh"
#include "k_common.cuh"
#include "k_kernels.h"

namespace uwie {
int uwie_fail(int code, const std::string& msg);   // uwie_engine.cu
}
using namespace uwie;

namespace {

constexpr int NT = 256;
constexpr int kBinsAB = 2048, kBinsC = 1024;
constexpr int kTargets = 2;                 // low / high order statistic of a plane
constexpr int kFeat = 79;                   // vgg_16_UIE.py:462-466 pads to 79

struct SelState {                           // per (plane, target)
    long long rank;                         // rank still to resolve inside the current prefix
    unsigned prefix;                        // key bits resolved so far
    float value;                            // the order statistic, once all 32 bits are known
};

// monotone map float32 -> uint32 (negative values below positive ones, -0 < +0)
__device__ __forceinline__ unsigned sort_key(float x) {
    const unsigned u = __float_as_uint(x);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_value(unsigned k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// PASS 0: histogram of key >> 21 (shared by both targets)
// PASS 1: per target, keys with key >> 21 == prefix: histogram of (key >> 10) & 2047
// PASS 2: per target, keys with key >> 10 == prefix: histogram of key & 1023
// PASS 0 also asks whether the plane is made of bytes - at most one distinct float32 per byte bin round(255 x), which is
// what imread / 255 (use_trained_model.py:119) and every loader of the reference produce, whichever way the division
// was rounded - and counts the bins: then the order statistics follow from 256 bins at once (vgg_sel_scan_kernel<0> marks
// the targets resolved, rank = -1) and passes 1 and 2 return in their first instruction: the plane is read twice (here
// and by the enhancement), not four times.  x -> round(255 x) is monotone, so the bins are in value order.
constexpr int kByteHist = 264;              // 256 counts + [256] = "two different values met in one bin, or a value outside" + padding
constexpr int kByteRows = 2;                // two blocks of [planes][kByteHist]: the counts (cleared to 0), then the bins' float32 bit patterns (cleared to kNoValue)
constexpr unsigned kNoValue = 0xffffffffu;  // not a value this path accepts (a NaN): what memset 0xFF leaves
template <int PASS>
__global__ void __launch_bounds__(NT) vgg_sel_hist_kernel(const float* __restrict__ data, long long n, long long plane_stride,
                                                          int elem_stride, const SelState* __restrict__ state,
                                                          unsigned* __restrict__ hist, unsigned* __restrict__ bhist) {
    constexpr int BINS = PASS == 2 ? kBinsC : kBinsAB;
    constexpr int NH = PASS == 0 ? 1 : kTargets;
    __shared__ unsigned sh[NH][BINS];
    __shared__ unsigned sb[PASS == 0 ? 256 : 1];
    __shared__ unsigned s_rep[PASS == 0 ? 256 : 1];
    const int plane = blockIdx.y;
    if (PASS > 0 && state[plane * kTargets].rank < 0 && state[plane * kTargets + 1].rank < 0) return;   // resolved from the bytes
    for (int i = threadIdx.x; i < NH * BINS; i += NT) (&sh[0][0])[i] = 0;
    if (PASS == 0)
        for (int i = threadIdx.x; i < 256; i += NT) { sb[i] = 0; s_rep[i] = kNoValue; }
    bool odd = false;                           // this thread met a value that is not a byte
    unsigned pre[kTargets] = {0, 0};
    if (PASS > 0) {
#pragma unroll
        for (int t = 0; t < kTargets; ++t) pre[t] = state[plane * kTargets + t].prefix;
    }
    __syncthreads();
    const float* p = data + (size_t)plane * plane_stride;
    auto visit = [&](float x) {
        const unsigned k = sort_key(x);
        if (PASS == 0) {
            atomicAdd(&sh[0][k >> 21], 1u);
            const int v = __float2int_rn(__fmul_rn(x, 255.0f));
            const unsigned bits = __float_as_uint(x);
            if ((unsigned)v <= 255u) {
                unsigned rep = s_rep[v];
                if (rep == kNoValue) { rep = atomicCAS(&s_rep[v], kNoValue, bits); if (rep == kNoValue) rep = bits; }
                if (rep == bits) atomicAdd(&sb[v], 1u);
                else odd = true;
            } else {
                odd = true;
            }
        } else if (PASS == 1) {
#pragma unroll
            for (int t = 0; t < kTargets; ++t) if ((k >> 21) == pre[t]) atomicAdd(&sh[t][(k >> 10) & 2047u], 1u);
        } else {
#pragma unroll
            for (int t = 0; t < kTargets; ++t) if ((k >> 10) == pre[t]) atomicAdd(&sh[t][k & 1023u], 1u);
        }
    };
    if (elem_stride == 1 && (n & 3) == 0 && (plane_stride & 3) == 0 && (reinterpret_cast<uintptr_t>(data) & 15) == 0) {
        const float4* q = reinterpret_cast<const float4*>(p);
        for (long long i = (long long)blockIdx.x * NT + threadIdx.x; i < n / 4; i += (long long)gridDim.x * NT) {
            const float4 v = __ldg(q + i);
            visit(v.x); visit(v.y); visit(v.z); visit(v.w);
        }
    } else {
        for (long long i = (long long)blockIdx.x * NT + threadIdx.x; i < n; i += (long long)gridDim.x * NT)
            visit(p[i * elem_stride]);
    }
    int any_odd = __syncthreads_or(odd);
    unsigned* gh = hist + (size_t)plane * NH * BINS;
    for (int i = threadIdx.x; i < NH * BINS; i += NT) {
        const unsigned v = (&sh[0][0])[i];
        if (v) atomicAdd(&gh[i], v);
    }
    if (PASS == 0) {
        unsigned* gb = bhist + (size_t)plane * kByteHist;
        unsigned* grep = bhist + ((size_t)gridDim.y + plane) * kByteHist;
        bool clash = false;
        for (int i = threadIdx.x; i < 256; i += NT)
            if (sb[i]) {
                atomicAdd(&gb[i], sb[i]);
                const unsigned old = atomicCAS(&grep[i], kNoValue, s_rep[i]);          // the plane's value of this bin
                if (old != kNoValue && old != s_rep[i]) clash = true;
            }
        any_odd |= __syncthreads_or(clash);
        if (any_odd && threadIdx.x == 0) atomicOr(&gb[256], 1u);
    }
}

// one block per plane: walk the histogram of each target to the bin holding its rank
template <int PASS>
__global__ void __launch_bounds__(NT) vgg_sel_scan_kernel(SelState* state, unsigned* hist, unsigned* bhist) {
    constexpr int BINS = PASS == 2 ? kBinsC : kBinsAB;
    constexpr int NH = PASS == 0 ? 1 : kTargets;
    constexpr int PER = BINS / NT;
    const int plane = blockIdx.x;
    __shared__ unsigned long long part[NT];
    for (int t = 0; t < kTargets; ++t) {
        unsigned* h = hist + ((size_t)plane * NH + (PASS == 0 ? 0 : t)) * BINS;
        unsigned long long s = 0;
        for (int i = 0; i < PER; ++i) s += h[threadIdx.x * PER + i];
        part[threadIdx.x] = s;
        __syncthreads();
        if (threadIdx.x == 0 && state[plane * kTargets + t].rank >= 0) {
            SelState st = state[plane * kTargets + t];
            unsigned long long acc = 0;
            int seg = 0;
            while (seg < NT - 1 && acc + part[seg] <= (unsigned long long)st.rank) acc += part[seg++];
            int bin = seg * PER;
            while (bin < seg * PER + PER - 1 && acc + h[bin] <= (unsigned long long)st.rank) acc += h[bin++];
            st.rank -= (long long)acc;
            st.prefix = PASS == 0 ? (unsigned)bin : ((st.prefix << (PASS == 1 ? 11 : 10)) | (unsigned)bin);
            if (PASS == 2) st.value = key_value(st.prefix);
            if (PASS == 0 && bhist[(size_t)plane * kByteHist + 256] == 0) {
                // a plane of bytes: the order statistic is the value of the bin that holds the rank
                const unsigned* hb = bhist + (size_t)plane * kByteHist;
                const unsigned* hrep = bhist + ((size_t)gridDim.x + plane) * kByteHist;
                const long long rank = st.rank + (long long)acc;           // the rank in the whole plane
                long long c = 0;
                int v = 0;
                for (; v < 255; ++v) { c += hb[v]; if (c > rank) break; }
                st.value = __uint_as_float(hrep[v]);
                st.rank = -1;
            }
            state[plane * kTargets + t] = st;
        }
        __syncthreads();
    }
    // leave the histograms clean for the next pass / call
    for (int i = threadIdx.x; i < NH * BINS; i += NT) hist[(size_t)plane * NH * BINS + i] = 0;
}

struct EnhanceParams {       // per image, device copy
    float omega, gamma;
};

__device__ __forceinline__ float clamp01(float x) { return fminf(fmaxf(x, 0.f), 1.f); }

// (x + 1e-8) ** gamma for x in [0, 1] (VGG:119-128) as 2^(gamma * log2(x + 1e-8)) on the special-function unit: three
// instructions instead of powf's ~45.  The operand is a normal number in [1e-8, 1], so the flush-to-zero forms lose
// nothing.  Error: lg2.approx is within 2^-22 absolute on [0.5, 2] and 2 ulp elsewhere, ex2.approx within 2 ulp, which
// bounds the result's absolute error by 0.69 gamma 2^-22 + 2^-22 <= 6.2e-7 for gamma <= 3 (the range of
// use_trained_model.py:74-79 and of the network's head) -- the tests hold it to 1e-6 against torch's pow, three orders
// below north_star's 1/255.  x + 1e-8 == 1 gives exactly 1.
__device__ __forceinline__ float pow_unit(float x, float gamma) {
    float l, y;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(__fadd_rn(x, 1e-8f)));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(__fmul_rn(gamma, l)));
    return y;
}

// stretch of one value, rounded like the tensor ops (VGG:88-90)
__device__ __forceinline__ float stretch_value(float x, float lo, float rng) { return clamp01(__fdiv_rn(__fsub_rn(x, lo), rng)); }

// dehaze -> gamma -> clamp on one stretched pixel (r, g, b in [0, 1]); every step rounds to float32 like the tensor ops.
// The three divisions by t share one correctly rounded reciprocal (k_arith.cuh InvDiv: bit-identical to __fdiv_rn for
// |x| <= 1 and a divisor in [1e-6, 1]; here x = c - 0.6 and t is in [0.1, 1]; uwie_selftest_division checks the range).
__device__ __forceinline__ void dehaze_gamma_pixel(float& r, float& g, float& b, bool dehaze, float omega, bool gamma_on, float gamma) {
    if (dehaze) {                                              // VGG:94-117, A = 0.6
        const float d = fminf(fminf(r, g), b);
        float t = __fsub_rn(1.0f, __fmul_rn(omega, d));
        t = fminf(fmaxf(t, 0.1f), 1.0f);
        const float rt = InvDiv<float>::rcp(t);
        r = clamp01(__fadd_rn(InvDiv<float>::div(__fsub_rn(r, 0.6f), t, rt), 0.6f));
        g = clamp01(__fadd_rn(InvDiv<float>::div(__fsub_rn(g, 0.6f), t, rt), 0.6f));
        b = clamp01(__fadd_rn(InvDiv<float>::div(__fsub_rn(b, 0.6f), t, rt), 0.6f));
    }
    if (gamma_on) {                                            // VGG:119-128, then the clamp of VGG:55
        r = clamp01(pow_unit(r, gamma));
        g = clamp01(pow_unit(g, gamma));
        b = clamp01(pow_unit(b, gamma));
    }
}

// stretch -> dehaze -> gamma -> clamp on one pixel of a float image
__device__ __forceinline__ void enhance_pixel(float& r, float& g, float& b, const float* lo, const float* rng,
                                              bool dehaze, float omega, bool gamma_on, float gamma) {
    r = stretch_value(r, lo[0], rng[0]);
    g = stretch_value(g, lo[1], rng[1]);
    b = stretch_value(b, lo[2], rng[2]);
    dehaze_gamma_pixel(r, g, b, dehaze, omega, gamma_on, gamma);
}

__global__ void __launch_bounds__(NT) vgg_enhance_kernel(const float* __restrict__ img, float* __restrict__ out, long long n,
                                                         const SelState* __restrict__ state,
                                                         const EnhanceParams* __restrict__ prm, int dehaze, int gamma_on) {
    const int b = blockIdx.y;
    __shared__ float lo[3], rng[3];
    if (threadIdx.x < 3) {
        const float p_lo = state[(b * 3 + threadIdx.x) * kTargets + 0].value;
        const float p_hi = state[(b * 3 + threadIdx.x) * kTargets + 1].value;
        lo[threadIdx.x] = p_lo;
        rng[threadIdx.x] = __fadd_rn(__fsub_rn(p_hi, p_lo), 1e-8f);     // VGG:87
    }
    __syncthreads();
    const float omega = prm[b].omega, gamma = prm[b].gamma;
    const float* p = img + (size_t)b * 3 * n;
    float* o = out + (size_t)b * 3 * n;
    if ((n & 3) == 0 && ((reinterpret_cast<uintptr_t>(img) | reinterpret_cast<uintptr_t>(out)) & 15) == 0) {
        const float4* pr = reinterpret_cast<const float4*>(p);
        const float4* pg = reinterpret_cast<const float4*>(p + n);
        const float4* pb = reinterpret_cast<const float4*>(p + 2 * n);
        float4* qr = reinterpret_cast<float4*>(o);
        float4* qg = reinterpret_cast<float4*>(o + n);
        float4* qb = reinterpret_cast<float4*>(o + 2 * n);
        for (long long i = (long long)blockIdx.x * NT + threadIdx.x; i < n / 4; i += (long long)gridDim.x * NT) {
            float4 r = __ldg(pr + i), g = __ldg(pg + i), bl = __ldg(pb + i);
            enhance_pixel(r.x, g.x, bl.x, lo, rng, dehaze, omega, gamma_on, gamma);
            enhance_pixel(r.y, g.y, bl.y, lo, rng, dehaze, omega, gamma_on, gamma);
            enhance_pixel(r.z, g.z, bl.z, lo, rng, dehaze, omega, gamma_on, gamma);
            enhance_pixel(r.w, g.w, bl.w, lo, rng, dehaze, omega, gamma_on, gamma);
            qr[i] = r; qg[i] = g; qb[i] = bl;
        }
    } else {
        for (long long i = (long long)blockIdx.x * NT + threadIdx.x; i < n; i += (long long)gridDim.x * NT) {
            float r = p[i], g = p[n + i], bl = p[2 * n + i];
            enhance_pixel(r, g, bl, lo, rng, dehaze, omega, gamma_on, gamma);
            o[i] = r; o[n + i] = g; o[2 * n + i] = bl;
        }
    }
}

// ---------------------------------------------------------------- 18 statistics from the byte histograms
// img = frame / 255 in float32 (vgg_16_UIE.py:425-433); all statistics are functions of the three
// 256-bin histograms.  Sums are taken in float64 and rounded once (NumPy sums float32 pairwise; the
// difference is below 1e-6 relative, inside the 1e-4 tolerance of the float stages).
__global__ void __launch_bounds__(256) vgg_features_kernel(const unsigned* __restrict__ hist, long long n, float* __restrict__ feat) {
    const int b = blockIdx.x;
    const unsigned* h = hist + (size_t)b * 768;
    float* f = feat + (size_t)b * kFeat;
    __shared__ double s1[3], s2[3];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < kFeat; i += 256) f[i] = 0.f;
    if (w < 3) {
        double a = 0, q = 0;
        for (int v = lane; v < 256; v += 32) {
            const float x = __fdiv_rn((float)v, 255.0f);
            const double c = (double)h[w * 256 + v];
            a += c * (double)x;
            q += c * (double)__fmul_rn(x, x);          // img ** 2 is a float32 array (VGG:458)
        }
        a = warp_sum_f64(a); q = warp_sum_f64(q);
        if (lane == 0) { s1[w] = a; s2[w] = q; }
    }
    __syncthreads();
    if (w < 3) {
        const double mean = s1[w] / (double)n;
        double var = 0;
        for (int v = lane; v < 256; v += 32) {
            const double d = (double)__fdiv_rn((float)v, 255.0f) - mean;
            var += (double)h[w * 256 + v] * d * d;
        }
        var = warp_sum_f64(var);
        if (lane == 0) {
            int vmin = 0, vmax = 255;
            while (vmin < 255 && h[w * 256 + vmin] == 0) ++vmin;
            while (vmax > 0 && h[w * 256 + vmax] == 0) --vmax;
            // np.median: mean of the two middle order statistics (the same one when n is odd)
            const long long k0 = (n - 1) / 2, k1 = n / 2;
            long long acc = 0; int m0 = -1, m1 = -1;
            for (int v = 0; v < 256 && m1 < 0; ++v) {
                acc += h[w * 256 + v];
                if (m0 < 0 && acc > k0) m0 = v;
                if (acc > k1) m1 = v;
            }
            const float x0 = __fdiv_rn((float)m0, 255.0f), x1 = __fdiv_rn((float)m1, 255.0f);
            f[w * 5 + 0] = (float)mean;
            f[w * 5 + 1] = (float)sqrt(var / (double)n);
            f[w * 5 + 2] = __fdiv_rn((float)vmin, 255.0f);
            f[w * 5 + 3] = __fdiv_rn((float)vmax, 255.0f);
            f[w * 5 + 4] = m0 == m1 ? x0 : __fdiv_rn(__fadd_rn(x0, x1), 2.0f);
        }
    }
    __syncthreads();
    if (w == 3) {                                      // overall mean, std, second moment (VGG:455-459)
        const double tot = 3.0 * (double)n;
        const double mean = (s1[0] + s1[1] + s1[2]) / tot;
        double var = 0;
        for (int i = lane; i < 768; i += 32) {
            const double d = (double)__fdiv_rn((float)(i & 255), 255.0f) - mean;
            var += (double)h[i] * d * d;
        }
        var = warp_sum_f64(var);
        if (lane == 0) {
            f[15] = (float)mean;
            f[16] = (float)sqrt(var / tot);
            f[17] = (float)((s2[0] + s2[1] + s2[2]) / tot);
        }
    }
}

// ---------------------------------------------------------------- 224 x 224 network input
// cv2.resize(uint8, INTER_LINEAR) in OpenCV's 11-bit fixed point, / 255, ImageNet
// normalisation (torchvision Normalize: (x - mean) / std in float32), HWC -> CHW.
struct ResizeTap { int s0, s1; int c1; };
// horizontal taps: a source index outside the row is clamped AND its fraction zeroed; vertical taps keep
// the fraction and only clamp the two row indices (cv2's resize treats the two directions differently,
// which shows in the first / last rows of an upscale because the two products are truncated separately)
__device__ __forceinline__ ResizeTap resize_tap(int i, double scale, int n, bool vertical) {
    float f = (float)(((double)i + 0.5) * scale - 0.5);
    int s = (int)floorf(f);
    f -= (float)s;
    ResizeTap t;
    if (vertical) {
        t.s0 = min(max(s, 0), n - 1); t.s1 = min(max(s + 1, 0), n - 1);
    } else {
        if (s < 0) { s = 0; f = 0.f; }
        if (s >= n - 1) { s = n - 1; f = 0.f; }
        t.s0 = s; t.s1 = min(s + 1, n - 1);
    }
    t.c1 = __float2int_rn(f * 2048.0f);
    return t;
}

template <class In>
__device__ __forceinline__ int load_byte(const In* p);
template <> __device__ __forceinline__ int load_byte<uint8_t>(const uint8_t* p) { return *p; }
template <> __device__ __forceinline__ int load_byte<float>(const float* p) { return (int)(uint8_t)(int)__fmul_rn(*p, 255.0f); }   // (img * 255).astype(np.uint8)

template <class In>
__global__ void __launch_bounds__(NT) vgg_preprocess_kernel(const In* __restrict__ frames, int H, int W, int size,
                                                            float* __restrict__ out) {
    const int b = blockIdx.y;
    const int i = blockIdx.x * NT + threadIdx.x;
    if (i >= size * size) return;
    const int oy = i / size, ox = i - oy * size;
    if (H == 2 * size && W == 2 * size) {
        // cv2.resize turns INTER_LINEAR into INTER_AREA for an exact 2x decimation: rounded 2x2 mean
        const In* f2 = frames + (size_t)b * H * W * 3;
        const float mean2[3] = {0.485f, 0.456f, 0.406f}, std2[3] = {0.229f, 0.224f, 0.225f};
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const In* q = f2 + ((size_t)(2 * oy) * W + 2 * ox) * 3 + c;
            const int v = (load_byte<In>(q) + load_byte<In>(q + 3) + load_byte<In>(q + (size_t)W * 3) + load_byte<In>(q + (size_t)W * 3 + 3) + 2) >> 2;
            out[((size_t)b * 3 + c) * size * size + i] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)v, 255.0f), mean2[c]), std2[c]);
        }
        return;
    }
    const ResizeTap ty = resize_tap(oy, (double)H / (double)size, H, true);
    const ResizeTap tx = resize_tap(ox, (double)W / (double)size, W, false);
    const In* f = frames + (size_t)b * H * W * 3;
    const float mean[3] = {0.485f, 0.456f, 0.406f}, stdv[3] = {0.229f, 0.224f, 0.225f};
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const int p00 = load_byte<In>(f + ((size_t)ty.s0 * W + tx.s0) * 3 + c), p01 = load_byte<In>(f + ((size_t)ty.s0 * W + tx.s1) * 3 + c);
        const int p10 = load_byte<In>(f + ((size_t)ty.s1 * W + tx.s0) * 3 + c), p11 = load_byte<In>(f + ((size_t)ty.s1 * W + tx.s1) * 3 + c);
        const int s0 = p00 * (2048 - tx.c1) + p01 * tx.c1;
        const int s1 = p10 * (2048 - tx.c1) + p11 * tx.c1;
        const int v = ((((2048 - ty.c1) * (s0 >> 4)) >> 16) + ((ty.c1 * (s1 >> 4)) >> 16) + 2) >> 2;
        const float x = __fdiv_rn((float)min(max(v, 0), 255), 255.0f);
        out[((size_t)b * 3 + c) * size * size + i] = __fdiv_rn(__fsub_rn(x, mean[c]), stdv[c]);
    }
}

struct VggLayout { size_t state, prm, hist, bhist, hist_bytes, bhist_bytes, total; };
VggLayout vgg_layout(int batch) {
    VggLayout L;
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t at = off; off += (bytes + 255) & ~(size_t)255; return at; };
    L.state = take((size_t)batch * 3 * kTargets * sizeof(SelState));
    L.prm = take((size_t)batch * sizeof(EnhanceParams));
    // the widest use: pass 1 of the selection (planes x targets x 2048 bins); features need batch x 768
    L.hist = take((size_t)batch * 3 * kTargets * kBinsAB * 4);
    L.hist_bytes = (size_t)batch * 3 * kTargets * kBinsAB * 4;
    L.bhist_bytes = (size_t)batch * 3 * kByteRows * kByteHist * 4;
    L.bhist = take(L.bhist_bytes);
    L.total = off;
    return L;
}

#define CUV(call)                                                                                  \
    do {                                                                                           \
        cudaError_t _e = (call);                                                                   \
        if (_e != cudaSuccess) return uwie_fail(UWIE_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(_e)); \
    } while (0)

int grid_for(long long items) {
    long long g = (items + NT * 8 - 1) / (NT * 8);
    if (g < 1) g = 1;
    if (g > 148 * 8) g = 148 * 8;
    return (int)g;
}

// ---------------------------------------------------------------- backward of the enhancement (training-time use)
// d loss / d omega and d loss / d gamma of out = clamp(pow(clamp((s - .6) / clamp(1 - omega * min_c s, .1, 1) + .6, 0, 1) + 1e-8,
// gamma), 0, 1) with s the stretched pixel: what autograd derives for vgg_16_UIE.py:32-128.  L_low / L_high receive no gradient
// in the reference either (they pass through .item(), VGG:78-79).  torch.clamp passes the gradient on the closed interval.
__global__ void __launch_bounds__(NT) vgg_enhance_bwd_kernel(const float* __restrict__ img, const float* __restrict__ gout, long long n,
                                                             const SelState* __restrict__ state, const EnhanceParams* __restrict__ prm,
                                                             double* __restrict__ gparams /* [2][batch] */, int batch) {
    const int b = blockIdx.y;
    __shared__ float lo[3], rng[3];
    if (threadIdx.x < 3) {
        const float p_lo = state[(b * 3 + threadIdx.x) * kTargets + 0].value;
        const float p_hi = state[(b * 3 + threadIdx.x) * kTargets + 1].value;
        lo[threadIdx.x] = p_lo;
        rng[threadIdx.x] = __fadd_rn(__fsub_rn(p_hi, p_lo), 1e-8f);
    }
    __syncthreads();
    const float omega = prm[b].omega, gamma = prm[b].gamma;
    const float* p = img + (size_t)b * 3 * n;
    const float* g = gout + (size_t)b * 3 * n;
    double g_om = 0.0, g_ga = 0.0;
    for (long long i = (long long)blockIdx.x * NT + threadIdx.x; i < n; i += (long long)gridDim.x * NT) {
        float s[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) s[c] = clamp01(__fdiv_rn(__fsub_rn(p[c * n + i], lo[c]), rng[c]));
        const float d = fminf(fminf(s[0], s[1]), s[2]);
        const float traw = __fsub_rn(1.0f, __fmul_rn(omega, d));
        const float t = fminf(fmaxf(traw, 0.1f), 1.0f);
        float gt = 0.f;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float zraw = __fadd_rn(__fdiv_rn(__fsub_rn(s[c], 0.6f), t), 0.6f);
            const float z = clamp01(zraw);
            const float base = __fadd_rn(z, 1e-8f);
            const float y = powf(base, gamma);
            float gy = g[c * n + i];
            if (!(y >= 0.f && y <= 1.f)) gy = 0.f;                       // final clamp
            g_ga += (double)(gy * y * logf(base));
            float gz = gy * gamma * powf(base, gamma - 1.0f);
            if (!(zraw >= 0.f && zraw <= 1.f)) gz = 0.f;                  // clamp after the recovery
            gt += gz * (-(s[c] - 0.6f) / (t * t));
        }
        if (traw >= 0.1f && traw <= 1.0f) g_om += (double)(gt * (-d));    // clamp of t
    }
    __shared__ double red[NT / 32][2];
    g_om = warp_sum_f64(g_om); g_ga = warp_sum_f64(g_ga);
    if ((threadIdx.x & 31) == 0) { red[threadIdx.x >> 5][0] = g_om; red[threadIdx.x >> 5][1] = g_ga; }
    __syncthreads();
    if (threadIdx.x < 2) {
        double v = 0;
        for (int w = 0; w < NT / 32; ++w) v += red[w][threadIdx.x];
        atomicAdd(&gparams[threadIdx.x * batch + b], v);
    }
}

// ---------------------------------------------------------------- the same enhancement straight from the uint8 frames
// use_trained_model.py feeds imread / 255 (UTM:119): the float image takes 256 values per channel, so the order statistics
// sorted[int(L / 100 * n)] follow from the channel histograms and the frame is read twice as BYTES (histogram, apply)
// instead of four times as floats: 3 N + 3 N bytes in, 12 N out.
__global__ void vgg_ranks_u8_kernel(const unsigned* __restrict__ hist /* [B][3][256] */, int batch, SelState* __restrict__ state) {
    const int pc = blockIdx.x * blockDim.x + threadIdx.x;          // plane = b * 3 + c
    if (pc >= batch * 3) return;
    const unsigned* h = hist + (size_t)pc * 256;
    for (int tg = 0; tg < kTargets; ++tg) {
        SelState& st = state[(size_t)pc * kTargets + tg];
        long long acc = 0;
        int v = 0;
        for (; v < 255; ++v) { acc += h[v]; if (acc > st.rank) break; }
        st.value = __fdiv_rn((float)v, 255.0f);                    // img = frame.astype(float32) / 255.0
    }
}

__global__ void __launch_bounds__(NT) vgg_enhance_u8_kernel(const uint8_t* __restrict__ frames, float* __restrict__ out, long long n,
                                                            const SelState* __restrict__ state, const EnhanceParams* __restrict__ prm) {
    const int b = blockIdx.y;
    // the stretched value is a function of the byte: one table per channel, built with the per-pixel formula of the float
    // kernel (so the two entry points agree bit for bit), replaces a subtraction and an IEEE division per value
    __shared__ float lo[3], rng[3], slut[3][256];
    if (threadIdx.x < 3) {
        const float p_lo = state[(b * 3 + threadIdx.x) * kTargets + 0].value;
        const float p_hi = state[(b * 3 + threadIdx.x) * kTargets + 1].value;
        lo[threadIdx.x] = p_lo;
        rng[threadIdx.x] = __fadd_rn(__fsub_rn(p_hi, p_lo), 1e-8f);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 768; i += NT) {
        const int c = i >> 8;
        slut[c][i & 255] = stretch_value(__fdiv_rn((float)(i & 255), 255.0f), lo[c], rng[c]);
    }
    __syncthreads();
    const float omega = prm[b].omega, gamma = prm[b].gamma;
    const uint8_t* f = frames + (size_t)b * 3 * n;
    float* o = out + (size_t)b * 3 * n;
    const float* lr = slut[0];
    const float* lg = slut[1];
    const float* lb = slut[2];
    const bool vec = (n & 3) == 0 && (reinterpret_cast<uintptr_t>(f) & 3) == 0 && (reinterpret_cast<uintptr_t>(o) & 15) == 0;
    if (vec) {
        const uint32_t* w = reinterpret_cast<const uint32_t*>(f);
        const long long groups = n / 4, step = (long long)gridDim.x * NT;
        auto one = [&](long long g, uint32_t w0, uint32_t w1, uint32_t w2) {
            float4 r, gg, bl;
            r.x = lr[w0 & 255]; gg.x = lg[(w0 >> 8) & 255]; bl.x = lb[(w0 >> 16) & 255];
            r.y = lr[w0 >> 24]; gg.y = lg[w1 & 255]; bl.y = lb[(w1 >> 8) & 255];
            r.z = lr[(w1 >> 16) & 255]; gg.z = lg[w1 >> 24]; bl.z = lb[w2 & 255];
            r.w = lr[(w2 >> 8) & 255]; gg.w = lg[(w2 >> 16) & 255]; bl.w = lb[w2 >> 24];
            dehaze_gamma_pixel(r.x, gg.x, bl.x, true, omega, true, gamma);
            dehaze_gamma_pixel(r.y, gg.y, bl.y, true, omega, true, gamma);
            dehaze_gamma_pixel(r.z, gg.z, bl.z, true, omega, true, gamma);
            dehaze_gamma_pixel(r.w, gg.w, bl.w, true, omega, true, gamma);
            __stcs(reinterpret_cast<float4*>(o) + g, r);
            __stcs(reinterpret_cast<float4*>(o + n) + g, gg);
            __stcs(reinterpret_cast<float4*>(o + 2 * n) + g, bl);
        };
        long long g = (long long)blockIdx.x * NT + threadIdx.x;
        // two groups per trip: six loads in flight per thread before the first table lookup
        for (; g + step < groups; g += 2 * step) {
            const uint32_t a0 = __ldg(w + 3 * g), a1 = __ldg(w + 3 * g + 1), a2 = __ldg(w + 3 * g + 2);
            const long long h = g + step;
            const uint32_t b0 = __ldg(w + 3 * h), b1 = __ldg(w + 3 * h + 1), b2 = __ldg(w + 3 * h + 2);
            one(g, a0, a1, a2);
            one(h, b0, b1, b2);
        }
        if (g < groups) one(g, __ldg(w + 3 * g), __ldg(w + 3 * g + 1), __ldg(w + 3 * g + 2));
    } else {
        for (long long i = (long long)blockIdx.x * NT + threadIdx.x; i < n; i += (long long)gridDim.x * NT) {
            float r = lr[f[3 * i]], g = lg[f[3 * i + 1]], bl = lb[f[3 * i + 2]];
            dehaze_gamma_pixel(r, g, bl, true, omega, true, gamma);
            o[i] = r; o[n + i] = g; o[2 * n + i] = bl;
        }
    }
}

// device-resident parameters (the network's outputs): ranks int(L / 100.0 * n) in double arithmetic (VGG:78-82),
// clamped like use_trained_model.py:74-79 when `clamp_ranges` is set
__global__ void vgg_params_kernel(const float* __restrict__ p /* [4][batch]: L_low, L_high, omega, gamma */, int batch,
                                  long long n, int clamp_ranges, SelState* __restrict__ state, EnhanceParams* __restrict__ prm,
                                  float* __restrict__ used /* [4][batch] or null: the values actually applied */) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= batch) return;
    float llo = p[b], lhi = p[batch + b], om = p[2 * batch + b], ga = p[3 * batch + b];
    if (clamp_ranges) {
        llo = fminf(fmaxf(llo, 1.0f), 30.0f); lhi = fminf(fmaxf(lhi, 65.0f), 99.0f);
        om = fminf(fmaxf(om, 0.1f), 0.9f); ga = fminf(fmaxf(ga, 0.5f), 3.0f);
    }
    long long lo = (long long)(((double)llo / 100.0) * (double)n);
    long long hi = (long long)(((double)lhi / 100.0) * (double)n);
    lo = lo < 0 ? 0 : (lo > n - 1 ? n - 1 : lo);
    hi = hi < 0 ? 0 : (hi > n - 1 ? n - 1 : hi);
    for (int c = 0; c < 3; ++c) {
        state[(size_t)(b * 3 + c) * kTargets + 0] = SelState{lo, 0u, 0.f};
        state[(size_t)(b * 3 + c) * kTargets + 1] = SelState{hi, 0u, 0.f};
    }
    prm[b].omega = om; prm[b].gamma = ga;
    if (used) { used[b] = llo; used[batch + b] = lhi; used[2 * batch + b] = om; used[3 * batch + b] = ga; }
}

}  // namespace

extern "C" {

size_t uwie_vgg_workspace_bytes(int batch) { return batch < 1 ? 0 : vgg_layout(batch).total; }

int uwie_vgg_enhance(const float* img_dev, int batch, int height, int width, const float* l_low, const float* l_high,
                     const float* omega, const float* gamma, float* out_dev, void* workspace_dev,
                     size_t workspace_bytes, void* stream) {
    if (!img_dev || !out_dev || !l_low || !l_high || batch < 1 || height < 1 || width < 1)
        return uwie_fail(UWIE_ERR_ARG, "uwie_vgg_enhance: bad arguments");
    const VggLayout L = vgg_layout(batch);
    if (!workspace_dev || workspace_bytes < L.total) return uwie_fail(UWIE_ERR_ARG, "uwie_vgg_enhance: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)workspace_dev;
    const long long n = (long long)height * width;
    // ranks: int(L / 100.0 * n) clamped to [0, n - 1], in Python float (double) arithmetic (VGG:78-82)
    std::vector<SelState> hs((size_t)batch * 3 * kTargets);
    std::vector<EnhanceParams> hp(batch);
    for (int b = 0; b < batch; ++b) {
        long long lo = (long long)(((double)l_low[b] / 100.0) * (double)n);
        long long hi = (long long)(((double)l_high[b] / 100.0) * (double)n);
        lo = lo < 0 ? 0 : (lo > n - 1 ? n - 1 : lo);
        hi = hi < 0 ? 0 : (hi > n - 1 ? n - 1 : hi);
        for (int c = 0; c < 3; ++c) {
            hs[(size_t)(b * 3 + c) * kTargets + 0] = SelState{lo, 0u, 0.f};
            hs[(size_t)(b * 3 + c) * kTargets + 1] = SelState{hi, 0u, 0.f};
        }
        hp[b].omega = omega ? omega[b] : 0.f;
        hp[b].gamma = gamma ? gamma[b] : 1.f;
    }
    SelState* state = (SelState*)(ws + L.state);
    EnhanceParams* prm = (EnhanceParams*)(ws + L.prm);
    unsigned* hist = (unsigned*)(ws + L.hist);
    unsigned* bhist = (unsigned*)(ws + L.bhist);
    // pageable sources: the runtime stages them before returning, so the vectors may die with this frame
    CUV(cudaMemcpyAsync(state, hs.data(), hs.size() * sizeof(SelState), cudaMemcpyHostToDevice, st));
    CUV(cudaMemcpyAsync(prm, hp.data(), hp.size() * sizeof(EnhanceParams), cudaMemcpyHostToDevice, st));
    CUV(cudaMemsetAsync(hist, 0, L.hist_bytes, st));
    CUV(cudaMemsetAsync(bhist, 0, L.bhist_bytes / 2, st));
    CUV(cudaMemsetAsync(reinterpret_cast<char*>(bhist) + L.bhist_bytes / 2, 0xff, L.bhist_bytes / 2, st));
    const int planes = batch * 3;
    const dim3 grid(grid_for(n / 4 + 1), planes);
    vgg_sel_hist_kernel<0><<<grid, NT, 0, st>>>(img_dev, n, n, 1, state, hist, bhist);
    vgg_sel_scan_kernel<0><<<planes, NT, 0, st>>>(state, hist, bhist);
    vgg_sel_hist_kernel<1><<<grid, NT, 0, st>>>(img_dev, n, n, 1, state, hist, bhist);
    vgg_sel_scan_kernel<1><<<planes, NT, 0, st>>>(state, hist, bhist);
    vgg_sel_hist_kernel<2><<<grid, NT, 0, st>>>(img_dev, n, n, 1, state, hist, bhist);
    vgg_sel_scan_kernel<2><<<planes, NT, 0, st>>>(state, hist, bhist);
    vgg_enhance_kernel<<<dim3(grid_for(n / 4 + 1), batch), NT, 0, st>>>(img_dev, out_dev, n, state, prm, omega ? 1 : 0, gamma ? 1 : 0);
    CUV(cudaGetLastError());
    return UWIE_OK;
}

int uwie_vgg_enhance_dev(const float* img_dev, int batch, int height, int width, const float* params_dev, int clamp_ranges,
                         float* used_dev, float* out_dev, void* workspace_dev, size_t workspace_bytes, void* stream) {
    if (!img_dev || !out_dev || !params_dev || batch < 1 || height < 1 || width < 1)
        return uwie_fail(UWIE_ERR_ARG, "uwie_vgg_enhance_dev: bad arguments");
    const VggLayout L = vgg_layout(batch);
    if (!workspace_dev || workspace_bytes < L.total) return uwie_fail(UWIE_ERR_ARG, "uwie_vgg_enhance_dev: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)workspace_dev;
    const long long n = (long long)height * width;
    SelState* state = (SelState*)(ws + L.state);
    EnhanceParams* prm = (EnhanceParams*)(ws + L.prm);
    unsigned* hist = (unsigned*)(ws + L.hist);
    unsigned* bhist = (unsigned*)(ws + L.bhist);
    vgg_params_kernel<<<(batch + 127) / 128, 128, 0, st>>>(params_dev, batch, n, clamp_ranges, state, prm, used_dev);
    CUV(cudaMemsetAsync(hist, 0, L.hist_bytes, st));
    CUV(cudaMemsetAsync(bhist, 0, L.bhist_bytes / 2, st));
    CUV(cudaMemsetAsync(reinterpret_cast<char*>(bhist) + L.bhist_bytes / 2, 0xff, L.bhist_bytes / 2, st));
    const int planes = batch * 3;
    const dim3 grid(grid_for(n / 4 + 1), planes);
    vgg_sel_hist_kernel<0><<<grid, NT, 0, st>>>(img_dev, n, n, 1, state, hist, bhist);
    vgg_sel_scan_kernel<0><<<planes, NT, 0, st>>>(state, hist, bhist);
    vgg_sel_hist_kernel<1><<<grid, NT, 0, st>>>(img_dev, n, n, 1, state, hist, bhist);
    vgg_sel_scan_kernel<1><<<planes, NT, 0, st>>>(state, hist, bhist);
    vgg_sel_hist_kernel<2><<<grid, NT, 0, st>>>(img_dev, n, n, 1, state, hist, bhist);
    vgg_sel_scan_kernel<2><<<planes, NT, 0, st>>>(state, hist, bhist);
    vgg_enhance_kernel<<<dim3(grid_for(n / 4 + 1), batch), NT, 0, st>>>(img_dev, out_dev, n, state, prm, 1, 1);
    CUV(cudaGetLastError());
    return UWIE_OK;
}

int uwie_vgg_enhance_u8(const uint8_t* frames_dev, int batch, int height, int width, const float* params_dev, int clamp_ranges,
                        float* used_dev, float* out_dev, void* workspace_dev, size_t workspace_bytes, void* stream) {
    if (!frames_dev || !out_dev || !params_dev || batch < 1 || height < 1 || width < 1)
        return uwie_fail(UWIE_ERR_ARG, "uwie_vgg_enhance_u8: bad arguments");
    const VggLayout L = vgg_layout(batch);
    if (!workspace_dev || workspace_bytes < L.total) return uwie_fail(UWIE_ERR_ARG, "uwie_vgg_enhance_u8: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)workspace_dev;
    const long long n = (long long)height * width;
    SelState* state = (SelState*)(ws + L.state);
    EnhanceParams* prm = (EnhanceParams*)(ws + L.prm);
    unsigned* hist = (unsigned*)(ws + L.hist);
    vgg_params_kernel<<<(batch + 127) / 128, 128, 0, st>>>(params_dev, batch, n, clamp_ranges, state, prm, used_dev);
    CUV(cudaMemsetAsync(hist, 0, (size_t)batch * 768 * 4, st));
    launch_frame_hist(frames_dev, batch, height, width, hist, st);
    vgg_ranks_u8_kernel<<<(batch * 3 + 127) / 128, 128, 0, st>>>(hist, batch, state);
    vgg_enhance_u8_kernel<<<dim3(grid_for(n / 4 + 1), batch), NT, 0, st>>>(frames_dev, out_dev, n, state, prm);
    CUV(cudaGetLastError());
    return UWIE_OK;
}

int uwie_vgg_enhance_bwd(const float* img_dev, const float* grad_out_dev, int batch, int height, int width,
                         const void* workspace_dev, size_t workspace_bytes, double* grad_params_dev, void* stream) {
    if (!img_dev || !grad_out_dev || !grad_params_dev || batch < 1 || height < 1 || width < 1)
        return uwie_fail(UWIE_ERR_ARG, "uwie_vgg_enhance_bwd: bad arguments");
    const VggLayout L = vgg_layout(batch);
    if (!workspace_dev || workspace_bytes < L.total) return uwie_fail(UWIE_ERR_ARG, "uwie_vgg_enhance_bwd: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    const char* ws = (const char*)workspace_dev;
    const long long n = (long long)height * width;
    CUV(cudaMemsetAsync(grad_params_dev, 0, sizeof(double) * 2 * batch, st));
    vgg_enhance_bwd_kernel<<<dim3(grid_for(n), batch), NT, 0, st>>>(img_dev, grad_out_dev, n, (const SelState*)(ws + L.state),
                                                                  (const EnhanceParams*)(ws + L.prm), grad_params_dev, batch);
    CUV(cudaGetLastError());
    return UWIE_OK;
}

int uwie_vgg_features_u8(const uint8_t* frames_dev, int batch, int height, int width, float* feat_dev,
                         void* workspace_dev, size_t workspace_bytes, void* stream) {
    if (!frames_dev || !feat_dev || batch < 1 || height < 1 || width < 1)
        return uwie_fail(UWIE_ERR_ARG, "uwie_vgg_features_u8: bad arguments");
    const VggLayout L = vgg_layout(batch);
    if (!workspace_dev || workspace_bytes < L.total) return uwie_fail(UWIE_ERR_ARG, "uwie_vgg_features_u8: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    unsigned* hist = (unsigned*)((char*)workspace_dev + L.hist);
    CUV(cudaMemsetAsync(hist, 0, (size_t)batch * 768 * 4, st));
    launch_frame_hist(frames_dev, batch, height, width, hist, st);
    vgg_features_kernel<<<batch, 256, 0, st>>>(hist, (long long)height * width, feat_dev);
    CUV(cudaGetLastError());
    return UWIE_OK;
}

static int vgg_preprocess_check(const void* in, const void* out, int batch, int height, int width, int size) {
    if (!in || !out || batch < 1 || height < 1 || width < 1 || size < 1)
        return uwie_fail(UWIE_ERR_ARG, "uwie_vgg_preprocess: bad arguments");
    return UWIE_OK;
}

int uwie_vgg_preprocess_u8(const uint8_t* frames_dev, int batch, int height, int width, int size, float* out_dev, void* stream) {
    if (int rc = vgg_preprocess_check(frames_dev, out_dev, batch, height, width, size)) return rc;
    vgg_preprocess_kernel<uint8_t><<<dim3((size * size + NT - 1) / NT, batch), NT, 0, (cudaStream_t)stream>>>(frames_dev, height, width, size, out_dev);
    CUV(cudaGetLastError());
    return UWIE_OK;
}

int uwie_vgg_preprocess_f32(const float* img_dev, int batch, int height, int width, int size, float* out_dev, void* stream) {
    if (int rc = vgg_preprocess_check(img_dev, out_dev, batch, height, width, size)) return rc;
    vgg_preprocess_kernel<float><<<dim3((size * size + NT - 1) / NT, batch), NT, 0, (cudaStream_t)stream>>>(img_dev, height, width, size, out_dev);
    CUV(cudaGetLastError());
    return UWIE_OK;
}

}  // extern "C"
