// 8-bit integer colour arithmetic shared by the CUDA kernels and the host-side table builder.
//
// Everything here restates the fixed-point paths OpenCV 4.x takes for 8-bit images, which is
// what the reference calls through cv2.cvtColor at
//   enhancement_strategies.py:180,228,299,306   six_stadigy.py:149,177,204,207
//   quality_assessment.py:25,88,112,136,198
// (formulas: SURVEY.md appendix A.3-A.6).  The functions are __host__ __device__ so that the
// CPU-only test suite can check them exhaustively against cv2 without a GPU
// (tests/test_host_color.py through csrc/hostcheck.cpp); the product only ever calls them
// from kernels.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define UW_HD __host__ __device__ __forceinline__
#else
#define UW_HD inline
#endif

namespace uwie {

// Tables built once on the host (uwie_tables.cpp) and copied to device memory.
struct ColorTables {
    uint16_t gamma_in[256];    // round(2040 * srgb_to_linear(i/255))
    uint16_t cbrt_tab[2048 + 8];  // round(32768 * f(k/2040)), k <= 2040, two entries patched
    uint8_t  gamma_out[4096];  // round(255 * linear_to_srgb(i/4096))
    int32_t  sdiv[256];        // round((255 << 12) / i)
    int16_t  lab_yf[512];      // [2L] = y, [2L+1] = ify for Lab -> RGB
};

UW_HD int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// RGB -> GRAY, Q15
UW_HD int gray_q15(int r, int g, int b) {
    return (r * 9798 + g * 19235 + b * 3735 + 16384) >> 15;
}

// RGB -> HSV saturation channel only; V = max, S = (d * sdiv[V] + 2048) >> 12
UW_HD int hsv_saturation(const int32_t* sdiv, int r, int g, int b) {
    int v = r > g ? r : g; v = v > b ? v : b;
    int m = r < g ? r : g; m = m < b ? m : b;
    return ((v - m) * sdiv[v] + 2048) >> 12;
}

// RGB -> Lab 8-bit
UW_HD void rgb_to_lab(const uint16_t* gamma_in, const uint16_t* cbrt_tab,
                      int r, int g, int b, int& L, int& A, int& B) {
    const int lr = gamma_in[r], lg = gamma_in[g], lb = gamma_in[b];
    const int fx = cbrt_tab[(lr * 1777 + lg * 1541 + lb * 778 + 2048) >> 12];
    const int fy = cbrt_tab[(lr * 871 + lg * 2929 + lb * 296 + 2048) >> 12];
    const int fz = cbrt_tab[(lr * 73 + lg * 448 + lb * 3575 + 2048) >> 12];
    L = clampi((296 * fy - 1336934 + 16384) >> 15, 0, 255);
    A = clampi((500 * (fx - fy) + 128 * 32768 + 16384) >> 15, 0, 255);
    B = clampi((200 * (fy - fz) + 128 * 32768 + 16384) >> 15, 0, 255);
}

UW_HD int lab_L_only(const uint16_t* gamma_in, const uint16_t* cbrt_tab, int r, int g, int b) {
    const int fy = cbrt_tab[(gamma_in[r] * 871 + gamma_in[g] * 2929 + gamma_in[b] * 296 + 2048) >> 12];
    return clampi((296 * fy - 1336934 + 16384) >> 15, 0, 255);
}

// inverse of f() on OpenCV's 14-bit grid, C truncating division at every '/'
UW_HD int lab_finv(int i) {
    const int BASE = 16384;
    if (i <= 3390) return i * 108 / 841 - BASE * 16 / 116 * 108 / 841;
    return ((i * i / BASE) * i) / BASE;
}

// Lab -> RGB 8-bit
UW_HD void lab_to_rgb(const int16_t* lab_yf, const uint8_t* gamma_out,
                      int L, int A, int B, int& r, int& g, int& b) {
    const int BASE = 16384;
    const int y = lab_yf[2 * L], ify = lab_yf[2 * L + 1];
    const int adiv = ((5 * A * 53687 + 128) >> 13) - 128 * BASE / 500;
    const int bdiv = ((B * 41943 + 16) >> 9) - 128 * BASE / 200 + 1;
    const int x = lab_finv(ify + adiv);
    const int z = lab_finv(ify - bdiv);
    const int vr = clampi((12615 * x - 6296 * y - 2223 * z + 8192) >> 14, 0, 4095);
    const int vg = clampi((-3773 * x + 7684 * y + 185 * z + 8192) >> 14, 0, 4095);
    const int vb = clampi((217 * x - 836 * y + 4715 * z + 8192) >> 14, 0, 4095);
    r = gamma_out[vr]; g = gamma_out[vg]; b = gamma_out[vb];
}

}  // namespace uwie
