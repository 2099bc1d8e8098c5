// TEST-ONLY host build of the __host__ __device__ integer arithmetic (uwie_color.h, uwie_hd.h)
// and of the host planners.  Compiled by __graft_entry__.build() into tests/_hostcheck.so and
// loaded only by the CPU test-suite so that the exact device arithmetic can be compared with
// cv2 / NumPy without a GPU.  The product library (libuwie_b200.so) does not contain or call
// anything in this file: there is no CPU execution path in the product.
#include <stdint.h>
#include <string.h>

#include "uwie_color.h"
#include "uwie_hd.h"
#include "uwie_tables.h"

using namespace uwie;

static ColorTables g_tab;
static bool g_ready = false;
static const ColorTables& tabs() {
    if (!g_ready) { build_color_tables(&g_tab); g_ready = true; }
    return g_tab;
}

extern "C" {

void hc_tables(ColorTables* out) { *out = tabs(); }
int hc_tables_size(void) { return (int)sizeof(ColorTables); }

void hc_gray(const uint8_t* rgb, long long n, uint8_t* out) {
    for (long long i = 0; i < n; ++i) out[i] = (uint8_t)gray_q15(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]);
}

void hc_sat(const uint8_t* rgb, long long n, uint8_t* out) {
    const ColorTables& t = tabs();
    for (long long i = 0; i < n; ++i)
        out[i] = (uint8_t)hsv_saturation(t.sdiv, rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]);
}

void hc_rgb2lab(const uint8_t* rgb, long long n, uint8_t* out) {
    const ColorTables& t = tabs();
    for (long long i = 0; i < n; ++i) {
        int L, A, B;
        rgb_to_lab(t.gamma_in, t.cbrt_tab, rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2], L, A, B);
        out[3 * i] = (uint8_t)L; out[3 * i + 1] = (uint8_t)A; out[3 * i + 2] = (uint8_t)B;
    }
}

void hc_lab2rgb(const uint8_t* lab, long long n, uint8_t* out) {
    const ColorTables& t = tabs();
    for (long long i = 0; i < n; ++i) {
        int r, g, b;
        lab_to_rgb(t.lab_yf, t.gamma_out, lab[3 * i], lab[3 * i + 1], lab[3 * i + 2], r, g, b);
        out[3 * i] = (uint8_t)r; out[3 * i + 1] = (uint8_t)g; out[3 * i + 2] = (uint8_t)b;
    }
}

// CLAHE tile LUT from a 256-bin tile histogram (uwie_hd.h)
void hc_clahe_lut(const uint32_t* hist, int clip, int area, uint8_t* lut) {
    int h[256];
    for (int i = 0; i < 256; ++i) h[i] = (int)hist[i];
    clahe_clip_redistribute(h, clip);
    int acc = 0;
    const float scale = clahe_lut_scale(area);
    for (int i = 0; i < 256; ++i) { acc += h[i]; lut[i] = clahe_lut_entry(acc, scale); }
}

void hc_equalize_lut(const uint32_t* hist, long long total, uint8_t* lut) {
    equalize_lut(hist, total, lut);
}

int hc_canny_class(int dx, int dy, int m, int m_l, int m_r, int m_u, int m_d, int m_ul, int m_ur,
                   int m_dl, int m_dr) {
    return canny_nms_class(dx, dy, m, m_l, m_r, m_u, m_d, m_ul, m_ur, m_dl, m_dr, 50, 150);
}

void hc_percentile_plan(long long n, double pct, int f32, long long* k, double* gamma) {
    PercentilePlan p = f32 ? percentile_plan_f32(n, pct) : percentile_plan_f64(n, pct);
    *k = p.k; *gamma = p.gamma;
}

}  // extern "C"
