// Host-side builders for the lookup tables of uwie_color.h and for the percentile plan.
// Compile with -ffp-contract=off: the float expressions must round exactly like NumPy's.
#include "uwie_tables.h"

#include <cmath>

namespace uwie {

static double srgb_to_linear(double x) {
    return x <= 0.04045 ? x / 12.92 : std::pow((x + 0.055) / 1.055, 2.4);
}

static double linear_to_srgb(double x) {
    return x <= 0.0031308 ? x * 12.92 : 1.055 * std::pow(x, 1.0 / 2.4) - 0.055;
}

void build_color_tables(ColorTables* t) {
    for (int i = 0; i < 256; ++i)
        t->gamma_in[i] = (uint16_t)std::lround(2040.0 * srgb_to_linear(i / 255.0));
    for (int k = 0; k < 2048 + 8; ++k) {
        const double x = k / 2040.0;
        const double f = x < 216.0 / 24389.0 ? x * 841.0 / 108.0 + 16.0 / 116.0 : std::cbrt(x);
        t->cbrt_tab[k] = (uint16_t)std::lround(32768.0 * f);
    }
    // OpenCV builds this table through its soft-float cube root; two entries land on the
    // other side of a rounding tie (found by solving against cv2 on all 2^24 colours,
    // SURVEY.md A.5).
    t->cbrt_tab[49] = 9454;
    t->cbrt_tab[628] = 22126;
    for (int i = 0; i < 4096; ++i)
        t->gamma_out[i] = (uint8_t)std::lround(255.0 * linear_to_srgb(i / 4096.0));
    t->sdiv[0] = 0;
    for (int i = 1; i < 256; ++i)
        t->sdiv[i] = (int32_t)std::nearbyint((255 << 12) / (double)i);
    const double BASE = 16384.0;
    for (int L = 0; L < 256; ++L) {
        double y, ify;
        if (L <= 20) {
            y = std::round(L * BASE * 180.0 / (17.0 * 29.0 * 29.0 * 29.0));
            ify = std::round(BASE * (16.0 / 116.0 + 5.0 * L / (3.0 * 17.0 * 29.0)));
        } else {
            const double fy = L * 100.0 * BASE / (255.0 * 116.0) + 16.0 * BASE / 116.0;
            ify = std::round(fy);
            y = std::round(fy * fy * fy / (BASE * BASE));
        }
        t->lab_yf[2 * L] = (int16_t)y;
        t->lab_yf[2 * L + 1] = (int16_t)ify;
    }
}

// NumPy's 'linear' percentile of n values (numpy/lib/_function_base_impl.py,
// _QuantileMethods['linear'] + _get_gamma): virtual index (n - 1) * q with q = pct / 100, both
// evaluated in the data's own float type (NumPy 2.x divides by a.dtype.type(100)).
PercentilePlan percentile_plan_f32(long long n, double pct) {
    volatile float q = (float)pct / 100.0f;
    volatile float vi = (float)(n - 1) * q;      // 'linear': (n - 1) * quantiles, in the data's float type
    PercentilePlan p;
    if (vi >= (float)(n - 1)) { p.k = n - 1; p.gamma = 0.0; return p; }
    const float fl = std::floor(vi);
    p.k = (long long)fl;
    if (p.k < 0) { p.k = 0; p.gamma = 0.0; return p; }
    volatile float gm = vi - fl;
    p.gamma = gm;
    return p;
}

PercentilePlan percentile_plan_f64(long long n, double pct) {
    volatile double q = pct / 100.0;
    volatile double vi = (double)(n - 1) * q;
    PercentilePlan p;
    if (vi >= (double)(n - 1)) { p.k = n - 1; p.gamma = 0.0; return p; }
    const double fl = std::floor(vi);
    p.k = (long long)fl;
    if (p.k < 0) { p.k = 0; p.gamma = 0.0; return p; }
    volatile double gm = vi - fl;
    p.gamma = gm;
    return p;
}

}  // namespace uwie
