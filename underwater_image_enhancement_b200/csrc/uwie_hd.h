// Small pieces of integer / float32 arithmetic shared between kernels and the CPU host check:
// CLAHE tile LUT construction, cv2.equalizeHist LUT, Canny non-maximum suppression.
// Restated from the OpenCV behaviour the reference relies on;
// reference call sites: enhancement_strategies.py:181,302-303,342-343, six_stadigy.py:150,205-206,
// quality_assessment.py:139.
#pragma once
#include <stdint.h>
#include <math.h>

#include "uwie_color.h"

namespace uwie {

UW_HD int rint_to_int(float v) {
#if defined(__CUDA_ARCH__)
    return __float2int_rn(v);
#else
    return (int)nearbyintf(v);  // default rounding mode: half to even
#endif
}

// CLAHE: clip the tile histogram at `clip`, hand the excess back uniformly, then the residual
// one count at a time with stride max(256 / residual, 1).
UW_HD void clahe_clip_redistribute(int* h, int clip) {
    int excess = 0;
    for (int i = 0; i < 256; ++i) {
        if (h[i] > clip) { excess += h[i] - clip; h[i] = clip; }
    }
    const int batch = excess / 256;
    int residual = excess - batch * 256;
    for (int i = 0; i < 256; ++i) h[i] += batch;
    if (residual != 0) {
        int step = 256 / residual;
        if (step < 1) step = 1;
        for (int i = 0; i < 256 && residual > 0; i += step, --residual) h[i] += 1;
    }
}

UW_HD float clahe_lut_scale(int area) {
#if defined(__CUDA_ARCH__)
    return __fdiv_rn(255.0f, (float)area);
#else
    volatile float s = 255.0f / (float)area;
    return s;
#endif
}

UW_HD uint8_t clahe_lut_entry(int cumsum, float scale) {
#if defined(__CUDA_ARCH__)
    const float v = __fmul_rn((float)cumsum, scale);
#else
    volatile float v = (float)cumsum * scale;
#endif
    return (uint8_t)clampi(rint_to_int(v), 0, 255);
}

// clip limit in counts for a tile of `area` pixels: max(int(clipLimit * area / 256), 1)
inline int clahe_clip_count(double clip_limit, int area) {
    int c = (int)(clip_limit * area / 256);
    return c < 1 ? 1 : c;
}

// cv2.equalizeHist LUT; a constant plane maps to itself.
UW_HD void equalize_lut(const uint32_t* hist, long long total, uint8_t* lut) {
    int i0 = 0;
    while (i0 < 255 && hist[i0] == 0) ++i0;
    if ((long long)hist[i0] == total) {
        for (int i = 0; i < 256; ++i) lut[i] = (uint8_t)i;
        return;
    }
#if defined(__CUDA_ARCH__)
    const float scale = __fdiv_rn(255.0f, (float)(total - hist[i0]));
#else
    volatile float scale = 255.0f / (float)(total - hist[i0]);
#endif
    for (int i = 0; i <= i0; ++i) lut[i] = 0;
    long long sum = 0;
    for (int i = i0 + 1; i < 256; ++i) {
        sum += hist[i];
#if defined(__CUDA_ARCH__)
        const float v = __fmul_rn((float)sum, scale);
#else
        volatile float v = (float)sum * scale;
#endif
        lut[i] = (uint8_t)clampi(rint_to_int(v), 0, 255);
    }
}

// Canny NMS for one pixel: 0 = not an edge candidate, 1 = weak (m > low), 2 = strong (m > high).
// m_* are the L1 gradient magnitudes of the 8 neighbours (0 outside the image).
UW_HD int canny_nms_class(int dx, int dy, int m, int m_l, int m_r, int m_u, int m_d,
                          int m_ul, int m_ur, int m_dl, int m_dr, int low, int high) {
    if (m <= low) return 0;
    const int x = dx < 0 ? -dx : dx;
    const int y = (dy < 0 ? -dy : dy) << 15;
    const int tg22x = x * 13573;
    bool keep;
    if (y < tg22x) {
        keep = m > m_l && m >= m_r;
    } else {
        const int tg67x = tg22x + (x << 16);
        if (y > tg67x) {
            keep = m > m_u && m >= m_d;
        } else if ((dx ^ dy) < 0) {
            keep = m > m_ur && m > m_dl;
        } else {
            keep = m > m_ul && m > m_dr;
        }
    }
    if (!keep) return 0;
    return m > high ? 2 : 1;
}

}  // namespace uwie
