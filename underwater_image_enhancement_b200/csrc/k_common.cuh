// Device-side helpers shared by all kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "uwie_color.h"
#include "uwie_hd.h"

namespace uwie {

constexpr int kStrategiesMax = 6;
constexpr int kScoreCols = 9;        // total + 8 sub-scores
constexpr int kSelBinsA = 1056;      // first-level bins of the percentile selection: key >> 21 in [0, 1024], padded to 33 * 32
constexpr int kSelBinsB = 2048;      // second level: 11 bits
constexpr int kSelBinsC = 1024;      // third level: 10 bits
constexpr int kSelTargets = 24;      // 3 channels x 2 stretches x (lo, hi) x (k, k+1)
constexpr int kSelListCap = 32768;   // members of a target's level-A bin kept as a list (bit patterns); a longer bin takes the streaming levels B / C
constexpr int kSelTable = 256;       // distinct values ranked inside a final bin of width 2^-31 (hash table slots)
constexpr int kQuadLevelsMax = 24;   // log2 of the largest supported side, with slack

// cv2 BORDER_REFLECT_101: -1 -> 1, n -> n-2
__device__ __forceinline__ int reflect101(int i, int n) {
    if (n == 1) return 0;
    while (i < 0 || i >= n) {
        if (i < 0) i = -i;
        else i = 2 * n - 2 - i;
    }
    return i;
}

__device__ __forceinline__ int clamp_idx(int i, int n) { return i < 0 ? 0 : (i >= n ? n - 1 : i); }

__device__ __forceinline__ unsigned warp_sum_u32(unsigned v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Per-frame colour-cast state (six_stadigy.py:292-323): the corrected frame is a per-channel
// function of the input byte, so it is carried as two 256-entry tables instead of a float image.
struct CastTables {
    float curf[3][256];    // float32 value of the corrected frame for input byte v
    uint8_t curq[3][256];  // trunc(curf * 255): the byte every cv2 call on the corrected frame sees
};

// Rectangle of the quadtree descent (enhancement_strategies.py:99-139, six_stadigy.py:69-108)
struct QuadRect {
    int y0, y1, x0, x1;
    int active;  // 1 while rows > 1 and cols > 1
    int level;   // number of completed levels
};

// Per (frame, strategy) accumulators of the scoring pass (quality_assessment.py:15-212)
struct QaAccum {
    uint32_t gray_hist[256];
    uint32_t sat_hist[256];
    unsigned long long lum_sum;     // sum of 8-bit Lab L
    unsigned int edge_count;        // Canny(50,150) edge pixels
    unsigned int pad;
};

}  // namespace uwie
