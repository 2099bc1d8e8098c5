// A strategy of the reference, written down as data.  Both strategy libraries of the reference are
// instances of the same chain
//
//   frame -> [colour cast] -> [dehaze] -> pre-stretches -> [equalizeHist | CLAHE(Lab L)] ->
//            post-stretches -> [gamma] -> float image -> (x * 255) truncated to bytes
//
// six_stadigy.py:230-285 (six strategies, float32 buffers, guard 1e-6, x ** gamma, no clip) and
// enhancement_strategies.py:349-474 with config.py:28-75 (five strategies, float64 buffers,
// guard 1e-10, clip(x ** (1/gamma))).  The engine in uwie_engine.cu walks these records; nothing
// in the kernels is specific to one library except the float type they are instantiated on.
#pragma once
#include <stdint.h>

namespace uwie {

constexpr int kMaxStretch = 2;

struct RankPlan {        // np.percentile 'linear': lerp(sorted[k], sorted[min(k+1, n-1)], gamma)
    long long k;
    double gamma;        // already rounded to the data's float type
};

struct StretchPlan {     // clip((x - p_lo) / (p_hi - p_lo + guard), 0, 1)
    RankPlan lo, hi;
};

struct Recipe {
    int dehaze;              // 1: quadtree light -> transmission -> guided filter -> recover
    float omega;
    int radius;              // guided-filter box is radius x radius (NOT 2r+1)
    double eps;
    int n_pre;               // stretches applied before the histogram stage
    StretchPlan pre[kMaxStretch];
    int equalize;            // per-channel cv2.equalizeHist of the truncated frame (ES:331-345)
    int clahe;               // CLAHE on the 8-bit Lab L plane
    int tiles_x, tiles_y;    // cv2 tileGridSize (8, 8 everywhere in the reference's configs)
    int clahe_clip;          // clip limit in counts: max(int(clipLimit * tileArea / 256), 1)
    int n_post;              // stretches applied after CLAHE / equalizeHist
    StretchPlan post[kMaxStretch];
    int gamma_on;
    double gamma_exp;        // exponent actually applied (gamma for SS, 1/gamma for ES)
    int gamma_clip;          // ES clips to [0,1] after the power
};

struct Library {             // what differs between the two strategy libraries
    int f64;                 // 0: float32 image buffers (SS)   1: float64 (ES)
    int colour_cast;         // 1: detect_image_type + color_correction first (SS:292-323)
    double dark_guard;       // img / (A + guard): 1e-6 (SS:170) or 1e-10 (ES:221)
    int clip_t0;             // SS:174 clips the raw transmission to [.1,1] before the guided filter
    double stretch_guard;    // 1e-6 (SS:198,218) or 1e-10 (ES:268)
    int n_strategies;
    double weights[8];       // comprehensive_assessment weights in QA key order (QA:229-238)
};

}  // namespace uwie
