// Launch wrappers of the kernel files (internal; the public boundary is include/uwie.h).
// Every wrapper enqueues on `st`, never synchronises, and returns the number of kernels launched.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "k_common.cuh"
#include "uwie_recipe.h"

namespace uwie {

// ------------------------------------------------------------------ k_canny.cu
enum { CANNY_QUADRANTS = 0, CANNY_PLANES = 1 };
constexpr int kCannyTile = 64;       // tile side
constexpr int kCannyListCap = 128;   // components of a tile that can touch its rim (<= 126: rim pixels of
                                     // different components are never adjacent on the 252-pixel ring)
struct CannyArgs {
    int mode;
    int H, W;
    int low, high;
    // CANNY_QUADRANTS: instance = frame * 4 + quadrant of rects[frame]; gray is computed from the
    // frame bytes through the colour-cast byte table; per-quadrant channel histograms are produced.
    const uint8_t* frames;      // [B][H][W][3]
    const CastTables* cast;     // [B]
    const QuadRect* rects;      // [B]
    int level;
    unsigned* quad_hist;        // [B][kQuadLevelsMax][4][3][256]
    // CANNY_PLANES: instance = one gray plane
    const uint8_t* gray;        // [instances][H][W]
    int planes_per_img;
    // scratch + result
    int* labels;                // canny_label_bytes(slots): rim labels [tile slots][252], then union-find nodes [tile slots][kCannyListCap]
    int2* tile_list;            // [tile slots][kCannyListCap]   {node id, size}
    int* tile_n;                // [tile slots]
    unsigned* counts;           // QUADRANTS: [B][kQuadLevelsMax][4]; PLANES: [instances]
};
int launch_canny(const CannyArgs& a, int max_rows, int max_cols, int instances, cudaStream_t st);
size_t canny_tile_slots(int max_rows, int max_cols, int instances);   // tile slots one launch uses
size_t canny_label_bytes(size_t slots);                                 // bytes of CannyArgs::labels for that many slots

// ------------------------------------------------------------------ k_stats.cu
// per-channel 256-bin histograms of uint8 HWC frames: hist[B][3][256] (must be zeroed)
int launch_frame_hist(const uint8_t* frames, int B, int H, int W, unsigned* hist, cudaStream_t st);

// ------------------------------------------------------------------ k_quadtree.cu
struct QuadArgs {
    int B, H, W;
    const uint8_t* frames;
    const CastTables* cast;
    QuadRect* rects;
    const unsigned* quad_hist;   // [B][L][4][3][256]
    const unsigned* counts;      // [B][L][4]
    double* q_trace;             // [B][L][4]  Q of every visited quadrant (diagnostics / tests)
    float* light;                // [B][3]
};
int launch_quad_init(const QuadArgs& a, cudaStream_t st);
int launch_quad_descend(const QuadArgs& a, int level, cudaStream_t st);
int launch_quad_brightest(const QuadArgs& a, cudaStream_t st);

// ------------------------------------------------------------------ k_setup.cu
// Per (frame, strategy) tables.  Values are stored as double but always hold numbers of the
// library's float type (float32 values are exact in double).
struct StratTables {
    uint8_t preq[3][256];     // no-dehaze strategies: byte entering CLAHE, by frame byte
    double postf[3][256];     // final float value, by CLAHE-output byte (or by frame byte if no CLAHE)
    uint8_t postq[3][256];    // its truncation to a byte
    double st_lo[kMaxStretch][3];   // dehaze strategies: float-domain stretches  (x - lo) / den
    double st_den[kMaxStretch][3];
};
struct SetupArgs {
    int B, S;
    long long n;                 // pixels per frame
    Library lib;
    const Recipe* recipes;       // [S] device
    const unsigned* frame_hist;  // [B][3][256]
    const unsigned* out_hist;    // [B][S][3][256] histograms of the CLAHE output (post-stretch strategies)
    int* types;                  // [B]
    CastTables* cast;            // [B]
    StratTables* tables;         // [B][S]
};
int launch_setup_cast(const SetupArgs& a, cudaStream_t st);
int launch_setup_frame_luts(const SetupArgs& a, cudaStream_t st);
int launch_setup_post_luts(const SetupArgs& a, cudaStream_t st);

// ------------------------------------------------------------------ k_guided.cu / k_select.cu
enum { UWIE_STATUS_SELECT_OVERFLOW = 1, UWIE_STATUS_SELECT_INTERNAL = 2 };
struct GuidedArgs {
    int B, S, H, W;
    Library lib;
    const Recipe* recipes;       // [S] device
    int n_dehaze;                // number of dehaze strategies
    int dehaze_idx[kStrategiesMax];   // their strategy indices
    int rows_per_block;
    int eps_safe;                // every dehaze strategy has eps >= 1e-9: a = cov / (var + eps) needs no range check
    uint2* gd_plane;             // [B][H][W]   {R | G << 8 | B << 16 | gray << 24 of the (colour-corrected) frame,
                                 //              float32 bits of min_c img / (A + guard)}
    const uint8_t* frames;
    const CastTables* cast;
    const float* light;          // [B][3]
    double2* ab;                 // [B * n_dehaze][H][W] {a, b}
    void* recovered;             // [B * n_dehaze][H][W][3] in the library's float type
    unsigned* sel_hist_a;        // [B * n_dehaze][3][kSelBinsA]
};
int launch_guided(const GuidedArgs& a, int max_radius, cudaStream_t st);
int selftest_division(unsigned long long count, unsigned long long seed, unsigned long long* mismatches_host);

struct SelTarget {
    long long rank;      // rank still to resolve inside the current bin
    unsigned prefix;     // digits of the key resolved so far (11, 22, then all 32 bits)
    int active;
    int rep;             // target of the same channel whose histogram / table / list this one shares
    int pad;
    double val;          // active == 3: the order statistic itself (resolved from the candidate list)
};
struct SelectArgs {
    int B, S;
    long long n;
    Library lib;
    const Recipe* recipes;
    int n_dehaze;
    int dehaze_idx[kStrategiesMax];
    const void* recovered;
    const unsigned* sel_hist_a;  // [inst][3][kSelBinsA]
    unsigned* sel_hist_b;        // [inst][kSelTargets][kSelBinsB]
    unsigned* sel_hist_c;        // [inst][kSelTargets][kSelBinsC]
    SelTarget* targets;          // [inst][kSelTargets]
    unsigned long long* tab_keys;   // [inst][kSelTargets][kSelTable], 0 = empty, else bit pattern + 1
    unsigned* tab_counts;           // [inst][kSelTargets][kSelTable]
    void* cand;                     // [inst][kSelTargets][kSelListCap] bit patterns (32 / 64 bits) of the members of a target's level-A bin
    unsigned* cand_n;               // [inst][kSelTargets] members seen (may exceed kSelListCap: list unusable)
    unsigned* stats;                // [2] targets resolved from lists / targets sent down the streaming levels (diagnostics)
    StratTables* tables;         // [B][S]
    int* status;
};
int launch_select(const SelectArgs& a, cudaStream_t st);

// ------------------------------------------------------------------ k_apply.cu / k_score.cu
constexpr int kMaxTiles = 256;   // CLAHE tile grids up to 16 x 16
struct ApplyArgs {
    int B, S, H, W;
    Library lib;
    const Recipe* recipes;
    int n_dehaze;
    int dehaze_slot[kStrategiesMax];   // strategy -> slot in the recovered buffer, -1 if none
    int var_n[4];                      // strategies per kernel variant (0 tables only, 1 dehaze, 2 CLAHE) and their indices:
    int var_s[4][kStrategiesMax];      // every variant is its own instantiation of apply_kernel
    int var_hist[4];                   // strategies of the variant that need the histogram pass (CLAHE + post-stretches)
    int fin_n[4];                      // the same lists for the FINAL pass: a CLAHE strategy with post-stretches left its CLAHE
    int fin_s[4][kStrategiesMax];      // output bytes in `out` during the histogram pass and finishes as variant 0 (tables only)
    int from_out[kStrategiesMax];      // 1: the final pass reads `out` (in place) instead of the frame
    const uint8_t* frames;
    const StratTables* tables;
    const void* recovered;
    const ColorTables* ctab;
    uint8_t* tile_lut;           // [B][S][kMaxTiles][256]
    unsigned* out_hist;          // [B][S][3][256]
    uint8_t* out;                // [B][S][H][W][3]
    uint8_t* gray;               // [B][S][H][W]
    QaAccum* qa;                 // [B][S]
    double* cf_part;             // [B][S][final_blocks][4]
    int pix_per_block;
    int final_blocks;
};
int launch_clahe_luts(const ApplyArgs& a, int max_tiles, cudaStream_t st);
int launch_apply(const ApplyArgs& a, bool final_pass, cudaStream_t st);

struct ScoreArgs {
    int B, S;
    long long n;
    Library lib;
    const QaAccum* qa;
    const double* cf_part;
    int final_blocks;
    const unsigned* edge_counts; // [B][S]
    double* scores;              // [B][S][kScoreCols]
    int* best;                   // [B]
};
int launch_scores(const ScoreArgs& a, cudaStream_t st);

}  // namespace uwie
