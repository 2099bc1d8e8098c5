/* uwie.h — C ABI of the B200-native enhance + score hot path.
 *
 * The reference (submarine0418/underwater_image_enhancement) is pure Python and has no FFI layer;
 * its boundary for this path is the static-method API
 *     six_stadigy.py:22-323        EnhancementStrategies.strategy1..6, detect_image_type, color_correction
 *     enhancement_strategies.py:13-508   EnhancementStrategies.apply_strategy(img, name, params)
 *     quality_assessment.py:215-286      QualityAssessment.comprehensive_assessment(img, weights)
 *     main.py:116-155              per-image loop: every strategy, score each, keep the first maximum
 *     six_stadigy.py:392-447       per-image loop: colour cast, six strategies, write all six
 * The entry points below are what a ctypes binding of those loops calls (INTEGRATION.md shows the
 * stubs).  Conventions: plain pointers and sizes only; `*_dev` pointers are CUDA device memory,
 * `*_host` pointers are host memory; every call returns 0 on success or a negative uwie_status;
 * nothing throws across the boundary; device entry points are asynchronous on `stream` (pass the
 * cudaStream_t as a void*; NULL is the default stream) and never synchronise; there is NO CPU
 * implementation behind this header — without a CUDA device every compute call fails with
 * UWIE_ERR_CUDA.
 */
#ifndef UWIE_H
#define UWIE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define UWIE_VERSION 100          /* 0.1.0 */
#define UWIE_MAX_STRATEGIES 6
#define UWIE_SCORE_COLS 9         /* total, contrast, sharpness, entropy, saturation, brightness,
                                     edge_density, colorfulness, naturalness (QA:243-281 order) */

typedef enum {
    UWIE_OK = 0,
    UWIE_ERR_ARG = -1,            /* bad argument (the Python shims raise ValueError) */
    UWIE_ERR_CUDA = -2,           /* CUDA runtime error, see uwie_last_error() */
    UWIE_ERR_NOMEM = -3,
    UWIE_ERR_SELECT = -4          /* exact percentile selection overflowed its candidate buffer */
} uwie_status;

/* The two strategy libraries of the reference, as presets. */
typedef enum {
    UWIE_LIB_SIX = 0,             /* six_stadigy.py: colour cast + strategy1..6, float32 buffers */
    UWIE_LIB_ES = 1               /* enhancement_strategies.py + config.py:28-75, float64 buffers */
} uwie_library_id;

/* One strategy, as data (see csrc/uwie_recipe.h for the chain it describes). */
typedef struct {
    int dehaze;                   /* quadtree light + transmission + guided filter + recover */
    double omega;
    int radius;
    double eps;
    int n_pre;                    /* percentile stretches before the histogram stage (<= 2) */
    double pre_lo[2], pre_hi[2];
    int equalize;                 /* per-channel cv2.equalizeHist (ES:331-345) */
    int clahe;                    /* CLAHE on the 8-bit Lab L plane */
    double clip_limit;
    int tiles_x, tiles_y;
    int n_post;                   /* percentile stretches after CLAHE / equalizeHist (<= 2) */
    double post_lo[2], post_hi[2];
    int gamma_on;
    double gamma;                 /* the reference's gamma argument (SS: x**gamma, ES: x**(1/gamma)) */
} uwie_strategy;

typedef struct {
    int f64;                      /* float type of the image buffers: 0 float32 (SS), 1 float64 (ES) */
    int colour_cast;              /* detect_image_type + color_correction first (SS:292-323) */
    double dark_guard;            /* 1e-6 (SS:170) / 1e-10 (ES:221) */
    int clip_t0;                  /* clip raw transmission to [.1,1] before the guided filter (SS:174) */
    double stretch_guard;         /* 1e-6 (SS:198) / 1e-10 (ES:268) */
    int gamma_inverse;            /* 1: exponent is 1/gamma and the result is clipped (ES:284-285) */
    double weights[8];            /* comprehensive_assessment weights, QA key order */
} uwie_library;

typedef struct uwie_engine uwie_engine;

int uwie_version(void);
const char* uwie_last_error(void);

/* Fill `lib` and `strategies` (room for UWIE_MAX_STRATEGIES) with a preset; returns the count.
 * UWIE_LIB_SIX: the constants hard-coded in strategy1..6 (six_stadigy.py:230-285) and the default weights of
 * comprehensive_assessment (quality_assessment.py:229-238); UWIE_LIB_ES: Config.STRATEGIES / Config.QUALITY_WEIGHTS
 * (config.py:28-85) with the defaults of enhancement_strategies.py:349-474. */
int uwie_preset(int library_id, uwie_library* lib, uwie_strategy* strategies);

/* Device memory an engine needs for `max_batch` frames of H x W (0 on bad arguments). */
size_t uwie_workspace_bytes(const uwie_library* lib, const uwie_strategy* strategies, int n_strategies,
                            int max_batch, int height, int width);

/* Create an engine.  `workspace_dev` may be NULL (the engine allocates with cudaMalloc and owns
 * it) or caller-owned device memory of at least uwie_workspace_bytes(...). */
int uwie_engine_create(uwie_engine** out, const uwie_library* lib, const uwie_strategy* strategies,
                       int n_strategies, int max_batch, int height, int width,
                       void* workspace_dev, size_t workspace_bytes);
void uwie_engine_destroy(uwie_engine* e);

/* Enhance + score `batch` (<= max_batch) uint8 RGB frames [batch][H][W][3] resident on the device.
 * Any output pointer may be NULL.
 *   out_all_dev   uint8  [batch][S][H][W][3]   every strategy's output       (SS:430-447)
 *   out_best_dev  uint8  [batch][H][W][3]      the selected strategy's output (main.py:150-155)
 *   scores_dev    double [batch][S][9]         total + 8 sub-scores          (QA:215-286)
 *   best_dev      int32  [batch]               first maximum of the totals   (main.py:144-146)
 *   type_dev      int32  [batch]               0 normal / 1 greenish / 2 bluish (SS:292-302)
 *   light_dev     float  [batch][3]            atmospheric light             (ES:77-144, SS:48-113)
 * Asynchronous on `stream`. */
int uwie_enhance_score(uwie_engine* e, const uint8_t* frames_dev, int batch,
                       uint8_t* out_all_dev, uint8_t* out_best_dev, double* scores_dev,
                       int32_t* best_dev, int32_t* type_dev, float* light_dev, void* stream);

/* Same call with HOST buffers (any count of frames, processed in chunks of max_batch): copies the
 * frames to the device, runs, copies the requested results back, synchronises.  Pinned host
 * memory makes the copies asynchronous with respect to the kernels of the neighbouring chunk.
 * The call allocates nothing: uwie_engine_host_prepare must have been called once with
 * max_count >= count and the want_* flags of the outputs that will be asked for (it owns the
 * device slots, the pinned landing zone of the small results, the copy streams and, when max_count >
 * max_batch, one further engine + stream per chunk of the largest call, up to four lanes in all: chunk c
 * runs on lane c % lanes, so the latency-bound stretches of one chunk overlap the wide kernels of the
 * others; UWIE_HOST_LANES=n caps the lanes); otherwise UWIE_ERR_ARG.
 * This is the loop body of main.py:88-155 / six_stadigy.py:392-447 without the file I/O. */
int uwie_engine_host_prepare(uwie_engine* e, int max_count, int want_all, int want_best);
int uwie_enhance_score_host(uwie_engine* e, const uint8_t* frames_host, int count,
                            uint8_t* out_all_host, uint8_t* out_best_host, double* scores_host,
                            int32_t* best_host, int32_t* type_host, float* light_host, void* stream);

/* detect_image_type (six_stadigy.py:292-302) of `batch` frames on the device: 0 normal / 1 greenish / 2 bluish.
 * Runs the channel histograms and the colour-cast set-up only.  Asynchronous on `stream`. */
int uwie_engine_image_types(uwie_engine* e, const uint8_t* frames_dev, int batch, int32_t* type_dev, void* stream);

/* Device-side status of the engine's calls since the last check (synchronises `stream`, read-and-clear):
 * 0 or a negative uwie_status.  uwie_enhance_score is asynchronous and cannot report a device-side
 * failure itself (an exact percentile selection that overflowed, UWIE_ERR_SELECT): callers of the
 * device entry point must check before they trust the results; the host entry point does it itself. */
int uwie_engine_check(uwie_engine* e, void* stream);

/* Kernels launched by the last uwie_enhance_score call on this engine. */
int uwie_engine_last_launches(const uwie_engine* e);

/* uwie_enhance_score[_host] replays its kernel sequence as ONE CUDA graph launch once it has seen the same batch size and
 * buffer pointers twice on a capturable stream (not the legacy / per-thread default stream; UWIE_NO_GRAPH=1 disables it).
 * Number of calls served by a graph replay so far (a diagnostic; the kernels and their results are the same). */
long long uwie_engine_graph_replays(const uwie_engine* e);

/* Per-stage device timing.  uwie_engine_profile(e, 1) clears the record and makes every following
 * uwie_enhance_score[_host] call bracket each stage of its kernel sequence with CUDA events on the
 * call's stream; (e, 0) stops.  uwie_engine_profile_read synchronises those events and returns, per
 * stage name (UWIE_STAGE_NAME_LEN bytes each, NUL-terminated), the summed milliseconds, the number
 * of kernels launched and the number of times the stage ran.  Returns the number of stages (<= cap)
 * or a negative uwie_status. */
#define UWIE_STAGE_NAME_LEN 32
int uwie_engine_profile(uwie_engine* e, int enable);
int uwie_engine_profile_read(uwie_engine* e, char* names, double* ms, int* kernels, int* runs, int cap);

/* Test hook: device pointer and byte size of a named intermediate of the last run
 * ("frame_hist", "cast", "tables", "rects", "q_trace", "quad_counts", "light", "ab", "recovered",
 *  "tile_lut", "out_hist", "gray", "qa", "edge_counts", "targets").  NULL if unknown. */
const void* uwie_engine_debug(const uwie_engine* e, const char* name, size_t* bytes);

/* ---------------------------------------------------------------------------------------------
 * VGG-guided path (use_trained_model.py): the per-pixel pieces around the network, which itself
 * stays on PyTorch.  Stateless calls on caller-owned device memory; `workspace_dev` must hold
 * uwie_vgg_workspace_bytes(batch) bytes; all are asynchronous on `stream`.
 * ------------------------------------------------------------------------------------------- */
#define UWIE_VGG_FEATURES 79      /* vgg_16_UIE.py:462-466: 18 statistics zero-padded to 79 */

size_t uwie_vgg_workspace_bytes(int batch);

/* DifferentiableEnhancement.forward (vgg_16_UIE.py:32-55): colour stretch between the order
 * statistics sorted[int(L/100 * n)] of each plane (VGG:57-92), dark-channel dehaze with A = 0.6
 * (VGG:94-117) when `omega` is given, pow(x + 1e-8, gamma) (VGG:119-128) when `gamma` is given,
 * clamp.  img / out: float32 NCHW [batch][3][H][W] on the device; l_low, l_high, omega, gamma:
 * HOST arrays of `batch` floats (omega and gamma may be NULL = step skipped, like a missing key). */
int uwie_vgg_enhance(const float* img_dev, int batch, int height, int width,
                     const float* l_low, const float* l_high, const float* omega, const float* gamma,
                     float* out_dev, void* workspace_dev, size_t workspace_bytes, void* stream);

/* The same enhancement with DEVICE-resident parameters — the outputs of the parameter network, so that a batch
 * needs no per-image host round trip (use_trained_model.py:63-66 does .cpu().item() per parameter).
 * params_dev: float32 [4][batch] = L_low, L_high, omega, gamma rows.  clamp_ranges: apply the safety ranges of
 * use_trained_model.py:74-79 (L_low 1..30, L_high 65..99, omega .1...9, gamma .5..3) on the device first;
 * used_dev (may be NULL): float32 [4][batch], the values actually applied. */
int uwie_vgg_enhance_dev(const float* img_dev, int batch, int height, int width, const float* params_dev, int clamp_ranges,
                         float* used_dev, float* out_dev, void* workspace_dev, size_t workspace_bytes, void* stream);

/* The same with the frames still uint8 [batch][H][W][3] (imread bytes; use_trained_model.py:119 divides them by 255): the
 * order statistics come from the channel histograms, so the frame is read twice as bytes instead of four times as floats.
 * out_dev: float32 [batch][3][H][W].  Bit-identical to uwie_vgg_enhance_dev on frame / 255. */
int uwie_vgg_enhance_u8(const uint8_t* frames_dev, int batch, int height, int width, const float* params_dev, int clamp_ranges,
                        float* used_dev, float* out_dev, void* workspace_dev, size_t workspace_bytes, void* stream);

/* Backward of the enhancement for the training loops (vgg_16_UIE.py:529, deep_learning_parameters.py:281): given
 * d loss / d out (float32 NCHW), the input image and the WORKSPACE of the forward uwie_vgg_enhance[_dev] call on the same
 * batch (it still holds the order statistics and parameters), writes d loss / d omega and d loss / d gamma as float64
 * [2][batch].  L_low / L_high get no gradient in the reference either (they pass through .item(), VGG:78-79). */
int uwie_vgg_enhance_bwd(const float* img_dev, const float* grad_out_dev, int batch, int height, int width,
                         const void* workspace_dev, size_t workspace_bytes, double* grad_params_dev, void* stream);

/* extract_all_features (vgg_16_UIE.py:435-466) of uint8 RGB frames [batch][H][W][3] (the reference
 * divides by 255 first, VGG:425-433): per channel mean, std, min, max, median; overall mean, std,
 * mean of squares; zeros up to 79.  feat_dev: float32 [batch][79]. */
int uwie_vgg_features_u8(const uint8_t* frames_dev, int batch, int height, int width, float* feat_dev,
                         void* workspace_dev, size_t workspace_bytes, void* stream);

/* EnhancementPredictor._preprocess_for_vgg (use_trained_model.py:39-46): truncate to bytes,
 * cv2.resize(size x size, INTER_LINEAR) in 8-bit fixed point, / 255, ImageNet normalisation,
 * HWC -> CHW.  Input uint8 or float32 [0,1] HWC frames; out_dev: float32 [batch][3][size][size]. */
int uwie_vgg_preprocess_u8(const uint8_t* frames_dev, int batch, int height, int width, int size,
                           float* out_dev, void* stream);
int uwie_vgg_preprocess_f32(const float* img_dev, int batch, int height, int width, int size,
                            float* out_dev, void* stream);

/* ---------------------------------------------------------------------------------------------
 * quality_assessment.py: QualityAssessment.comprehensive_assessment(img, weights) (QA:215-286) of
 * images that did not come out of an engine.  img_dev: [batch][H][W][3] on the device, dtype 0 = uint8
 * (the image is bytes / 255 in float32), 1 = float32, 2 = float64 (the enhanced image in [0, 1]);
 * weights: HOST array of 8 doubles in QA key order; scores_dev: double [batch][UWIE_SCORE_COLS].
 * Stateless, asynchronous on `stream`.
 * ------------------------------------------------------------------------------------------- */
size_t uwie_quality_workspace_bytes(int batch, int height, int width);
int uwie_quality_scores(const void* img_dev, int dtype, int batch, int height, int width, const double* weights,
                        double* scores_dev, void* workspace_dev, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------
 * feature_extraction.py: FeatureExtractor.extract_all_features (FE:249-297) of uint8 RGB frames
 * [batch][H][W][3] on the device -> float64 [batch][79]:
 *   colour 35 (FE:17-77) | texture 22 (FE:80-120) | frequency 5 (FE:123-158) | edge 7 (FE:161-200) |
 *   quality 10 (FE:203-246).
 * Stateless, asynchronous on `stream`; `workspace_dev` must hold uwie_features79_workspace_bytes(...).
 * The DCT of the frequency group is hand-written (csrc/k_dct.cu: float32 GEMMs with the even / odd butterfly, the
 * coefficient plane reduced in the epilogue); its cosine matrices are built once per (device, length) and cached.
 * ------------------------------------------------------------------------------------------- */
#define UWIE_FEATURES 79
size_t uwie_features79_workspace_bytes(int batch, int height, int width);
int uwie_features79(const uint8_t* frames_dev, int batch, int height, int width, double* feat_dev,
                    void* workspace_dev, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Per-primitive entry points: the public static methods of the reference's two
 * EnhancementStrategies classes as stand-alone operators.  Stateless, caller-owned device memory and
 * workspace (`*_workspace_bytes`), asynchronous on `stream` unless stated.  Float images are
 * [H][W][3] with dtype 1 = float32, 2 = float64 (the codes of uwie_quality_scores); they need NOT be
 * made from bytes (example_usage.py:27 feeds np.random.rand images).
 * ------------------------------------------------------------------------------------------- */

/* Transmission head on uint8 frames [batch][H][W][3] (BASELINE.json configs[2], 7 N bytes per frame):
 * t0 = 1 - omega * min_c((v_c / 255) / (A_c + guard)) in float32, clipped to [.1, 1] when clip_t0
 * (six_stadigy.py:170-174; enhancement_strategies.py:221-225 with clip_t0 = 0).  light_dev: float32
 * [batch][3]; t_dev: float32 [batch][H][W]. */
int uwie_prim_transmission_head_u8(const uint8_t* frames_dev, int batch, int height, int width, const float* light_dev,
                                   float omega, float guard, int clip_t0, float* t_dev, void* stream);

/* cv2.createCLAHE(clip_limit, (tiles_x, tiles_y)).apply on uint8 planes [batch][H][W] — the Lab L plane of
 * enhancement_strategies.py:302-303 / six_stadigy.py:205-206 (configs[2], 2 N bytes per plane). Bit-exact. */
size_t uwie_prim_clahe_lab_workspace_bytes(int batch, int tiles_x, int tiles_y);
int uwie_prim_clahe_lab(const uint8_t* plane_dev, int batch, int height, int width, double clip_limit, int tiles_x,
                        int tiles_y, uint8_t* out_dev, void* workspace_dev, size_t workspace_bytes, void* stream);

/* cv2.equalizeHist per channel of uint8 frames [batch][H][W][3] (enhancement_strategies.py:339-343).
 * Workspace: batch * 3840 bytes. */
int uwie_prim_equalize_hist(const uint8_t* frames_dev, int batch, int height, int width, uint8_t* out_dev,
                            void* workspace_dev, size_t workspace_bytes, void* stream);

/* Allocation-free variant of uwie_prim_equalize_hist.  It has NO counterpart in the reference: it was added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  Same arguments and same result as uwie_prim_equalize_hist, minus
 * the workspace: the histograms and byte tables live in a static device array inside the library
 * (UWIE_EQ_STATIC_MAX_BATCH x 3840 bytes = 240 KB on every GPU that loads the library, until the process exits).
 * batch > UWIE_EQ_STATIC_MAX_BATCH returns UWIE_ERR_ARG (split the batch).  The call allocates nothing (no cudaMalloc /
 * cudaFree, no allocating library routine); the device of `stream` must be current.  All calls on a device share the
 * array: calls whose GPU work can overlap (different streams without ordering) race and must be serialised by the
 * caller; calls on one stream are safe back to back. */
#define UWIE_EQ_STATIC_MAX_BATCH 64
int uwie_prim_equalize_hist_static(const uint8_t* frames_dev, int batch, int height, int width, uint8_t* out_dev,
                                   void* stream);

/* (img * 255).astype(uint8) of `count` float values (ES:298, SS:204, QA:25). */
int uwie_prim_quantise(const void* img_dev, int dtype, size_t count, uint8_t* out_dev, void* stream);

/* estimate_atmospheric_light(image, min_size=1) (ES:77-206, SS:48-165): quadtree descent on a float image;
 * light_dev receives the (3,) pixel in the image's dtype. */
size_t uwie_prim_atmospheric_light_workspace_bytes(int height, int width);
int uwie_prim_atmospheric_light(const void* img_dev, int dtype, int height, int width, void* light_dev,
                                void* workspace_dev, size_t workspace_bytes, void* stream);

/* compute_Q(block) (ES:147-188, SS:115-157) of a whole float image -> one float64 on the device, and
 * get_brightest_pixel(block) (ES:191-206, SS:159-165) -> 3 values of the image's dtype (workspace: 256 bytes). */
size_t uwie_prim_region_score_workspace_bytes(int height, int width);
int uwie_prim_region_score(const void* img_dev, int dtype, int height, int width, double* q_dev, void* workspace_dev,
                           size_t workspace_bytes, void* stream);
int uwie_prim_brightest_pixel(const void* img_dev, int dtype, int height, int width, void* light_dev, void* workspace_dev,
                              size_t workspace_bytes, void* stream);

/* guided_filter(I, p, r, eps) on float64 planes [H][W] (ES:16-46, SS:25-46); clip_01_1: clip the result to [.1, 1]. */
size_t uwie_prim_guided_filter_workspace_bytes(int height, int width);
int uwie_prim_guided_filter(const double* guide_dev, const double* src_dev, int height, int width, int radius, double eps,
                            int clip_01_1, double* out_dev, void* workspace_dev, size_t workspace_bytes, void* stream);

/* estimate_transmission(img, A, omega, r, eps) -> float64 [H][W] in [.1, 1].  light_dev: 3 values of the image's
 * dtype.  (guard, clip_t0) = (1e-6, 1) is six_stadigy.py:167-180, (1e-10, 0) enhancement_strategies.py:209-234. */
size_t uwie_prim_transmission_workspace_bytes(int height, int width);
int uwie_prim_transmission(const void* img_dev, int dtype, int height, int width, const void* light_dev, double omega,
                           int radius, double eps, double guard, int clip_t0, double* t_dev, void* workspace_dev,
                           size_t workspace_bytes, void* stream);

/* recover_image / restore_image: clip((img - A) / t + A, 0, 1) stored as out_dtype (ES:237-249 float64, SS:182-188 float32). */
int uwie_prim_recover(const void* img_dev, int dtype, int height, int width, const double* t_dev, const void* light_dev,
                      void* out_dev, int out_dtype, void* stream);

/* color_enhancement / enhance_contrast / white_balance: per channel clip((x - p_lo) / (p_hi - p_lo + guard), 0, 1) with the
 * exact np.percentile (ES:252-273, SS:190-219).  The image must lie in [0, 1], else UWIE_ERR_ARG.  Synchronises `stream`. */
size_t uwie_prim_percentile_stretch_workspace_bytes(int dtype);
int uwie_prim_percentile_stretch(const void* img_dev, int dtype, int height, int width, double lo_pct, double hi_pct,
                                 double guard, void* out_dev, void* workspace_dev, size_t workspace_bytes, void* stream);

/* gamma_correction: img ** exponent, clipped to [0, 1] when `clip` (ES:276-285: 1 / gamma, clip; SS:221-224: gamma, no clip). */
int uwie_prim_gamma(const void* img_dev, int dtype, size_t count, double exponent, int clip, void* out_dev, void* stream);

/* color_correction (SS:305-323): clip(img * scale3[c], 0, 1); scale3 is a HOST array. */
int uwie_prim_scale_clip(const void* img_dev, int dtype, int height, int width, const double* scale3, void* out_dev, void* stream);

/* Channel means of a float image (detect_image_type, SS:292-302): 3 float64 on the device.  Workspace: 4096 bytes. */
int uwie_prim_channel_means(const void* img_dev, int dtype, int height, int width, double* means_dev, void* workspace_dev,
                            size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Self-test of the guided filter's branch-free float64 division and reciprocal (k_arith.cuh) against the
 * compiler's IEEE-754 routines: `count` seeded operand pairs in the ranges the kernels guarantee
 * (divisor in [1e-9, 2], dividend 0 or |x| in [1e-18, 2]; reciprocal of t in [0.1, 1]), and of the percentile stretch's
 * invariant-divisor division x / den = RN(q + (x - q * den) * r), r = RN(1 / den) (float32 and float64; den in [guard, 1 + guard],
 * dividend 0 or 2^-30 <= |x| <= 1).
 * *mismatches_host receives the number of results whose bits differ (must be 0).  Synchronous.
 * ------------------------------------------------------------------------------------------- */
int uwie_selftest_division(unsigned long long count, unsigned long long seed, unsigned long long* mismatches_host);

#ifdef __cplusplus
}
#endif
#endif /* UWIE_H */
