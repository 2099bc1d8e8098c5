// Engine: owns the workspace layout and enqueues the kernel sequence of one batch.
// C ABI in include/uwie.h.  No CPU path: every entry point needs a CUDA device.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <array>
#include <cstring>
#include <map>
#include <set>
#include <string>
#include <vector>

#include "../../include/uwie.h"
#include "k_kernels.h"
#include "uwie_tables.h"

using namespace uwie;

namespace {

thread_local std::string g_err;

int fail(int code, const std::string& msg) { g_err = msg; return code; }
}  // namespace
namespace uwie {
int uwie_fail(int code, const std::string& msg) { return fail(code, msg); }   // for the other C-ABI files
}
namespace {

#define CU(call)                                                                          \
    do {                                                                                  \
        cudaError_t _e = (call);                                                          \
        if (_e != cudaSuccess)                                                            \
            return fail(UWIE_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(_e)); \
    } while (0)

struct Bump {
    size_t off = 0;
    size_t take(size_t bytes) {
        const size_t at = off;
        off += (bytes + 255) & ~(size_t)255;
        return at;
    }
};

struct Layout {
    // zeroed every run (contiguous)
    size_t zero_begin, frame_hist, quad_hist, quad_counts, sel_hist_a, sel_hist_b, sel_hist_c, tab_keys,
        tab_counts, cand_n, sel_stats, out_hist, qa,
        edge_counts, zero_end;
    size_t recipes, ctab, status, types, cast, tables, rects, q_trace, light, tile_list, tile_n, labels, ab,
        recovered, targets, cand, gf_gray, gf_dark, tile_lut, out, gray, cf_part, scores, best;
    size_t total;
};

}  // namespace

struct uwie_engine {
    Library lib;
    std::vector<Recipe> recipes;
    int S = 0, B = 0, H = 0, W = 0;
    int n_dehaze = 0;
    int dehaze_idx[kStrategiesMax];
    int dehaze_slot[kStrategiesMax];
    int max_radius = 0, max_tiles = 0;
    int eps_safe = 1;            // every dehaze eps >= 1e-9 (k_guided.cu: division without a range check)
    bool need_hist_pass = false;
    int final_blocks = 0, pix_per_block = 0;
    std::vector<std::pair<int, int>> quad_dims;   // per level: max quadrant rows, cols
    Layout L;
    char* ws = nullptr;
    bool owns_ws = false;
    int last_launches = 0;
    // per-stage timing (uwie_engine_profile): events[i] opens record i, events[i + 1] closes it
    bool prof_on = false;
    struct Mark { const char* name; cudaEvent_t ev; int kernels; };
    std::vector<Mark> marks;
    std::vector<cudaEvent_t> ev_pool;
    // two-slot pipeline of the host entry point
    struct HostPipe {
        static constexpr int kLanes = 4;             // slots = lanes: chunk c uses slot c % slots and runs on lane c % lanes
        bool ready = false;
        int slots = 0;                               // slots allocated so far (2 for a one-lane pipeline: double buffering)
        cudaStream_t s_in = nullptr, s_out = nullptr;
        cudaEvent_t ev_in[kLanes] = {}, ev_run[kLanes] = {}, ev_out[kLanes] = {}, ev_start = nullptr;
        uint8_t* frames[kLanes] = {}; uint8_t* best_img[kLanes] = {}; uint8_t* out_all[kLanes] = {}; char* small[kLanes] = {};
        // pinned landing zone of the small results: a D2H copy into PAGEABLE memory blocks the host
        // until the stream gets there, which would serialise the whole pipeline
        char* h_small = nullptr; size_t h_small_frames = 0;
        // more lanes: chunk c runs on twin engine c % lanes (own workspace, own stream; lane 0 is this engine on the
        // caller's stream), so that the launch- and latency-bound stretches of one chunk (quadtree levels, table
        // set-up, the guided filter) overlap the wide kernels of the others - the same four lanes the device-resident
        // bench runs
        int lanes = 1;
        uwie_engine* twin[kLanes] = {}; bool twin_failed = false;
        cudaStream_t s_run[kLanes] = {}; cudaEvent_t ev_twin[kLanes] = {};
    } pipe;
    // CUDA graphs of the kernel sequence, one per (batch, pointer set): a step of the bench or of the host pipeline
    // presents the same few pointer sets again and again (chunk slices / the two slots), so the ~60 launches of a
    // call are replayed as ONE graph launch.  First sight of a pointer set runs eagerly, the second captures.
    typedef std::array<uintptr_t, 9> GraphKey;
    struct GraphExec { cudaGraphExec_t exec; int launches; };
    std::map<GraphKey, GraphExec> graphs;
    std::set<GraphKey> seen;
    bool use_graphs = true;
    long long graph_replays = 0;
    uwie_library pub_lib;                         // what the engine was created from (for the twin)
    std::vector<uwie_strategy> pub_strategies;
    template <class T> T* at(size_t off) const { return reinterpret_cast<T*>(ws + off); }
};

namespace {

int build_recipes(const uwie_library* lib, const uwie_strategy* st, int n, int H, int W, Library* L,
                  std::vector<Recipe>* out) {
    if (!lib || !st || n < 1 || n > kStrategiesMax || H < 1 || W < 1) return fail(UWIE_ERR_ARG, "bad engine arguments");
    std::memset(L, 0, sizeof(*L));
    L->f64 = lib->f64 ? 1 : 0;
    L->colour_cast = lib->colour_cast ? 1 : 0;
    L->dark_guard = lib->dark_guard;
    L->clip_t0 = lib->clip_t0 ? 1 : 0;
    L->stretch_guard = lib->stretch_guard;
    L->n_strategies = n;
    for (int i = 0; i < 8; ++i) L->weights[i] = lib->weights[i];
    const long long npix = (long long)H * W;
    auto plan = [&](double pct) {
        PercentilePlan p = L->f64 ? percentile_plan_f64(npix, pct) : percentile_plan_f32(npix, pct);
        RankPlan r; r.k = p.k; r.gamma = p.gamma; return r;
    };
    out->clear();
    for (int i = 0; i < n; ++i) {
        const uwie_strategy& s = st[i];
        Recipe r;
        std::memset(&r, 0, sizeof(r));
        if (s.n_pre < 0 || s.n_pre > kMaxStretch || s.n_post < 0 || s.n_post > kMaxStretch)
            return fail(UWIE_ERR_ARG, "at most two stretches before and after the histogram stage");
        if (s.dehaze && (s.radius < 1 || s.radius > 64)) return fail(UWIE_ERR_ARG, "guided radius must be in [1, 64]");
        if (s.dehaze && (s.equalize || s.n_post)) return fail(UWIE_ERR_ARG, "unsupported chain: dehaze + equalize/post-stretch");
        if (s.clahe && s.equalize) return fail(UWIE_ERR_ARG, "unsupported chain: clahe + equalize");
        if (!s.clahe && !s.equalize && !s.dehaze && s.n_post) return fail(UWIE_ERR_ARG, "post-stretch needs a histogram stage");
        r.dehaze = s.dehaze ? 1 : 0;
        r.omega = (float)s.omega;
        r.radius = s.radius;
        r.eps = s.eps;
        r.n_pre = s.n_pre;
        for (int k = 0; k < s.n_pre; ++k) { r.pre[k].lo = plan(s.pre_lo[k]); r.pre[k].hi = plan(s.pre_hi[k]); }
        r.equalize = s.equalize ? 1 : 0;
        r.clahe = s.clahe ? 1 : 0;
        if (r.clahe) {
            if (s.tiles_x < 1 || s.tiles_y < 1 || s.tiles_x * s.tiles_y > kMaxTiles)
                return fail(UWIE_ERR_ARG, "CLAHE tile grid must have between 1 and 256 tiles");
            r.tiles_x = s.tiles_x; r.tiles_y = s.tiles_y;
            int tw, th;
            if (W % r.tiles_x == 0 && H % r.tiles_y == 0) { tw = W / r.tiles_x; th = H / r.tiles_y; }
            else { tw = (W + (r.tiles_x - W % r.tiles_x)) / r.tiles_x; th = (H + (r.tiles_y - H % r.tiles_y)) / r.tiles_y; }
            r.clahe_clip = clahe_clip_count(s.clip_limit, tw * th);
        }
        r.n_post = s.n_post;
        for (int k = 0; k < s.n_post; ++k) { r.post[k].lo = plan(s.post_lo[k]); r.post[k].hi = plan(s.post_hi[k]); }
        r.gamma_on = s.gamma_on ? 1 : 0;
        if (lib->gamma_inverse) { r.gamma_exp = 1.0 / s.gamma; r.gamma_clip = 1; }
        else { r.gamma_exp = s.gamma; r.gamma_clip = 0; }
        if (!L->f64) r.gamma_exp = (double)(float)r.gamma_exp;   // a Python float meets a float32 array
        out->push_back(r);
    }
    return UWIE_OK;
}

void plan_layout(uwie_engine* e) {
    Bump b;
    Layout& L = e->L;
    const size_t B = e->B, S = e->S, N = (size_t)e->H * e->W, nd = e->n_dehaze;
    const size_t tsz = e->lib.f64 ? 8 : 4;
    L.zero_begin = b.off;
    L.frame_hist = b.take(B * 768 * 4);
    L.quad_hist = b.take(nd ? B * kQuadLevelsMax * 4 * 768 * 4 : 0);
    L.quad_counts = b.take(nd ? B * kQuadLevelsMax * 4 * 4 : 0);
    L.sel_hist_a = b.take(B * nd * 3 * kSelBinsA * 4);
    L.sel_hist_b = b.take(B * nd * kSelTargets * kSelBinsB * 4);
    L.sel_hist_c = b.take(B * nd * kSelTargets * kSelBinsC * 4);
    L.tab_keys = b.take(B * nd * kSelTargets * kSelTable * 8);
    L.tab_counts = b.take(B * nd * kSelTargets * kSelTable * 4);
    L.cand_n = b.take(B * nd * kSelTargets * 4);
    L.sel_stats = b.take(256);
    L.out_hist = b.take(B * S * 768 * 4);
    L.qa = b.take(B * S * sizeof(QaAccum));
    L.edge_counts = b.take(B * S * 4);
    L.zero_end = b.off;
    L.recipes = b.take(S * sizeof(Recipe));
    L.ctab = b.take(sizeof(ColorTables));
    L.status = b.take(256);
    L.types = b.take(B * 4);
    L.cast = b.take(B * sizeof(CastTables));
    L.tables = b.take(B * S * sizeof(StratTables));
    L.rects = b.take(B * sizeof(QuadRect));
    L.q_trace = b.take(B * kQuadLevelsMax * 4 * 8);
    L.light = b.take(B * 3 * 4);
    {
        size_t slots = canny_tile_slots(e->H, e->W, (int)(B * S));
        if (nd && !e->quad_dims.empty())
            slots = std::max(slots, canny_tile_slots(e->quad_dims[0].first, e->quad_dims[0].second, (int)(B * 4)));
        L.tile_list = b.take(slots * kCannyListCap * sizeof(int2));
        L.tile_n = b.take(slots * 4);
        L.labels = b.take(canny_label_bytes(slots));
    }
    L.ab = b.take(B * nd * 2 * N * 8);
    L.recovered = b.take(B * nd * N * 3 * tsz);
    L.targets = b.take(B * nd * kSelTargets * sizeof(SelTarget));
    L.cand = b.take(B * nd * kSelTargets * (size_t)kSelListCap * tsz);
    L.gf_gray = b.take(nd ? B * N * 8 : 0);
    L.tile_lut = b.take(e->max_tiles ? B * S * kMaxTiles * 256 : 0);
    L.out = b.take(B * S * N * 3);
    L.gray = b.take(B * S * N);
    L.cf_part = b.take(B * S * (size_t)e->final_blocks * 4 * 8);
    L.scores = b.take(B * S * kScoreCols * 8);
    L.best = b.take(B * 4);
    L.total = b.off;
}

int init_engine_shape(uwie_engine* e, const uwie_library* lib, const uwie_strategy* st, int n, int B, int H, int W) {
    if (B < 1) return fail(UWIE_ERR_ARG, "max_batch must be positive");
    if ((long long)H * W >= (1LL << 30)) return fail(UWIE_ERR_ARG, "frame too large");
    int rc = build_recipes(lib, st, n, H, W, &e->lib, &e->recipes);
    if (rc) return rc;
    e->S = n; e->B = B; e->H = H; e->W = W;
    e->n_dehaze = 0; e->max_radius = 0; e->max_tiles = 0; e->need_hist_pass = false; e->eps_safe = 1;
    for (int i = 0; i < kStrategiesMax; ++i) { e->dehaze_idx[i] = 0; e->dehaze_slot[i] = -1; }
    for (int i = 0; i < n; ++i) {
        const Recipe& r = e->recipes[i];
        if (r.dehaze) {
            e->dehaze_slot[i] = e->n_dehaze;
            e->dehaze_idx[e->n_dehaze++] = i;
            if (r.radius > e->max_radius) e->max_radius = r.radius;
            if (!(r.eps >= 1e-9 && r.eps <= 1e9)) e->eps_safe = 0;
        }
        if (r.clahe) {
            if (r.tiles_x * r.tiles_y > e->max_tiles) e->max_tiles = r.tiles_x * r.tiles_y;
            if (r.n_post > 0) e->need_hist_pass = true;
        }
    }
    const size_t N = (size_t)H * W;
    // multiple of 4: apply_kernel moves 4-pixel groups as words.  32768 (was 8192): the block prologue (13 KB of colour and
    // strategy tables into shared memory) was 12 % of apply_kernel's stall samples with 32 pixels per thread
    e->pix_per_block = 32768;
    e->final_blocks = (int)((N + e->pix_per_block - 1) / e->pix_per_block);
    // quadtree: upper bound of the quadrant size per level (ceil halves), enhancement_strategies.py:104-119
    e->quad_dims.clear();
    int rows = H, cols = W;
    while (rows > 1 && cols > 1 && (int)e->quad_dims.size() < kQuadLevelsMax) {
        rows = rows - rows / 2; cols = cols - cols / 2;
        e->quad_dims.push_back({rows, cols});
    }
    plan_layout(e);
    return UWIE_OK;
}

}  // namespace

// =====================================================================================
// C ABI
// =====================================================================================
extern "C" {

int uwie_version(void) { return UWIE_VERSION; }

int uwie_selftest_division(unsigned long long count, unsigned long long seed, unsigned long long* mismatches_host) {
    if (!mismatches_host) return fail(UWIE_ERR_ARG, "null mismatch counter");
    if (uwie::selftest_division(count, seed, mismatches_host)) return fail(UWIE_ERR_CUDA, "division self-test did not run");
    return UWIE_OK;
}
const char* uwie_last_error(void) { return g_err.c_str(); }

int uwie_preset(int library_id, uwie_library* lib, uwie_strategy* st) {
    if (!lib || !st) return fail(UWIE_ERR_ARG, "null preset output");
    std::memset(lib, 0, sizeof(*lib));
    std::memset(st, 0, sizeof(uwie_strategy) * UWIE_MAX_STRATEGIES);
    auto dehaze = [](uwie_strategy& s, double omega, int r, double eps) { s.dehaze = 1; s.omega = omega; s.radius = r; s.eps = eps; };
    auto pre = [](uwie_strategy& s, double lo, double hi) { s.pre_lo[s.n_pre] = lo; s.pre_hi[s.n_pre] = hi; ++s.n_pre; };
    auto post = [](uwie_strategy& s, double lo, double hi) { s.post_lo[s.n_post] = lo; s.post_hi[s.n_post] = hi; ++s.n_post; };
    auto clahe = [](uwie_strategy& s, double clip) { s.clahe = 1; s.clip_limit = clip; s.tiles_x = 8; s.tiles_y = 8; };
    auto gamma = [](uwie_strategy& s, double g) { s.gamma_on = 1; s.gamma = g; };
    if (library_id == UWIE_LIB_SIX) {
        // six_stadigy.py:230-285; white_balance(p) is the stretch (p, 100 - p) (SS:210-219)
        lib->f64 = 0; lib->colour_cast = 1; lib->dark_guard = 1e-6; lib->clip_t0 = 1;
        lib->stretch_guard = 1e-6; lib->gamma_inverse = 0;
        const double w[8] = {0.20, 0.20, 0.15, 0.15, 0.10, 0.10, 0.05, 0.05};   // QA:229-238
        std::memcpy(lib->weights, w, sizeof(w));
        dehaze(st[0], 0.3, 20, 0.5); pre(st[0], 5, 98); clahe(st[0], 3.0); gamma(st[0], 1.5);
        dehaze(st[1], 0.5, 15, 0.5); pre(st[1], 15, 95); clahe(st[1], 2.0);
        dehaze(st[2], 0.7, 10, 0.1); pre(st[2], 20, 85); pre(st[2], 2, 98);
        clahe(st[3], 4.0); post(st[3], 10, 95); post(st[3], 3, 97); gamma(st[3], 1.3);
        pre(st[4], 2, 98); pre(st[4], 15, 90); clahe(st[4], 1.5); gamma(st[4], 1.2);
        pre(st[5], 5, 98); clahe(st[5], 3.5); gamma(st[5], 1.4);
        return 6;
    }
    if (library_id == UWIE_LIB_ES) {
        // config.py:28-85 through enhancement_strategies.py:349-474 (eps stays at its 1e-3 default)
        lib->f64 = 1; lib->colour_cast = 0; lib->dark_guard = 1e-10; lib->clip_t0 = 0;
        lib->stretch_guard = 1e-10; lib->gamma_inverse = 1;
        const double w[8] = {0.25, 0.20, 0.15, 0.15, 0.15, 0.10, 0.0, 0.0};     // config.py:78-85
        std::memcpy(lib->weights, w, sizeof(w));
        dehaze(st[0], 0.5, 15, 1e-3); pre(st[0], 10, 95); gamma(st[0], 1.2);
        dehaze(st[1], 0.6, 20, 1e-3); pre(st[1], 15, 92); gamma(st[1], 1.2);
        dehaze(st[2], 0.4, 10, 1e-3); pre(st[2], 15, 95);
        clahe(st[3], 2.0); post(st[3], 20, 85);
        st[4].equalize = 1; post(st[4], 10, 95);
        return 5;
    }
    return fail(UWIE_ERR_ARG, "unknown library id");
}

size_t uwie_workspace_bytes(const uwie_library* lib, const uwie_strategy* st, int n, int B, int H, int W) {
    uwie_engine tmp;
    if (init_engine_shape(&tmp, lib, st, n, B, H, W)) return 0;
    return tmp.L.total;
}

int uwie_engine_create(uwie_engine** out, const uwie_library* lib, const uwie_strategy* st, int n, int B,
                       int H, int W, void* workspace_dev, size_t workspace_bytes) {
    if (!out) return fail(UWIE_ERR_ARG, "null engine output");
    *out = nullptr;
    uwie_engine* e = new uwie_engine();
    int rc = init_engine_shape(e, lib, st, n, B, H, W);
    if (rc) { delete e; return rc; }
    e->pub_lib = *lib;
    e->pub_strategies.assign(st, st + n);
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) {
        delete e;
        return fail(UWIE_ERR_CUDA, "no CUDA device: this library has no CPU path");
    }
    if (workspace_dev) {
        if (workspace_bytes < e->L.total) { delete e; return fail(UWIE_ERR_NOMEM, "workspace too small"); }
        e->ws = static_cast<char*>(workspace_dev);
    } else {
        if (cudaMalloc(&e->ws, e->L.total) != cudaSuccess) {
            cudaGetLastError();
            delete e;
            return fail(UWIE_ERR_NOMEM, "cudaMalloc of the workspace failed");
        }
        e->owns_ws = true;
    }
    ColorTables tabs;
    build_color_tables(&tabs);
    cudaError_t ce = cudaMemcpy(e->ws + e->L.ctab, &tabs, sizeof(tabs), cudaMemcpyHostToDevice);
    if (ce == cudaSuccess) ce = cudaMemcpy(e->ws + e->L.recipes, e->recipes.data(), sizeof(Recipe) * e->S, cudaMemcpyHostToDevice);
    if (ce == cudaSuccess) ce = cudaMemset(e->ws + e->L.status, 0, 256);
    if (ce != cudaSuccess) {
        std::string msg = std::string("engine upload: ") + cudaGetErrorString(ce);
        uwie_engine_destroy(e);
        return fail(UWIE_ERR_CUDA, msg);
    }
    *out = e;
    return UWIE_OK;
}

void uwie_engine_destroy(uwie_engine* e) {
    if (!e) return;
    if (e->owns_ws && e->ws) cudaFree(e->ws);
    {
        auto& hp = e->pipe;
        for (int k = 0; k < uwie_engine::HostPipe::kLanes; ++k) {
            if (hp.frames[k]) cudaFree(hp.frames[k]);
            if (hp.best_img[k]) cudaFree(hp.best_img[k]);
            if (hp.out_all[k]) cudaFree(hp.out_all[k]);
            if (hp.small[k]) cudaFree(hp.small[k]);
            if (hp.ev_in[k]) cudaEventDestroy(hp.ev_in[k]);
            if (hp.ev_run[k]) cudaEventDestroy(hp.ev_run[k]);
            if (hp.ev_out[k]) cudaEventDestroy(hp.ev_out[k]);
            if (hp.twin[k]) uwie_engine_destroy(hp.twin[k]);
            if (hp.s_run[k]) cudaStreamDestroy(hp.s_run[k]);
            if (hp.ev_twin[k]) cudaEventDestroy(hp.ev_twin[k]);
        }
        if (hp.ev_start) cudaEventDestroy(hp.ev_start);
        if (hp.h_small) cudaFreeHost(hp.h_small);
        if (hp.s_in) cudaStreamDestroy(hp.s_in);
        if (hp.s_out) cudaStreamDestroy(hp.s_out);
    }
    for (auto& m : e->marks) cudaEventDestroy(m.ev);
    for (auto& ev : e->ev_pool) cudaEventDestroy(ev);
    for (auto& g : e->graphs) cudaGraphExecDestroy(g.second.exec);
    delete e;
}

int uwie_engine_last_launches(const uwie_engine* e) { return e ? e->last_launches : 0; }
long long uwie_engine_graph_replays(const uwie_engine* e) {
    if (!e) return 0;
    long long n = e->graph_replays;
    for (int k = 1; k < uwie_engine::HostPipe::kLanes; ++k) if (e->pipe.twin[k]) n += e->pipe.twin[k]->graph_replays;
    return n;
}

}  // extern "C"

// =====================================================================================
// the kernel sequence of one batch
// =====================================================================================
namespace {

__global__ void gather_best_kernel(const uint8_t* __restrict__ all, const int* __restrict__ best, int S,
                                   size_t bytes_per_image, uint8_t* __restrict__ out) {
    const int b = blockIdx.y;
    const uint8_t* src = all + ((size_t)b * S + best[b]) * bytes_per_image;
    uint8_t* dst = out + (size_t)b * bytes_per_image;
    const size_t n16 = bytes_per_image / 16;
    const bool vec = ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0;
    if (vec) {
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x)
            reinterpret_cast<uint4*>(dst)[i] = reinterpret_cast<const uint4*>(src)[i];
        for (size_t i = n16 * 16 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < bytes_per_image;
             i += (size_t)gridDim.x * blockDim.x)
            dst[i] = src[i];
    } else {
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < bytes_per_image; i += (size_t)gridDim.x * blockDim.x)
            dst[i] = src[i];
    }
}

// Stage bracket: closes the previous stage (kernel count = launches since its mark) and opens `name`.
// name == nullptr only closes.  No-op unless profiling is on.
void stage(uwie_engine* e, cudaStream_t st, const char* name, int launches_so_far) {
    if (!e->prof_on) return;
    if (!e->marks.empty() && e->marks.back().name) e->marks.back().kernels = launches_so_far - e->marks.back().kernels;
    cudaEvent_t ev;
    if (!e->ev_pool.empty()) { ev = e->ev_pool.back(); e->ev_pool.pop_back(); }
    else if (cudaEventCreate(&ev) != cudaSuccess) { e->prof_on = false; return; }
    cudaEventRecord(ev, st);
    e->marks.push_back({name, ev, launches_so_far});
}

int enqueue_batch(uwie_engine* e, const uint8_t* frames, int batch, uint8_t* out_all, uint8_t* out_best,
                  double* scores, int32_t* best, int32_t* types, float* light, cudaStream_t st) {
    const Layout& L = e->L;
    const int S = e->S, H = e->H, W = e->W;
    const size_t N = (size_t)H * W;
    int launches = 0;
    stage(e, st, "zero", launches);
    CU(cudaMemsetAsync(e->ws + L.zero_begin, 0, L.zero_end - L.zero_begin, st));

    uint8_t* out_buf = out_all ? out_all : e->at<uint8_t>(L.out);

    stage(e, st, "frame_hist", launches);
    launches += launch_frame_hist(frames, batch, H, W, e->at<unsigned>(L.frame_hist), st);

    SetupArgs sa;
    sa.B = batch; sa.S = S; sa.n = (long long)N; sa.lib = e->lib;
    sa.recipes = e->at<Recipe>(L.recipes);
    sa.frame_hist = e->at<unsigned>(L.frame_hist);
    sa.out_hist = e->at<unsigned>(L.out_hist);
    sa.types = e->at<int>(L.types);
    sa.cast = e->at<CastTables>(L.cast);
    sa.tables = e->at<StratTables>(L.tables);
    stage(e, st, "setup_luts", launches);
    launches += launch_setup_cast(sa, st);
    launches += launch_setup_frame_luts(sa, st);

    if (e->n_dehaze > 0) {
        QuadArgs qa;
        qa.B = batch; qa.H = H; qa.W = W; qa.frames = frames; qa.cast = sa.cast;
        qa.rects = e->at<QuadRect>(L.rects);
        qa.quad_hist = e->at<unsigned>(L.quad_hist);
        qa.counts = e->at<unsigned>(L.quad_counts);
        qa.q_trace = e->at<double>(L.q_trace);
        qa.light = e->at<float>(L.light);
        stage(e, st, "quadtree_canny", launches);
        launches += launch_quad_init(qa, st);
        CannyArgs ca;
        std::memset(&ca, 0, sizeof(ca));
        ca.mode = CANNY_QUADRANTS; ca.H = H; ca.W = W; ca.low = 50; ca.high = 150;
        ca.frames = frames; ca.cast = sa.cast; ca.rects = qa.rects;
        ca.quad_hist = e->at<unsigned>(L.quad_hist);
        ca.labels = e->at<int>(L.labels);
        ca.tile_list = e->at<int2>(L.tile_list); ca.tile_n = e->at<int>(L.tile_n);
        ca.counts = e->at<unsigned>(L.quad_counts);
        for (size_t lv = 0; lv < e->quad_dims.size(); ++lv) {
            ca.level = (int)lv;
            launches += launch_canny(ca, e->quad_dims[lv].first, e->quad_dims[lv].second, batch * 4, st);
            launches += launch_quad_descend(qa, (int)lv, st);
        }
        launches += launch_quad_brightest(qa, st);

        GuidedArgs ga;
        std::memset(&ga, 0, sizeof(ga));
        ga.B = batch; ga.S = S; ga.H = H; ga.W = W; ga.lib = e->lib; ga.recipes = sa.recipes;
        ga.n_dehaze = e->n_dehaze;
        for (int i = 0; i < kStrategiesMax; ++i) ga.dehaze_idx[i] = e->dehaze_idx[i];
        ga.rows_per_block = 0;                                // chosen by launch_guided (strip and chunk geometry live there)
        ga.gd_plane = e->at<uint2>(L.gf_gray);
        ga.eps_safe = e->eps_safe;
        ga.frames = frames; ga.cast = sa.cast; ga.light = qa.light;
        ga.ab = e->at<double2>(L.ab); ga.recovered = e->at<void>(L.recovered);
        ga.sel_hist_a = e->at<unsigned>(L.sel_hist_a);
        stage(e, st, "guided_recover", launches);
        launches += launch_guided(ga, e->max_radius, st);

        SelectArgs se;
        std::memset(&se, 0, sizeof(se));
        se.B = batch; se.S = S; se.n = (long long)N; se.lib = e->lib; se.recipes = sa.recipes;
        se.n_dehaze = e->n_dehaze;
        for (int i = 0; i < kStrategiesMax; ++i) se.dehaze_idx[i] = e->dehaze_idx[i];
        se.recovered = ga.recovered; se.sel_hist_a = ga.sel_hist_a;
        se.sel_hist_b = e->at<unsigned>(L.sel_hist_b);
        se.targets = e->at<SelTarget>(L.targets);
        se.sel_hist_c = e->at<unsigned>(L.sel_hist_c);
        se.tab_keys = e->at<unsigned long long>(L.tab_keys);
        se.tab_counts = e->at<unsigned>(L.tab_counts);
        se.cand = e->at<void>(L.cand);
        se.cand_n = e->at<unsigned>(L.cand_n);
        se.stats = e->at<unsigned>(L.sel_stats);
        se.tables = sa.tables;
        se.status = e->at<int>(L.status);
        stage(e, st, "percentile_select", launches);
        launches += launch_select(se, st);
    }

    ApplyArgs aa;
    std::memset(&aa, 0, sizeof(aa));
    aa.B = batch; aa.S = S; aa.H = H; aa.W = W; aa.lib = e->lib; aa.recipes = sa.recipes;
    aa.n_dehaze = e->n_dehaze;
    for (int i = 0; i < kStrategiesMax; ++i) aa.dehaze_slot[i] = e->dehaze_slot[i];
    for (int i = 0; i < S; ++i) {
        // a CLAHE strategy continues from the Lab bytes of clahe_lut_kernel, with or without a dehaze step before: variant 2
        const int v = e->recipes[i].clahe ? 2 : (e->recipes[i].dehaze ? 1 : 0);
        aa.var_s[v][aa.var_n[v]++] = i;
        const bool two_pass = e->recipes[i].clahe && e->recipes[i].n_post > 0;
        if (two_pass) ++aa.var_hist[v];
        const int fv = two_pass ? 0 : v;                   // the histogram pass leaves the CLAHE output in `out`
        aa.fin_s[fv][aa.fin_n[fv]++] = i;
        aa.from_out[i] = two_pass ? 1 : 0;
    }
    aa.frames = frames; aa.tables = sa.tables; aa.recovered = e->at<void>(L.recovered);
    aa.ctab = e->at<ColorTables>(L.ctab);
    aa.tile_lut = e->at<uint8_t>(L.tile_lut);
    aa.out_hist = e->at<unsigned>(L.out_hist);
    aa.out = out_buf; aa.gray = e->at<uint8_t>(L.gray);
    aa.qa = e->at<QaAccum>(L.qa); aa.cf_part = e->at<double>(L.cf_part);
    aa.pix_per_block = e->pix_per_block; aa.final_blocks = e->final_blocks;
    stage(e, st, "clahe_tile_luts", launches);
    launches += launch_clahe_luts(aa, e->max_tiles, st);
    if (e->need_hist_pass) {
        stage(e, st, "clahe_out_hist", launches);
        launches += launch_apply(aa, false, st);
        launches += launch_setup_post_luts(sa, st);
    }
    stage(e, st, "final_apply", launches);
    launches += launch_apply(aa, true, st);

    CannyArgs cq;
    std::memset(&cq, 0, sizeof(cq));
    cq.mode = CANNY_PLANES; cq.H = H; cq.W = W; cq.low = 50; cq.high = 150;
    cq.gray = aa.gray; cq.planes_per_img = S;
    cq.labels = e->at<int>(L.labels);
    cq.tile_list = e->at<int2>(L.tile_list); cq.tile_n = e->at<int>(L.tile_n);
    cq.counts = e->at<unsigned>(L.edge_counts);
    stage(e, st, "score_canny", launches);
    launches += launch_canny(cq, H, W, batch * S, st);

    ScoreArgs sc;
    sc.B = batch; sc.S = S; sc.n = (long long)N; sc.lib = e->lib;
    sc.qa = aa.qa; sc.cf_part = aa.cf_part; sc.final_blocks = e->final_blocks;
    sc.edge_counts = e->at<unsigned>(L.edge_counts);
    sc.scores = e->at<double>(L.scores); sc.best = e->at<int>(L.best);
    stage(e, st, "scores", launches);
    launches += launch_scores(sc, st);

    stage(e, st, "gather_copy", launches);
    if (out_best) {
        gather_best_kernel<<<dim3(64, batch), 256, 0, st>>>(out_buf, sc.best, S, N * 3, out_best);
        ++launches;
    }
    if (scores) CU(cudaMemcpyAsync(scores, sc.scores, sizeof(double) * batch * S * kScoreCols, cudaMemcpyDeviceToDevice, st));
    if (best) CU(cudaMemcpyAsync(best, sc.best, sizeof(int) * batch, cudaMemcpyDeviceToDevice, st));
    if (types) CU(cudaMemcpyAsync(types, sa.types, sizeof(int) * batch, cudaMemcpyDeviceToDevice, st));
    if (light) CU(cudaMemcpyAsync(light, e->at<float>(L.light), sizeof(float) * batch * 3, cudaMemcpyDeviceToDevice, st));
    stage(e, st, nullptr, launches);
    CU(cudaGetLastError());
    e->last_launches = launches;
    return UWIE_OK;
}

// enqueue_batch through the graph cache (see uwie_engine::graphs).  Streams that cannot be captured (the legacy and
// per-thread default streams), profiling runs and UWIE_NO_GRAPH=1 take the eager path.
int run_batch(uwie_engine* e, const uint8_t* frames, int batch, uint8_t* out_all, uint8_t* out_best, double* scores,
              int32_t* best, int32_t* types, float* light, cudaStream_t st) {
    static const bool disabled = std::getenv("UWIE_NO_GRAPH") != nullptr;
    const bool capturable = st != nullptr && st != cudaStreamLegacy && st != cudaStreamPerThread;
    if (disabled || !e->use_graphs || e->prof_on || !capturable)
        return enqueue_batch(e, frames, batch, out_all, out_best, scores, best, types, light, st);
    const uwie_engine::GraphKey key = {(uintptr_t)frames, (uintptr_t)out_all, (uintptr_t)out_best, (uintptr_t)scores,
                                       (uintptr_t)best, (uintptr_t)types, (uintptr_t)light, (uintptr_t)batch, 0};
    auto it = e->graphs.find(key);
    if (it != e->graphs.end()) {
        CU(cudaGraphLaunch(it->second.exec, st));
        e->last_launches = it->second.launches;
        ++e->graph_replays;
        return UWIE_OK;
    }
    if (e->seen.insert(key).second || e->graphs.size() >= 64)          // first sight (or cache full): eager
        return enqueue_batch(e, frames, batch, out_all, out_best, scores, best, types, light, st);
    if (cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
        cudaGetLastError();
        e->use_graphs = false;
        return enqueue_batch(e, frames, batch, out_all, out_best, scores, best, types, light, st);
    }
    const int rc = enqueue_batch(e, frames, batch, out_all, out_best, scores, best, types, light, st);
    cudaGraph_t g = nullptr;
    const cudaError_t ce = cudaStreamEndCapture(st, &g);
    cudaGraphExec_t exec = nullptr;
    if (rc == UWIE_OK && ce == cudaSuccess && g && cudaGraphInstantiate(&exec, g, 0) == cudaSuccess) {
        cudaGraphDestroy(g);
        e->graphs[key] = {exec, e->last_launches};
        CU(cudaGraphLaunch(exec, st));
        ++e->graph_replays;
        return UWIE_OK;
    }
    if (g) cudaGraphDestroy(g);
    cudaGetLastError();
    e->use_graphs = false;                                            // capture is not possible here: stay eager
    return enqueue_batch(e, frames, batch, out_all, out_best, scores, best, types, light, st);
}

}  // namespace

extern "C" {

int uwie_enhance_score(uwie_engine* e, const uint8_t* frames_dev, int batch, uint8_t* out_all_dev,
                       uint8_t* out_best_dev, double* scores_dev, int32_t* best_dev, int32_t* type_dev,
                       float* light_dev, void* stream) {
    if (!e || !frames_dev) return fail(UWIE_ERR_ARG, "null engine or frames");
    if (batch < 1 || batch > e->B) return fail(UWIE_ERR_ARG, "batch must be in [1, max_batch]");
    return run_batch(e, frames_dev, batch, out_all_dev, out_best_dev, scores_dev, best_dev, type_dev,
                     light_dev, static_cast<cudaStream_t>(stream));
}

// detect_image_type (six_stadigy.py:292-302) of a batch: channel histograms + the colour-cast set-up kernel only.
int uwie_engine_image_types(uwie_engine* e, const uint8_t* frames_dev, int batch, int32_t* type_dev, void* stream) {
    if (!e || !frames_dev || !type_dev) return fail(UWIE_ERR_ARG, "null engine, frames or output");
    if (batch < 1 || batch > e->B) return fail(UWIE_ERR_ARG, "batch must be in [1, max_batch]");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const Layout& L = e->L;
    CU(cudaMemsetAsync(e->ws + L.frame_hist, 0, (size_t)batch * 768 * 4, st));
    launch_frame_hist(frames_dev, batch, e->H, e->W, e->at<unsigned>(L.frame_hist), st);
    SetupArgs sa;
    sa.B = batch; sa.S = e->S; sa.n = (long long)e->H * e->W; sa.lib = e->lib;
    sa.lib.colour_cast = 1;                                  // the type is wanted even by an engine that does not correct
    sa.recipes = e->at<Recipe>(L.recipes);
    sa.frame_hist = e->at<unsigned>(L.frame_hist);
    sa.out_hist = e->at<unsigned>(L.out_hist);
    sa.types = e->at<int>(L.types);
    sa.cast = e->at<CastTables>(L.cast);
    sa.tables = e->at<StratTables>(L.tables);
    launch_setup_cast(sa, st);
    CU(cudaMemcpyAsync(type_dev, sa.types, sizeof(int) * batch, cudaMemcpyDeviceToDevice, st));
    CU(cudaGetLastError());
    return UWIE_OK;
}

int uwie_engine_check(uwie_engine* e, void* stream) {
    if (!e) return fail(UWIE_ERR_ARG, "null engine");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int status = 0;
    CU(cudaMemcpyAsync(&status, e->ws + e->L.status, sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (status) {
        // read-and-clear: the failure is reported once, for the calls since the last check
        CU(cudaMemsetAsync(e->ws + e->L.status, 0, sizeof(int), st));
        char buf[128];
        std::snprintf(buf, sizeof(buf), "percentile selection failed on the device (status bits %d)", status);
        return fail(UWIE_ERR_SELECT, buf);
    }
    return UWIE_OK;
}

// Host entry point: a two-slot pipeline.  Slot k owns device copies of a chunk's frames and of
// every result; the H2D copy of chunk i+1 and the D2H copies of chunk i-1 run on their own streams
// while the kernels of chunk i run on the caller's stream (the workspace is single, so kernels of
// different chunks never overlap each other).  With pinned host buffers the copies are truly
// asynchronous; with pageable ones CUDA stages them and the result is the same, only slower.
//
// Every allocation of that pipeline (device slots, the pinned landing zone of the small results,
// streams, events and the second lane's engine) happens in uwie_engine_host_prepare; the call
// itself allocates nothing and fails with UWIE_ERR_ARG when the engine was not prepared for it.
int uwie_engine_host_prepare(uwie_engine* e, int max_count, int want_all, int want_best) {
    if (!e) return fail(UWIE_ERR_ARG, "null engine");
    if (max_count < 1) return fail(UWIE_ERR_ARG, "max_count must be positive");
    const size_t N = (size_t)e->H * e->W, S = e->S, B = e->B;
    uwie_engine::HostPipe& hp = e->pipe;
    // lanes: one per chunk of the largest call, at most four (UWIE_HOST_LANES=n overrides, UWIE_HOST_ONE_LANE=1 forces one)
    int want_lanes = (max_count + (int)B - 1) / (int)B;
    if (want_lanes > uwie_engine::HostPipe::kLanes) want_lanes = uwie_engine::HostPipe::kLanes;
    if (const char* v = std::getenv("UWIE_HOST_LANES")) {
        const int n = std::atoi(v);
        if (n >= 1 && n < want_lanes) want_lanes = n;
    }
    if (std::getenv("UWIE_HOST_ONE_LANE")) want_lanes = 1;
    const int want_slots = want_lanes > 2 ? want_lanes : 2;
    if (!hp.ready) {
        CU(cudaStreamCreateWithFlags(&hp.s_in, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&hp.s_out, cudaStreamNonBlocking));
        CU(cudaEventCreateWithFlags(&hp.ev_start, cudaEventDisableTiming));
        hp.ready = true;
    }
    for (int k = hp.slots; k < want_slots; ++k) {
        CU(cudaEventCreateWithFlags(&hp.ev_in[k], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&hp.ev_run[k], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&hp.ev_out[k], cudaEventDisableTiming));
        CU(cudaMalloc(&hp.frames[k], B * N * 3));
        CU(cudaMalloc(&hp.small[k], B * (S * kScoreCols * 8 + 4 + 4 + 12) + 256));
        hp.slots = k + 1;
    }
    for (int k = 0; k < hp.slots; ++k) {
        if (want_best && !hp.best_img[k]) CU(cudaMalloc(&hp.best_img[k], B * N * 3));
        if (want_all && !hp.out_all[k]) CU(cudaMalloc(&hp.out_all[k], B * S * N * 3));   // S times a chunk: only if somebody asks for it
    }
    const size_t rec_bytes = S * kScoreCols * 8 + 4 + 4 + 12;
    if (hp.h_small_frames < (size_t)max_count) {
        if (hp.h_small) cudaFreeHost(hp.h_small);
        hp.h_small = nullptr; hp.h_small_frames = 0;
        CU(cudaMallocHost(&hp.h_small, (size_t)max_count * rec_bytes + 64));
        hp.h_small_frames = max_count;
    }
    // twin engines (see HostPipe): only worth their workspaces when there is more than one chunk
    for (int k = hp.lanes; k < want_lanes && !hp.twin_failed; ++k) {
        if (uwie_engine_create(&hp.twin[k], &e->pub_lib, e->pub_strategies.data(), (int)S, (int)B, e->H, e->W, nullptr, 0) != UWIE_OK ||
            cudaStreamCreateWithFlags(&hp.s_run[k], cudaStreamNonBlocking) != cudaSuccess ||
            cudaEventCreateWithFlags(&hp.ev_twin[k], cudaEventDisableTiming) != cudaSuccess) {
            cudaGetLastError();
            if (hp.twin[k]) { uwie_engine_destroy(hp.twin[k]); hp.twin[k] = nullptr; }
            hp.twin_failed = true;                   // e.g. no memory for another workspace: the lanes so far it is
        } else {
            hp.lanes = k + 1;
        }
    }
    return UWIE_OK;
}

int uwie_enhance_score_host(uwie_engine* e, const uint8_t* frames_host, int count, uint8_t* out_all_host,
                            uint8_t* out_best_host, double* scores_host, int32_t* best_host,
                            int32_t* type_host, float* light_host, void* stream) {
    if (!e || !frames_host) return fail(UWIE_ERR_ARG, "null engine or frames");
    if (count < 1) return fail(UWIE_ERR_ARG, "count must be positive");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t N = (size_t)e->H * e->W, S = e->S, B = e->B;
    uwie_engine::HostPipe& hp = e->pipe;
    if (!hp.ready || hp.slots < 2 || hp.h_small_frames < (size_t)count || (out_all_host && !hp.out_all[hp.slots - 1]) ||
        (out_best_host && !hp.best_img[hp.slots - 1]))
        return fail(UWIE_ERR_ARG, "uwie_enhance_score_host: call uwie_engine_host_prepare(engine, max_count, want_all, "
                                  "want_best) first (the call itself allocates nothing)");
    double* p_scores = reinterpret_cast<double*>(hp.h_small);
    int32_t* p_best = reinterpret_cast<int32_t*>(p_scores + (size_t)count * S * kScoreCols);
    int32_t* p_types = p_best + count;
    float* p_light = reinterpret_cast<float*>(p_types + count);
    const int lanes = (!e->prof_on && count > (int)B) ? hp.lanes : 1;
    const int slots = lanes > 2 ? lanes : 2;
    // the copy streams must not start before work already queued on the caller's stream
    CU(cudaEventRecord(hp.ev_start, st));
    CU(cudaStreamWaitEvent(hp.s_in, hp.ev_start, 0));
    CU(cudaStreamWaitEvent(hp.s_out, hp.ev_start, 0));
    for (int l = 1; l < lanes; ++l) CU(cudaStreamWaitEvent(hp.s_run[l], hp.ev_start, 0));
    int launches_total = 0;
    int chunk = 0;
    for (int done = 0; done < count; done += (int)B, ++chunk) {
        const int k = chunk % slots, ln = chunk % lanes;
        const int nb = count - done < (int)B ? count - done : (int)B;
        double* d_scores = reinterpret_cast<double*>(hp.small[k]);
        int32_t* d_best = reinterpret_cast<int32_t*>(d_scores + B * S * kScoreCols);
        int32_t* d_types = d_best + B;
        float* d_light = reinterpret_cast<float*>(d_types + B);
        // H2D: the slot's frames are free once the kernels of chunk - slots are done
        if (chunk >= slots) CU(cudaStreamWaitEvent(hp.s_in, hp.ev_run[k], 0));
        CU(cudaMemcpyAsync(hp.frames[k], frames_host + (size_t)done * N * 3, (size_t)nb * N * 3,
                           cudaMemcpyHostToDevice, hp.s_in));
        CU(cudaEventRecord(hp.ev_in[k], hp.s_in));
        // kernels: need the frames, and the slot's result buffers drained (D2H of chunk - slots)
        uwie_engine* lane = ln ? hp.twin[ln] : e;
        cudaStream_t ls = ln ? hp.s_run[ln] : st;
        CU(cudaStreamWaitEvent(ls, hp.ev_in[k], 0));
        if (chunk >= slots) CU(cudaStreamWaitEvent(ls, hp.ev_out[k], 0));
        int rc = run_batch(lane, hp.frames[k], nb, out_all_host ? hp.out_all[k] : nullptr,
                           out_best_host ? hp.best_img[k] : nullptr, d_scores, d_best, d_types, d_light, ls);
        if (rc) return rc;
        launches_total += lane->last_launches;
        CU(cudaEventRecord(hp.ev_run[k], ls));
        // D2H
        CU(cudaStreamWaitEvent(hp.s_out, hp.ev_run[k], 0));
        if (out_all_host)
            CU(cudaMemcpyAsync(out_all_host + (size_t)done * S * N * 3, hp.out_all[k], (size_t)nb * S * N * 3, cudaMemcpyDeviceToHost, hp.s_out));
        if (out_best_host)
            CU(cudaMemcpyAsync(out_best_host + (size_t)done * N * 3, hp.best_img[k], (size_t)nb * N * 3, cudaMemcpyDeviceToHost, hp.s_out));
        if (scores_host)
            CU(cudaMemcpyAsync(p_scores + (size_t)done * S * kScoreCols, d_scores, sizeof(double) * nb * S * kScoreCols, cudaMemcpyDeviceToHost, hp.s_out));
        if (best_host) CU(cudaMemcpyAsync(p_best + done, d_best, sizeof(int) * nb, cudaMemcpyDeviceToHost, hp.s_out));
        if (type_host) CU(cudaMemcpyAsync(p_types + done, d_types, sizeof(int) * nb, cudaMemcpyDeviceToHost, hp.s_out));
        if (light_host) CU(cudaMemcpyAsync(p_light + (size_t)done * 3, d_light, sizeof(float) * nb * 3, cudaMemcpyDeviceToHost, hp.s_out));
        CU(cudaEventRecord(hp.ev_out[k], hp.s_out));
    }
    for (int l = 1; l < lanes; ++l) {                 // the caller's stream is done when all lanes are
        CU(cudaEventRecord(hp.ev_twin[l], hp.s_run[l]));
        CU(cudaStreamWaitEvent(st, hp.ev_twin[l], 0));
    }
    CU(cudaStreamSynchronize(hp.s_out));
    e->last_launches = launches_total;
    for (int l = 1; l < lanes; ++l) {
        const int rc2 = uwie_engine_check(hp.twin[l], hp.s_run[l]);
        if (rc2) return rc2;
    }
    if (scores_host) std::memcpy(scores_host, p_scores, sizeof(double) * (size_t)count * S * kScoreCols);
    if (best_host) std::memcpy(best_host, p_best, sizeof(int) * (size_t)count);
    if (type_host) std::memcpy(type_host, p_types, sizeof(int) * (size_t)count);
    if (light_host) std::memcpy(light_host, p_light, sizeof(float) * (size_t)count * 3);
    return uwie_engine_check(e, stream);
}

int uwie_engine_profile(uwie_engine* e, int enable) {
    if (!e) return fail(UWIE_ERR_ARG, "null engine");
    for (auto& m : e->marks) e->ev_pool.push_back(m.ev);
    e->marks.clear();
    e->prof_on = enable != 0;
    return UWIE_OK;
}

int uwie_engine_profile_read(uwie_engine* e, char* names, double* ms, int* kernels, int* runs, int cap) {
    if (!e || !names || !ms || !kernels || !runs || cap < 1) return fail(UWIE_ERR_ARG, "bad profile_read arguments");
    int n = 0;
    for (size_t i = 0; i + 1 < e->marks.size(); ++i) {
        const auto& m = e->marks[i];
        if (!m.name) continue;                       // closing mark of a call
        CU(cudaEventSynchronize(e->marks[i + 1].ev));
        float t = 0.f;
        CU(cudaEventElapsedTime(&t, m.ev, e->marks[i + 1].ev));
        int k = 0;
        while (k < n && std::strncmp(names + (size_t)k * UWIE_STAGE_NAME_LEN, m.name, UWIE_STAGE_NAME_LEN - 1) != 0) ++k;
        if (k == n) {
            if (n == cap) continue;
            std::memset(names + (size_t)k * UWIE_STAGE_NAME_LEN, 0, UWIE_STAGE_NAME_LEN);
            std::strncpy(names + (size_t)k * UWIE_STAGE_NAME_LEN, m.name, UWIE_STAGE_NAME_LEN - 1);
            ms[k] = 0.0; kernels[k] = 0; runs[k] = 0;
            ++n;
        }
        ms[k] += t; kernels[k] += m.kernels; runs[k] += 1;
    }
    return n;
}

const void* uwie_engine_debug(const uwie_engine* e, const char* name, size_t* bytes) {
    if (!e || !name) return nullptr;
    const Layout& L = e->L;
    const size_t B = e->B, S = e->S, N = (size_t)e->H * e->W, nd = e->n_dehaze;
    const size_t tsz = e->lib.f64 ? 8 : 4;
    struct Item { const char* n; size_t off, sz; };
    const Item items[] = {
        {"frame_hist", L.frame_hist, B * 768 * 4},
        {"cast", L.cast, B * sizeof(CastTables)},
        {"tables", L.tables, B * S * sizeof(StratTables)},
        {"rects", L.rects, B * sizeof(QuadRect)},
        {"q_trace", L.q_trace, B * kQuadLevelsMax * 4 * 8},
        {"quad_counts", L.quad_counts, nd ? B * kQuadLevelsMax * 4 * 4 : 0},
        {"quad_hist", L.quad_hist, nd ? B * kQuadLevelsMax * 4 * 768 * 4 : 0},
        {"light", L.light, B * 3 * 4},
        {"ab", L.ab, B * nd * 2 * N * 8},
        {"recovered", L.recovered, B * nd * N * 3 * tsz},
        {"tile_lut", L.tile_lut, e->max_tiles ? B * S * kMaxTiles * 256 : 0},
        {"out_hist", L.out_hist, B * S * 768 * 4},
        {"out", L.out, B * S * N * 3},
        {"gray", L.gray, B * S * N},
        {"qa", L.qa, B * S * sizeof(QaAccum)},
        {"edge_counts", L.edge_counts, B * S * 4},
        {"targets", L.targets, B * nd * kSelTargets * sizeof(SelTarget)},
        {"sel_stats", L.sel_stats, 8},
        {"scores", L.scores, B * S * kScoreCols * 8},
    };
    for (const Item& it : items) {
        if (std::strcmp(it.n, name) == 0) {
            if (bytes) *bytes = it.sz;
            return e->ws + it.off;
        }
    }
    return nullptr;
}

}  // extern "C"
