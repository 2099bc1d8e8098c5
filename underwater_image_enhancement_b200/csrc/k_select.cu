// Exact np.percentile on float data (the recovered frame of the dehaze strategies):
//   enhance_contrast / white_balance   six_stadigy.py:190-199, 210-219   (float32 data)
//   color_enhancement                  enhancement_strategies.py:252-273 (float64 data)
// np.percentile(method='linear') needs the order statistics k and k+1.  They are found exactly by
// a most-significant-digit radix selection over the IEEE bit pattern (monotone for values in [0,1]):
// a 32-bit prefix p of every value is cut into digits of 12 / 10 / 10 bits,
//   float32: p = the bit pattern itself            -> three levels resolve the value completely
//   float64: p = bit pattern >> 30 (exponent + 22 mantissa bits); the members of the final bin
//            (normally 1-3 distinct values, or one value repeated many times) are put into a small
//            hash table of (full bit pattern, count) and ranked there.
// Level A is histogrammed inside gf_recover_kernel; levels B and C cost one read of the data each.
// Heavy ties are the normal case, not a corner case: every pixel whose channel byte equals the
// airlight's byte recovers to exactly A_c, whatever its transmission.
// A second stretch applied on top of the first is a monotone map of already ranked data, so its
// order statistics are the mapped order statistics of the recovered frame: no further pass.
#include "k_arith.cuh"
#include "k_kernels.h"

namespace uwie {

namespace {

constexpr int NT = 256;

__device__ __forceinline__ long long target_rank(const Recipe& rc, long long n, int t) {
    // t = stretch * 4 + endpoint * 2 + j
    const int i = (t >> 2) & 1, e = (t >> 1) & 1, j = t & 1;
    if (i >= rc.n_pre) return -1;
    const RankPlan& p = e ? rc.pre[i].hi : rc.pre[i].lo;
    const long long k = p.k + j;
    return k < n ? k : n - 1;
}

// ---- scans: one warp per target walks the histogram of the current level
// LEVEL 0: level-A histogram [inst][3][kSelBinsA]; 1: sel_hist_b; 2: sel_hist_c
template <int LEVEL>
__global__ void __launch_bounds__(kSelTargets * 32) sel_scan_kernel(SelectArgs a) {
    const int inst = blockIdx.x, tg = threadIdx.x >> 5, lane = threadIdx.x & 31;
    SelTarget& T = a.targets[(size_t)inst * kSelTargets + tg];
    long long rank;
    if (LEVEL == 0) {
        const Recipe& rc = a.recipes[a.dehaze_idx[inst % a.n_dehaze]];
        rank = target_rank(rc, a.n, tg & 7);
        if (lane == 0) { T.active = rank >= 0; T.prefix = 0; T.rank = rank; }
        if (rank < 0) return;
    } else {
        if (!T.active) return;
        rank = T.rank;
    }
    __syncwarp();
    const int nbins = LEVEL == 0 ? kSelBinsA : kSelSub;
    const unsigned* h = LEVEL == 0 ? a.sel_hist_a + ((size_t)inst * 3 + (tg >> 3)) * kSelBinsA
                                   : (LEVEL == 1 ? a.sel_hist_b : a.sel_hist_c) + ((size_t)inst * kSelTargets + tg) * kSelSub;
    const int per = nbins / 32;
    long long mine = 0;
    for (int i = 0; i < per; ++i) mine += h[lane * per + i];
    long long incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const long long v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    const long long excl = incl - mine;
    const bool here = excl <= rank && rank < incl;
    const unsigned who = __ballot_sync(0xffffffffu, here);
    if (who == 0) {             // cannot happen (ranks are < n); keep the run alive and flag it
        if (lane == 0) { atomicOr(a.status, UWIE_STATUS_SELECT_INTERNAL); T.active = 0; }
        return;
    }
    if (here) {
        long long acc = excl;
        int bin = lane * per;
        for (int i = 0; i < per; ++i, ++bin) {
            const long long nx = acc + h[bin];
            if (nx > rank) break;
            acc = nx;
        }
        T.rank = rank - acc;
        T.prefix = LEVEL == 0 ? (unsigned)bin : ((T.prefix << 10) | (unsigned)bin);
    }
}

// ---- level B / C histograms: members of each target's current bin, by their next 10 bits
template <class T, int LEVEL>
__global__ void __launch_bounds__(NT) sel_hist_kernel(SelectArgs a) {
    extern __shared__ unsigned sh[];          // [kSelTargets][kSelSub]
    const int inst = blockIdx.y;
    __shared__ unsigned s_pre[kSelTargets];
    __shared__ int s_on[kSelTargets];
    for (int i = threadIdx.x; i < kSelTargets * kSelSub; i += NT) sh[i] = 0;
    if (threadIdx.x < kSelTargets) {
        const SelTarget& t = a.targets[(size_t)inst * kSelTargets + threadIdx.x];
        // a target that shares its bin with the previous one of its channel rides on that histogram
        s_pre[threadIdx.x] = t.prefix;
        s_on[threadIdx.x] = t.active;
    }
    __syncthreads();
    constexpr int SHIFT = LEVEL == 1 ? 20 : 10;
    const T* src = reinterpret_cast<const T*>(a.recovered) + (size_t)inst * a.n * 3;
    const long long total = a.n * 3;
    for (long long i = (long long)blockIdx.x * NT + threadIdx.x; i < total; i += (long long)gridDim.x * NT) {
        const unsigned p = sel_prefix<T>(src[i]);
        const int c = (int)(i % 3);
        const unsigned hi = p >> SHIFT, sub = (p >> (SHIFT - 10)) & 1023u;
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            const int tg = c * 8 + t;
            if (s_on[tg] && s_pre[tg] == hi) atomicAdd(&sh[tg * kSelSub + sub], 1u);
        }
    }
    __syncthreads();
    unsigned* gh = (LEVEL == 1 ? a.sel_hist_b : a.sel_hist_c) + (size_t)inst * kSelTargets * kSelSub;
    for (int i = threadIdx.x; i < kSelTargets * kSelSub; i += NT) {
        const unsigned v = sh[i];
        if (v) atomicAdd(&gh[i], v);
    }
}

// ---- float64 only: (bit pattern, count) table of the members of each target's final bin
__device__ __forceinline__ void table_insert(unsigned long long* keys, unsigned* counts, unsigned long long key,
                                             unsigned n, int* status) {
    const unsigned long long stored = key + 1;          // 0 marks an empty slot
    unsigned h = (unsigned)((key * 0x9E3779B97F4A7C15ull) >> 54);   // 10 bits
    for (int probe = 0; probe < kSelTable; ++probe) {
        const unsigned long long prev = atomicCAS(&keys[h], 0ull, stored);
        if (prev == 0ull || prev == stored) { atomicAdd(&counts[h], n); return; }
        h = (h + 1) & (kSelTable - 1);
    }
    atomicOr(status, UWIE_STATUS_SELECT_OVERFLOW);
}

__global__ void __launch_bounds__(NT) sel_table_kernel(SelectArgs a) {
    const int inst = blockIdx.y;
    __shared__ unsigned s_pre[kSelTargets];
    __shared__ int s_on[kSelTargets];
    if (threadIdx.x < kSelTargets) {
        const SelTarget& t = a.targets[(size_t)inst * kSelTargets + threadIdx.x];
        s_pre[threadIdx.x] = t.prefix;
        s_on[threadIdx.x] = t.active;
    }
    __syncthreads();
    const double* src = reinterpret_cast<const double*>(a.recovered) + (size_t)inst * a.n * 3;
    unsigned long long* keys = a.tab_keys + (size_t)inst * kSelTargets * kSelTable;
    unsigned* counts = a.tab_counts + (size_t)inst * kSelTargets * kSelTable;
    const long long total = a.n * 3;
    const long long stride = (long long)gridDim.x * NT;
    const long long rounds = (total + stride - 1) / stride;
    for (long long rd = 0; rd < rounds; ++rd) {          // whole warps stay in the loop (match_any below)
        const long long i = rd * stride + (long long)blockIdx.x * NT + threadIdx.x;
        const bool in = i < total;
        const double x = in ? src[i] : 0.0;
        const unsigned p = sel_prefix<double>(x);
        const unsigned long long key = sel_key64(x);
        const int c = (int)(i % 3);
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            const int tg = c * 8 + t;
            const bool hit = in && s_on[tg] && s_pre[tg] == p;
            const unsigned hits = __ballot_sync(0xffffffffu, hit);
            if (hits == 0) continue;
            if (hit) {
                // lanes of one warp hold consecutive elements: same t means same slot only if same channel,
                // so aggregate on (target, key)
                const unsigned long long tag = key ^ ((unsigned long long)tg << 58);
                const unsigned same = __match_any_sync(hits, tag);
                if ((int)(__ffs(same) - 1) == (int)(threadIdx.x & 31))
                    table_insert(keys + (size_t)tg * kSelTable, counts + (size_t)tg * kSelTable, key,
                                 (unsigned)__popc(same), a.status);
            }
        }
    }
}

// ---- one block per instance: value of every target, then the stretch parameters
template <class T>
__global__ void __launch_bounds__(NT) sel_finish_kernel(SelectArgs a) {
    const int inst = blockIdx.x;
    const int b = inst / a.n_dehaze, k = inst % a.n_dehaze;
    const int s = a.dehaze_idx[k];
    const Recipe& rc = a.recipes[s];
    __shared__ double s_val[kSelTargets];
    __shared__ unsigned long long l_key[kSelTable];
    __shared__ unsigned l_cnt[kSelTable];
    __shared__ int l_n;
    for (int tg = 0; tg < kSelTargets; ++tg) {
        const SelTarget t = a.targets[(size_t)inst * kSelTargets + tg];
        if (!t.active) continue;                       // uniform per block
        if (sizeof(T) == 4) {
            if (threadIdx.x == 0) s_val[tg] = (double)__uint_as_float(t.prefix);
            continue;
        }
        if (threadIdx.x == 0) l_n = 0;
        __syncthreads();
        const unsigned long long* keys = a.tab_keys + ((size_t)inst * kSelTargets + tg) * kSelTable;
        const unsigned* counts = a.tab_counts + ((size_t)inst * kSelTargets + tg) * kSelTable;
        for (int i = threadIdx.x; i < kSelTable; i += NT) {
            const unsigned long long kv = keys[i];
            if (kv) { const int pos = atomicAdd(&l_n, 1); l_key[pos] = kv - 1; l_cnt[pos] = counts[i]; }
        }
        __syncthreads();
        const int m = l_n;
        bool found = false;
        for (int i = threadIdx.x; i < m; i += NT) {
            long long less = 0;
            for (int j = 0; j < m; ++j) if (l_key[j] < l_key[i]) less += l_cnt[j];
            if (less <= t.rank && t.rank < less + (long long)l_cnt[i]) {
                s_val[tg] = __longlong_as_double((long long)l_key[i]);
                found = true;
            }
        }
        if (__syncthreads_or(found) == 0 && threadIdx.x == 0) {
            atomicOr(a.status, UWIE_STATUS_SELECT_INTERNAL);
            s_val[tg] = 0.0;
        }
        __syncthreads();
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        const int c = threadIdx.x;
        StratTables& st = a.tables[(size_t)b * a.S + s];
        T lo0 = 0, den0 = 1;
        for (int i = 0; i < rc.n_pre; ++i) {
            T ord[4];
            for (int e = 0; e < 4; ++e) {
                T x = (T)s_val[c * 8 + i * 4 + e];
                if (i == 1) x = stretch_apply<T>(x, lo0, den0);   // rank statistics survive a monotone map
                ord[e] = x;
            }
            const T lo = np_lerp<T>(ord[0], ord[1], (T)rc.pre[i].lo.gamma);
            const T hi = np_lerp<T>(ord[2], ord[3], (T)rc.pre[i].hi.gamma);
            const T den = stretch_den<T>(lo, hi, a.lib.stretch_guard);
            st.st_lo[i][c] = (double)lo;
            st.st_den[i][c] = (double)den;
            if (i == 0) { lo0 = lo; den0 = den; }
        }
    }
}

template <class T>
int launch_select_t(const SelectArgs& a, int inst, int nblk, cudaStream_t st) {
    const size_t sm = (size_t)kSelTargets * kSelSub * sizeof(unsigned);
    static bool attr_done = false;     // per template instance
    if (!attr_done) {
        cudaFuncSetAttribute(sel_hist_kernel<T, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        cudaFuncSetAttribute(sel_hist_kernel<T, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        attr_done = true;
    }
    int launches = 0;
    sel_scan_kernel<0><<<inst, kSelTargets * 32, 0, st>>>(a); ++launches;
    sel_hist_kernel<T, 1><<<dim3(nblk, inst), NT, sm, st>>>(a); ++launches;
    sel_scan_kernel<1><<<inst, kSelTargets * 32, 0, st>>>(a); ++launches;
    sel_hist_kernel<T, 2><<<dim3(nblk, inst), NT, sm, st>>>(a); ++launches;
    sel_scan_kernel<2><<<inst, kSelTargets * 32, 0, st>>>(a); ++launches;
    if (sizeof(T) == 8) { sel_table_kernel<<<dim3(nblk, inst), NT, 0, st>>>(a); ++launches; }
    sel_finish_kernel<T><<<inst, NT, 0, st>>>(a); ++launches;
    return launches;
}

}  // namespace

int launch_select(const SelectArgs& a, cudaStream_t st) {
    if (a.n_dehaze == 0) return 0;
    const int inst = a.B * a.n_dehaze;
    long long per = (a.n * 3 + NT * 16 - 1) / (NT * 16);
    int nblk = (int)(per < 1 ? 1 : (per > 296 ? 296 : per));
    return a.lib.f64 ? launch_select_t<double>(a, inst, nblk, st) : launch_select_t<float>(a, inst, nblk, st);
}

}  // namespace uwie
