// Exact np.percentile on float data (the recovered frame of the dehaze strategies):
//   enhance_contrast / white_balance   six_stadigy.py:190-199, 210-219   (float32 data)
//   color_enhancement                  enhancement_strategies.py:252-273 (float64 data)
// np.percentile(method='linear') needs the order statistics k and k+1.  They are found exactly by a
// most-significant-digit radix selection over key = floor(x * 2^31) (k_arith.cuh: monotone, uniform
// in value, exact), cut into digits of 11 / 11 / 10 bits:
//   level A  histogram of key >> 21, accumulated inside gf_recover_kernel (no extra pass); its first and
//            last bins hold nothing but exact zeros / exact ones (the frame is clipped to [0, 1])
//   level B  one streaming read: every member of a target's A bin appends its bit pattern to that
//            target's candidate list (kSelListCap entries) and counts in a histogram of its next 11 bits
//   lists    sel_small_kernel: the order statistic inside the list by a radix selection over the bit
//            patterns (8-bit digits) - no further read of the frame.  This is the normal end.
// Only a target whose A bin held more members than a list goes on:
//   level C  one streaming read: members of the target's A.B bin, by their last 10 bits
//   After level C a float32 value >= 2^-8 is known exactly (key * 2^-31).  Values that can share a
//   key (float64 data; float32 below 2^-8) are told apart by a small hash table of (bit pattern, count)
//   filled in one more streaming read that exits at once when no target needs it.
// Heavy ties are the normal case, not a corner case: every pixel whose channel byte equals the
// airlight's byte recovers to exactly A_c, whatever its transmission - hence the route through
// counts (levels B / C, table) stays for the bins a list cannot hold.
// Because a level-A bin is 1/1024 of the value range, few elements match a target after level A;
// the streaming passes therefore test one shared-memory byte per element (does my A bin hold any
// target of my channel?) and only the rare hits do warp-aggregated global atomics.
// A second stretch applied on top of the first is a monotone map of already ranked data, so its
// order statistics are the mapped order statistics of the recovered frame: no further pass.
#include "k_arith.cuh"
#include "k_kernels.h"

namespace uwie {

namespace {

constexpr int NT = 256;
constexpr unsigned kExactKeyMin = 1u << 23;     // float32: key >= 2^23  <=>  x >= 2^-8  <=>  key is injective

__device__ __forceinline__ long long target_rank(const Recipe& rc, long long n, int t) {
    // t = stretch * 4 + endpoint * 2 + j
    const int i = (t >> 2) & 1, e = (t >> 1) & 1, j = t & 1;
    if (i >= rc.n_pre) return -1;
    const RankPlan& p = e ? rc.pre[i].hi : rc.pre[i].lo;
    const long long k = p.k + j;
    return k < n ? k : n - 1;
}

// ---- scans: one warp per target walks the histogram of the current level
template <int LEVEL>
__global__ void __launch_bounds__(kSelTargets * 32) sel_scan_kernel(SelectArgs a) {
    const int inst = blockIdx.x, tg = threadIdx.x >> 5, lane = threadIdx.x & 31;
    SelTarget& T = a.targets[(size_t)inst * kSelTargets + tg];
    long long rank;
    if (LEVEL == 0) {
        const Recipe& rc = a.recipes[a.dehaze_idx[inst % a.n_dehaze]];
        rank = target_rank(rc, a.n, tg & 7);
        if (lane == 0) { T.active = rank >= 0; T.prefix = 0; T.rank = rank; T.rep = tg; }
    } else {
        rank = T.active == 1 ? T.rank : -1;
    }
    __syncwarp();
    if (rank >= 0) {
    constexpr int nbins = LEVEL == 0 ? kSelBinsA : (LEVEL == 1 ? kSelBinsB : kSelBinsC);
    constexpr int bits = LEVEL == 1 ? 11 : 10;
    const unsigned* h = LEVEL == 0 ? a.sel_hist_a + ((size_t)inst * 3 + (tg >> 3)) * kSelBinsA
                      : LEVEL == 1 ? a.sel_hist_b + ((size_t)inst * kSelTargets + T.rep) * kSelBinsB
                                   : a.sel_hist_c + ((size_t)inst * kSelTargets + T.rep) * kSelBinsC;
    constexpr int per = nbins / 32;
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
        if (LEVEL == 0) {
            // bin 0 of level A holds the exact zeros (clipping makes them very common): resolved at once
            // and the last bin (key >> 21 == 1024) holds nothing but exact ones: the recovered frame is clipped to [0, 1]
            if (bin == 0) T.active = 2;
            else if (bin == 1025) T.active = 4;
            else T.prefix = (unsigned)(bin - 1);
        } else {
            T.prefix = (T.prefix << bits) | (unsigned)bin;
        }
    }
    }
    // targets of one channel that sit in the same bin share one histogram at the next level:
    // rep = the first such target (k and k+1 of a percentile almost always coincide)
    __syncthreads();
    if (lane == 0) {
        int rep = tg;
        if (T.active == 1) {
            for (int o = (tg & ~7); o < tg; ++o) {
                const SelTarget& O = a.targets[(size_t)inst * kSelTargets + o];
                if (O.active == 1 && O.prefix == T.prefix) { rep = o; break; }
            }
        }
        T.rep = rep;
    }
}

template <class T> struct SelBits;
template <> struct SelBits<float> { using type = unsigned; };
template <> struct SelBits<double> { using type = unsigned long long; };

// Shared prologue of the streaming passes: the targets of this instance and, per channel, which
// level-A bins hold a target at all.
struct PassState {
    unsigned pre[kSelTargets];
    int on[kSelTargets];
    int list[3][8];              // per channel: the targets that are on, compacted
    int cnt[3];
    uint8_t hot[3][kSelBinsA];
};
template <int LEVEL>      // 1, 2: prefixes hold 11 / 22 bits; 3: all 32
__device__ __forceinline__ bool pass_prologue(const SelectArgs& a, int inst, PassState& ps, bool only_inexact) {
    for (int i = threadIdx.x; i < 3 * kSelBinsA; i += NT) (&ps.hot[0][0])[i] = 0;
    __syncthreads();
    if (threadIdx.x < kSelTargets) {
        const SelTarget& t = a.targets[(size_t)inst * kSelTargets + threadIdx.x];
        int on = t.active == 1;
        if (only_inexact && t.prefix >= kExactKeyMin) on = 0;     // float32 value already exact
        if (t.rep != (int)threadIdx.x) on = 0;                     // rides on its representative's histogram
        ps.pre[threadIdx.x] = t.prefix;
        ps.on[threadIdx.x] = on;
        if (on) ps.hot[threadIdx.x >> 3][LEVEL == 1 ? t.prefix : (LEVEL == 2 ? t.prefix >> 11 : t.prefix >> 21)] = 1;
    }
    const bool any = __syncthreads_or(threadIdx.x < kSelTargets && ps.on[threadIdx.x]) != 0;
    if (threadIdx.x < 3) {
        int n = 0;
        for (int t = 0; t < 8; ++t) if (ps.on[threadIdx.x * 8 + t]) ps.list[threadIdx.x][n++] = threadIdx.x * 8 + t;
        ps.cnt[threadIdx.x] = n;
    }
    __syncthreads();
    return any;
}

// warp-aggregated histogram increment for the lanes in `hits` (all of them call this)
__device__ __forceinline__ void agg_add(unsigned* h, unsigned bin, unsigned hits) {
    const unsigned same = __match_any_sync(hits, bin);
    if ((int)(__ffs(same) - 1) == (int)(threadIdx.x & 31)) atomicAdd(&h[bin], (unsigned)__popc(same));
}

// ---- level B / C: one streaming read of the recovered frame of an instance
// 12 consecutive values (4 pixels) with 16-byte loads
__device__ __forceinline__ void load12(const float* p, float (&v)[12]) {
    const float4* q = reinterpret_cast<const float4*>(p);
#pragma unroll
    for (int i = 0; i < 3; ++i) { const float4 f = __ldg(q + i); v[4 * i] = f.x; v[4 * i + 1] = f.y; v[4 * i + 2] = f.z; v[4 * i + 3] = f.w; }
}
__device__ __forceinline__ void load12(const double* p, double (&v)[12]) {
    const double2* q = reinterpret_cast<const double2*>(p);
#pragma unroll
    for (int i = 0; i < 6; ++i) { const double2 f = __ldg(q + i); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
}

// VEC: every thread takes 4 pixels per round (n % 4 == 0, so every instance starts 16-byte aligned)
template <class T, int LEVEL, bool VEC>
__global__ void __launch_bounds__(NT, 4) sel_hist_kernel(SelectArgs a) {
    __shared__ PassState ps;
    const int inst = blockIdx.y;
    if (!pass_prologue<LEVEL>(a, inst, ps, false)) return;
    constexpr int SHIFT = LEVEL == 1 ? 21 : 10;                 // key >> SHIFT is what the targets resolved so far
    constexpr unsigned SUBMASK = LEVEL == 1 ? 2047u : 1023u;
    constexpr int SUBSHIFT = LEVEL == 1 ? 10 : 0;
    constexpr int BINS = LEVEL == 1 ? kSelBinsB : kSelBinsC;
    constexpr int PER = VEC ? 4 : 1;                            // pixels per thread and round
    const T* src = reinterpret_cast<const T*>(a.recovered) + (size_t)inst * a.n * 3;
    unsigned* gh = (LEVEL == 1 ? a.sel_hist_b : a.sel_hist_c) + (size_t)inst * kSelTargets * BINS;
    const long long units = a.n / PER;
    const long long stride = (long long)gridDim.x * NT;
    const long long rounds = (units + stride - 1) / stride;     // whole warps stay in the loop
    using U = typename SelBits<T>::type;
    __shared__ unsigned s_qkey[NT / 32][64];
    __shared__ U s_qbits[NT / 32][64];
    __shared__ uint8_t s_qch[NT / 32][64];
    const int lane = threadIdx.x & 31;
    unsigned* qkey = s_qkey[threadIdx.x >> 5];
    U* qbits = s_qbits[threadIdx.x >> 5];
    uint8_t* qch = s_qch[threadIdx.x >> 5];
    int qn = 0;                                                 // warp-uniform
    // entries 0 .. cnt-1 of the queue, one per lane: histogram of the next digit and, at level 1, the candidate list
    auto drain = [&](int cnt) {
        int tg = -1;
        unsigned sub = 0;
        U bits = 0;
        if (lane < cnt) {
            const unsigned k = qkey[lane];
            const int c = qch[lane];
            const unsigned hi = k >> SHIFT;
            sub = (k >> SUBSHIFT) & SUBMASK;
            bits = qbits[lane];
            const int nt = ps.cnt[c];
            for (int t = 0; t < nt; ++t) {                      // the targets that are on have distinct prefixes: at most one matches
                const int cand_t = ps.list[c][t];
                if (ps.pre[cand_t] == hi) tg = cand_t;
            }
        }
        const unsigned valid = __ballot_sync(0xffffffffu, tg >= 0);
        if (tg >= 0) {
            const unsigned peers = __match_any_sync(valid, (unsigned)tg);                   // same target
            const unsigned same = __match_any_sync(peers, sub);                             // ... and same bin: one atomic
            if ((int)(__ffs(same) - 1) == lane) atomicAdd(gh + (size_t)tg * BINS + sub, (unsigned)__popc(same));
            if (LEVEL == 1) {
                // the member itself goes to the target's candidate list (one atomic per target present)
                const int lead = __ffs(peers) - 1;
                unsigned base = 0;
                if (lane == lead) base = atomicAdd(&a.cand_n[(size_t)inst * kSelTargets + tg], (unsigned)__popc(peers));
                base = __shfl_sync(peers, base, lead) + (unsigned)__popc(peers & ((1u << lane) - 1u));
                if (base < (unsigned)kSelListCap)
                    reinterpret_cast<U*>(a.cand)[((size_t)inst * kSelTargets + tg) * kSelListCap + base] = bits;
            }
        }
        __syncwarp();
    };
    for (long long rd = 0; rd < rounds; ++rd) {
        const long long u = rd * stride + (long long)blockIdx.x * NT + threadIdx.x;
        const bool in = u < units;
        T x[3 * PER];
        if (VEC) {
            if (in) load12(src + u * 12, reinterpret_cast<T(&)[12]>(x));
            else {
#pragma unroll
                for (int j = 0; j < 3 * PER; ++j) x[j] = (T)0;
            }
        } else {
#pragma unroll
            for (int c = 0; c < 3; ++c) x[c] = in ? src[u * 3 + c] : (T)0;
        }
        unsigned key[3 * PER];
        unsigned hot = 0;
#pragma unroll
        for (int j = 0; j < 3 * PER; ++j) {
            key[j] = x[j] != (T)0 ? sel_key<T>(x[j]) : 0xffffffffu;   // exact zeros live in level-A bin 0, not here
            const unsigned ka = key[j] >> 21;
            if (ka <= 1024u && ps.hot[j % 3][ka] != 0) hot |= 1u << j;
        }
        // A hot value is rare per LANE (the targets' level-A bins hold ~1 % of the data) but not per warp (384 values a
        // round).  Hot values are first compacted into the warp's queue in shared memory - every lane pushes one of its
        // own per trip, positions from a ballot - and handled 32 at a time, one per lane.
        unsigned m = hot;
        while (true) {
            const unsigned has = __ballot_sync(0xffffffffu, m != 0);
            if (has == 0) break;
            if (m != 0) {
                const int j = __ffs(m) - 1;
                m &= m - 1;
                unsigned kj = 0;
                U bj = 0;
#pragma unroll
                for (int jj = 0; jj < 3 * PER; ++jj)
                    if (jj == j) { kj = key[jj]; bj = (U)sel_bits<T>(x[jj]); }
                const int at = qn + __popc(has & ((1u << lane) - 1u));
                qkey[at] = kj; qbits[at] = bj; qch[at] = (uint8_t)(j % 3);
            }
            qn += __popc(has);
            __syncwarp();
            if (qn >= 32) {
                drain(32);
                // the (at most 31) entries behind the drained ones move to the front
                const int left = qn - 32;
                unsigned k2 = 0; U b2 = 0; uint8_t c2 = 0;
                if (lane < left) { k2 = qkey[32 + lane]; b2 = qbits[32 + lane]; c2 = qch[32 + lane]; }
                __syncwarp();
                if (lane < left) { qkey[lane] = k2; qbits[lane] = b2; qch[lane] = c2; }
                qn = left;
                __syncwarp();
            }
        }
    }
    if (qn > 0) drain(qn);
}

// ---- candidate lists: a target whose level-A bin fitted its list is resolved here, with no further
// streaming read: radix selection (8-bit digits, most significant first) over the bit patterns, which
// order like the values because the recovered frame is >= 0.  One block per target.
template <class T>
__global__ void __launch_bounds__(NT) sel_small_kernel(SelectArgs a) {
    using U = typename SelBits<T>::type;
    constexpr int BITS = sizeof(U) * 8;
    const int tg = blockIdx.x, inst = blockIdx.y;
    SelTarget& Tt = a.targets[(size_t)inst * kSelTargets + tg];
    if (Tt.active != 1) return;                                   // block-uniform
    const int rep = Tt.rep;
    const unsigned m = a.cand_n[(size_t)inst * kSelTargets + rep];
    if (m > (unsigned)kSelListCap) {                              // list unusable: levels B / C resolve this one
        if (threadIdx.x == 0) atomicAdd(&a.stats[1], 1u);
        return;
    }
    const U* list = reinterpret_cast<const U*>(a.cand) + ((size_t)inst * kSelTargets + rep) * kSelListCap;
    __shared__ unsigned hist[256];
    __shared__ long long s_rank;
    __shared__ int s_digit;
    U prefix = 0;
    if (threadIdx.x == 0) s_rank = Tt.rank;
    const unsigned rounds = (m + NT - 1) / NT;
    for (int shift = BITS - 8; shift >= 0; shift -= 8) {
        hist[threadIdx.x] = 0;
        __syncthreads();
        for (unsigned rd = 0; rd < rounds; ++rd) {
            const unsigned i = rd * NT + threadIdx.x;
            bool in = i < m;
            U v = 0;
            if (in) { v = list[i]; in = shift == BITS - 8 || (v >> (shift + 8)) == prefix; }
            const unsigned who = __ballot_sync(0xffffffffu, in);
            if (in) agg_add(hist, (unsigned)(v >> shift) & 255u, who);
        }
        __syncthreads();
        if (threadIdx.x < 32) {
            const int lane = threadIdx.x;
            const long long rank = s_rank;
            long long mine = 0;
            for (int i = 0; i < 8; ++i) mine += hist[lane * 8 + i];
            long long incl = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const long long v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += v;
            }
            const long long excl = incl - mine;
            const bool here = excl <= rank && rank < incl;
            if (__ballot_sync(0xffffffffu, here) == 0 && lane == 0) {     // cannot happen; keep the run alive and flag it
                atomicOr(a.status, UWIE_STATUS_SELECT_INTERNAL);
                s_digit = 0;
            }
            if (here) {
                long long acc = excl;
                int bin = lane * 8;
                for (int i = 0; i < 8; ++i, ++bin) {
                    const long long nx = acc + hist[bin];
                    if (nx > rank) break;
                    acc = nx;
                }
                s_rank = rank - acc;
                s_digit = bin;
            }
        }
        __syncthreads();
        prefix = (prefix << 8) | (U)s_digit;
    }
    if (threadIdx.x == 0) {
        Tt.val = (double)sel_from_bits<T>((unsigned long long)prefix);
        Tt.active = 3;
        atomicAdd(&a.stats[0], 1u);
    }
}

// ---- values that share a key: (bit pattern, count) table of the members of each target's final bin
__device__ __forceinline__ void table_insert(unsigned long long* keys, unsigned* counts, unsigned long long bits,
                                             unsigned n, int* status) {
    const unsigned long long stored = bits + 1;          // 0 marks an empty slot
    unsigned h = (unsigned)((bits * 0x9E3779B97F4A7C15ull) >> 56) & (kSelTable - 1);
    for (int probe = 0; probe < kSelTable; ++probe) {
        const unsigned long long prev = atomicCAS(&keys[h], 0ull, stored);
        if (prev == 0ull || prev == stored) { atomicAdd(&counts[h], n); return; }
        h = (h + 1) & (kSelTable - 1);
    }
    atomicOr(status, UWIE_STATUS_SELECT_OVERFLOW);
}

template <class T>
__global__ void __launch_bounds__(NT) sel_table_kernel(SelectArgs a) {
    __shared__ PassState ps;
    const int inst = blockIdx.y;
    if (!pass_prologue<3>(a, inst, ps, sizeof(T) == 4)) return;  // float32: normally nothing to do
    const T* src = reinterpret_cast<const T*>(a.recovered) + (size_t)inst * a.n * 3;
    unsigned long long* keys = a.tab_keys + (size_t)inst * kSelTargets * kSelTable;
    unsigned* counts = a.tab_counts + (size_t)inst * kSelTargets * kSelTable;
    const long long stride = (long long)gridDim.x * NT;
    const long long rounds = (a.n + stride - 1) / stride;
    for (long long rd = 0; rd < rounds; ++rd) {
        const long long px = rd * stride + (long long)blockIdx.x * NT + threadIdx.x;
        const bool in = px < a.n;
        T x[3];
        unsigned key[3];
        bool hot = false;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            x[c] = in ? src[px * 3 + c] : (T)0;
            key[c] = x[c] != (T)0 ? sel_key<T>(x[c]) : 0xffffffffu;
            hot |= (key[c] >> 21) <= 1024u ? ps.hot[c][key[c] >> 21] != 0 : false;
        }
        if (__ballot_sync(0xffffffffu, hot) == 0) continue;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const bool hc = (key[c] >> 21) <= 1024u ? ps.hot[c][key[c] >> 21] != 0 : false;
            if (__ballot_sync(0xffffffffu, hc) == 0) continue;
            const unsigned long long bits = sel_bits<T>(x[c]);
            const int nt = ps.cnt[c];                            // uniform
#pragma unroll 1
            for (int t = 0; t < nt; ++t) {
                const int tg = ps.list[c][t];
                const bool hit = hc && ps.pre[tg] == key[c];
                const unsigned hits = __ballot_sync(0xffffffffu, hit);
                if (hit) {
                    const unsigned same = __match_any_sync(hits, bits);
                    if ((int)(__ffs(same) - 1) == (int)(threadIdx.x & 31))
                        table_insert(keys + (size_t)tg * kSelTable, counts + (size_t)tg * kSelTable, bits,
                                     (unsigned)__popc(same), a.status);
                }
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
    __shared__ unsigned long long l_bits[kSelTable];
    __shared__ unsigned l_cnt[kSelTable];
    __shared__ int l_n;
    for (int tg = 0; tg < kSelTargets; ++tg) {
        const SelTarget t = a.targets[(size_t)inst * kSelTargets + tg];
        if (!t.active) continue;                       // uniform per block
        if (t.active == 2) { if (threadIdx.x == 0) s_val[tg] = 0.0; continue; }
        if (t.active == 4) { if (threadIdx.x == 0) s_val[tg] = 1.0; continue; }
        if (t.active == 3) { if (threadIdx.x == 0) s_val[tg] = t.val; continue; }
        if (sizeof(T) == 4 && t.prefix >= kExactKeyMin) {
            // key = x * 2^31 exactly, with at most 24 significant bits
            if (threadIdx.x == 0) s_val[tg] = (double)__fmul_rn(__uint2float_rn(t.prefix), 4.656612873077393e-10f);
            continue;
        }
        if (threadIdx.x == 0) l_n = 0;
        __syncthreads();
        const unsigned long long* keys = a.tab_keys + ((size_t)inst * kSelTargets + t.rep) * kSelTable;
        const unsigned* counts = a.tab_counts + ((size_t)inst * kSelTargets + t.rep) * kSelTable;
        for (int i = threadIdx.x; i < kSelTable; i += NT) {
            const unsigned long long kv = keys[i];
            if (kv) { const int pos = atomicAdd(&l_n, 1); l_bits[pos] = kv - 1; l_cnt[pos] = counts[i]; }
        }
        __syncthreads();
        const int m = l_n;
        bool found = false;
        for (int i = threadIdx.x; i < m; i += NT) {
            long long less = 0;
            for (int j = 0; j < m; ++j) if (l_bits[j] < l_bits[i]) less += l_cnt[j];
            if (less <= t.rank && t.rank < less + (long long)l_cnt[i]) {
                s_val[tg] = (double)sel_from_bits<T>(l_bits[i]);
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
int launch_select_t(const SelectArgs& a, int inst, cudaStream_t st) {
    // ONE wave of blocks across all instances: every block does the same number of rounds, so a second
    // partial wave would only be a tail.  Resident blocks per SM come from the occupancy calculator.
    static const int resident = [] {
        int b1 = 0, b2 = 0, sms = 0, dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b1, sel_hist_kernel<T, 1, true>, NT, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b2, sel_hist_kernel<T, 2, true>, NT, 0);
        const int b = b1 < b2 ? b1 : b2;
        return (b > 0 ? b : 4) * (sms > 0 ? sms : 148);
    }();
    const long long per = (a.n + NT * 8 - 1) / (NT * 8);
    const int cap = resident / inst < 1 ? 1 : resident / inst;
    const int nblk = (int)(per < 1 ? 1 : (per > cap ? cap : per));
    int launches = 0;
    sel_scan_kernel<0><<<inst, kSelTargets * 32, 0, st>>>(a); ++launches;
    const bool vec = a.n % 4 == 0 && reinterpret_cast<uintptr_t>(a.recovered) % 16 == 0;
    if (vec) sel_hist_kernel<T, 1, true><<<dim3(nblk, inst), NT, 0, st>>>(a);
    else sel_hist_kernel<T, 1, false><<<dim3(nblk, inst), NT, 0, st>>>(a);
    ++launches;
    sel_small_kernel<T><<<dim3(kSelTargets, inst), NT, 0, st>>>(a); ++launches;
    // from here on only for targets whose list overflowed; otherwise every block exits in its prologue
    sel_scan_kernel<1><<<inst, kSelTargets * 32, 0, st>>>(a); ++launches;
    if (vec) sel_hist_kernel<T, 2, true><<<dim3(nblk, inst), NT, 0, st>>>(a);
    else sel_hist_kernel<T, 2, false><<<dim3(nblk, inst), NT, 0, st>>>(a);
    ++launches;
    sel_scan_kernel<2><<<inst, kSelTargets * 32, 0, st>>>(a); ++launches;
    sel_table_kernel<T><<<dim3(nblk, inst), NT, 0, st>>>(a); ++launches;
    sel_finish_kernel<T><<<inst, NT, 0, st>>>(a); ++launches;
    return launches;
}

}  // namespace

int launch_select(const SelectArgs& a, cudaStream_t st) {
    if (a.n_dehaze == 0) return 0;
    const int inst = a.B * a.n_dehaze;
    return a.lib.f64 ? launch_select_t<double>(a, inst, st) : launch_select_t<float>(a, inst, st);
}

}  // namespace uwie
