// Exact np.percentile on float data (the recovered frame of the dehaze strategies):
//   enhance_contrast / white_balance   six_stadigy.py:190-199, 210-219   (float32 data)
//   color_enhancement                  enhancement_strategies.py:252-273 (float64 data)
// np.percentile(method='linear') needs the order statistics k and k+1; they are found exactly with
// a fixed number of launches: level-A histogram (fused into gf_recover_kernel), scan, level-B
// histogram, scan, gather of the (few) values sharing the level-B bin, exact selection in one block.
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

// one block per instance, one thread per target
__global__ void sel_scan_a_kernel(SelectArgs a) {
    const int inst = blockIdx.x, tg = threadIdx.x;
    if (tg >= kSelTargets) return;
    const int k = inst % a.n_dehaze;
    const Recipe& rc = a.recipes[a.dehaze_idx[k]];
    SelTarget& T = a.targets[(size_t)inst * kSelTargets + tg];
    const int c = tg >> 3;
    long long rank = target_rank(rc, a.n, tg & 7);
    T.active = rank >= 0;
    T.count = 0; T.bin_a = -1; T.bin_b = -1; T.rank = rank;
    if (rank < 0) return;
    const unsigned* h = a.sel_hist_a + ((size_t)inst * 3 + c) * kSelBins;
    long long acc = 0;
    for (int i = 0; i < kSelBins; ++i) {
        const long long nx = acc + h[i];
        if (nx > rank) { T.bin_a = i; T.rank = rank - acc; return; }
        acc = nx;
    }
    T.bin_a = kSelBins - 1; T.rank = 0;
}

template <class T>
__global__ void __launch_bounds__(NT) sel_hist_b_kernel(SelectArgs a) {
    extern __shared__ unsigned sh[];          // [kSelTargets][kSelSub]
    const int inst = blockIdx.y;
    __shared__ int s_bin[kSelTargets];
    for (int i = threadIdx.x; i < kSelTargets * kSelSub; i += NT) sh[i] = 0;
    if (threadIdx.x < kSelTargets) {
        const SelTarget& t = a.targets[(size_t)inst * kSelTargets + threadIdx.x];
        s_bin[threadIdx.x] = t.active ? t.bin_a : -1;
    }
    __syncthreads();
    const T* src = reinterpret_cast<const T*>(a.recovered) + (size_t)inst * a.n * 3;
    const long long total = a.n * 3;
    for (long long i = (long long)blockIdx.x * NT + threadIdx.x; i < total; i += (long long)gridDim.x * NT) {
        const T x = src[i];
        const int c = (int)(i % 3);
        const int ba = sel_bin_a<T>(x);
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            if (s_bin[c * 8 + t] == ba) atomicAdd(&sh[(c * 8 + t) * kSelSub + sel_bin_b<T>(x, ba)], 1u);
        }
    }
    __syncthreads();
    unsigned* gh = a.sel_hist_b + (size_t)inst * kSelTargets * kSelSub;
    for (int i = threadIdx.x; i < kSelTargets * kSelSub; i += NT) {
        const unsigned v = sh[i];
        if (v) atomicAdd(&gh[i], v);
    }
}

__global__ void sel_scan_b_kernel(SelectArgs a) {
    const int inst = blockIdx.x, tg = threadIdx.x;
    if (tg >= kSelTargets) return;
    SelTarget& T = a.targets[(size_t)inst * kSelTargets + tg];
    if (!T.active) return;
    const unsigned* h = a.sel_hist_b + ((size_t)inst * kSelTargets + tg) * kSelSub;
    long long acc = 0;
    const long long rank = T.rank;
    for (int i = 0; i < kSelSub; ++i) {
        const long long nx = acc + h[i];
        if (nx > rank) { T.bin_b = i; T.rank = rank - acc; return; }
        acc = nx;
    }
    T.bin_b = kSelSub - 1; T.rank = 0;
}

template <class T>
__global__ void __launch_bounds__(NT) sel_gather_kernel(SelectArgs a) {
    const int inst = blockIdx.y;
    __shared__ int s_a[kSelTargets], s_b[kSelTargets];
    if (threadIdx.x < kSelTargets) {
        const SelTarget& t = a.targets[(size_t)inst * kSelTargets + threadIdx.x];
        // exact 0 / exact 1 bins hold one distinct value: nothing to gather
        const bool need = t.active && t.bin_a != 0 && t.bin_a != kSelBins - 1;
        s_a[threadIdx.x] = need ? t.bin_a : -1;
        s_b[threadIdx.x] = t.bin_b;
    }
    __syncthreads();
    const T* src = reinterpret_cast<const T*>(a.recovered) + (size_t)inst * a.n * 3;
    T* buf = reinterpret_cast<T*>(a.sel_buf) + (size_t)inst * kSelTargets * kSelCap;
    const long long total = a.n * 3;
    for (long long i = (long long)blockIdx.x * NT + threadIdx.x; i < total; i += (long long)gridDim.x * NT) {
        const T x = src[i];
        const int c = (int)(i % 3);
        const int ba = sel_bin_a<T>(x);
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            const int tg = c * 8 + t;
            if (s_a[tg] == ba && s_b[tg] == sel_bin_b<T>(x, ba)) {
                const int pos = atomicAdd(&a.targets[(size_t)inst * kSelTargets + tg].count, 1);
                if (pos < kSelCap) buf[(size_t)tg * kSelCap + pos] = x;
            }
        }
    }
}

// one block per instance: exact selection inside each gathered set, then the stretch parameters
template <class T>
__global__ void __launch_bounds__(NT) sel_finish_kernel(SelectArgs a) {
    const int inst = blockIdx.x;
    const int b = inst / a.n_dehaze, k = inst % a.n_dehaze;
    const int s = a.dehaze_idx[k];
    const Recipe& rc = a.recipes[s];
    __shared__ double s_val[kSelTargets];
    __shared__ int s_found;
    const T* buf = reinterpret_cast<const T*>(a.sel_buf) + (size_t)inst * kSelTargets * kSelCap;
    for (int tg = 0; tg < kSelTargets; ++tg) {
        const SelTarget t = a.targets[(size_t)inst * kSelTargets + tg];
        if (!t.active) continue;                       // uniform per block
        if (t.bin_a == 0) { if (threadIdx.x == 0) s_val[tg] = 0.0; continue; }
        if (t.bin_a == kSelBins - 1) { if (threadIdx.x == 0) s_val[tg] = 1.0; continue; }
        if (threadIdx.x == 0) s_found = 0;
        __syncthreads();
        const int m = t.count;
        if (m > kSelCap) {
            if (threadIdx.x == 0) { atomicOr(a.status, UWIE_STATUS_SELECT_OVERFLOW); s_val[tg] = 0.0; }
            __syncthreads();
            continue;
        }
        const T* v = buf + (size_t)tg * kSelCap;
        for (int i = threadIdx.x; i < m; i += NT) {
            const T x = v[i];
            int less = 0, eq = 0;
            for (int j = 0; j < m; ++j) { const T y = v[j]; less += y < x; eq += y == x; }
            if (less <= t.rank && t.rank < less + eq) { s_val[tg] = (double)x; s_found = 1; }
        }
        __syncthreads();
        if (threadIdx.x == 0 && !s_found) atomicOr(a.status, UWIE_STATUS_SELECT_INTERNAL);
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

}  // namespace

int launch_select(const SelectArgs& a, cudaStream_t st) {
    if (a.n_dehaze == 0) return 0;
    const int inst = a.B * a.n_dehaze;
    long long per = (a.n * 3 + NT * 16 - 1) / (NT * 16);
    int nblk = (int)(per < 1 ? 1 : (per > 592 ? 592 : per));
    const size_t sm = (size_t)kSelTargets * kSelSub * sizeof(unsigned);
    sel_scan_a_kernel<<<inst, 32, 0, st>>>(a);
    if (a.lib.f64) {
        cudaFuncSetAttribute(sel_hist_b_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        sel_hist_b_kernel<double><<<dim3(nblk, inst), NT, sm, st>>>(a);
    } else {
        cudaFuncSetAttribute(sel_hist_b_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        sel_hist_b_kernel<float><<<dim3(nblk, inst), NT, sm, st>>>(a);
    }
    sel_scan_b_kernel<<<inst, 32, 0, st>>>(a);
    if (a.lib.f64) {
        sel_gather_kernel<double><<<dim3(nblk, inst), NT, 0, st>>>(a);
        sel_finish_kernel<double><<<inst, NT, 0, st>>>(a);
    } else {
        sel_gather_kernel<float><<<dim3(nblk, inst), NT, 0, st>>>(a);
        sel_finish_kernel<float><<<inst, NT, 0, st>>>(a);
    }
    return 5;
}

}  // namespace uwie
