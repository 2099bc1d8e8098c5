// Per-channel 256-bin histograms of uint8 HWC frames.
// Everything the reference derives from channel statistics of the raw / colour-corrected frame
// (six_stadigy.py:294 channel means, the np.percentile calls of six_stadigy.py:194-198,214-218
// on byte-valued data) is a function of these 768 counters.
#include "k_common.cuh"
#include "k_kernels.h"

namespace uwie {

namespace {

constexpr int NT = 256;
constexpr int WARPS = NT / 32;
constexpr int GROUP = 16;  // pixels per thread step (48 bytes = three 128-bit loads)

__device__ __forceinline__ void add_word(unsigned* h, unsigned w, int phase) {
    // phase = channel of the lowest byte of w
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int c = (phase + k) % 3;
        atomicAdd(&h[c * 256 + ((w >> (8 * k)) & 255u)], 1u);
    }
}

__global__ void __launch_bounds__(NT) frame_hist_kernel(const uint8_t* __restrict__ frames, int H, int W,
                                                       unsigned* __restrict__ hist) {
    __shared__ unsigned sh[WARPS][768];
    const int b = blockIdx.y;
    const size_t npix = (size_t)H * W;
    const uint8_t* f = frames + (size_t)b * npix * 3;
    for (int i = threadIdx.x; i < WARPS * 768; i += NT) (&sh[0][0])[i] = 0;
    __syncthreads();
    unsigned* h = sh[threadIdx.x >> 5];

    const size_t ngroups = npix / GROUP;
    const bool aligned = ((reinterpret_cast<uintptr_t>(f) & 15) == 0);
    for (size_t g = (size_t)blockIdx.x * NT + threadIdx.x; g < ngroups; g += (size_t)gridDim.x * NT) {
        const uint8_t* p = f + g * (GROUP * 3);
        if (aligned) {
            const uint4* q = reinterpret_cast<const uint4*>(p);
            const uint4 a = __ldg(q), bq = __ldg(q + 1), c = __ldg(q + 2);
            // byte offsets 0,4,8,...,44 -> channel phase = offset % 3
            add_word(h, a.x, 0); add_word(h, a.y, 1); add_word(h, a.z, 2); add_word(h, a.w, 0);
            add_word(h, bq.x, 1); add_word(h, bq.y, 2); add_word(h, bq.z, 0); add_word(h, bq.w, 1);
            add_word(h, c.x, 2); add_word(h, c.y, 0); add_word(h, c.z, 1); add_word(h, c.w, 2);
        } else {
            for (int k = 0; k < GROUP * 3; ++k) atomicAdd(&h[(k % 3) * 256 + p[k]], 1u);
        }
    }
    // tail pixels (fewer than GROUP), done by block 0
    if (blockIdx.x == 0) {
        for (size_t i = ngroups * GROUP + threadIdx.x; i < npix; i += NT) {
            const uint8_t* p = f + i * 3;
            atomicAdd(&h[p[0]], 1u); atomicAdd(&h[256 + p[1]], 1u); atomicAdd(&h[512 + p[2]], 1u);
        }
    }
    __syncthreads();
    unsigned* gh = hist + (size_t)b * 768;
    for (int i = threadIdx.x; i < 768; i += NT) {
        unsigned v = 0;
#pragma unroll
        for (int w = 0; w < WARPS; ++w) v += sh[w][i];
        if (v) atomicAdd(&gh[i], v);
    }
}

}  // namespace

int launch_frame_hist(const uint8_t* frames, int B, int H, int W, unsigned* hist, cudaStream_t st) {
    const size_t ngroups = ((size_t)H * W) / GROUP;
    int nblk = (int)((ngroups + (size_t)NT * 8 - 1) / ((size_t)NT * 8));
    if (nblk < 1) nblk = 1;
    if (nblk > 148 * 4) nblk = 148 * 4;
    frame_hist_kernel<<<dim3(nblk, B), NT, 0, st>>>(frames, H, W, hist);
    return 1;
}

}  // namespace uwie
