// Order statistics of a per-path fp32 array by radix select (the VaR quantiles the
// reference prints at every time step, np.quantile(VaR, .98 / .99), Replicating_Portfolio.py:120-122).
//
// HBM-bound integer/byte work: each pass streams the array once (4 B per path) and histograms one 8-bit digit
// of the order-preserving key among the elements that match the digits chosen so far; four passes pin the k-th
// smallest element exactly, one more pass finds its successor.  All accumulation is integer (atomicAdd / atomicMin),
// so the result does not depend on scheduling.  Between passes the host picks the digit (and, on several GPUs,
// adds the ranks' 257-entry histograms first), which keeps the device side free of grid-wide synchronisation.
#include "orp_common.h"

namespace orp {

constexpr int kSelThreads = 256;

// fp32 -> uint32 whose unsigned order is the float order (-inf < ... < -0 < +0 < ... < +inf); NaN reported apart
__device__ __forceinline__ uint32_t order_key(float f) {
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__global__ void __launch_bounds__(kSelThreads) select_hist_kernel(const float* __restrict__ x, uint64_t n, int shift,
                                                                  uint32_t prefix, uint32_t mask,
                                                                  unsigned long long* __restrict__ hist) {
    __shared__ unsigned int s_h[ORP_SELECT_BINS];
    for (int i = threadIdx.x; i < ORP_SELECT_BINS; i += kSelThreads) s_h[i] = 0u;
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * kSelThreads * 4;
    for (uint64_t i0 = ((uint64_t)blockIdx.x * kSelThreads + threadIdx.x) * 4; i0 < n; i0 += stride) {
        float v[4];
        if (i0 + 3 < n && (reinterpret_cast<uintptr_t>(x) & 15u) == 0) {
            const float4 q = *reinterpret_cast<const float4*>(x + i0);
            v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = i0 + j < n ? x[i0 + j] : 0.f;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (i0 + j >= n) break;
            if (v[j] != v[j]) { atomicAdd(&s_h[256], 1u); continue; }          // NaN: np.quantile then returns NaN
            const uint32_t key = order_key(v[j]);
            if (((key ^ prefix) & mask) == 0u) atomicAdd(&s_h[(key >> shift) & 0xffu], 1u);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < ORP_SELECT_BINS; i += kSelThreads)
        if (s_h[i]) atomicAdd(&hist[i], (unsigned long long)s_h[i]);
}

// out[0] += #{key <= K}; out[1] = min(out[1], min{key : key > K}) (0xffffffff when there is none)
__global__ void __launch_bounds__(kSelThreads) select_next_kernel(const float* __restrict__ x, uint64_t n, uint32_t K,
                                                                  unsigned long long* __restrict__ out) {
    unsigned long long le = 0;
    uint32_t mn = 0xffffffffu;
    for (uint64_t i = (uint64_t)blockIdx.x * kSelThreads + threadIdx.x; i < n; i += (uint64_t)gridDim.x * kSelThreads) {
        const float v = x[i];
        if (v != v) continue;
        const uint32_t key = order_key(v);
        if (key <= K) ++le; else mn = min(mn, key);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        le += __shfl_down_sync(0xffffffffu, le, o);
        mn = min(mn, __shfl_down_sync(0xffffffffu, mn, o));
    }
    if ((threadIdx.x & 31) == 0) {
        if (le) atomicAdd(&out[0], le);
        atomicMin(&out[1], (unsigned long long)mn);
    }
}

static int select_grid(uint64_t n, int per_thread) {
    uint64_t want = (n + (uint64_t)kSelThreads * per_thread - 1) / ((uint64_t)kSelThreads * per_thread);
    if (want < 1) want = 1;
    const uint64_t cap = (uint64_t)sm_count() * 8;
    return (int)(want < cap ? want : cap);
}

}  // namespace orp

using namespace orp;

extern "C" int orp_select_hist(const float* d_x, uint64_t n, int shift, uint32_t prefix, uint32_t mask,
                               unsigned long long* d_hist, void* stream) {
    ORP_REQUIRE(d_hist && (d_x || n == 0), "null pointer");
    ORP_REQUIRE(shift == 0 || shift == 8 || shift == 16 || shift == 24, "shift must be 0, 8, 16 or 24");
    if (n == 0) return ORP_OK;
    select_hist_kernel<<<select_grid(n, 4), kSelThreads, 0, (cudaStream_t)stream>>>(d_x, n, shift, prefix, mask, d_hist);
    ORP_LAUNCH_CHECK("select_hist_kernel");
    return ORP_OK;
}

extern "C" int orp_select_next(const float* d_x, uint64_t n, uint32_t key, unsigned long long* d_out, void* stream) {
    ORP_REQUIRE(d_out && (d_x || n == 0), "null pointer");
    if (n == 0) return ORP_OK;
    select_next_kernel<<<select_grid(n, 1), kSelThreads, 0, (cudaStream_t)stream>>>(d_x, n, key, d_out);
    ORP_LAUNCH_CHECK("select_next_kernel");
    return ORP_OK;
}
