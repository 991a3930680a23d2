// Order statistics of a per-path fp32 array by radix select (the VaR quantiles the
// reference prints at every time step, np.quantile(VaR, .98 / .99), Replicating_Portfolio.py:120-122).
//
// HBM-bound integer/byte work: each pass streams the array once (4 B per path) and histograms one 8-bit digit
// of the order-preserving key among the elements that match the digits chosen so far; four passes pin the k-th
// smallest element exactly, one more pass finds its successor.  All accumulation is integer (atomicAdd / atomicMin),
// so the result does not depend on scheduling.  Two forms: the digit is picked on the HOST between passes
// (orp_select_hist / orp_select_next: five round trips per quantile), or on the DEVICE (orp_select_hist_multi ->
// [sum of the ranks' histograms] -> orp_select_pick, four times, then orp_select_next_multi): several quantiles share
// every pass over the array, the selection state stays in device memory and the host reads the result once.
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

// ---- device-side selection state: one entry per requested order statistic ----------------------------------------
// (layout = orp_select_state in include/orp_b200.h)
struct SelState { unsigned long long k; uint32_t prefix, mask, nan, bad; unsigned long long reserved; };
static_assert(sizeof(SelState) == sizeof(orp_select_state), "orp_select_state layout");

template <int Q>
__global__ void __launch_bounds__(kSelThreads) select_hist_multi_kernel(const float* __restrict__ x, uint64_t n, int shift,
                                                                        const SelState* __restrict__ st,
                                                                        unsigned long long* __restrict__ hist) {
    __shared__ unsigned int s_h[Q][ORP_SELECT_BINS];
    __shared__ uint32_t s_prefix[Q], s_mask[Q];
    for (int i = threadIdx.x; i < Q * ORP_SELECT_BINS; i += kSelThreads) (&s_h[0][0])[i] = 0u;
    if (threadIdx.x < Q) { s_prefix[threadIdx.x] = st[threadIdx.x].prefix; s_mask[threadIdx.x] = st[threadIdx.x].mask; }
    __syncthreads();
    uint32_t prefix[Q], mask[Q];
#pragma unroll
    for (int q = 0; q < Q; ++q) { prefix[q] = s_prefix[q]; mask[q] = s_mask[q]; }
    const bool vec = (reinterpret_cast<uintptr_t>(x) & 15u) == 0;
    const uint64_t stride = (uint64_t)gridDim.x * kSelThreads * 4;
    for (uint64_t i0 = ((uint64_t)blockIdx.x * kSelThreads + threadIdx.x) * 4; i0 < n; i0 += stride) {
        float v[4];
        if (i0 + 3 < n && vec) {
            const float4 f = *reinterpret_cast<const float4*>(x + i0);
            v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = i0 + j < n ? x[i0 + j] : 0.f;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (i0 + j >= n) break;
            if (v[j] != v[j]) { atomicAdd(&s_h[0][256], 1u); continue; }       // NaNs: one counter, copied to every entry below
            const uint32_t key = order_key(v[j]);
            const uint32_t digit = (key >> shift) & 0xffu;
#pragma unroll
            for (int q = 0; q < Q; ++q)
                if (((key ^ prefix[q]) & mask[q]) == 0u) atomicAdd(&s_h[q][digit], 1u);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < Q * ORP_SELECT_BINS; i += kSelThreads) {
        const unsigned int c = (i % ORP_SELECT_BINS) == 256 ? s_h[0][256] : (&s_h[0][0])[i];     // every entry carries the NaN count
        if (c) atomicAdd(&hist[i], (unsigned long long)c);
    }
}

// One block per entry: find the digit whose cumulative count passes k, descend into it, clear the histogram.
__global__ void __launch_bounds__(256) select_pick_kernel(unsigned long long* __restrict__ hist, int shift,
                                                          SelState* __restrict__ st) {
    __shared__ unsigned long long s_cum[256];
    __shared__ unsigned long long s_warp[8];
    const int q = blockIdx.x, t = threadIdx.x, lane = t & 31, w = t >> 5;
    unsigned long long* h = hist + (size_t)q * ORP_SELECT_BINS;
    const unsigned long long c = h[t];
    unsigned long long inc = c;                                       // inclusive scan over the 256 bins
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += u;
    }
    if (lane == 31) s_warp[w] = inc;
    __syncthreads();
    unsigned long long base = 0;
    for (int i = 0; i < w; ++i) base += s_warp[i];
    inc += base;
    s_cum[t] = inc;
    const unsigned long long nan = h[256];
    const unsigned long long k = st[q].k;                             // read by all before the barrier, written by one after it
    __syncthreads();
    const unsigned long long before = t ? s_cum[t - 1] : 0ull;
    if (before <= k && k < inc) {                                     // exactly one thread (bins are disjoint ranges)
        st[q].k = k - before;
        st[q].prefix |= (uint32_t)t << shift;
        st[q].mask |= 0xffu << shift;
    }
    if (t == 0) {
        if (k >= s_cum[255]) st[q].bad = 1u;                          // rank beyond the number of elements
        if (nan) st[q].nan = 1u;
    }
    __syncthreads();
    h[t] = 0ull;
    if (t == 0) h[256] = 0ull;
}

// out[q][0] += #{key <= K_q}; out[q][1] = min(out[q][1], min{key : key > K_q}), K_q = st[q].prefix
template <int Q>
__global__ void __launch_bounds__(kSelThreads) select_next_multi_kernel(const float* __restrict__ x, uint64_t n,
                                                                        const SelState* __restrict__ st,
                                                                        unsigned long long* __restrict__ out) {
    uint32_t K[Q];
    unsigned long long le[Q];
    uint32_t mn[Q];
#pragma unroll
    for (int q = 0; q < Q; ++q) { K[q] = st[q].prefix; le[q] = 0; mn[q] = 0xffffffffu; }
    for (uint64_t i = (uint64_t)blockIdx.x * kSelThreads + threadIdx.x; i < n; i += (uint64_t)gridDim.x * kSelThreads) {
        const float v = x[i];
        if (v != v) continue;
        const uint32_t key = order_key(v);
#pragma unroll
        for (int q = 0; q < Q; ++q) { if (key <= K[q]) ++le[q]; else mn[q] = min(mn[q], key); }
    }
#pragma unroll
    for (int q = 0; q < Q; ++q) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            le[q] += __shfl_down_sync(0xffffffffu, le[q], o);
            mn[q] = min(mn[q], __shfl_down_sync(0xffffffffu, mn[q], o));
        }
        if ((threadIdx.x & 31) == 0) {
            if (le[q]) atomicAdd(&out[2 * q], le[q]);
            atomicMin(&out[2 * q + 1], (unsigned long long)mn[q]);
        }
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

// ---- device-side digit pick: no host round trip between the passes -------------------------------------------------

template <int Q>
static void launch_hist_multi(const float* d_x, uint64_t n, int shift, const SelState* st, unsigned long long* hist, cudaStream_t s) {
    select_hist_multi_kernel<Q><<<select_grid(n, 4), kSelThreads, 0, s>>>(d_x, n, shift, st, hist);
}
template <int Q>
static void launch_next_multi(const float* d_x, uint64_t n, const SelState* st, unsigned long long* out, cudaStream_t s) {
    select_next_multi_kernel<Q><<<select_grid(n, 1), kSelThreads, 0, s>>>(d_x, n, st, out);
}

extern "C" int orp_select_hist_multi(const float* d_x, uint64_t n, int shift, int n_q, const orp_select_state* d_state,
                                     unsigned long long* d_hist, void* stream) {
    ORP_REQUIRE(d_hist && d_state && (d_x || n == 0), "null pointer");
    ORP_REQUIRE(shift == 0 || shift == 8 || shift == 16 || shift == 24, "shift must be 0, 8, 16 or 24");
    ORP_REQUIRE(n_q >= 1 && n_q <= ORP_SELECT_MAX_Q, "n_q must be in [1, %d]", ORP_SELECT_MAX_Q);
    if (n == 0) return ORP_OK;
    const SelState* st = reinterpret_cast<const SelState*>(d_state);
    cudaStream_t s = (cudaStream_t)stream;
    switch (n_q) {
        case 1: launch_hist_multi<1>(d_x, n, shift, st, d_hist, s); break;
        case 2: launch_hist_multi<2>(d_x, n, shift, st, d_hist, s); break;
        case 3: launch_hist_multi<3>(d_x, n, shift, st, d_hist, s); break;
        default: launch_hist_multi<4>(d_x, n, shift, st, d_hist, s); break;
    }
    ORP_LAUNCH_CHECK("select_hist_multi_kernel");
    return ORP_OK;
}

extern "C" int orp_select_pick(unsigned long long* d_hist, int shift, int n_q, orp_select_state* d_state, void* stream) {
    ORP_REQUIRE(d_hist && d_state, "null pointer");
    ORP_REQUIRE(shift == 0 || shift == 8 || shift == 16 || shift == 24, "shift must be 0, 8, 16 or 24");
    ORP_REQUIRE(n_q >= 1 && n_q <= ORP_SELECT_MAX_Q, "n_q must be in [1, %d]", ORP_SELECT_MAX_Q);
    select_pick_kernel<<<n_q, 256, 0, (cudaStream_t)stream>>>(d_hist, shift, reinterpret_cast<SelState*>(d_state));
    ORP_LAUNCH_CHECK("select_pick_kernel");
    return ORP_OK;
}

extern "C" int orp_select_next_multi(const float* d_x, uint64_t n, int n_q, const orp_select_state* d_state,
                                     unsigned long long* d_out, void* stream) {
    ORP_REQUIRE(d_out && d_state && (d_x || n == 0), "null pointer");
    ORP_REQUIRE(n_q >= 1 && n_q <= ORP_SELECT_MAX_Q, "n_q must be in [1, %d]", ORP_SELECT_MAX_Q);
    if (n == 0) return ORP_OK;
    const SelState* st = reinterpret_cast<const SelState*>(d_state);
    cudaStream_t s = (cudaStream_t)stream;
    switch (n_q) {
        case 1: launch_next_multi<1>(d_x, n, st, d_out, s); break;
        case 2: launch_next_multi<2>(d_x, n, st, d_out, s); break;
        case 3: launch_next_multi<3>(d_x, n, st, d_out, s); break;
        default: launch_next_multi<4>(d_x, n, st, d_out, s); break;
    }
    ORP_LAUNCH_CHECK("select_next_multi_kernel");
    return ORP_OK;
}
