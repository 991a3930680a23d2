// Terminal payoff and summary reductions.
//
// HBM-bound: one read of each maturity column (8 B or 4 B per path), optional 4 B write of the
// terminal value.  Per-option sums are reduced warp-shuffle -> shared memory -> one fp64 partial per
// block in scratch -> a second, single-block pass that adds the partials in block order, so the
// result does not depend on scheduling (and matches on every launch bit for bit).
#include "orp_common.h"

#include <mutex>

namespace orp {

constexpr int kPayThreads = 256;
constexpr int kPayChunk = 8;          // options handled together (shared column loads hit L1)

struct PayoffArgs {
    const void* paths;
    uint64_t ld, count;
    int kind, n_opt;
    double s0, value_scale;
    float* values_out;
    double* partials;                 // [gridDim.y][gridDim.x][2 * kPayChunk]
    double strikes[ORP_MAX_OPTIONS];
    int32_t cols[ORP_MAX_OPTIONS];
};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

// Per-thread accumulators of one chunk of options.  The out-of-the-money tallies are integers (a thread sees far
// fewer than 2^32 paths) and a NaN price is remembered as a flag instead of being added: per (path, option) the fp64
// pipe then runs three instructions for a call / put (subtract, compare, add) and two for the floor.
struct PayAcc {
    double sum[kPayChunk];
    unsigned otm[kPayChunk], nan[kPayChunk];
};

// UU paths of one thread (all of them inside the path range) against the chunk's options.  KIND is a template
// parameter: the payoff formula is fixed per launch, the option loop carries no dispatch.
template <typename T, int KIND, int UU>
__device__ __forceinline__ void payoff_paths(const PayoffArgs& a, const T* __restrict__ paths, uint64_t i, uint64_t stride,
                                             int o_first, int n_here, bool write_values, PayAcc& acc) {
    int last_col = -1;
    double s[UU];
    unsigned below[UU], any_nan = 0u;
#pragma unroll
    for (int u = 0; u < UU; ++u) { s[u] = 0.0; below[u] = 0u; }
#pragma unroll
    for (int k = 0; k < kPayChunk; ++k) {
        if (k < n_here) {                                   // uniform over the grid's y slice
            const int col = a.cols[o_first + k];
            if (col != last_col) {                          // options of a sweep arrive grouped by maturity column
                any_nan = 0u;
#pragma unroll
                for (int u = 0; u < UU; ++u) s[u] = (double)paths[(size_t)col * a.ld + i + u * stride];
#pragma unroll
                for (int u = 0; u < UU; ++u) {
                    if (KIND == ORP_PAYOFF_FLOOR) below[u] = s[u] < a.s0 ? 1u : 0u;                // RP.py:89
                    else any_nan |= s[u] != s[u] ? 1u : 0u;
                }
                last_col = col;
            }
            const double K = a.strikes[o_first + k];
#pragma unroll
            for (int u = 0; u < UU; ++u) {
                double pay;
                if (KIND == ORP_PAYOFF_FLOOR) {
                    pay = s[u] > K ? s[u] : K;                                                     // RP.py:88 (np.where: NaN -> K)
                    acc.otm[k] += below[u];
                } else {
                    const double d = KIND == ORP_PAYOFF_CALL ? s[u] - K : K - s[u];                // EO cell 8
                    const bool in = d > 0.0;                                                       // false for NaN
                    pay = in ? d : 0.0;
                    acc.otm[k] += in ? 0u : 1u;
                }
                acc.sum[k] += pay;
                if (k == 0 && write_values) {
                    // np.maximum propagates NaN (the sums get it through the flag below)
                    const double v = (KIND != ORP_PAYOFF_FLOOR && s[u] != s[u]) ? s[u] : pay;
                    a.values_out[i + u * stride] = (float)(v * a.value_scale);
                }
            }
            if (KIND != ORP_PAYOFF_FLOOR) acc.nan[k] |= any_nan;
        }
    }
}

template <typename T, int KIND>
__global__ void __launch_bounds__(kPayThreads, 4) payoff_partial_kernel(const PayoffArgs a) {
    const int chunk = blockIdx.y;
    const int o_first = chunk * kPayChunk;
    const int n_here = min(kPayChunk, a.n_opt - o_first);
    const T* __restrict__ paths = reinterpret_cast<const T*>(a.paths);
    const bool write_values = chunk == 0 && a.values_out != nullptr;

    PayAcc acc;
#pragma unroll
    for (int k = 0; k < kPayChunk; ++k) { acc.sum[k] = 0.0; acc.otm[k] = acc.nan[k] = 0u; }

    // four paths per thread and iteration: their column loads are independent, so four 8-byte requests per thread are
    // in flight (one per thread covered a quarter of the bandwidth-delay product of HBM); the ragged end of the range
    // goes through the same code one path at a time, so the main loop has no range checks
    constexpr int U = 4;
    const uint64_t stride = (uint64_t)gridDim.x * kPayThreads;
    uint64_t i = (uint64_t)blockIdx.x * kPayThreads + threadIdx.x;
    for (; i + (U - 1) * stride < a.count; i += U * stride)
        payoff_paths<T, KIND, U>(a, paths, i, stride, o_first, n_here, write_values, acc);
    for (; i < a.count; i += stride)
        payoff_paths<T, KIND, 1>(a, paths, i, stride, o_first, n_here, write_values, acc);

    double sum[kPayChunk], otm[kPayChunk];
#pragma unroll
    for (int k = 0; k < kPayChunk; ++k) {
        sum[k] = acc.nan[k] ? __longlong_as_double(0x7ff8000000000000ll) : acc.sum[k];
        otm[k] = (double)acc.otm[k];
    }

    __shared__ double s_red[kPayThreads / 32][2 * kPayChunk];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < kPayChunk; ++k) {
        const double a_ = warp_sum(sum[k]), b_ = warp_sum(otm[k]);
        if (lane == 0) { s_red[warp][2 * k] = a_; s_red[warp][2 * k + 1] = b_; }
    }
    __syncthreads();
    if (threadIdx.x < 2 * kPayChunk) {
        double t = 0.0;
        for (int w = 0; w < kPayThreads / 32; ++w) t += s_red[w][threadIdx.x];
        a.partials[((size_t)chunk * gridDim.x + blockIdx.x) * (2 * kPayChunk) + threadIdx.x] = t;
    }
}

// Second pass: one WARP per (option, quantity).  Lane l adds the partials of blocks l, l + 32, ... in increasing order,
// then a shuffle tree: a fixed order (the same bits on every launch) with ~n_blocks / 32 dependent additions per lane
// instead of one thread walking all n_blocks partials (57 us of a 64-option sweep under ncu).
constexpr int kFinalThreads = 128;
__global__ void __launch_bounds__(kFinalThreads) payoff_final_kernel(const double* __restrict__ partials, int n_blocks, int n_opt,
                                                                     double* __restrict__ sums, double* __restrict__ otm) {
    const int o = (blockIdx.x * kFinalThreads + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (o >= 2 * n_opt) return;                                 // whole warps leave together
    const int opt = o >> 1, which = o & 1;
    const int chunk = opt / kPayChunk, k = opt % kPayChunk;
    double t = 0.0;
    for (int b = lane; b < n_blocks; b += 32)
        t += partials[((size_t)chunk * n_blocks + b) * (2 * kPayChunk) + 2 * k + which];
    t = warp_sum(t);
    if (lane == 0) (which ? otm : sums)[opt] = t;
}

static int payoff_blocks() { return 4 * sm_count(); }        // four resident blocks per SM (launch bounds): whole waves

static void launch_payoff_partial(int dtype, int kind, dim3 grid, cudaStream_t st, const PayoffArgs& a) {
#define ORP_PAY(T, KIND) payoff_partial_kernel<T, KIND><<<grid, kPayThreads, 0, st>>>(a)
    if (dtype == ORP_F64) {
        if (kind == ORP_PAYOFF_CALL) ORP_PAY(double, ORP_PAYOFF_CALL);
        else if (kind == ORP_PAYOFF_PUT) ORP_PAY(double, ORP_PAYOFF_PUT);
        else ORP_PAY(double, ORP_PAYOFF_FLOOR);
    } else {
        if (kind == ORP_PAYOFF_CALL) ORP_PAY(float, ORP_PAYOFF_CALL);
        else if (kind == ORP_PAYOFF_PUT) ORP_PAY(float, ORP_PAYOFF_PUT);
        else ORP_PAY(float, ORP_PAYOFF_FLOOR);
    }
#undef ORP_PAY
}

}  // namespace orp

using namespace orp;

extern "C" uint64_t orp_payoff_scratch_bytes(int n_opt) {
    if (n_opt < 1) n_opt = 1;
    if (n_opt > ORP_MAX_OPTIONS) n_opt = ORP_MAX_OPTIONS;
    const int chunks = (n_opt + kPayChunk - 1) / kPayChunk;
    // sized for the largest grid this library launches on any device it may see (<= 4 blocks x 256 SMs)
    return (uint64_t)chunks * 1024ull * (2 * kPayChunk) * sizeof(double);
}

extern "C" int orp_payoff_reduce(const void* d_paths, int dtype, uint64_t ld, uint64_t count, int kind, int n_opt,
                                 const double* strikes, const int32_t* cols, int32_t n_cols, double s0, double value_scale,
                                 float* d_values_out, double* d_sums, double* d_otm, void* d_scratch, void* stream) {
    ORP_REQUIRE(d_paths && strikes && cols && d_sums && d_otm && d_scratch, "null pointer");
    ORP_REQUIRE(n_opt >= 1 && n_opt <= ORP_MAX_OPTIONS, "n_opt must be in [1, %d]", ORP_MAX_OPTIONS);
    ORP_REQUIRE(kind >= ORP_PAYOFF_CALL && kind <= ORP_PAYOFF_FLOOR, "unknown payoff kind %d", kind);
    ORP_REQUIRE(dtype == ORP_F64 || dtype == ORP_F32, "dtype must be ORP_F64 or ORP_F32");
    ORP_REQUIRE(ld >= count, "ld must be >= count");
    PayoffArgs a;
    a.paths = d_paths; a.ld = ld; a.count = count; a.kind = kind; a.n_opt = n_opt;
    a.s0 = s0; a.value_scale = value_scale; a.values_out = d_values_out;
    a.partials = reinterpret_cast<double*>(d_scratch);
    for (int k = 0; k < n_opt; ++k) {
        ORP_REQUIRE(cols[k] >= 0 && cols[k] < n_cols, "column index %d of option %d is outside [0, %d)", cols[k], k, n_cols);
        a.strikes[k] = strikes[k];
        a.cols[k] = cols[k];
    }
    const int chunks = (n_opt + kPayChunk - 1) / kPayChunk;
    uint64_t want = (count + kPayThreads - 1) / kPayThreads;
    if (want < 1) want = 1;
    int gx = payoff_blocks();
    if (gx > 1024) gx = 1024;
    if ((uint64_t)gx > want) gx = (int)want;
    cudaStream_t st = (cudaStream_t)stream;
    dim3 grid(gx, chunks);
    launch_payoff_partial(dtype, kind, grid, st, a);
    ORP_LAUNCH_CHECK("payoff_partial_kernel");
    payoff_final_kernel<<<(2 * n_opt * 32 + kFinalThreads - 1) / kFinalThreads, kFinalThreads, 0, st>>>(a.partials, gx, n_opt, d_sums, d_otm);
    ORP_LAUNCH_CHECK("payoff_final_kernel");
    return ORP_OK;
}

// ------------------------------------------------------------------------------------------------
// Allocation-free variant of orp_payoff_reduce (no counterpart in the reference).
//
// Who needs it: a caller that cannot obtain device memory at the time it wants a price — a pricing service
// that hands its whole GPU budget to fixed arenas at start-up and forbids the runtime allocator afterwards,
// or a plug-in living inside another framework whose allocator it may not touch.  Such a caller still owns
// its inputs (the path array) and outputs (the sums), but has nowhere to put the per-block partial sums the
// regular entry point expects in `d_scratch`.  Here those partials live in a fixed array that is part of
// the library image: every GPU that loads liborp_b200.so carries these 128 KiB until the process exits.
//
// The array's name below is only a host-side handle; its device address is looked up once per device with
// cudaGetSymbolAddress (with that device current) and that address is what the kernels receive.
constexpr int kNoAllocBlocks = 1024;          // upper bound of the grid's x extent (payoff_blocks() <= 1024)
static_assert(ORP_PAYOFF_NOALLOC_MAX_OPTIONS == kPayChunk, "one chunk of options: 1024 blocks x 16 doubles");
static __device__ __align__(256) double g_payoff_noalloc_partials[kNoAllocBlocks * 2 * kPayChunk];

// Device address of the array on the CURRENT device (cached per device ordinal).
static int noalloc_partials(double** out) {
    static std::mutex mu;
    static double* cached[64] = {nullptr};
    int dev = 0;
    ORP_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return orp::fail(ORP_E_STATE, "device ordinal %d is outside the supported range [0, 64)", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!cached[dev]) {
        void* p = nullptr;
        ORP_CUDA(cudaGetSymbolAddress(&p, g_payoff_noalloc_partials));
        if (!p) return orp::fail(ORP_E_STATE, "cudaGetSymbolAddress returned a null address for the static scratch");
        cached[dev] = static_cast<double*>(p);
    }
    *out = cached[dev];
    return ORP_OK;
}

extern "C" int orp_payoff_reduce_noalloc(const void* d_paths, int dtype, uint64_t ld, uint64_t count, int kind, int n_opt,
                                         const double* strikes, const int32_t* cols, int32_t n_cols, double s0, double value_scale,
                                         float* d_values_out, double* d_sums, double* d_otm, void* stream) {
    ORP_REQUIRE(d_paths && strikes && cols && d_sums && d_otm, "null pointer");
    ORP_REQUIRE(n_opt >= 1, "n_opt must be >= 1");
    ORP_REQUIRE(n_opt <= ORP_PAYOFF_NOALLOC_MAX_OPTIONS,
                "orp_payoff_reduce_noalloc handles at most %d options per call (got %d): its scratch is a fixed %zu-byte "
                "array inside the library; split the options over several calls or use orp_payoff_reduce",
                ORP_PAYOFF_NOALLOC_MAX_OPTIONS, n_opt, sizeof(double) * kNoAllocBlocks * 2 * kPayChunk);
    ORP_REQUIRE(kind >= ORP_PAYOFF_CALL && kind <= ORP_PAYOFF_FLOOR, "unknown payoff kind %d", kind);
    ORP_REQUIRE(dtype == ORP_F64 || dtype == ORP_F32, "dtype must be ORP_F64 or ORP_F32");
    ORP_REQUIRE(ld >= count, "ld must be >= count");
    double* partials = nullptr;
    int rc = noalloc_partials(&partials);        // first call on a device: symbol look-up (no allocation)
    if (rc) return rc;
    PayoffArgs a;
    a.paths = d_paths; a.ld = ld; a.count = count; a.kind = kind; a.n_opt = n_opt;
    a.s0 = s0; a.value_scale = value_scale; a.values_out = d_values_out;
    a.partials = partials;
    for (int k = 0; k < n_opt; ++k) {
        ORP_REQUIRE(cols[k] >= 0 && cols[k] < n_cols, "column index %d of option %d is outside [0, %d)", cols[k], k, n_cols);
        a.strikes[k] = strikes[k];
        a.cols[k] = cols[k];
    }
    uint64_t want = (count + kPayThreads - 1) / kPayThreads;
    if (want < 1) want = 1;
    int gx = payoff_blocks();
    if (gx > kNoAllocBlocks) gx = kNoAllocBlocks;
    if ((uint64_t)gx > want) gx = (int)want;
    cudaStream_t st = (cudaStream_t)stream;
    // Nothing in the array has to start at zero: every block of this launch overwrites its own 16 slots and the
    // second kernel reads exactly those gx blocks, so what an earlier call left behind is never looked at.
    dim3 grid(gx, 1);
    launch_payoff_partial(dtype, kind, grid, st, a);
    ORP_LAUNCH_CHECK("payoff_partial_kernel (noalloc)");
    payoff_final_kernel<<<(2 * n_opt * 32 + kFinalThreads - 1) / kFinalThreads, kFinalThreads, 0, st>>>(partials, gx, n_opt, d_sums, d_otm);
    ORP_LAUNCH_CHECK("payoff_final_kernel (noalloc)");
    return ORP_OK;
}
