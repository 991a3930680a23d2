// Terminal payoff and summary reductions.
//
// HBM-bound: one read of each maturity column (8 B or 4 B per path), optional 4 B write of the
// terminal value.  Per-option sums are reduced warp-shuffle -> shared memory -> one fp64 partial per
// block in scratch -> a second, single-block pass that adds the partials in block order, so the
// result does not depend on scheduling (and matches on every launch bit for bit).
#include "orp_common.h"

#include <algorithm>
#include <cstdlib>
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
    // bucket form only: strikes[] is sorted inside every chunk; slot[o] = position of that option in the caller's chunk
    int8_t slot[ORP_MAX_OPTIONS];
    double strike0;                   // the caller's option 0 (values_out)
};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

// Per-thread accumulators of one chunk of options.  Per (path, option) the loop runs ONE fp64 compare and, for the
// paths in the money, one fp64 add and one integer add: with d = S - K_ref (K_ref = the chunk's first
// strike, one subtraction per path and column)
//     call:  sum max(S - K, 0) = sum_{S > K} d - (K - K_ref) n_in          put:  (K - K_ref) n_in - sum_{S < K} d
//     floor: sum (S > K ? S : K) = sum_{S > K} d + K_ref n_in + K (n - n_in)
// are formed once per thread at the end; the in-the-money test is on S and K themselves (exact), the reference point
// keeps the sums as small as the payoffs (no cancellation against the price level), and for the chunk's first option
// (K = K_ref: the single-option pass of the backward loop) every bit equals the plain sum of S - K.  A NaN price is
// remembered as a flag instead of being added.
struct PayAcc {
    double sum[kPayChunk];
    unsigned n_in[kPayChunk], nan[kPayChunk];
    unsigned below[kPayChunk];        // floor only: paths with S < s0 at the option's column (RP.py:89)
    unsigned n;                       // paths seen
};

// One (path, option) pair.  (ptxas turns the conditional fp64 add into add + compare + two selects — also when it is
// written as a predicated add in PTX — so a pair costs five to six instructions; strike ladders avoid the per-option
// work altogether, see payoff_bucket_kernel.)
template <int KIND>
__device__ __forceinline__ void pay_pair(double s, double K, double d, double& sum, unsigned& n_in) {
    const bool in = KIND == ORP_PAYOFF_PUT ? s < K : s > K;
    if (in) { sum += d; n_in += 1u; }
}

// UU paths of one thread (all of them inside the path range) against the chunk's options.  KIND is a template
// parameter: the payoff formula is fixed per launch, the option loop carries no dispatch.  s_K: the chunk's strikes in
// shared memory; new_col: bit k set when option k reads another column than option k - 1 (both fixed per block — read
// from the launch arguments per option and iteration they cost twelve uniform-datapath instructions each).
template <typename T, int KIND, int UU>
__device__ __forceinline__ void payoff_paths(const PayoffArgs& a, const T* __restrict__ paths, uint64_t i, uint64_t stride,
                                             int o_first, int n_here, bool write_values, const double* s_K, unsigned new_col,
                                             PayAcc& acc) {
    double s[UU], d[UU];
    unsigned below[UU], any_nan = 0u;
    const double k_ref = s_K[0];
#pragma unroll
    for (int u = 0; u < UU; ++u) { s[u] = d[u] = 0.0; below[u] = 0u; }
    acc.n += UU;
#pragma unroll
    for (int k = 0; k < kPayChunk; ++k) {
        if (k < n_here) {                                   // uniform over the grid's y slice
            if (new_col & (1u << k)) {                      // options of a sweep arrive grouped by maturity column
                const int col = a.cols[o_first + k];
                any_nan = 0u;
#pragma unroll
                for (int u = 0; u < UU; ++u) s[u] = (double)paths[(size_t)col * a.ld + i + u * stride];
#pragma unroll
                for (int u = 0; u < UU; ++u) {
                    d[u] = s[u] - k_ref;
                    if (KIND == ORP_PAYOFF_FLOOR) below[u] = s[u] < a.s0 ? 1u : 0u;                // RP.py:89
                    else any_nan |= s[u] != s[u] ? 1u : 0u;
                }
            }
            const double K = s_K[k];
#pragma unroll
            for (int u = 0; u < UU; ++u) {
                // in the money (false for NaN): call S > K, put S < K (EO cell 8); floor: S > K keeps S (RP.py:88, np.where: NaN -> K)
                pay_pair<KIND>(s[u], K, d[u], acc.sum[k], acc.n_in[k]);
                if (KIND == ORP_PAYOFF_FLOOR) acc.below[k] += below[u];
                if (k == 0 && write_values) {
                    // option 0 is the chunk's reference: d = S - K.  np.maximum propagates NaN (the sums get it through the flag)
                    const bool in = KIND == ORP_PAYOFF_PUT ? s[u] < K : s[u] > K;
                    double v;
                    if (KIND == ORP_PAYOFF_FLOOR) v = in ? s[u] : K;
                    else v = s[u] != s[u] ? s[u] : (in ? (KIND == ORP_PAYOFF_CALL ? d[u] : -d[u]) : 0.0);
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

    __shared__ double s_K[kPayChunk];
    if (threadIdx.x < kPayChunk) s_K[threadIdx.x] = threadIdx.x < n_here ? a.strikes[o_first + threadIdx.x] : 0.0;
    unsigned new_col = 1u;
    for (int k = 1; k < n_here; ++k) new_col |= a.cols[o_first + k] != a.cols[o_first + k - 1] ? 1u << k : 0u;
    __syncthreads();

    PayAcc acc;
#pragma unroll
    for (int k = 0; k < kPayChunk; ++k) { acc.sum[k] = 0.0; acc.n_in[k] = acc.nan[k] = acc.below[k] = 0u; }
    acc.n = 0u;

    // four paths per thread and iteration: their column loads are independent, so four 8-byte requests per thread are
    // in flight (one per thread covered a quarter of the bandwidth-delay product of HBM); the ragged end of the range
    // goes through the same code one path at a time, so the main loop has no range checks
    constexpr int U = 4;
    const uint64_t stride = (uint64_t)gridDim.x * kPayThreads;
    uint64_t i = (uint64_t)blockIdx.x * kPayThreads + threadIdx.x;
    for (; i + (U - 1) * stride < a.count; i += U * stride)
        payoff_paths<T, KIND, U>(a, paths, i, stride, o_first, n_here, write_values, s_K, new_col, acc);
    for (; i < a.count; i += stride)
        payoff_paths<T, KIND, 1>(a, paths, i, stride, o_first, n_here, write_values, s_K, new_col, acc);

    double sum[kPayChunk], otm[kPayChunk];
    const double k_ref = s_K[0];
#pragma unroll
    for (int k = 0; k < kPayChunk; ++k) {
        const double K = s_K[k], n_in = (double)acc.n_in[k], dk = K - k_ref;
        double t;
        if (KIND == ORP_PAYOFF_CALL) t = acc.sum[k] - dk * n_in;
        else if (KIND == ORP_PAYOFF_PUT) t = dk * n_in - acc.sum[k];
        else t = acc.sum[k] + k_ref * n_in + K * (double)(acc.n - acc.n_in[k]);
        sum[k] = acc.nan[k] ? __longlong_as_double(0x7ff8000000000000ll) : t;
        otm[k] = KIND == ORP_PAYOFF_FLOOR ? (double)acc.below[k] : (double)(acc.n - acc.n_in[k]);
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

// ------------------------------------------------------------------------------------------------
// Sweep form: every chunk's options share ONE maturity column (a strike ladder per maturity, C4).  The strikes of a
// chunk are sorted (calls and floors ascending, puts descending; the host does it and passes the original positions in
// `slot`), so the options in the money for a path are exactly the first idx of them, idx = #{k : S > K_k} (puts: S <
// K_k).  A path then costs eight compares and ONE accumulation — d = S - K_ref and a count into bucket idx of the
// thread's own nine buckets in shared memory — instead of eight compare/add/select groups (ncu on the pair loop: 395
// instructions per four paths and eight options, issue slots 77 % busy, DRAM 27 %).  At the end
//     sum_{S > K_k} d = sum_{j > k} bucket[j],   n_in(k) = sum_{j > k} count[j]
// and the per-option sums follow by the formulas of PayAcc.  The in-the-money sets are exact (compares on S and K
// themselves), so the out-of-the-money tallies equal those of the pair loop; the sums differ from it by fp64 rounding.
constexpr int kBuckets = kPayChunk + 1;

template <typename T, int KIND>
__global__ void __launch_bounds__(kPayThreads, 4) payoff_bucket_kernel(const PayoffArgs a) {
    const int chunk = blockIdx.y;
    const int o_first = chunk * kPayChunk;
    const int n_here = min(kPayChunk, a.n_opt - o_first);
    const T* __restrict__ col = reinterpret_cast<const T*>(a.paths) + (size_t)a.cols[o_first] * a.ld;
    const bool write_values = chunk == 0 && a.values_out != nullptr;

    __shared__ double s_bd[kBuckets][kPayThreads];
    __shared__ unsigned s_bc[kBuckets][kPayThreads];
#pragma unroll
    for (int j = 0; j < kBuckets; ++j) { s_bd[j][threadIdx.x] = 0.0; s_bc[j][threadIdx.x] = 0u; }
    // never-in-the-money padding behind the last option of a short chunk
    const double pad = __longlong_as_double(KIND == ORP_PAYOFF_PUT ? 0xfff0000000000000ll : 0x7ff0000000000000ll);
    double K[kPayChunk];
#pragma unroll
    for (int k = 0; k < kPayChunk; ++k) K[k] = k < n_here ? a.strikes[o_first + k] : pad;
    const double k_ref = K[0];
    unsigned n = 0u, below = 0u, any_nan = 0u;

    auto one_path = [&](double sv, uint64_t i) {
        unsigned idx = 0u;
#pragma unroll
        for (int k = 0; k < kPayChunk; ++k) idx += (KIND == ORP_PAYOFF_PUT ? sv < K[k] : sv > K[k]) ? 1u : 0u;
        s_bd[idx][threadIdx.x] += sv - k_ref;
        s_bc[idx][threadIdx.x] += 1u;
        if (KIND == ORP_PAYOFF_FLOOR) below += sv < a.s0 ? 1u : 0u;                            // RP.py:89
        else any_nan |= sv != sv ? 1u : 0u;
        if (write_values) {                                 // the caller's option 0; np.maximum propagates NaN
            const double K0 = a.strike0;
            const bool in = KIND == ORP_PAYOFF_PUT ? sv < K0 : sv > K0;
            double v;
            if (KIND == ORP_PAYOFF_FLOOR) v = in ? sv : K0;
            else v = sv != sv ? sv : (in ? (KIND == ORP_PAYOFF_CALL ? sv - K0 : K0 - sv) : 0.0);
            a.values_out[i] = (float)(v * a.value_scale);
        }
    };

    constexpr int U = 4;                                    // four independent column loads per thread in flight
    const uint64_t stride = (uint64_t)gridDim.x * kPayThreads;
    uint64_t i = (uint64_t)blockIdx.x * kPayThreads + threadIdx.x;
    for (; i + (U - 1) * stride < a.count; i += U * stride) {
        double sv[U];
#pragma unroll
        for (int u = 0; u < U; ++u) sv[u] = (double)col[i + u * stride];
#pragma unroll
        for (int u = 0; u < U; ++u) one_path(sv[u], i + u * stride);
        n += U;
    }
    for (; i < a.count; i += stride) { one_path((double)col[i], i); n += 1u; }

    // buckets -> per-option sums (sorted positions), then the same block reduction as the pair form
    double sum[kPayChunk], otm[kPayChunk];
    double tail_d = 0.0;
    unsigned tail_c = 0u;
#pragma unroll
    for (int k = kPayChunk - 1; k >= 0; --k) {
        tail_d += s_bd[k + 1][threadIdx.x];
        tail_c += s_bc[k + 1][threadIdx.x];
        const double n_in = (double)tail_c, dk = K[k] - k_ref;
        double t;
        if (KIND == ORP_PAYOFF_CALL) t = tail_d - dk * n_in;
        else if (KIND == ORP_PAYOFF_PUT) t = dk * n_in - tail_d;
        else t = tail_d + k_ref * n_in + K[k] * (double)(n - tail_c);
        if (k >= n_here) t = 0.0;                           // padding: keep inf * 0 out of the partials
        sum[k] = (KIND != ORP_PAYOFF_FLOOR && any_nan) ? __longlong_as_double(0x7ff8000000000000ll) : t;
        otm[k] = KIND == ORP_PAYOFF_FLOOR ? (double)below : (double)(n - tail_c);
    }

    __shared__ double s_red[kPayThreads / 32][2 * kPayChunk];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < kPayChunk; ++k) {
        const double a_ = warp_sum(sum[k]), b_ = warp_sum(otm[k]);
        if (lane == 0) { s_red[warp][2 * k] = a_; s_red[warp][2 * k + 1] = b_; }
    }
    __syncthreads();
    if (threadIdx.x < 2 * kPayChunk && (threadIdx.x >> 1) < n_here) {
        double t = 0.0;
        for (int w = 0; w < kPayThreads / 32; ++w) t += s_red[w][threadIdx.x];
        const int pos = a.slot[o_first + (threadIdx.x >> 1)];                  // back to the caller's order inside the chunk
        a.partials[((size_t)chunk * gridDim.x + blockIdx.x) * (2 * kPayChunk) + 2 * pos + (threadIdx.x & 1)] = t;
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

static void launch_payoff_partial(int dtype, int kind, dim3 grid, cudaStream_t st, const PayoffArgs& a, bool buckets) {
#define ORP_PAY(T, KIND) do { if (buckets) payoff_bucket_kernel<T, KIND><<<grid, kPayThreads, 0, st>>>(a); \
                              else payoff_partial_kernel<T, KIND><<<grid, kPayThreads, 0, st>>>(a); } while (0)
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

// Bucket form (payoff_bucket_kernel) for a call of several options whose chunks each read ONE column: sorts the strikes
// inside every chunk (calls / floors ascending, puts descending) and records where each came from.
static bool prepare_buckets(PayoffArgs& a) {
    if (a.n_opt < 2 || getenv("ORP_B200_NO_BUCKETS")) return false;
    for (int k = 1; k < a.n_opt; ++k)
        if (k % kPayChunk != 0 && a.cols[k] != a.cols[k - 1]) return false;
    a.strike0 = a.strikes[0];
    for (int o = 0; o < a.n_opt; o += kPayChunk) {
        const int n = std::min(kPayChunk, a.n_opt - o);
        int order[kPayChunk];
        for (int k = 0; k < n; ++k) order[k] = k;
        std::stable_sort(order, order + n, [&](int x, int y) {
            return a.kind == ORP_PAYOFF_PUT ? a.strikes[o + x] > a.strikes[o + y] : a.strikes[o + x] < a.strikes[o + y];
        });
        double sorted[kPayChunk];
        for (int k = 0; k < n; ++k) { sorted[k] = a.strikes[o + order[k]]; a.slot[o + k] = (int8_t)order[k]; }
        for (int k = 0; k < n; ++k) a.strikes[o + k] = sorted[k];
    }
    return true;
}

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
    for (int k = 0; k < n_opt; ++k) ORP_REQUIRE(strikes[k] == strikes[k], "strike %d is NaN", k);
    const bool buckets = prepare_buckets(a);
    // strike ladders with short blocks (< 2^16 paths each on a full grid per chunk): ONE wave of blocks over all chunks
    // together — each block zeroes and folds 18 bucket columns and runs sixteen warp reductions, fixed costs that a full
    // grid per chunk pays eight times per SM slot (64 options: 298 -> 288 us at 2^24 paths; at 2^26 the full grid stays,
    // 1032 against 1072 us)
    if (buckets && count / (uint64_t)gx < (1ull << 16) && !getenv("ORP_B200_BUCKET_FULL_GRID"))
        gx = std::max(1, std::min(gx, (payoff_blocks() + chunks - 1) / chunks));
    dim3 grid(gx, chunks);
    launch_payoff_partial(dtype, kind, grid, st, a, buckets);
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
    launch_payoff_partial(dtype, kind, grid, st, a, false);
    ORP_LAUNCH_CHECK("payoff_partial_kernel (noalloc)");
    payoff_final_kernel<<<(2 * n_opt * 32 + kFinalThreads - 1) / kFinalThreads, kFinalThreads, 0, st>>>(partials, gx, n_opt, d_sums, d_otm);
    ORP_LAUNCH_CHECK("payoff_final_kernel (noalloc)");
    return ORP_OK;
}
