// Sobol points, inverse normal and the fused path kernels (SURVEY.md §8 rows a-2 .. a-5, a-8).
//
// One persistent thread block per SM keeps the 116 KB inverse-normal coefficient table in shared
// memory; a warp owns 64 consecutive Sobol indices (two per lane) and walks the time axis with the
// whole path state in registers, storing the kept columns time-major (256 contiguous bytes per warp
// store).  Nothing but the kept columns ever touches HBM: 8 B (fp64) or 4 B (fp32) per stored path-step.
#include "orp_common.h"
#include "orp_device.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <utility>
#include <vector>

namespace orp {

constexpr int kPathThreads = 1024;   // 32 warps: 2048 paths per block tile
constexpr int kWarpPaths = 64;

// ------------------------------------------------------------------------------------------------
// a-3 test entry: z[i] = ndtri(x[i] * 2^-30)
__global__ void __launch_bounds__(256) ndtri_u32_kernel(const uint32_t* __restrict__ x, uint64_t n,
                                                        double* __restrict__ z) {
    extern __shared__ __align__(128) char s_tab[];
    load_ndtri_table(reinterpret_cast<double*>(s_tab), threadIdx.x, blockDim.x);
    __syncthreads();
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        z[i] = ndtri_bits(x[i], (uint32_t)__cvta_generic_to_shared(s_tab));
}

// ------------------------------------------------------------------------------------------------
// a-2: integer points.  layout 1: out[t][count] (native), layout 0: out[count][d] (SciPy's) through a
// per-warp shared-memory transpose so both are written with full 128-byte lines.
constexpr int kPointsWarps = 4;

template <int LAYOUT>
__global__ void __launch_bounds__(kPointsWarps * 32)
sobol_points_kernel(const uint32_t* __restrict__ sv32, const uint2* __restrict__ tlane, int d, uint64_t start,
                    uint64_t count, uint32_t* __restrict__ out) {
    __shared__ uint32_t s_tile[LAYOUT == 0 ? kPointsWarps : 1][64][33];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint64_t end = start + count;
    const uint64_t w_first = start / kWarpPaths, w_last = (end - 1) / kWarpPaths;
    const uint64_t n_tiles = w_last - w_first + 1;
    for (uint64_t tile = (uint64_t)blockIdx.x * kPointsWarps + warp; tile < n_tiles;
         tile += (uint64_t)gridDim.x * kPointsWarps) {
        const uint64_t w = w_first + tile;
        const uint64_t i0 = w * kWarpPaths + lane, i1 = i0 + 32;
        const bool ok0 = i0 >= start && i0 < end, ok1 = i1 >= start && i1 < end;
        for (int t0 = 0; t0 < d; t0 += 32) {
            const int my_dim = t0 + lane;
            const uint32_t bl = my_dim < d ? warp_base(sv32, ORP_SV_DPAD(d), my_dim, w) : 0u;
            const int nc = min(32, d - t0);
            for (int s = 0; s < nc; ++s) {
                const int t = t0 + s;
                const uint32_t base = __shfl_sync(kFull, bl, s);
                const uint2 tl = __ldg(tlane + (size_t)t * 32 + lane);
                const uint32_t x0 = base ^ tl.x, x1 = base ^ tl.y;
                if (LAYOUT == 1) {
                    if (ok0) out[(size_t)t * count + (i0 - start)] = x0;
                    if (ok1) out[(size_t)t * count + (i1 - start)] = x1;
                } else {
                    s_tile[warp][lane][s] = x0;
                    s_tile[warp][lane + 32][s] = x1;
                }
            }
            if (LAYOUT == 0) {
                __syncwarp();
                for (int r = 0; r < 64; ++r) {
                    const uint64_t i = w * kWarpPaths + r;
                    if (i >= start && i < end && lane < nc)
                        out[(size_t)(i - start) * d + t0 + lane] = s_tile[warp][r][lane];
                }
                __syncwarp();
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// a-4 / a-5 / a-8: fused Sobol -> ndtri -> step -> store-kept-columns.

struct PathArgs {
    const uint32_t* sv32_a;
    const uint2* tlane_a;
    const uint32_t* sv32_b;   // SV only: the volatility engine (seed 1234 in the reference)
    const uint2* tlane_b;
    void* out;
    uint64_t ld, start, count;
    int n_steps, keep_every, dpad;
    double s0;                // Y / S0
    double c_drift;           // GBM_LOG: (mu - sigma^2/2) dt;  GBM_ARITH: mu dt;  SV: mu
    double c_vol;             // GBM: sigma sqrt(dt);           SV: sqrt(dt)
    double dt, v0, a, b, c;
};


// models whose state is the level itself (no exp at the kept columns)
template <int MODEL> constexpr bool kLevel = MODEL == ORP_MODEL_GBM_ARITH || MODEL == ORP_MODEL_LINEAR;

// exp(x) for the one-step log-increments of the multiplicative form (MULT): degree-9 Taylor series inside
// |x| < 1/8 (remainder x^10 / 10! < 3e-16 relative), the library exp outside — including -inf, +inf and NaN,
// which fail the comparison.  The host only selects MULT when six standard deviations stay inside the fast range.
// (coefficients in constant memory: as literals each use costs two moves into uniform registers, and the stochastic-volatility
//  loop — 234 executed instructions per step of two paths, issue slots 75 % busy under ncu — is bound by issue, not by the fp64 pipe)
__constant__ double c_exp_series[8] = {1.0 / 362880.0, 1.0 / 40320.0, 1.0 / 5040.0, 1.0 / 720.0, 1.0 / 120.0, 1.0 / 24.0, 1.0 / 6.0, 0.5};
__device__ __forceinline__ double exp_series(double x) {
    double p = c_exp_series[0];
#pragma unroll
    for (int k = 1; k < 8; ++k) p = fma(p, x, c_exp_series[k]);
    p = fma(p, x, 1.0);
    return fma(p, x, 1.0);
}
// (one range test per path: a shared test for the two paths of a lane with the library code behind a call measured
//  19 % slower on the bridge kernel)
__device__ __forceinline__ double exp_step(double x) { return fabs(x) < 0.125 ? exp_series(x) : exp(x); }

// One time step for the two paths of a lane.  z: normals of the price engine, y: of the volatility engine.
// MULT (log-models only): the state is the price itself and is multiplied by exp_step(increment), so that storing
// a column costs nothing; otherwise the state is the log-price and exp() is taken at the kept columns.
template <int MODEL, bool MULT = false>
__device__ __forceinline__ void advance(const PathArgs& a, double& s_0, double& s_1, double& v_0, double& v_1,
                                        double z0, double z1, double y0, double y1) {
    if (MODEL == ORP_MODEL_GBM_LOG) {
        // lnS_t = lnS_{t-1} + ((r - sigma^2/2) dt + sigma sqrt(dt) W_t)                      EO cell 6
        if (MULT) { s_0 *= exp_step(fma(a.c_vol, z0, a.c_drift)); s_1 *= exp_step(fma(a.c_vol, z1, a.c_drift)); }
        else { s_0 += fma(a.c_vol, z0, a.c_drift); s_1 += fma(a.c_vol, z1, a.c_drift); }
    } else if (MODEL == ORP_MODEL_GBM_ARITH) {
        // Y_t = Y_{t-1} + Y_{t-1} (mu dt + sigma sqrt(dt) W_t)                                RP.py:65
        s_0 = fma(s_0, fma(a.c_vol, z0, a.c_drift), s_0);
        s_1 = fma(s_1, fma(a.c_vol, z1, a.c_drift), s_1);
    } else if (MODEL == ORP_MODEL_LINEAR) {
        // l_t = l_{t-1} + (c l_{t-1} dt + ita sqrt(dt) W_t)                                    RP.py:76
        s_0 += fma(a.c_vol, z0, s_0 * a.c_drift);
        s_1 += fma(a.c_vol, z1, s_1 * a.c_drift);
    } else {
        // v_t = v + a (b - v) + c sqrt(v dt) Wv_t;  lnY_t = lnY + ((mu - v_t^2/2) dt + v_t sqrt(dt) W_t)   RP.py:285-287
        v_0 = (v_0 + a.a * (a.b - v_0)) + (a.c * sqrt(v_0 * a.dt)) * y0;
        v_1 = (v_1 + a.a * (a.b - v_1)) + (a.c * sqrt(v_1 * a.dt)) * y1;
        const double i0 = (a.c_drift - 0.5 * (v_0 * v_0)) * a.dt + (v_0 * a.c_vol) * z0;
        const double i1 = (a.c_drift - 0.5 * (v_1 * v_1)) * a.dt + (v_1 * a.c_vol) * z1;
        if (MULT) { s_0 *= exp_step(i0); s_1 *= exp_step(i1); } else { s_0 += i0; s_1 += i1; }
    }
}

__device__ __forceinline__ uint2 lds_u32x2(uint32_t saddr) {
    uint2 r;
    asm("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "r"(saddr));
    return r;
}

// Bytes of one engine's lane table in shared memory (one spare row: the loop prefetches one step ahead).
__host__ __device__ inline size_t tlane_smem_bytes(int n_steps) { return (size_t)(n_steps + 2) * 256; }

// FUSED (log-Euler GBM only): the table is transformed so that one Horner chain yields exp(drift + vol z);
// the state is then the price itself, S_t = S_{t-1} * factor — no exp, no separate normal.
// TLS: the lane tables (256 B per dimension and engine) are staged in shared memory behind the coefficient table
// when they fit (D <= ~300 for two engines); otherwise they are read through L1.
// ALL: keep_every == 1, every step is stored (no keep test in the loop).
// MULT: log-models with every column stored and small steps — price state, exp_step() per step, no exp() per column.
template <int MODEL, typename OutT, bool FUSED, bool TLS, bool ALL, bool MULT>
__global__ void __launch_bounds__(kPathThreads, 1) paths_kernel(const PathArgs a) {
    static_assert(!MULT || (ALL && !FUSED && !kLevel<MODEL>), "MULT is the keep-every-1 form of the unfused log-models");
    static_assert(!FUSED || MODEL == ORP_MODEL_GBM_LOG, "only the log-Euler GBM step fuses into the table");
    extern __shared__ __align__(128) char s_tab[];
    load_ndtri_table(reinterpret_cast<double*>(s_tab), threadIdx.x, blockDim.x);
    const int D = a.n_steps;
    if (TLS) {
        // rows 0 .. D of each engine's table, plus one zero row
        uint2* dst_a = reinterpret_cast<uint2*>(s_tab + kNdtriTableBytes);
        const int n = (D + 1) * 32;
        for (int i = threadIdx.x; i < n + 32; i += blockDim.x) dst_a[i] = i < n ? __ldg(a.tlane_a + i) : make_uint2(0u, 0u);
        if (MODEL == ORP_MODEL_SV) {
            uint2* dst_b = reinterpret_cast<uint2*>(s_tab + kNdtriTableBytes + tlane_smem_bytes(D));
            for (int i = threadIdx.x; i < n + 32; i += blockDim.x) dst_b[i] = i < n ? __ldg(a.tlane_b + i) : make_uint2(0u, 0u);
        }
    }
    __syncthreads();
    if (FUSED) {
        fuse_gbm_table(s_tab, a.c_drift, a.c_vol, threadIdx.x, blockDim.x);
        __syncthreads();
    }

    const uint32_t tab = (uint32_t)__cvta_generic_to_shared(s_tab);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t tls_a = tab + kNdtriTableBytes + lane * 8u;
    const uint32_t tls_b = tls_a + (uint32_t)tlane_smem_bytes(D);
    constexpr int kWarps = kPathThreads / 32;
    const uint64_t end = a.start + a.count;
    const uint64_t w_first = a.start / kWarpPaths, w_last = (end - 1) / kWarpPaths;
    const uint64_t n_tiles = w_last - w_first + 1;

    // block-contiguous tiling: the 32 warps of a block work on 2048 consecutive paths
    for (uint64_t tile = (uint64_t)blockIdx.x * kWarps + warp; tile < n_tiles; tile += (uint64_t)gridDim.x * kWarps) {
        const uint64_t w = w_first + tile;
        const uint64_t i0 = w * kWarpPaths + lane, i1 = i0 + 32;
        const bool ok0 = i0 >= a.start && i0 < end, ok1 = i1 >= a.start && i1 < end;
        const uint64_t j0 = i0 - a.start;

        // state: GBM_LOG ln S; GBM_ARITH S; SV ln Y and v
        double s_0, s_1, v_0 = a.v0, v_1 = a.v0;
        if (kLevel<MODEL> || FUSED || MULT) s_0 = s_1 = a.s0; else s_0 = s_1 = log(a.s0);
        OutT* __restrict__ optr = reinterpret_cast<OutT*>(a.out) + j0;       // running column pointer (+32 = second path)
        if (ok0) optr[0] = (OutT)a.s0;
        if (ok1) optr[32] = (OutT)a.s0;
        int next_keep = a.keep_every;

        for (int t0 = 1; t0 <= D; t0 += 32) {
            const int my_dim = t0 + lane;
            uint32_t bl_a = 0u, bl_b = 0u;
            if (my_dim <= D) {
                bl_a = warp_base(a.sv32_a, a.dpad, my_dim, w);
                if (MODEL == ORP_MODEL_SV) bl_b = warp_base(a.sv32_b, a.dpad, my_dim, w);
            }
            const int nc = min(32, D - t0 + 1);
            // lane-table entries are fetched one step ahead: their latency hides behind the current step
            uint32_t tp_a = tls_a + (uint32_t)t0 * 256u, tp_b = tls_b + (uint32_t)t0 * 256u;
            uint2 tl_next = TLS ? lds_u32x2(tp_a) : __ldg(a.tlane_a + (size_t)t0 * 32 + lane);
            uint2 tb_next = make_uint2(0u, 0u);
            if (MODEL == ORP_MODEL_SV) tb_next = TLS ? lds_u32x2(tp_b) : __ldg(a.tlane_b + (size_t)t0 * 32 + lane);
            for (int s = 0; s < nc; ++s) {
                const int t = t0 + s;
                const uint32_t base = __shfl_sync(kFull, bl_a, s);
                const uint2 tl = tl_next;
                if (TLS) { tp_a += 256u; tl_next = lds_u32x2(tp_a); }                      // row D + 1 exists (zeros)
                else tl_next = __ldg(a.tlane_a + (size_t)min(t + 1, D) * 32 + lane);
                if (FUSED) {
                    // exp(-inf) = 0 at the exact-zero coordinate, like np.exp of the reference's -inf log-price
                    s_0 *= table_eval(base ^ tl.x, tab);
                    s_1 *= table_eval(base ^ tl.y, tab);
                } else {
                    const double z0 = ndtri_bits(base ^ tl.x, tab);
                    const double z1 = ndtri_bits(base ^ tl.y, tab);
                    double y0 = 0.0, y1 = 0.0;
                    if (MODEL == ORP_MODEL_SV) {
                        const uint32_t base_b = __shfl_sync(kFull, bl_b, s);
                        const uint2 tb = tb_next;
                        if (TLS) { tp_b += 256u; tb_next = lds_u32x2(tp_b); }
                        else tb_next = __ldg(a.tlane_b + (size_t)min(t + 1, D) * 32 + lane);
                        y0 = ndtri_bits(base_b ^ tb.x, tab);
                        y1 = ndtri_bits(base_b ^ tb.y, tab);
                    }
                    advance<MODEL, MULT>(a, s_0, s_1, v_0, v_1, z0, z1, y0, y1);
                }
                if (ALL || t == next_keep) {
                    double o0 = s_0, o1 = s_1;
                    if (!kLevel<MODEL> && !FUSED && !MULT) { o0 = exp(s_0); o1 = exp(s_1); }
                    optr += a.ld;
                    if (ok0) optr[0] = (OutT)o0;
                    if (ok1) optr[32] = (OutT)o1;
                    if (!ALL) next_keep += a.keep_every;
                }
            }
        }
    }
}


// ------------------------------------------------------------------------------------------------
// Fused generate + payoff (north_star: "keeps them in registers/shared memory when the payoff is fused"; SURVEY 7.1
// K3 "fused as K2's epilogue so paths are never written").  Same tile shape, skip-ahead and step arithmetic as
// paths_kernel, but nothing is stored: at the time steps the options mature the two prices of a lane are paid off
// against that step's strikes (groups of <= 8 options sharing a maturity), the group's sums are reduced over the warp
// (reduce-scatter of the 8 doubles + two butterfly stages: 18 shuffles; the out-of-the-money tallies by integer warp
// reduction) and added to the warp's own accumulators in shared memory — a fixed order, so every launch gives the
// same bits.  Payoff formulas and NaN rules are those of payoff_partial_kernel (EO cell 8 raw 328-329, 344; RP.py:88-89).
// Algorithmic HBM traffic: none (a few KB of tables); the kernel is bound by the fp64 / shared-memory work of the
// path generation itself.
constexpr int kFusedGroup = 8;

struct FusedPayoff {
    int n_opt, n_groups, kind;
    double s0;                                   // FLOOR: out of the money means S < s0 (RP.py:89)
    double* partials;                            // [gridDim.x][n_opt][2]
    int32_t grp_step[ORP_MAX_OPTIONS];           // maturity (time step) of group g, ascending
    int16_t grp_first[ORP_MAX_OPTIONS], grp_n[ORP_MAX_OPTIONS];
    double strikes[ORP_MAX_OPTIONS];             // in group order
};

// (not inlined: it runs at a handful of time steps per tile; inlined it bloats the step loop and its register budget)
__device__ __noinline__ void fused_group(const FusedPayoff& f, int g, double p0, double p1, bool ok0, bool ok1, int lane,
                                            double* __restrict__ w_sum, unsigned* __restrict__ w_otm, unsigned long long* __restrict__ w_nan) {
    const int first = f.grp_first[g], n = f.grp_n[g];
    const int KIND = f.kind;                     // uniform over the launch
    double a[kFusedGroup];
    unsigned nan_bits = 0u;
    const bool nan0 = ok0 && p0 != p0, nan1 = ok1 && p1 != p1;
#pragma unroll
    for (int k = 0; k < kFusedGroup; ++k) {
        double pay0 = 0.0, pay1 = 0.0;
        unsigned out = 0u;
        if (k < n) {
            const double K = f.strikes[first + k];
            if (KIND == ORP_PAYOFF_FLOOR) {
                pay0 = p0 > K ? p0 : K;                                            // RP.py:88 (np.where: NaN -> K)
                pay1 = p1 > K ? p1 : K;
                out = (ok0 && p0 < f.s0 ? 1u : 0u) + (ok1 && p1 < f.s0 ? 1u : 0u);  // RP.py:89
            } else {
                const double d0 = KIND == ORP_PAYOFF_CALL ? p0 - K : K - p0, d1 = KIND == ORP_PAYOFF_CALL ? p1 - K : K - p1;
                const bool in0 = d0 > 0.0, in1 = d1 > 0.0;                         // false for NaN
                pay0 = in0 ? d0 : 0.0;
                pay1 = in1 ? d1 : 0.0;
                out = (ok0 && !in0 ? 1u : 0u) + (ok1 && !in1 ? 1u : 0u);
                if (nan0 || nan1) nan_bits |= 1u << k;                             // np.maximum propagates NaN into the mean
            }
        }
        a[k] = (ok0 ? pay0 : 0.0) + (ok1 ? pay1 : 0.0);
        const unsigned tot = __reduce_add_sync(kFull, out);
        if (lane == k && k < n) w_otm[first + k] += tot;
    }
    nan_bits = __reduce_or_sync(kFull, nan_bits);
    if (lane == 0 && nan_bits) *w_nan |= (unsigned long long)nan_bits << first;
    // reduce-scatter of the eight sums: after the stages with offsets 16, 8, 4 a lane holds the 8-lane partial of sum
    // idx = 4 b4 + 2 b3 + b2 (bN: bit N of the lane index); two butterfly stages finish it
    int len = kFusedGroup;
#pragma unroll
    for (int off = 16; off >= 4; off >>= 1) {
        const int half = len / 2;
        const bool up = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < kFusedGroup / 2; ++i)
            if (i < half) {
                const double lo = a[i], hi = a[i + half];
                a[i] = (up ? hi : lo) + __shfl_xor_sync(kFull, up ? lo : hi, off);
            }
        len = half;
    }
    a[0] += __shfl_xor_sync(kFull, a[0], 2);
    a[0] += __shfl_xor_sync(kFull, a[0], 1);
    const int idx = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
    if ((lane & 3) == 0 && idx < n) w_sum[first + idx] += a[0];
}

template <int MODEL, bool FUSED, bool TLS>
__global__ void __launch_bounds__(kPathThreads, 1) paths_payoff_kernel(const PathArgs a, const __grid_constant__ FusedPayoff f) {
    static_assert(!FUSED || MODEL == ORP_MODEL_GBM_LOG, "only the log-Euler GBM step fuses into the table");
    extern __shared__ __align__(128) char s_tab[];
    constexpr int kWarps = kPathThreads / 32;
    __shared__ double s_sum[kWarps][ORP_MAX_OPTIONS];
    __shared__ unsigned s_otm[kWarps][ORP_MAX_OPTIONS];
    __shared__ unsigned long long s_nan[kWarps];
    load_ndtri_table(reinterpret_cast<double*>(s_tab), threadIdx.x, blockDim.x);
    const int D = a.n_steps;
    if (TLS) {
        uint2* dst_a = reinterpret_cast<uint2*>(s_tab + kNdtriTableBytes);
        const int n = (D + 1) * 32;
        for (int i = threadIdx.x; i < n + 32; i += blockDim.x) dst_a[i] = i < n ? __ldg(a.tlane_a + i) : make_uint2(0u, 0u);
        if (MODEL == ORP_MODEL_SV) {
            uint2* dst_b = reinterpret_cast<uint2*>(s_tab + kNdtriTableBytes + tlane_smem_bytes(D));
            for (int i = threadIdx.x; i < n + 32; i += blockDim.x) dst_b[i] = i < n ? __ldg(a.tlane_b + i) : make_uint2(0u, 0u);
        }
    }
    for (int i = threadIdx.x; i < kWarps * ORP_MAX_OPTIONS; i += blockDim.x) { (&s_sum[0][0])[i] = 0.0; (&s_otm[0][0])[i] = 0u; }
    if (threadIdx.x < kWarps) s_nan[threadIdx.x] = 0ull;
    __syncthreads();
    if (FUSED) {
        fuse_gbm_table(s_tab, a.c_drift, a.c_vol, threadIdx.x, blockDim.x);
        __syncthreads();
    }

    const uint32_t tab = (uint32_t)__cvta_generic_to_shared(s_tab);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t tls_a = tab + kNdtriTableBytes + lane * 8u;
    const uint32_t tls_b = tls_a + (uint32_t)tlane_smem_bytes(D);
    const uint64_t end = a.start + a.count;
    const uint64_t w_first = a.start / kWarpPaths, w_last = (end - 1) / kWarpPaths;
    const uint64_t n_tiles = w_last - w_first + 1;
    double* const w_sum = s_sum[warp];
    unsigned* const w_otm = s_otm[warp];

    for (uint64_t tile = (uint64_t)blockIdx.x * kWarps + warp; tile < n_tiles; tile += (uint64_t)gridDim.x * kWarps) {
        const uint64_t w = w_first + tile;
        const uint64_t i0 = w * kWarpPaths + lane, i1 = i0 + 32;
        const bool ok0 = i0 >= a.start && i0 < end, ok1 = i1 >= a.start && i1 < end;
        double s_0, s_1, v_0 = a.v0, v_1 = a.v0;
        if (kLevel<MODEL> || FUSED) s_0 = s_1 = a.s0; else s_0 = s_1 = log(a.s0);
        int g = 0;
        while (g < f.n_groups && f.grp_step[g] == 0) { fused_group(f, g, a.s0, a.s0, ok0, ok1, lane, w_sum, w_otm, &s_nan[warp]); ++g; }
        int next_step = g < f.n_groups ? f.grp_step[g] : -1;

        for (int t0 = 1; t0 <= D && g < f.n_groups; t0 += 32) {          // nothing matures later: stop walking
            const int my_dim = t0 + lane;
            uint32_t bl_a = 0u, bl_b = 0u;
            if (my_dim <= D) {
                bl_a = warp_base(a.sv32_a, a.dpad, my_dim, w);
                if (MODEL == ORP_MODEL_SV) bl_b = warp_base(a.sv32_b, a.dpad, my_dim, w);
            }
            const int nc = min(32, D - t0 + 1);
            uint32_t tp_a = tls_a + (uint32_t)t0 * 256u, tp_b = tls_b + (uint32_t)t0 * 256u;
            uint2 tl_next = TLS ? lds_u32x2(tp_a) : __ldg(a.tlane_a + (size_t)t0 * 32 + lane);
            uint2 tb_next = make_uint2(0u, 0u);
            if (MODEL == ORP_MODEL_SV) tb_next = TLS ? lds_u32x2(tp_b) : __ldg(a.tlane_b + (size_t)t0 * 32 + lane);
            int s = 0;
            while (s < nc) {
                // steps up to and including the next maturity of this chunk run without a test per step
                const int s_end = next_step >= t0 + s && next_step - t0 < nc ? next_step - t0 + 1 : nc;
#pragma unroll 2
                for (; s < s_end; ++s) {
                    const int t = t0 + s;
                    const uint32_t base = __shfl_sync(kFull, bl_a, s);
                    const uint2 tl = tl_next;
                    if (TLS) { tp_a += 256u; tl_next = lds_u32x2(tp_a); }
                    else tl_next = __ldg(a.tlane_a + (size_t)min(t + 1, D) * 32 + lane);
                    if (FUSED) {
                        s_0 *= table_eval(base ^ tl.x, tab);
                        s_1 *= table_eval(base ^ tl.y, tab);
                    } else {
                        const double z0 = ndtri_bits(base ^ tl.x, tab);
                        const double z1 = ndtri_bits(base ^ tl.y, tab);
                        double y0 = 0.0, y1 = 0.0;
                        if (MODEL == ORP_MODEL_SV) {
                            const uint32_t base_b = __shfl_sync(kFull, bl_b, s);
                            const uint2 tb = tb_next;
                            if (TLS) { tp_b += 256u; tb_next = lds_u32x2(tp_b); }
                            else tb_next = __ldg(a.tlane_b + (size_t)min(t + 1, D) * 32 + lane);
                            y0 = ndtri_bits(base_b ^ tb.x, tab);
                            y1 = ndtri_bits(base_b ^ tb.y, tab);
                        }
                        advance<MODEL, false>(a, s_0, s_1, v_0, v_1, z0, z1, y0, y1);
                    }
                }
                const int t = t0 + s - 1;
                if (t == next_step) {                                       // warp-uniform
                    double p0 = s_0, p1 = s_1;
                    if (!kLevel<MODEL> && !FUSED) { p0 = exp(s_0); p1 = exp(s_1); }
                    do {
                        fused_group(f, g, p0, p1, ok0, ok1, lane, w_sum, w_otm, &s_nan[warp]);
                        ++g;
                    } while (g < f.n_groups && f.grp_step[g] == t);
                    next_step = g < f.n_groups ? f.grp_step[g] : -1;
                }
            }
        }
    }
    __syncthreads();
    // block totals: the warps' accumulators in warp order
    for (int o = threadIdx.x; o < f.n_opt; o += blockDim.x) {
        double t = 0.0;
        unsigned c = 0u;
        bool bad = false;
        for (int wq = 0; wq < kWarps; ++wq) { t += s_sum[wq][o]; c += s_otm[wq][o]; bad = bad || ((s_nan[wq] >> o) & 1ull); }
        f.partials[((size_t)blockIdx.x * f.n_opt + o) * 2] = bad ? __longlong_as_double(0x7ff8000000000000ll) : t;
        f.partials[((size_t)blockIdx.x * f.n_opt + o) * 2 + 1] = (double)c;
    }
}

struct FusedPerm { int16_t to_caller[ORP_MAX_OPTIONS]; };          // group-order index -> the caller's option index

// second pass: one warp per (option, quantity); lane l adds blocks l, l + 32, ... in increasing order, then a shuffle tree
__global__ void __launch_bounds__(128) paths_payoff_final_kernel(const double* __restrict__ partials, int n_blocks, int n_opt,
                                                                 const FusedPerm perm, double* __restrict__ sums, double* __restrict__ otm) {
    const int o = (blockIdx.x * 128 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (o >= 2 * n_opt) return;
    const int opt = o >> 1, which = o & 1;
    double t = 0.0;
    for (int b = lane; b < n_blocks; b += 32) t += partials[((size_t)b * n_opt + opt) * 2 + which];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) t += __shfl_down_sync(kFull, t, off);
    if (lane == 0) (which ? otm : sums)[perm.to_caller[opt]] = t;
}


// ------------------------------------------------------------------------------------------------
// Brownian-bridge ordering (NEW feature; specification in oracle/bridge.py, no reference counterpart).
// Sobol coordinates are random-access in the DIMENSION as well as in the index, so the bridge needs no
// buffering of a path's D normals: the bisection tree is walked depth-first in time order with a stack
// of at most ceil(log2 D) + 1 pending right end-points per path.  The walk is the same for every path,
// so the host compiles it into a program of D ops (orp_bridge_schedule): op o builds one node from
// Sobol dimension ops[o].dim and is followed by ops[o].emits emissions (pop + time step).
struct BridgeOp { int32_t dim; int32_t emits; double wl, wr, sd; };
constexpr int kBridgeMaxDepth = 20;
// the schedule's bytes in shared memory (a multiple of 128 so that the stack columns behind it stay aligned)
__host__ __device__ inline size_t bridge_ops_smem_bytes(int n_steps) { return ((size_t)n_steps * sizeof(BridgeOp) + 127) / 128 * 128; }

// SSTK: the stack lives in shared memory behind the coefficient table and the schedule (level-major, one 8-byte
// column per thread: conflict-free), its TOP in registers — a push stores the old top and needs no load at all, a pop
// loads the entry below for later use.  Chosen when (depth - 1) levels fit (D <= 512 for the one-engine models);
// otherwise the stack is a per-thread array in local memory.  With it the schedule is staged in shared memory too
// and the lane-table entry of the next op is requested one op ahead (its Sobol dimension comes from the lanes that
// evaluated the chunk's bases), so no step waits on two dependent global loads.
template <int MODEL, typename OutT, bool MULT, bool SSTK>
__global__ void __launch_bounds__(kPathThreads, 1) paths_bridge_kernel(const PathArgs a, const BridgeOp* __restrict__ ops) {
    extern __shared__ __align__(128) char s_tab[];
    load_ndtri_table(reinterpret_cast<double*>(s_tab), threadIdx.x, blockDim.x);
    const int D = a.n_steps;
    BridgeOp* s_ops = reinterpret_cast<BridgeOp*>(s_tab + kNdtriTableBytes);
    if (SSTK) {
        const ulonglong2* src = reinterpret_cast<const ulonglong2*>(ops);
        ulonglong2* dst = reinterpret_cast<ulonglong2*>(s_ops);
        for (int i = threadIdx.x; i < D * 2; i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();

    const uint32_t tab = (uint32_t)__cvta_generic_to_shared(s_tab);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int kWarps = kPathThreads / 32;
    constexpr int NS = MODEL == ORP_MODEL_SV ? 4 : 2;      // stacks: price (2 paths) [+ volatility (2 paths)]
    const uint64_t end = a.start + a.count;
    const uint64_t w_first = a.start / kWarpPaths, w_last = (end - 1) / kWarpPaths;
    const uint64_t n_tiles = w_last - w_first + 1;
    const BridgeOp* __restrict__ prog = SSTK ? s_ops : ops;
    // level l of stack q of this thread: s_stk[(l * NS + q) * kPathThreads]
    double* s_stk = reinterpret_cast<double*>(s_tab + kNdtriTableBytes + bridge_ops_smem_bytes(D)) + threadIdx.x;

    for (uint64_t tile = (uint64_t)blockIdx.x * kWarps + warp; tile < n_tiles; tile += (uint64_t)gridDim.x * kWarps) {
        const uint64_t w = w_first + tile;
        const uint64_t i0 = w * kWarpPaths + lane, i1 = i0 + 32;
        const bool ok0 = i0 >= a.start && i0 < end, ok1 = i1 >= a.start && i1 < end;
        const uint64_t j0 = i0 - a.start;

        double s_0, s_1, v_0 = a.v0, v_1 = a.v0;
        if (kLevel<MODEL> || MULT) s_0 = s_1 = a.s0; else s_0 = s_1 = log(a.s0);
        OutT* __restrict__ optr = reinterpret_cast<OutT*>(a.out) + j0;       // running column pointer (+32 = second path)
        if (ok0) optr[0] = (OutT)a.s0;
        if (ok1) optr[32] = (OutT)a.s0;
        int next_keep = a.keep_every, t = 0, sp = -1;
        double stk[NS][SSTK ? 1 : kBridgeMaxDepth];
        double top[NS];                                      // SSTK: the stack's top entry (valid while sp >= 0)
        double cur[NS];
#pragma unroll
        for (int q = 0; q < NS; ++q) { cur[q] = 0.0; top[q] = 0.0; }           // W(0) = 0

        for (int o0 = 0; o0 < D; o0 += 32) {
            uint32_t bl_a = 0u, bl_b = 0u;
            int my_dim = 0;
            if (o0 + lane < D) {
                my_dim = prog[o0 + lane].dim;
                bl_a = warp_base(a.sv32_a, a.dpad, my_dim, w);
                if (MODEL == ORP_MODEL_SV) bl_b = warp_base(a.sv32_b, a.dpad, my_dim, w);
            }
            const int nc = min(32, D - o0);
            uint2 tl_next = make_uint2(0u, 0u), tb_next = make_uint2(0u, 0u);
            if (SSTK) {
                const int d0 = __shfl_sync(kFull, my_dim, 0);
                tl_next = __ldg(a.tlane_a + (size_t)d0 * 32 + lane);
                if (MODEL == ORP_MODEL_SV) tb_next = __ldg(a.tlane_b + (size_t)d0 * 32 + lane);
            }
            for (int s = 0; s < nc; ++s) {
                const BridgeOp op = prog[o0 + s];
                double zz[NS];
                uint2 tl, tb = make_uint2(0u, 0u);
                if (SSTK) {
                    tl = tl_next; tb = tb_next;
                    const int dn = __shfl_sync(kFull, my_dim, (s + 1) & 31);       // past the chunk: a valid row, value unused
                    tl_next = __ldg(a.tlane_a + (size_t)dn * 32 + lane);
                    if (MODEL == ORP_MODEL_SV) tb_next = __ldg(a.tlane_b + (size_t)dn * 32 + lane);
                } else {
                    tl = __ldg(a.tlane_a + (size_t)op.dim * 32 + lane);
                    if (MODEL == ORP_MODEL_SV) tb = __ldg(a.tlane_b + (size_t)op.dim * 32 + lane);
                }
                {
                    const uint32_t base = __shfl_sync(kFull, bl_a, s);
                    zz[0] = ndtri_bits(base ^ tl.x, tab);
                    zz[1] = ndtri_bits(base ^ tl.y, tab);
                }
                if (MODEL == ORP_MODEL_SV) {
                    const uint32_t base_b = __shfl_sync(kFull, bl_b, s);
                    zz[NS - 2] = ndtri_bits(base_b ^ tb.x, tab);
                    zz[NS - 1] = ndtri_bits(base_b ^ tb.y, tab);
                }
                // W_m = wl W_left + wr W_right + sd z   (left = last emitted node, right = stack top)
#pragma unroll
                for (int q = 0; q < NS; ++q) {
                    if (SSTK) {
                        // (the first op has no right end-point: its wr is 0 and top still holds its initial 0)
                        if (sp >= 0) s_stk[(size_t)(sp * NS + q) * kPathThreads] = top[q];
                        top[q] = op.wl * cur[q] + op.wr * top[q] + op.sd * zz[q];
                    } else {
                        const double right = sp >= 0 ? stk[q][sp] : 0.0;
                        stk[q][sp + 1] = op.wl * cur[q] + op.wr * right + op.sd * zz[q];
                    }
                }
                ++sp;
                for (int e = 0; e < op.emits; ++e) {
                    double inc[NS];
#pragma unroll
                    for (int q = 0; q < NS; ++q) {
                        const double wj = SSTK ? top[q] : stk[q][sp];
                        inc[q] = wj - cur[q];
                        cur[q] = wj;
                    }
                    --sp;
                    if (SSTK && sp >= 0) {
#pragma unroll
                        for (int q = 0; q < NS; ++q) top[q] = s_stk[(size_t)(sp * NS + q) * kPathThreads];
                    }
                    ++t;
                    advance<MODEL, MULT>(a, s_0, s_1, v_0, v_1, inc[0], inc[1], inc[NS - 2], inc[NS - 1]);
                    if (t == next_keep) {
                        double q0 = s_0, q1 = s_1;
                        if (!kLevel<MODEL> && !MULT) { q0 = exp(s_0); q1 = exp(s_1); }
                        optr += a.ld;
                        if (ok0) optr[0] = (OutT)q0;
                        if (ok1) optr[32] = (OutT)q1;
                        next_keep += a.keep_every;
                    }
                }
            }
        }
    }
}

template <typename K>
static int set_table_smem(K kernel, size_t bytes = kNdtriTableBytes) {
    ORP_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return ORP_OK;
}

constexpr size_t kMaxDynSmem = 227 * 1024;       // per-block opt-in limit on sm_100

// Six standard deviations of a one-step log-increment stay inside exp_step()'s series range (|x| < 1/8)?
static bool small_steps(const PathArgs& a, int model) {
    if (model == ORP_MODEL_GBM_LOG) return 6.0 * fabs(a.c_vol) + fabs(a.c_drift) < 0.125;
    const double v = fmax(fabs(a.v0), fabs(a.b)) * 1.5;            // SV: a bound on the volatility level (beyond it exp_step falls back to exp)
    return 6.0 * v * fabs(a.c_vol) + fabs(a.c_drift - 0.5 * v * v) * fabs(a.dt) < 0.125;
}

template <int MODEL, typename OutT, bool FUSED, bool TLS, bool ALL, bool MULT = false>
static int launch_paths_v(const PathArgs& a, size_t smem, cudaStream_t st) {
    int rc = set_table_smem(paths_kernel<MODEL, OutT, FUSED, TLS, ALL, MULT>, smem);
    if (rc) return rc;
    const uint64_t n_tiles = (a.start + a.count - 1) / kWarpPaths - a.start / kWarpPaths + 1;
    const uint64_t want = (n_tiles + 31) / 32;
    const int grid = (int)(want < (uint64_t)sm_count() ? want : (uint64_t)sm_count());
    paths_kernel<MODEL, OutT, FUSED, TLS, ALL, MULT><<<grid, kPathThreads, smem, st>>>(a);
    ORP_LAUNCH_CHECK("paths_kernel");
    return ORP_OK;
}

template <int MODEL, typename OutT, bool FUSED = false>
static int launch_paths(const PathArgs& a, cudaStream_t st) {
    const size_t engines = MODEL == ORP_MODEL_SV ? 2 : 1;
    const size_t with_tl = kNdtriTableBytes + engines * tlane_smem_bytes(a.n_steps);
    const bool tls = with_tl <= kMaxDynSmem && !getenv("ORP_B200_NO_TLS");
    const bool all = a.keep_every == 1;
    if constexpr (!FUSED && (MODEL == ORP_MODEL_GBM_LOG || MODEL == ORP_MODEL_SV)) {
        if (all && small_steps(a, MODEL) && !getenv("ORP_B200_NO_MULT"))
            return tls ? launch_paths_v<MODEL, OutT, false, true, true, true>(a, with_tl, st)
                       : launch_paths_v<MODEL, OutT, false, false, true, true>(a, kNdtriTableBytes, st);
    }
    if (tls) return all ? launch_paths_v<MODEL, OutT, FUSED, true, true>(a, with_tl, st)
                        : launch_paths_v<MODEL, OutT, FUSED, true, false>(a, with_tl, st);
    return all ? launch_paths_v<MODEL, OutT, FUSED, false, true>(a, kNdtriTableBytes, st)
               : launch_paths_v<MODEL, OutT, FUSED, false, false>(a, kNdtriTableBytes, st);
}


// Deepest stack of the depth-first bridge walk over D steps (the walk of orp_bridge_schedule without the weights).
static int bridge_depth(int D) {
    std::vector<int> stack{D};
    int left = 0, depth = 1;
    for (int j = 1; j <= D; ++j) {
        while (stack.back() > j) {
            stack.push_back((left + stack.back()) / 2);
            if ((int)stack.size() > depth) depth = (int)stack.size();
        }
        stack.pop_back();
        left = j;
    }
    return depth;
}

template <int MODEL, typename OutT, bool MULT, bool SSTK>
static int launch_paths_bridge_v(const PathArgs& a, const void* d_bridge, size_t smem, cudaStream_t st) {
    int rc = set_table_smem(paths_bridge_kernel<MODEL, OutT, MULT, SSTK>, smem);
    if (rc) return rc;
    const uint64_t n_tiles = (a.start + a.count - 1) / kWarpPaths - a.start / kWarpPaths + 1;
    const uint64_t want = (n_tiles + 31) / 32;
    const int grid = (int)(want < (uint64_t)sm_count() ? want : (uint64_t)sm_count());
    paths_bridge_kernel<MODEL, OutT, MULT, SSTK><<<grid, kPathThreads, smem, st>>>(a, reinterpret_cast<const BridgeOp*>(d_bridge));
    ORP_LAUNCH_CHECK("paths_bridge_kernel");
    return ORP_OK;
}

template <int MODEL, typename OutT>
static int launch_paths_bridge(const PathArgs& a, const void* d_bridge, cudaStream_t st) {
    constexpr size_t n_stacks = MODEL == ORP_MODEL_SV ? 4 : 2;
    // the top entry stays in registers: depth - 1 levels of one 8-byte column per thread and stack
    const size_t with_stack = kNdtriTableBytes + bridge_ops_smem_bytes(a.n_steps)
                              + (size_t)(bridge_depth(a.n_steps) - 1) * n_stacks * kPathThreads * sizeof(double);
    const bool sstk = with_stack <= kMaxDynSmem && !getenv("ORP_B200_NO_SMEM_STACK");
    if constexpr (MODEL == ORP_MODEL_GBM_LOG || MODEL == ORP_MODEL_SV) {
        if (a.keep_every <= 2 && small_steps(a, MODEL) && !getenv("ORP_B200_NO_MULT"))
            return sstk ? launch_paths_bridge_v<MODEL, OutT, true, true>(a, d_bridge, with_stack, st)
                        : launch_paths_bridge_v<MODEL, OutT, true, false>(a, d_bridge, kNdtriTableBytes, st);
    }
    return sstk ? launch_paths_bridge_v<MODEL, OutT, false, true>(a, d_bridge, with_stack, st)
                : launch_paths_bridge_v<MODEL, OutT, false, false>(a, d_bridge, kNdtriTableBytes, st);
}

// PathArgs of a launch from the caller's parameter block (shared by orp_paths and orp_paths_payoff)
static int build_path_args(const orp_path_params* p, const uint32_t* d_sv32, const uint32_t* d_tlane, const uint32_t* d_sv32_b,
                           const uint32_t* d_tlane_b, uint64_t start, uint64_t count, PathArgs& a) {
    a.sv32_a = d_sv32;  a.tlane_a = reinterpret_cast<const uint2*>(d_tlane);
    a.sv32_b = d_sv32_b; a.tlane_b = reinterpret_cast<const uint2*>(d_tlane_b);
    a.out = nullptr; a.ld = 0; a.start = start; a.count = count;
    a.n_steps = p->n_steps; a.keep_every = p->keep_every;
    const int sobol_dim = p->sobol_dim > 0 ? p->sobol_dim : p->n_steps + 1;
    ORP_REQUIRE(sobol_dim >= p->n_steps + 1 && sobol_dim <= ORP_SOBOL_MAXDIM, "sobol_dim must be in [n_steps + 1, %d]", ORP_SOBOL_MAXDIM);
    a.dpad = ORP_SV_DPAD(sobol_dim);
    a.s0 = p->s0; a.dt = p->dt; a.v0 = p->v0; a.a = p->a; a.b = p->b; a.c = p->c;
    const double sqdt = sqrt(p->dt);
    switch (p->model) {
        case ORP_MODEL_GBM_LOG:   a.c_drift = (p->mu - 0.5 * p->sigma * p->sigma) * p->dt; a.c_vol = p->sigma * sqdt; break;
        case ORP_MODEL_GBM_ARITH: a.c_drift = p->mu * p->dt;                              a.c_vol = p->sigma * sqdt; break;
        case ORP_MODEL_SV:        a.c_drift = p->mu;                                      a.c_vol = sqdt;            break;
        case ORP_MODEL_LINEAR:    a.c_drift = p->mu * p->dt;                              a.c_vol = p->sigma * sqdt; break;
        default: return fail(ORP_E_VALUE, "unknown model %d", p->model);
    }
    return ORP_OK;
}

template <int MODEL, bool FUSED>
static int launch_paths_payoff(const PathArgs& a, const FusedPayoff& f, int grid, cudaStream_t st) {
    const size_t engines = MODEL == ORP_MODEL_SV ? 2 : 1;
    const size_t with_tl = kNdtriTableBytes + engines * tlane_smem_bytes(a.n_steps);
    // static shared memory of the kernel (the warps' accumulators) counts against the same 227 KB
    const bool tls = with_tl + 26 * 1024 <= kMaxDynSmem && !getenv("ORP_B200_NO_TLS");
    if (tls) {
        int rc = set_table_smem(paths_payoff_kernel<MODEL, FUSED, true>, with_tl);
        if (rc) return rc;
        paths_payoff_kernel<MODEL, FUSED, true><<<grid, kPathThreads, with_tl, st>>>(a, f);
    } else {
        int rc = set_table_smem(paths_payoff_kernel<MODEL, FUSED, false>, kNdtriTableBytes);
        if (rc) return rc;
        paths_payoff_kernel<MODEL, FUSED, false><<<grid, kPathThreads, kNdtriTableBytes, st>>>(a, f);
    }
    ORP_LAUNCH_CHECK("paths_payoff_kernel");
    return ORP_OK;
}

}  // namespace orp

using namespace orp;

extern "C" uint64_t orp_bridge_schedule_bytes(int n_steps) { return n_steps > 0 ? (uint64_t)n_steps * sizeof(BridgeOp) : 0; }

// Breadth-first bisection numbers the nodes (Sobol dimension k drives the k-th node built); the ops are
// then emitted in depth-first time order.  Same tree, same weights as oracle/bridge.py::bridge_schedule.
extern "C" int orp_bridge_schedule(int n_steps, void* sched_out) {
    ORP_REQUIRE(sched_out, "null pointer");
    ORP_REQUIRE(n_steps >= 1 && n_steps + 1 <= ORP_SOBOL_MAXDIM, "n_steps must be in [1, %d]", ORP_SOBOL_MAXDIM - 1);
    const int D = n_steps;
    std::vector<int> dim_of(D + 1, 0);
    {
        std::vector<std::pair<int, int>> q;
        q.reserve(2 * (size_t)D + 2);
        dim_of[D] = 1;
        int next_dim = 2;
        q.emplace_back(0, D);
        for (size_t h = 0; h < q.size(); ++h) {
            const int l = q[h].first, r = q[h].second;
            if (r - l >= 2) {
                const int m = (l + r) / 2;
                dim_of[m] = next_dim++;
                q.emplace_back(l, m);
                q.emplace_back(m, r);
            }
        }
    }
    BridgeOp* ops = reinterpret_cast<BridgeOp*>(sched_out);
    int n_ops = 0, depth_max = 0;
    std::vector<int> stack;                      // node indices, top = nearest pending right end-point
    ops[n_ops++] = BridgeOp{dim_of[D], 0, 0.0, 0.0, sqrt((double)D)};
    stack.push_back(D);
    int left = 0;
    for (int j = 1; j <= D; ++j) {
        while (stack.back() > j) {
            const int r = stack.back(), m = (left + r) / 2;
            ops[n_ops++] = BridgeOp{dim_of[m], 0, (double)(r - m) / (double)(r - left), (double)(m - left) / (double)(r - left),
                                    sqrt((double)(m - left) * (double)(r - m) / (double)(r - left))};
            stack.push_back(m);
            if ((int)stack.size() > depth_max) depth_max = (int)stack.size();
        }
        ops[n_ops - 1].emits += 1;               // emission of node j follows the last op executed so far
        stack.pop_back();
        left = j;
    }
    ORP_REQUIRE(n_ops == D && depth_max < kBridgeMaxDepth, "bridge schedule inconsistent (ops %d, depth %d)", n_ops, depth_max);
    return ORP_OK;
}

extern "C" int orp_ndtri_u32(const uint32_t* d_x, uint64_t n, double* d_z, void* stream) {
    ORP_REQUIRE(d_x && d_z, "null pointer");
    if (n == 0) return ORP_OK;
    int rc = set_table_smem(ndtri_u32_kernel);
    if (rc) return rc;
    const uint64_t blocks = (n + 255) / 256;
    const int grid = (int)(blocks < (uint64_t)sm_count() ? blocks : (uint64_t)sm_count());
    ndtri_u32_kernel<<<grid, 256, kNdtriTableBytes, (cudaStream_t)stream>>>(d_x, n, d_z);
    ORP_LAUNCH_CHECK("ndtri_u32_kernel");
    return ORP_OK;
}

extern "C" int orp_sobol_points(const uint32_t* d_sv32, const uint32_t* d_tlane, int d, uint64_t start, uint64_t count,
                                int layout, uint32_t* d_out, void* stream) {
    ORP_REQUIRE(d_sv32 && d_tlane && (d_out || count == 0), "null pointer");
    ORP_REQUIRE(d >= 1 && d <= ORP_SOBOL_MAXDIM, "Maximum supported dimensionality is %d.", ORP_SOBOL_MAXDIM);
    ORP_REQUIRE(layout == 0 || layout == 1, "layout must be 0 (points-major) or 1 (dimension-major)");
    ORP_REQUIRE(start + count <= (1ull << ORP_SOBOL_BITS), "At most 2**%d=%llu distinct points can be generated.",
                ORP_SOBOL_BITS, 1ull << ORP_SOBOL_BITS);
    if (count == 0) return ORP_OK;
    const uint64_t n_tiles = (start + count - 1) / kWarpPaths - start / kWarpPaths + 1;
    const uint64_t want = (n_tiles + kPointsWarps - 1) / kPointsWarps;
    const uint64_t cap = (uint64_t)sm_count() * 8;
    const int grid = (int)(want < cap ? want : cap);
    const uint2* tl = reinterpret_cast<const uint2*>(d_tlane);
    if (layout == 0)
        sobol_points_kernel<0><<<grid, kPointsWarps * 32, 0, (cudaStream_t)stream>>>(d_sv32, tl, d, start, count, d_out);
    else
        sobol_points_kernel<1><<<grid, kPointsWarps * 32, 0, (cudaStream_t)stream>>>(d_sv32, tl, d, start, count, d_out);
    ORP_LAUNCH_CHECK("sobol_points_kernel");
    return ORP_OK;
}

extern "C" int orp_paths(const orp_path_params* p, const uint32_t* d_sv32, const uint32_t* d_tlane,
                         const uint32_t* d_sv32_b, const uint32_t* d_tlane_b, const void* d_bridge, uint64_t start,
                         uint64_t count, void* d_out, uint64_t ld, void* stream) {
    ORP_REQUIRE(p && d_sv32 && d_tlane && (d_out || count == 0), "null pointer");
    ORP_REQUIRE(p->n_steps >= 1 && p->n_steps + 1 <= ORP_SOBOL_MAXDIM, "n_steps must be in [1, %d]", ORP_SOBOL_MAXDIM - 1);
    ORP_REQUIRE(p->keep_every >= 1, "keep_every must be >= 1");
    ORP_REQUIRE(start + count <= (1ull << ORP_SOBOL_BITS), "At most 2**%d=%llu distinct points can be generated.",
                ORP_SOBOL_BITS, 1ull << ORP_SOBOL_BITS);
    ORP_REQUIRE(ld >= count, "ld (%llu) must be >= count (%llu)", (unsigned long long)ld, (unsigned long long)count);
    ORP_REQUIRE(p->out_dtype == ORP_F64 || p->out_dtype == ORP_F32, "out_dtype must be ORP_F64 or ORP_F32");
    ORP_REQUIRE(p->model != ORP_MODEL_SV || (d_sv32_b && d_tlane_b), "the SV model needs a second Sobol engine");
    ORP_REQUIRE(p->ordering == ORP_ORDER_INCREMENTAL || (p->ordering == ORP_ORDER_BRIDGE && d_bridge),
                "ordering must be ORP_ORDER_INCREMENTAL, or ORP_ORDER_BRIDGE with a schedule");
    if (count == 0) return ORP_OK;

    PathArgs a;
    int rc = build_path_args(p, d_sv32, d_tlane, d_sv32_b, d_tlane_b, start, count, a);
    if (rc) return rc;
    a.out = d_out; a.ld = ld;
    if (p->ordering == ORP_ORDER_BRIDGE) {
        switch (p->model) {
            case ORP_MODEL_GBM_LOG:   return p->out_dtype == ORP_F32 ? launch_paths_bridge<ORP_MODEL_GBM_LOG, float>(a, d_bridge, (cudaStream_t)stream)   : launch_paths_bridge<ORP_MODEL_GBM_LOG, double>(a, d_bridge, (cudaStream_t)stream);
            case ORP_MODEL_GBM_ARITH: return p->out_dtype == ORP_F32 ? launch_paths_bridge<ORP_MODEL_GBM_ARITH, float>(a, d_bridge, (cudaStream_t)stream) : launch_paths_bridge<ORP_MODEL_GBM_ARITH, double>(a, d_bridge, (cudaStream_t)stream);
            case ORP_MODEL_LINEAR:    return p->out_dtype == ORP_F32 ? launch_paths_bridge<ORP_MODEL_LINEAR, float>(a, d_bridge, (cudaStream_t)stream)    : launch_paths_bridge<ORP_MODEL_LINEAR, double>(a, d_bridge, (cudaStream_t)stream);
            default:                  return p->out_dtype == ORP_F32 ? launch_paths_bridge<ORP_MODEL_SV, float>(a, d_bridge, (cudaStream_t)stream)        : launch_paths_bridge<ORP_MODEL_SV, double>(a, d_bridge, (cudaStream_t)stream);
        }
    }
    cudaStream_t st = (cudaStream_t)stream;
    const bool f32 = p->out_dtype == ORP_F32;
    switch (p->model) {
        case ORP_MODEL_GBM_LOG:
            // |vol| < 1/4 per step keeps the table transform's truncation below 1e-15; larger steps take the generic path
            if (fabs(a.c_vol) < 0.25 && fabs(a.c_drift) < 1.0 && !getenv("ORP_B200_NO_FUSE"))
                return f32 ? launch_paths<ORP_MODEL_GBM_LOG, float, true>(a, st) : launch_paths<ORP_MODEL_GBM_LOG, double, true>(a, st);
            return f32 ? launch_paths<ORP_MODEL_GBM_LOG, float>(a, st) : launch_paths<ORP_MODEL_GBM_LOG, double>(a, st);
        case ORP_MODEL_GBM_ARITH: return f32 ? launch_paths<ORP_MODEL_GBM_ARITH, float>(a, st) : launch_paths<ORP_MODEL_GBM_ARITH, double>(a, st);
        case ORP_MODEL_LINEAR:    return f32 ? launch_paths<ORP_MODEL_LINEAR, float>(a, st)    : launch_paths<ORP_MODEL_LINEAR, double>(a, st);
        default:                  return f32 ? launch_paths<ORP_MODEL_SV, float>(a, st)        : launch_paths<ORP_MODEL_SV, double>(a, st);
    }
}

static int fused_grid(uint64_t start, uint64_t count) {
    const uint64_t n_tiles = (start + count - 1) / kWarpPaths - start / kWarpPaths + 1;
    const uint64_t want = (n_tiles + 31) / 32;
    return (int)(want < (uint64_t)sm_count() ? want : (uint64_t)sm_count());
}

extern "C" uint64_t orp_paths_payoff_scratch_bytes(int n_opt) {
    if (n_opt < 1) n_opt = 1;
    if (n_opt > ORP_MAX_OPTIONS) n_opt = ORP_MAX_OPTIONS;
    return 1024ull * (uint64_t)n_opt * 2 * sizeof(double);       // one partial pair per option and block; <= 1024 blocks
}

extern "C" int orp_paths_payoff(const orp_path_params* p, const uint32_t* d_sv32, const uint32_t* d_tlane,
                                const uint32_t* d_sv32_b, const uint32_t* d_tlane_b, uint64_t start, uint64_t count,
                                int kind, int n_opt, const double* strikes, const int32_t* steps, double s0,
                                double* d_sums, double* d_otm, void* d_scratch, void* stream) {
    ORP_REQUIRE(p && d_sv32 && d_tlane && strikes && steps && d_sums && d_otm && d_scratch, "null pointer");
    ORP_REQUIRE(p->n_steps >= 1 && p->n_steps + 1 <= ORP_SOBOL_MAXDIM, "n_steps must be in [1, %d]", ORP_SOBOL_MAXDIM - 1);
    ORP_REQUIRE(count >= 1 && start + count <= (1ull << ORP_SOBOL_BITS), "At most 2**%d=%llu distinct points can be generated.",
                ORP_SOBOL_BITS, 1ull << ORP_SOBOL_BITS);
    ORP_REQUIRE(p->ordering == ORP_ORDER_INCREMENTAL, "the fused kernel walks the time axis in incremental order only");
    ORP_REQUIRE(p->model != ORP_MODEL_SV || (d_sv32_b && d_tlane_b), "the SV model needs a second Sobol engine");
    ORP_REQUIRE(n_opt >= 1 && n_opt <= ORP_MAX_OPTIONS, "n_opt must be in [1, %d]", ORP_MAX_OPTIONS);
    ORP_REQUIRE(kind >= ORP_PAYOFF_CALL && kind <= ORP_PAYOFF_FLOOR, "unknown payoff kind %d", kind);
    PathArgs a;
    int rc = build_path_args(p, d_sv32, d_tlane, d_sv32_b, d_tlane_b, start, count, a);
    if (rc) return rc;
    // options grouped by maturity step (ascending, stable), at most kFusedGroup per group
    FusedPayoff f;
    FusedPerm perm;
    int order[ORP_MAX_OPTIONS];
    for (int k = 0; k < n_opt; ++k) {
        ORP_REQUIRE(steps[k] >= 0 && steps[k] <= p->n_steps, "maturity step %d of option %d is outside [0, %d]", steps[k], k, p->n_steps);
        order[k] = k;
    }
    std::stable_sort(order, order + n_opt, [&](int x, int y) { return steps[x] < steps[y]; });
    f.n_opt = n_opt; f.kind = kind; f.s0 = s0; f.n_groups = 0;
    f.partials = reinterpret_cast<double*>(d_scratch);
    for (int k = 0; k < n_opt; ++k) {
        const int o = order[k];
        f.strikes[k] = strikes[o];
        perm.to_caller[k] = (int16_t)o;
        if (f.n_groups == 0 || f.grp_step[f.n_groups - 1] != steps[o] || f.grp_n[f.n_groups - 1] == kFusedGroup) {
            f.grp_step[f.n_groups] = steps[o];
            f.grp_first[f.n_groups] = (int16_t)k;
            f.grp_n[f.n_groups] = 0;
            ++f.n_groups;
        }
        f.grp_n[f.n_groups - 1] += 1;
    }
    int grid = fused_grid(start, count);
    if (grid > 1024) grid = 1024;
    cudaStream_t st = (cudaStream_t)stream;
    switch (p->model) {
        case ORP_MODEL_GBM_LOG:
            if (fabs(a.c_vol) < 0.25 && fabs(a.c_drift) < 1.0 && !getenv("ORP_B200_NO_FUSE")) rc = launch_paths_payoff<ORP_MODEL_GBM_LOG, true>(a, f, grid, st);
            else rc = launch_paths_payoff<ORP_MODEL_GBM_LOG, false>(a, f, grid, st);
            break;
        case ORP_MODEL_GBM_ARITH: rc = launch_paths_payoff<ORP_MODEL_GBM_ARITH, false>(a, f, grid, st); break;
        case ORP_MODEL_LINEAR:    rc = launch_paths_payoff<ORP_MODEL_LINEAR, false>(a, f, grid, st); break;
        default:                  rc = launch_paths_payoff<ORP_MODEL_SV, false>(a, f, grid, st); break;
    }
    if (rc) return rc;
    paths_payoff_final_kernel<<<(2 * n_opt * 32 + 127) / 128, 128, 0, st>>>(f.partials, grid, n_opt, perm, d_sums, d_otm);
    ORP_LAUNCH_CHECK("paths_payoff_final_kernel");
    return ORP_OK;
}
