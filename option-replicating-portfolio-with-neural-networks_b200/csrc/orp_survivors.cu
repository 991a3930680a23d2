// Binomial survivors N_t on the device: the reference draws
//   np.random.seed(1234 + t); N[:, t] = np.random.binomial(N[:, t-1], exp(-lambda[:, t] dt))      Replicating_Portfolio.py:78-84
// with NumPy's legacy generator, a sequential variable-rejection stream that cannot be reproduced in parallel.
// This kernel keeps the MODEL (binomial thinning of every path's population by its own mortality intensity) and
// replaces the stream by a counter-based one: Philox4x32-10 keyed by the seed and indexed by (path, step), so the
// draw of a path does not depend on how many GPUs or threads there are.  Parity with the reference is therefore
// DISTRIBUTIONAL (tests compare moments of N_T with the host NumPy draws), not bit-wise.
// One thread per path walks the time axis; lambda is read time-major (coalesced), N_t / N is written at the kept
// columns as fp32 (the feature the fit reads).  HBM-bound: 8 B read per path-step, 4 B written per kept column.
#include "orp_common.h"

namespace orp {

__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
    const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
    c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
}

// Philox4x32-10 (Salmon et al. 2011): counter (path, step, sub) -> one uniform in (0, 1) with 53 random bits
__device__ __forceinline__ double philox_uniform(uint64_t path, uint32_t step, uint64_t seed, uint32_t sub = 0u) {
    uint32_t c[4] = {(uint32_t)path, (uint32_t)(path >> 32), step, sub};
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        philox_round(c, k0, k1);
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    const uint64_t bits = ((uint64_t)c[0] << 21) | ((uint64_t)c[1] >> 11);            // 53 bits
    return ((double)bits + 0.5) * (1.0 / 9007199254740992.0);
}

// log(k!) - [ (k + 1/2) log(k + 1) - (k + 1) + log(2 pi) / 2 ]: the tail of Stirling's series (Hoermann 1993, table + series)
__device__ __forceinline__ double stirling_tail(double k) {
    const double t[10] = {0.0810614667953272, 0.0413406959554092, 0.0276779256849983, 0.02079067210376509,
                          0.0166446911898211, 0.0138761288230707, 0.0118967099458917, 0.0104112652619720,
                          0.00925546218271273, 0.00833056343336287};
    if (k <= 9.0) return t[(int)k];
    const double kp = k + 1.0, k2 = kp * kp;
    return (1.0 / 12.0 - (1.0 / 360.0 - 1.0 / 1260.0 / k2) / k2) / kp;
}

// D ~ Binomial(n, q).  Which sampler runs is decided by the mean UP FRONT (np.random.binomial does the same, and
// handles any n): inversion from 0 while n min(q, 1-q) <= 30 — then P(D = 0) >= e^-31 cannot underflow and the
// walk ends after a few dozen terms — and Hoermann's transformed rejection (BTRS, exact) beyond, with fresh
// uniforms from the same counter-based stream (sub-counter = attempt).  q > 1/2 is drawn through its complement.
__device__ __forceinline__ double binomial_deaths(double n, double q, uint64_t path, uint32_t step, uint64_t seed) {
    if (!(q > 0.0) || n <= 0.0) return 0.0;
    if (q >= 1.0) return n;
    const bool flip = q > 0.5;
    const double pq = flip ? 1.0 - q : q, pp = 1.0 - pq;
    double k;
    if (n * pq <= 30.0) {
        const double u = philox_uniform(path, step, seed), r = pq / pp;
        double pk = exp(n * log1p(-pq)), cdf = pk;                     // P(D = 0) = (1 - q)^n
        k = 0.0;
        for (int it = 0; it < 400 && u > cdf && k < n; ++it) {         // 400 terms: > 40 standard deviations past the mean
            pk *= r * (n - k) / (k + 1.0);
            k += 1.0;
            cdf += pk;
        }
    } else {
        const double spq = sqrt(n * pq * pp), b = 1.15 + 2.53 * spq, a = -0.0873 + 0.0248 * b + 0.01 * pq;
        const double c = n * pq + 0.5, v_r = 0.92 - 4.2 / b, r = pq / pp, alpha = (2.83 + 5.1 / b) * spq;
        const double m = floor((n + 1.0) * pq);
        k = rint(n * pq);                                              // only if 32 attempts all reject (p < 1e-20)
        for (uint32_t att = 0; att < 32u; ++att) {
            const double u = philox_uniform(path, step, seed, 2u * att) - 0.5;
            double v = philox_uniform(path, step, seed, 2u * att + 1u);
            const double us = 0.5 - fabs(u), kk = floor((2.0 * a / us + b) * u + c);
            if (us >= 0.07 && v <= v_r) { k = kk; break; }
            if (kk < 0.0 || kk > n) continue;
            v = log(v * alpha / (a / (us * us) + b));
            const double upper = (m + 0.5) * log((m + 1.0) / (r * (n - m + 1.0))) + (n + 1.0) * log((n - m + 1.0) / (n - kk + 1.0))
                               + (kk + 0.5) * log(r * (n - kk + 1.0) / (kk + 1.0))
                               + stirling_tail(m) + stirling_tail(n - m) - stirling_tail(kk) - stirling_tail(n - kk);
            if (v <= upper) { k = kk; break; }
        }
    }
    return flip ? n - k : k;
}

template <typename T>
__global__ void __launch_bounds__(256) survivors_kernel(const T* __restrict__ lam, uint64_t ld, uint64_t count, uint64_t start,
                                                       int n_steps, int keep_every, double n0, double dt, uint64_t seed,
                                                       float* __restrict__ out, uint64_t ld_out) {
    for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < count; j += (uint64_t)gridDim.x * blockDim.x) {
        double n = n0;
        out[j] = 1.f;
        int next_keep = keep_every, col = 1;
        for (int t = 1; t <= n_steps; ++t) {
            const double l = (double)lam[(size_t)t * ld + j];
            const double q = -expm1(-l * dt);                              // 1 - exp(-lambda_t dt)
            n -= binomial_deaths(n, q, start + j, (uint32_t)t, seed);
            if (t == next_keep) {
                out[(size_t)col * ld_out + j] = (float)(n / n0);
                next_keep += keep_every;
                ++col;
            }
        }
    }
}

}  // namespace orp

using namespace orp;

extern "C" int orp_survivors(const void* d_lambda, int dtype, uint64_t ld, uint64_t start, uint64_t count, int n_steps,
                             int keep_every, double n0, double dt, uint64_t seed, float* d_out, uint64_t ld_out, void* stream) {
    ORP_REQUIRE(d_lambda && d_out, "null pointer");
    ORP_REQUIRE(dtype == ORP_F64 || dtype == ORP_F32, "dtype must be ORP_F64 or ORP_F32");
    ORP_REQUIRE(n_steps >= 1 && keep_every >= 1, "n_steps and keep_every must be >= 1");
    ORP_REQUIRE(ld >= count && ld_out >= count, "leading dimensions must be >= count");
    ORP_REQUIRE(n0 >= 1.0 && n0 < 9.0e15, "n0 must be a population size");
    if (count == 0) return ORP_OK;
    uint64_t want = (count + 255) / 256;
    const uint64_t cap = (uint64_t)sm_count() * 8;
    const int grid = (int)(want < cap ? want : cap);
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == ORP_F64)
        survivors_kernel<double><<<grid, 256, 0, st>>>(reinterpret_cast<const double*>(d_lambda), ld, count, start, n_steps,
                                                       keep_every, n0, dt, seed, d_out, ld_out);
    else
        survivors_kernel<float><<<grid, 256, 0, st>>>(reinterpret_cast<const float*>(d_lambda), ld, count, start, n_steps,
                                                      keep_every, n0, dt, seed, d_out, ld_out);
    ORP_LAUNCH_CHECK("survivors_kernel");
    return ORP_OK;
}
