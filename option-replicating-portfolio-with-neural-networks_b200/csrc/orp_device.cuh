// Device-side building blocks shared by the path kernels: table-driven inverse normal on the 30-bit
// Sobol grid, and the warp decomposition of the Sobol skip-ahead.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

#include "../../include/orp_b200.h"

// The coefficient table is defined in the one translation unit that includes this header
// (orp_paths.cu); no relocatable device code is needed.
// raw 64-bit words: a row holds fp64 and fp32 coefficients
#define ORP_NDTRI_TABLE_DECL static __device__ const unsigned long long g_orp_ndtri_table[ORP_NDTRI_SLABS * 16]
#define ORP_NDTRI_OWNER_DECL static __device__ const unsigned char g_orp_ndtri_owner[ORP_NDTRI_SLABS]
#include "ndtri_table.inc"

namespace orp {

constexpr int kNdtriTableBytes = ORP_NDTRI_SLABS * 128;   // 61,568 B of shared memory
constexpr unsigned kFull = 0xffffffffu;

// Cooperative copy of the coefficient table into shared memory (16-byte units).
__device__ __forceinline__ void load_ndtri_table(double* s_tab, int tid, int nthreads) {
    const ulonglong2* src = reinterpret_cast<const ulonglong2*>(g_orp_ndtri_table);
    ulonglong2* dst = reinterpret_cast<ulonglong2*>(s_tab);
    for (int i = tid; i < ORP_NDTRI_SLABS * 8; i += nthreads) dst[i] = src[i];
}

__device__ __forceinline__ double2 lds_f64x2(uint32_t saddr) {
    double2 r;
    asm("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(r.x), "=d"(r.y) : "r"(saddr));
    return r;
}
__device__ __forceinline__ float4 lds_f32x4(uint32_t saddr) {
    float4 r;
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(saddr));
    return r;
}

// (a & b) | c in one LOP3 (the constants sit in registers: an instruction carries one immediate at most)
__device__ __forceinline__ uint32_t and_or(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}

// z = ndtri(x * 2^-30) for an integer Sobol point x in [0, 2^30); ndtri(0) = -inf like scipy.special.ndtri
// (Replicating_Portfolio.py:57).  `tab` is the shared-space address of the table copy (128-byte aligned;
// the table is read-only once the block has synchronised after filling it, hence the plain asm loads).
// Uniform control flow: fold to the nearer end, let ONE int->fp64 conversion normalise the argument (its
// exponent and top four mantissa bits ARE the row index), subtract the row centre (exact), then Horner:
// c5..c2 in fp32 (they carry < 1e-6 of the value), c1, c0 in fp64.  Two conflict-free LDS.128 per point.
// x = 0 needs no branch: (double)0 has a zero high word, which clamps to slab 0, a constant -inf row.
// Layout and accuracy: tools/gen_ndtri_table.py (9e-12 against mpmath on the 30-bit grid).
// With a transformed table (fuse_gbm_table) the same routine returns exp(a + b ndtri(u)).
constexpr uint32_t kNdtriBase = 1023u * 16u - 1u;          // (hi32(1.0) >> 16) - 1: v = 1 lands in slab 1
__device__ __forceinline__ double table_eval(uint32_t x, uint32_t tab) {
    const uint32_t v = min(x, 0x40000000u - x);           // fold to the nearer end
    const double d = __uint2double_rn(v);                  // exact: v < 2^30
    const uint32_t hi = (uint32_t)__double2hiint(d);
    const uint32_t slab = max(hi >> 16, kNdtriBase);       // (exponent, top four mantissa bits); v = 0 -> kNdtriBase
    const uint32_t g = x >> 27;                            // top three digits: distinct inside a quarter-warp
    const uint32_t a0 = tab - kNdtriBase * 128u + slab * 128u + g * 16u;     // (c0, c1) in unit g, (c2..c5) in unit g ^ 2
    const double2 c01 = lds_f64x2(a0);
    const float4 c25 = lds_f32x4(a0 ^ 32u);
    // mantissa m in [1, 2) and the centre of its sixteenth, 1 + (j + 1/2) / 16: same exponent, so m - centre is exact
    const double m = __hiloint2double((int)and_or(hi, 0x000fffffu, 0x3ff00000u), __double2loint(d));
    const double c = __hiloint2double((int)and_or(hi, 0x000f0000u, 0x3ff08000u), 0);
    const double s = m - c;
    const float sf = (float)s;
    float q = fmaf(c25.w, sf, c25.z);
    q = fmaf(q, sf, c25.y);
    q = fmaf(q, sf, c25.x);
    const double z = fma((double)q, s, c01.y);
    return fma(z, s, c01.x);
}

__device__ __forceinline__ double ndtri_bits(uint32_t x, uint32_t tab) { return table_eval(x, tab); }

// In-place transform of the shared-memory table: every row z(s) = sum c_k s^k becomes the degree-5 Taylor
// polynomial of y(s) = exp(a + b z(s)) about the sub-interval centre (recurrence n y_n = sum_k k p_k y_{n-k}
// with p = b z).  The neglected terms are O((b c_1 / 32)^6 / 720): below 1e-20 for every step size the kernels see.
// After this, table_eval() returns the log-Euler GBM growth factor directly: S_t = S_{t-1} * y(u_t) — the
// inverse normal, the drift/volatility FMA and the exponential of EO cell 6 collapse into one Horner chain.
__device__ __forceinline__ void fuse_gbm_table(char* __restrict__ tab, double a, double b, int tid, int nthreads) {
    for (int r = tid; r < ORP_NDTRI_SLABS * 8; r += nthreads) {
        const int slab = r >> 3, g = r & 7;
        if (!((g_orp_ndtri_owner[slab] >> g) & 1)) continue;           // no row of group g in this slab
        double2* u0 = reinterpret_cast<double2*>(tab + slab * 128 + g * 16);
        float4* u1 = reinterpret_cast<float4*>(tab + slab * 128 + (g ^ 2) * 16);
        const double2 c01 = *u0;
        const float4 c25 = *u1;
        const double p[6] = {0.0, b * c01.y, b * (double)c25.x, b * (double)c25.y, b * (double)c25.z, b * (double)c25.w};
        double y[6];
        y[0] = exp(fma(b, c01.x, a));
#pragma unroll
        for (int n = 1; n < 6; ++n) {
            double acc = 0.0;
#pragma unroll
            for (int k = 1; k <= n; ++k) acc = fma((double)k * p[k], y[n - k], acc);
            y[n] = acc / (double)n;
        }
        *u0 = make_double2(y[0], y[1]);
        *u1 = make_float4((float)y[2], (float)y[3], (float)y[4], (float)y[5]);
    }
}

// ---- Sobol skip-ahead, warp decomposition --------------------------------------------------------
// Path index i = 64 w + 32 p + l (w: warp tile, p in {0,1}, l: lane).  With gray(i) = i ^ (i >> 1):
//   x_i[t] = base(w, t) ^ tlane[t][l][p]
//   base(w, t) = shift[t] ^ (w_0 ? sv[t][5] : 0) ^ XOR_{b in gray(w)} sv[t][6 + b]          (warp-uniform)
//   tlane[t][l][p] = XOR_{b<4} (l_b ^ l_{b+1}) sv[t][b] ^ (l_4 ? sv[t][4] : 0) ^ (p ? sv[t][4] ^ sv[t][5] : 0)
// Each lane evaluates base(w, .) for ONE dimension of a 32-dimension chunk; the step loop broadcasts it
// with a shuffle, so the per-point cost is one SHFL + one 8-byte table load + two XORs for two paths.

// sv32 is word-major (word b of dimension k at sv32[b * dpad + k]): the 32 lanes of a warp evaluate 32
// consecutive dimensions, so every load below is one coalesced 128-byte line.
__device__ __forceinline__ uint32_t warp_base(const uint32_t* __restrict__ sv32, int dpad, int dim, uint64_t w) {
    const uint32_t* col = sv32 + dim;
    uint32_t b = __ldg(col + (size_t)31 * dpad);
    if (w & 1ull) b ^= __ldg(col + (size_t)5 * dpad);
    uint32_t g = (uint32_t)(w ^ (w >> 1));                // w < 2^24 because i < 2^30
    while (g) {
        const int bit = __ffs((int)g) - 1;
        b ^= __ldg(col + (size_t)(6 + bit) * dpad);
        g &= g - 1;
    }
    return b;
}

}  // namespace orp
