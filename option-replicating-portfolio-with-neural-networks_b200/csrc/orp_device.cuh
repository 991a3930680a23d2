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
#define ORP_NDTRI_QMAP_DECL static __device__ const unsigned char g_orp_ndtri_qmap[ORP_NDTRI_SLABS * 2]
#include "ndtri_table.inc"

namespace orp {

constexpr int kNdtriTableBytes = ORP_NDTRI_SLABS * 128;   // 118,784 B of shared memory
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

// z = ndtri(x * 2^-30) for an integer Sobol point x in [0, 2^30); ndtri(0) = -inf like scipy.special.ndtri
// (Replicating_Portfolio.py:57).  `tab` is the shared-space address of the table copy (128-byte aligned;
// the table is read-only once the block has synchronised after filling it, hence the plain asm loads).
// Uniform control flow: fold to the nearer end, normalise, pick one of 16 sub-intervals of the octave, then
// Horner: c7..c4 in fp32 (they carry < 1e-7 of the value), c3..c0 in fp64.  Three conflict-free LDS.128 per
// point.  x = 0 needs no branch: clz(0) = 32 indexes one octave past the last, a constant -inf row.
// Layout and accuracy: tools/gen_ndtri_table.py (4e-15 against scipy.special.ndtri on the 30-bit grid).
// With a transformed table (fuse_gbm_table) the same routine returns exp(a + b ndtri(u)).
__device__ __forceinline__ double table_eval(uint32_t x, uint32_t tab) {
    const uint32_t v = min(x, 0x40000000u - x);           // fold to the nearer end
    const uint32_t lz = (uint32_t)__clz((int)v);
    uint32_t mant;                                         // (v << lz) << 1: leading one dropped; shl clamps, so v = 0 -> 0
    asm("shl.b32 %0, %1, %2;" : "=r"(mant) : "r"(v), "r"(lz + 1u));
    const uint32_t g = x >> 27;                           // top three digits: distinct inside a quarter-warp
    uint32_t offs16;                                       // byte g of the packed per-group slab offsets (byte 0 is 0)
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(offs16) : "r"(ORP_NDTRI_OFFS_LO), "r"(ORP_NDTRI_OFFS_HI), "r"(g));
    const uint32_t a0 = tab + (lz + offs16) * 2048u + ((mant >> 21) & 0x780u) + g * 16u
                        - (2u * 16u + ORP_NDTRI_OFFS_BIAS) * 128u;     // unit k lives at position (g&4)|((g&3)^k)
    const double2 c01 = lds_f64x2(a0);
    const double2 c23 = lds_f64x2(a0 ^ 16u);
    const float4 c47 = lds_f32x4(a0 ^ 32u);
    const uint32_t whi = 0x3ff00000u | ((mant >> 8) & 0x000fffffu);       // 1 + f, f = low 28 bits of mant
    const double t = __hiloint2double((int)whi, (int)(mant << 24)) - 1.5;
    const float tf = __uint_as_float(0x3f800000u | ((mant >> 5) & 0x007fffffu)) - 1.5f;
    float hi = fmaf(c47.w, tf, c47.z);
    hi = fmaf(hi, tf, c47.y);
    hi = fmaf(hi, tf, c47.x);
    double z = fma((double)hi, t, c23.y);
    z = fma(z, t, c23.x);
    z = fma(z, t, c01.y);
    return fma(z, t, c01.x);
}

__device__ __forceinline__ double ndtri_bits(uint32_t x, uint32_t tab) { return table_eval(x, tab); }

// In-place transform of the shared-memory table: every row z(t) = sum c_k t^k becomes the degree-7 Taylor
// polynomial of y(t) = exp(a + b z(t)) about the sub-interval centre (recurrence n y_n = sum_k k p_k y_{n-k}
// with p = b z).  The neglected terms are O((b c_1 / 2)^8): below 1e-17 for every step size the kernels see.
// After this, table_eval() returns the log-Euler GBM growth factor directly: S_t = S_{t-1} * y(u_t) — the
// inverse normal, the drift/volatility FMA and the exponential of EO cell 6 collapse into one Horner chain.
__device__ __forceinline__ void fuse_gbm_table(char* __restrict__ tab, double a, double b, int tid, int nthreads) {
    for (int hs = tid; hs < ORP_NDTRI_SLABS * 2; hs += nthreads) {
        const int slab = hs >> 1, half = hs & 1;
        const int q = g_orp_ndtri_qmap[hs];
        char* base = tab + slab * 128 + half * 64;
        double2* u0 = reinterpret_cast<double2*>(base + ((q ^ 0) << 4));
        double2* u1 = reinterpret_cast<double2*>(base + ((q ^ 1) << 4));
        float4* u2 = reinterpret_cast<float4*>(base + ((q ^ 2) << 4));
        const double2 c01 = *u0, c23 = *u1;
        const float4 c47 = *u2;
        double p[8] = {0.0, b * c01.y, b * c23.x, b * c23.y, b * (double)c47.x, b * (double)c47.y, b * (double)c47.z, b * (double)c47.w};
        double y[8];
        y[0] = exp(fma(b, c01.x, a));
#pragma unroll
        for (int n = 1; n < 8; ++n) {
            double acc = 0.0;
#pragma unroll
            for (int k = 1; k <= n; ++k) acc = fma((double)k * p[k], y[n - k], acc);
            y[n] = acc / (double)n;
        }
        *u0 = make_double2(y[0], y[1]);
        *u1 = make_double2(y[2], y[3]);
        *u2 = make_float4((float)y[4], (float)y[5], (float)y[6], (float)y[7]);
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
