// Device-side building blocks shared by the path kernels: table-driven inverse normal on the 30-bit
// Sobol grid, and the warp decomposition of the Sobol skip-ahead.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

#include "../../include/orp_b200.h"

// The coefficient table is defined in the one translation unit that includes this header
// (orp_paths.cu); no relocatable device code is needed.
#define ORP_NDTRI_TABLE_DECL static __device__ const double g_orp_ndtri_table[ORP_NDTRI_SLABS * 16]
#include "ndtri_table.inc"

namespace orp {

constexpr int kNdtriTableBytes = ORP_NDTRI_SLABS * 128;   // 118,784 B of shared memory
constexpr unsigned kFull = 0xffffffffu;

// Cooperative copy of the coefficient table into shared memory (16-byte units).
__device__ __forceinline__ void load_ndtri_table(double* s_tab, int tid, int nthreads) {
    const double2* src = reinterpret_cast<const double2*>(g_orp_ndtri_table);
    double2* dst = reinterpret_cast<double2*>(s_tab);
    for (int i = tid; i < ORP_NDTRI_SLABS * 8; i += nthreads) dst[i] = src[i];
}

// z = ndtri(x * 2^-30) for an integer Sobol point x in [0, 2^30); ndtri(0) = -inf like scipy.special.ndtri
// (Replicating_Portfolio.py:57).  `tab` points at the shared-memory copy of the table (128-byte aligned).
// Uniform control flow: fold to the nearer end, normalise, pick one of 16 sub-intervals of the octave,
// Horner degree 7.  Layout and accuracy: tools/gen_ndtri_table.py (max error 3e-16 vs mpmath).
__device__ __forceinline__ double ndtri_bits(uint32_t x, const char* __restrict__ tab) {
    const uint32_t side = x >> 29;
    const uint32_t v = side ? (0x40000000u - x) : x;
    const int lz = __clz((int)(v | 1u));
    const uint32_t mant = (v << lz) << 1;                 // leading one dropped
    const uint32_t g = x >> 27;                           // top three digits: distinct inside a quarter-warp
    const uint32_t offs16 = __byte_perm(ORP_NDTRI_OFFS_LO, ORP_NDTRI_OFFS_HI, g) & 0xffu;
    const uint32_t slab = (uint32_t)(lz - 2) * 16u + (mant >> 28) + offs16 * 16u - ORP_NDTRI_OFFS_BIAS;
    const uint32_t a0 = slab * 128u + g * 16u;             // pair k lives at unit (g&4)|((g&3)^k)
    const double2 c01 = *reinterpret_cast<const double2*>(tab + (a0));
    const double2 c23 = *reinterpret_cast<const double2*>(tab + (a0 ^ 16u));
    const double2 c45 = *reinterpret_cast<const double2*>(tab + (a0 ^ 32u));
    const double2 c67 = *reinterpret_cast<const double2*>(tab + (a0 ^ 48u));
    const uint32_t rest = mant << 4;
    const double t = __hiloint2double((int)(0x3ff00000u | (rest >> 12)), (int)(rest << 20)) - 1.5;
    double z = fma(c67.y, t, c67.x);
    z = fma(z, t, c45.y);
    z = fma(z, t, c45.x);
    z = fma(z, t, c23.y);
    z = fma(z, t, c23.x);
    z = fma(z, t, c01.y);
    z = fma(z, t, c01.x);
    return x == 0u ? __longlong_as_double(0xfff0000000000000ll) : z;
}

// ---- Sobol skip-ahead, warp decomposition --------------------------------------------------------
// Path index i = 64 w + 32 p + l (w: warp tile, p in {0,1}, l: lane).  With gray(i) = i ^ (i >> 1):
//   x_i[t] = base(w, t) ^ tlane[t][l][p]
//   base(w, t) = shift[t] ^ (w_0 ? sv[t][5] : 0) ^ XOR_{b in gray(w)} sv[t][6 + b]          (warp-uniform)
//   tlane[t][l][p] = XOR_{b<4} (l_b ^ l_{b+1}) sv[t][b] ^ (l_4 ? sv[t][4] : 0) ^ (p ? sv[t][4] ^ sv[t][5] : 0)
// Each lane evaluates base(w, .) for ONE dimension of a 32-dimension chunk; the step loop broadcasts it
// with a shuffle, so the per-point cost is one SHFL + one 8-byte table load + two XORs for two paths.

__device__ __forceinline__ uint32_t warp_base(const uint32_t* __restrict__ sv32, int dim, uint64_t w) {
    const uint32_t* row = sv32 + (size_t)dim * ORP_SV_STRIDE;
    uint32_t b = __ldg(row + 31);
    if (w & 1ull) b ^= __ldg(row + 5);
    uint32_t g = (uint32_t)(w ^ (w >> 1));                // w < 2^24 because i < 2^30
    while (g) {
        const int bit = __ffs((int)g) - 1;
        b ^= __ldg(row + 6 + bit);
        g &= g - 1;
    }
    return b;
}

}  // namespace orp
