// Host-side Sobol set-up: Joe-Kuo direction numbers, LMS + digital-shift scramble, device packing.
// Integer / GF(2) arithmetic only; bit-exact with scipy.stats.qmc.Sobol(d, scramble=True) given the
// same random bits (scipy/stats/_qmc.py:1793-1828).  O(d * bits^2) work — stays on the host.
#include "orp_common.h"

#include <vector>

namespace orp {

char* last_error_buffer() {
    static thread_local char buf[512] = {0};
    return buf;
}

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

}  // namespace orp

extern "C" int orp_version(void) { return 100; }

extern "C" const char* orp_last_error(void) { return orp::last_error_buffer(); }

extern "C" int orp_sobol_direction_numbers(const int64_t* poly, const int64_t* vinit, int d, int bits,
                                           uint32_t* v_out) {
    ORP_REQUIRE(poly && vinit && v_out, "null pointer");
    ORP_REQUIRE(d >= 1 && d <= ORP_SOBOL_MAXDIM, "Maximum supported dimensionality is %d.", ORP_SOBOL_MAXDIM);
    ORP_REQUIRE(bits >= 1 && bits <= 32, "bits must be in [1, 32]");
    std::vector<uint64_t> v(bits);
    for (int k = 0; k < d; ++k) {
        if (k == 0) {
            for (int j = 0; j < bits; ++j) v[j] = 1;
        } else {
            const uint64_t p = (uint64_t)poly[k];
            int deg = 63 - __builtin_clzll(p);          // degree of the primitive polynomial
            for (int j = 0; j < deg && j < bits; ++j) v[j] = (uint64_t)vinit[(size_t)k * 18 + j];
            for (int j = deg; j < bits; ++j) {
                uint64_t nv = v[j - deg];
                uint64_t pow2 = 1;
                for (int q = 0; q < deg; ++q) {
                    pow2 <<= 1;
                    if ((p >> (deg - 1 - q)) & 1) nv ^= pow2 * v[j - q - 1];
                }
                v[j] = nv;
            }
        }
        for (int j = 0; j < bits; ++j) v_out[(size_t)k * bits + j] = (uint32_t)(v[j] << (bits - 1 - j));
    }
    return ORP_OK;
}

extern "C" int orp_sobol_scramble(int d, int bits, const uint32_t* v, const uint32_t* shift_bits,
                                  const uint32_t* ltm_bits, uint32_t* sv_out, uint32_t* shift_out) {
    ORP_REQUIRE(v && shift_bits && ltm_bits && sv_out && shift_out, "null pointer");
    ORP_REQUIRE(d >= 1 && d <= ORP_SOBOL_MAXDIM, "Maximum supported dimensionality is %d.", ORP_SOBOL_MAXDIM);
    ORP_REQUIRE(bits >= 1 && bits <= 32, "bits must be in [1, 32]");
    std::vector<uint32_t> rows(bits);
    for (int k = 0; k < d; ++k) {
        uint32_t sh = 0;
        for (int j = 0; j < bits; ++j) sh |= (shift_bits[(size_t)k * bits + j] & 1u) << j;
        shift_out[k] = sh;
        // row p of the unit-lower-triangular matrix as a word: column q sits at bit (bits-1-q)
        const uint32_t* L = ltm_bits + (size_t)k * bits * bits;
        for (int p = 0; p < bits; ++p) {
            uint32_t r = 1u << (bits - 1 - p);
            for (int q = 0; q < p; ++q) r |= (L[(size_t)p * bits + q] & 1u) << (bits - 1 - q);
            rows[p] = r;
        }
        for (int j = 0; j < bits; ++j) {
            const uint32_t col = v[(size_t)k * bits + j];
            uint32_t out = 0;
            for (int p = 0; p < bits; ++p)
                out |= (uint32_t)(__builtin_popcount(rows[p] & col) & 1) << (bits - 1 - p);
            sv_out[(size_t)k * bits + j] = out;
        }
    }
    return ORP_OK;
}

extern "C" int orp_sobol_pack(int d, int bits, const uint32_t* sv, const uint32_t* shift, uint32_t* sv32_out,
                              uint32_t* tlane_out) {
    ORP_REQUIRE(sv && shift && sv32_out && tlane_out, "null pointer");
    ORP_REQUIRE(d >= 1 && d <= ORP_SOBOL_MAXDIM, "Maximum supported dimensionality is %d.", ORP_SOBOL_MAXDIM);
    ORP_REQUIRE(bits == ORP_SOBOL_BITS, "the device path is built for bits = %d (SciPy's default)", ORP_SOBOL_BITS);
    const size_t dpad = (size_t)ORP_SV_DPAD(d);
    for (size_t i = 0; i < dpad * ORP_SV_WORDS; ++i) sv32_out[i] = 0u;
    for (int k = 0; k < d; ++k) {
        const uint32_t* s = sv + (size_t)k * bits;
        for (int b = 0; b < bits; ++b) sv32_out[(size_t)b * dpad + k] = s[b];
        sv32_out[(size_t)31 * dpad + k] = shift[k];
        // Path index i = 64 w + 32 p + l  =>  gray bits 0..3 = l_b ^ l_{b+1}, bit 4 = l_4 ^ p, bit 5 = p ^ w_0.
        for (int l = 0; l < 32; ++l) {
            uint32_t t = 0;
            for (int b = 0; b < 4; ++b)
                if (((l >> b) ^ (l >> (b + 1))) & 1) t ^= s[b];
            if ((l >> 4) & 1) t ^= s[4];
            tlane_out[((size_t)k * 32 + l) * 2 + 0] = t;
            tlane_out[((size_t)k * 32 + l) * 2 + 1] = t ^ s[4] ^ s[5];
        }
    }
    return ORP_OK;
}
