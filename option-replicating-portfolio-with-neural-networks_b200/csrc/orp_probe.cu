// Instruction-throughput probes (measurement support, no counterpart in the reference): the fit kernel is bound by the
// fp32 FMA pipe, for which MEASURED_PEAKS.json holds no figure, so bench.py measures the packed-FMA peak of the GPU it
// runs on with this kernel, in the same process, and quotes the fit's arithmetic rate against it.
//
// Each probe is a register-only loop of independent dependency chains (enough of them to cover the pipe latency at
// full occupancy), one block of 1024 threads per SM x 2, timed with CUDA events by the caller.
#include "orp_common.h"

namespace orp {

constexpr int kProbeThreads = 512;
constexpr int kProbeChains = 8;            // independent accumulators per thread
constexpr int kProbeInner = 64;            // instructions per chain and outer iteration

template <int KIND>
__global__ void __launch_bounds__(kProbeThreads) probe_kernel(int iters, float seed, float* __restrict__ sink) {
    const float a = seed + 1e-7f * (float)threadIdx.x, b = 1.f - seed;
    if (KIND == ORP_PROBE_FFMA2) {
        unsigned long long acc[kProbeChains], x, y;
        asm("mov.b64 %0, {%1, %2};" : "=l"(x) : "f"(a), "f"(a));
        asm("mov.b64 %0, {%1, %2};" : "=l"(y) : "f"(b), "f"(b));
#pragma unroll
        for (int c = 0; c < kProbeChains; ++c) asm("mov.b64 %0, {%1, %2};" : "=l"(acc[c]) : "f"((float)c), "f"(a));
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int k = 0; k < kProbeInner; ++k)
#pragma unroll
                for (int c = 0; c < kProbeChains; ++c) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(acc[c]) : "l"(x), "l"(y));
        unsigned long long t = 0ull;
#pragma unroll
        for (int c = 0; c < kProbeChains; ++c) t ^= acc[c];
        if (t == 0x1234567ull) sink[0] = 1.f;
    } else if (KIND == ORP_PROBE_FFMA) {
        float acc[kProbeChains];
#pragma unroll
        for (int c = 0; c < kProbeChains; ++c) acc[c] = (float)c + a;
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int k = 0; k < kProbeInner; ++k)
#pragma unroll
                for (int c = 0; c < kProbeChains; ++c) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(acc[c]) : "f"(a), "f"(b));
        float t = 0.f;
#pragma unroll
        for (int c = 0; c < kProbeChains; ++c) t += acc[c];
        if (t == 12345.678f) sink[0] = t;
    } else if (KIND == ORP_PROBE_FADD) {
        float acc[kProbeChains];
#pragma unroll
        for (int c = 0; c < kProbeChains; ++c) acc[c] = (float)c + a;
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int k = 0; k < kProbeInner; ++k)
#pragma unroll
                for (int c = 0; c < kProbeChains; ++c) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(acc[c]) : "f"(b));
        float t = 0.f;
#pragma unroll
        for (int c = 0; c < kProbeChains; ++c) t += acc[c];
        if (t == 12345.678f) sink[0] = t;
    } else if (KIND == ORP_PROBE_LOP3) {
        unsigned acc[kProbeChains];
#pragma unroll
        for (int c = 0; c < kProbeChains; ++c) acc[c] = (unsigned)c + __float_as_uint(a);
        const unsigned m = __float_as_uint(b);
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int k = 0; k < kProbeInner; ++k)
#pragma unroll
                for (int c = 0; c < kProbeChains; ++c) asm volatile("lop3.b32 %0, %0, %1, 0x9e3779b9, 0x96;" : "+r"(acc[c]) : "r"(m));
        unsigned t = 0u;
#pragma unroll
        for (int c = 0; c < kProbeChains; ++c) t ^= acc[c];
        if (t == 0x1234567u) sink[0] = 1.f;
    } else {                                   // ORP_PROBE_MMA_TF32: mma.sync.m16n8k8, kProbeChains independent accumulators
        float acc[kProbeChains][4];
#pragma unroll
        for (int c = 0; c < kProbeChains; ++c)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[c][q] = (float)(c + q);
        const unsigned ua = __float_as_uint(a), ub = __float_as_uint(b);
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int k = 0; k < kProbeInner / 8; ++k)
#pragma unroll
                for (int c = 0; c < kProbeChains; ++c)
                    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0, %1, %2, %3}, {%4, %4, %5, %5}, {%5, %4}, {%0, %1, %2, %3};"
                                 : "+f"(acc[c][0]), "+f"(acc[c][1]), "+f"(acc[c][2]), "+f"(acc[c][3]) : "r"(ua), "r"(ub));
        float t = 0.f;
#pragma unroll
        for (int c = 0; c < kProbeChains; ++c) t += acc[c][0] + acc[c][3];
        if (t == 12345.678f) sink[0] = t;
    }
}

}  // namespace orp

using namespace orp;

extern "C" int orp_probe_throughput(int kind, int iters, float* d_sink, double* warp_instructions_out, void* stream) {
    ORP_REQUIRE(kind >= ORP_PROBE_FFMA2 && kind <= ORP_PROBE_MMA_TF32, "unknown probe %d", kind);
    ORP_REQUIRE(iters >= 1 && d_sink && warp_instructions_out, "bad argument");
    const int grid = 2 * sm_count();
    cudaStream_t st = (cudaStream_t)stream;
    switch (kind) {
        case ORP_PROBE_FFMA2: probe_kernel<ORP_PROBE_FFMA2><<<grid, kProbeThreads, 0, st>>>(iters, 0.25f, d_sink); break;
        case ORP_PROBE_FFMA: probe_kernel<ORP_PROBE_FFMA><<<grid, kProbeThreads, 0, st>>>(iters, 0.25f, d_sink); break;
        case ORP_PROBE_FADD: probe_kernel<ORP_PROBE_FADD><<<grid, kProbeThreads, 0, st>>>(iters, 0.25f, d_sink); break;
        case ORP_PROBE_LOP3: probe_kernel<ORP_PROBE_LOP3><<<grid, kProbeThreads, 0, st>>>(iters, 0.25f, d_sink); break;
        default: probe_kernel<ORP_PROBE_MMA_TF32><<<grid, kProbeThreads, 0, st>>>(iters, 0.25f, d_sink); break;
    }
    ORP_LAUNCH_CHECK("probe_kernel");
    const double per_thread = (double)iters * kProbeChains * (kind == ORP_PROBE_MMA_TF32 ? kProbeInner / 8 : kProbeInner);
    *warp_instructions_out = per_thread * (double)grid * (kProbeThreads / 32);
    return ORP_OK;
}
