// Replicating-portfolio fit: the 8-wide MLP of
// Replicating_Portfolio.py:149-161 / European Options.ipynb cell 12, trained per time step.
//
// Hidden width 8 is far too narrow for a tensor-core contraction, so this is a fused per-path SIMT
// kernel: thread = sample, weights broadcast from shared memory, forward + hand-derived backward in
// registers, per-thread gradient accumulators, then warp-shuffle -> block -> fixed-order fp64 reduction.
// Bound: fp32 FMA issue (~315 FMA-class ops per sample against 12-24 B of HBM reads).
//
// Device-resident control block (`state`, doubles) so that the host can enqueue epochs back to back:
#include "orp_common.h"

#include <cstring>

namespace orp {

// ---- state layout (indices into the double array) ------------------------------------------------
enum {
    ST_LR = 0, ST_ADAM_T, ST_EPOCH, ST_EPOCHS, ST_PATIENCE, ST_USE_SCHED, ST_WAIT, ST_BEST, ST_STOP,
    ST_STOPPED_EARLY, ST_LAST_LOSS, ST_EPOCHS_RUN, ST_HAVE_BEST, ST_TOTAL_EPOCHS, ST_LOSS_ACC, ST_COUNT_ACC,
    ST_ADAM_M = 16, ST_ADAM_V = ST_ADAM_M + ORP_FIT_MAX_PARAMS, ST_BEST_W = ST_ADAM_V + ORP_FIT_MAX_PARAMS,
    ST_TICKET = ST_BEST_W + ORP_FIT_MAX_PARAMS,      // block-arrival counter of the single-launch step (used as uint32)
    ST_SPARE,                                        // (unused)
    ST_ABORT,                                        // persistent kernel: non-zero when a hand-over poll timed out (uint32)
    ST_END
};
static_assert(ST_END <= ORP_FIT_STATE_DOUBLES, "state block too small");

constexpr float kLeaky = 0.3f;          // Keras LeakyReLU default alpha
constexpr int kH = 8;

template <int NET> struct Net {
    static constexpr int IN = NET == ORP_NET_EO ? 1 : 3;
    static constexpr int OUT = NET == ORP_NET_EO ? 1 : 2;
    static constexpr int W1 = 0, B1 = W1 + IN * kH, W2 = B1 + kH, B2 = W2 + kH * kH, W3 = B2 + kH,
                         B3 = W3 + kH * OUT, P = B3 + OUT;
};
static_assert(Net<ORP_NET_EO>::P == 97 && Net<ORP_NET_RP>::P == 122, "parameter counts of the reference nets");

__device__ __forceinline__ float lrelu(float x) { return fmaxf(x, kLeaky * x); }
__device__ __forceinline__ float dlrelu(float pre) { return pre > 0.f ? 1.f : kLeaky; }

// ---- packed fp32 pairs ----------------------------------------------------------------------------
// Blackwell issues a 3-register FFMA every other cycle per scheduler; the packed FFMA2 (PTX fma.rn.f32x2)
// does two fp32 FMAs in the same slot.  One sample per thread, so the pairs run along the OUTPUT index of
// each layer: (j, j+1) share the scalar input, which FFMA2 takes as a broadcast operand.  Per-lane results
// are the IEEE fp32 FMA results of the scalar code.
typedef float2 f2;
__device__ __forceinline__ unsigned long long pk(f2 v) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(v.x), "f"(v.y));
    return r;
}
__device__ __forceinline__ f2 upk(unsigned long long r) {
    f2 v;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(v.x), "=f"(v.y) : "l"(r));
    return v;
}
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) {
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(pk(a)), "l"(pk(b)), "l"(pk(c)));
    return upk(d);
}
__device__ __forceinline__ f2 fma2s(float a, f2 b, f2 c) { return fma2(make_float2(a, a), b, c); }
__device__ __forceinline__ f2 mul2(f2 a, f2 b) {
    unsigned long long d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(pk(a)), "l"(pk(b)));
    return upk(d);
}
__device__ __forceinline__ f2 add2(f2 a, f2 b) {
    unsigned long long d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(pk(a)), "l"(pk(b)));
    return upk(d);
}
// Weight pairs k, k+1 (k even) of the shared-memory image, as one LDS.128.  `volatile` keeps the compiler from
// hoisting the ~160 loop-invariant weights into registers (which spills the gradient accumulators): they are
// re-read every sample as broadcast loads, which is what shared memory is good at.
struct f2x2 { f2 a, b; };
__device__ __forceinline__ f2x2 ldw(uint32_t s_w, int pair) {
    f2x2 r;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(r.a.x), "=f"(r.a.y), "=f"(r.b.x), "=f"(r.b.y) : "r"(s_w + 8u * (uint32_t)pair));
    return r;
}
__device__ __forceinline__ float ldw1(uint32_t s_w, int k) {
    float r;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(r) : "r"(s_w + 4u * (uint32_t)k));
    return r;
}
__device__ __forceinline__ f2 lrelu2(f2 v) {
    const f2 s = mul2(make_float2(kLeaky, kLeaky), v);
    return make_float2(fmaxf(v.x, s.x), fmaxf(v.y, s.y));
}
__device__ __forceinline__ f2 dlrelu2(f2 pre) { return make_float2(dlrelu(pre.x), dlrelu(pre.y)); }
__device__ __forceinline__ float elem(const f2 (&v)[kH / 2], int i) { return (i & 1) ? v[i >> 1].y : v[i >> 1].x; }

// Shared-memory image of the weights: the parameter vector (every block of it starts at an even offset, so
// it is also an array of pairs) followed by W2 transposed, which the backward pass reads pairwise over the
// INPUT index.
constexpr int kW2T = ORP_FIT_MAX_PARAMS;                 // float offset of W2^T[j][i]
constexpr int kSmemW = ORP_FIT_MAX_PARAMS + kH * kH;

template <int NET>
__device__ __forceinline__ void load_weights(float* __restrict__ s_w, const float* __restrict__ params, int tid, int nthreads) {
    using N = Net<NET>;
    for (int k = tid; k < N::P; k += nthreads) s_w[k] = params[k];
    for (int k = tid; k < kH * kH; k += nthreads) s_w[kW2T + (k & 7) * kH + (k >> 3)] = params[N::W2 + k];   // k = i*8+j -> [j][i]
}

// S samples are carried through the network TOGETHER by one thread: every weight pair is read from shared
// memory once and used S times (halving the LDS traffic, which otherwise runs at 60 % of the LSU pipe), and the
// S independent dependency chains fill the FFMA2 latency that 8 warps per SM cannot hide by switching.
// Only the post-activation values are kept: LeakyReLU preserves the sign, so the slope is read off h.
template <int NET, int S> struct Acts {
    f2 h1[S][kH / 2], h2[S][kH / 2];
};

// Forward pass: V[s] = phi[s] p0[s] + psi[s] p1.  `w` is the shared-space address of the weight image.
template <int NET, int S>
__device__ __forceinline__ void net_forward(uint32_t w, const float (&x)[S][Net<NET>::IN], const float (&p0)[S], float p1,
                                            Acts<NET, S>& a, float (&phi)[S], float (&psi)[S], float (&v)[S]) {
    using N = Net<NET>;
    static_assert(N::W1 % 4 == 0 && N::B1 % 4 == 0 && N::W2 % 4 == 0 && N::B2 % 4 == 0 && N::W3 % 4 == 0 && N::B3 % 2 == 0,
                  "parameter blocks must be 16-byte aligned for the paired loads");
    f2 pre[S][kH / 2];
#pragma unroll
    for (int q = 0; q < kH / 4; ++q) {
        const f2x2 b = ldw(w, N::B1 / 2 + 2 * q);
#pragma unroll
        for (int s = 0; s < S; ++s) { pre[s][2 * q] = b.a; pre[s][2 * q + 1] = b.b; }
    }
#pragma unroll
    for (int i = 0; i < N::IN; ++i)
#pragma unroll
        for (int q = 0; q < kH / 4; ++q) {
            const f2x2 c = ldw(w, (N::W1 + i * kH) / 2 + 2 * q);
#pragma unroll
            for (int s = 0; s < S; ++s) {
                pre[s][2 * q] = fma2s(x[s][i], c.a, pre[s][2 * q]);
                pre[s][2 * q + 1] = fma2s(x[s][i], c.b, pre[s][2 * q + 1]);
            }
        }
#pragma unroll
    for (int s = 0; s < S; ++s)
#pragma unroll
        for (int jp = 0; jp < kH / 2; ++jp) a.h1[s][jp] = lrelu2(pre[s][jp]);
#pragma unroll
    for (int q = 0; q < kH / 4; ++q) {
        const f2x2 b = ldw(w, N::B2 / 2 + 2 * q);
#pragma unroll
        for (int s = 0; s < S; ++s) { pre[s][2 * q] = b.a; pre[s][2 * q + 1] = b.b; }
    }
#pragma unroll
    for (int i = 0; i < kH; ++i)
#pragma unroll
        for (int q = 0; q < kH / 4; ++q) {
            const f2x2 c = ldw(w, (N::W2 + i * kH) / 2 + 2 * q);
#pragma unroll
            for (int s = 0; s < S; ++s) {
                const float hi = elem(a.h1[s], i);
                pre[s][2 * q] = fma2s(hi, c.a, pre[s][2 * q]);
                pre[s][2 * q + 1] = fma2s(hi, c.b, pre[s][2 * q + 1]);
            }
        }
#pragma unroll
    for (int s = 0; s < S; ++s)
#pragma unroll
        for (int jp = 0; jp < kH / 2; ++jp) a.h2[s][jp] = lrelu2(pre[s][jp]);
    if (NET == ORP_NET_EO) {
        const f2x2 c0 = ldw(w, N::W3 / 2), c1 = ldw(w, N::W3 / 2 + 2);        // phi = b3 + sum_i h2[i] w3[i], pairs over i
        const float b3 = ldw1(w, N::B3);
#pragma unroll
        for (int s = 0; s < S; ++s) {
            f2 t = mul2(a.h2[s][0], c0.a);
            t = fma2(a.h2[s][1], c0.b, t);
            t = fma2(a.h2[s][2], c1.a, t);
            t = fma2(a.h2[s][3], c1.b, t);
            phi[s] = b3 + (t.x + t.y);
            psi[s] = 1.f - phi[s];
        }
    } else {
        f2 o[S];
        const f2 b3 = make_float2(ldw1(w, N::B3), ldw1(w, N::B3 + 1));        // (phi, psi) = b3 + sum_i h2[i] W3[i][0:2]
#pragma unroll
        for (int s = 0; s < S; ++s) o[s] = b3;
#pragma unroll
        for (int q = 0; q < kH / 2; ++q) {
            const f2x2 c = ldw(w, N::W3 / 2 + 2 * q);
#pragma unroll
            for (int s = 0; s < S; ++s) {
                o[s] = fma2s(elem(a.h2[s], 2 * q), c.a, o[s]);
                o[s] = fma2s(elem(a.h2[s], 2 * q + 1), c.b, o[s]);
            }
        }
#pragma unroll
        for (int s = 0; s < S; ++s) { phi[s] = o[s].x; psi[s] = o[s].y; }
    }
#pragma unroll
    for (int s = 0; s < S; ++s) v[s] = fmaf(phi[s], p0[s], psi[s] * p1);
}

// Gradient accumulators of one thread, as pairs in the order of the parameter vector.
template <int NET> struct Grads {
    static_assert(Net<NET>::P % 2 == (NET == ORP_NET_EO ? 1 : 0), "pair layout");
    f2 v[Net<NET>::P / 2];
    float last;                                      // EO: the single output bias (P is odd)
    __device__ __forceinline__ void zero() {
#pragma unroll
        for (int k = 0; k < Net<NET>::P / 2; ++k) v[k] = make_float2(0.f, 0.f);
        last = 0.f;
    }
    __device__ __forceinline__ float get(int k) const {      // k is a compile-time constant after unrolling
        return k == 2 * (Net<NET>::P / 2) ? last : ((k & 1) ? v[k >> 1].y : v[k >> 1].x);
    }
};

// Backward pass of V = phi p0 + psi p1 for S samples, accumulating into g.  dV[s] = 0 switches a sample off.
template <int NET, int S>
__device__ __forceinline__ void net_backward(uint32_t w, const float (&x)[S][Net<NET>::IN], const float (&p0)[S], float p1,
                                             const float (&dV)[S], const Acts<NET, S>& a, Grads<NET>& g) {
    using N = Net<NET>;
    f2 d2[S][kH / 2];
    if (NET == ORP_NET_EO) {
        f2 dd[S];
#pragma unroll
        for (int s = 0; s < S; ++s) {
            const float dphi = dV[s] * (p0[s] - p1);                          // psi = 1 - phi
            dd[s] = make_float2(dphi, dphi);
            g.last += dphi;
        }
#pragma unroll
        for (int q = 0; q < kH / 4; ++q) {
            const f2x2 c = ldw(w, N::W3 / 2 + 2 * q);
#pragma unroll
            for (int s = 0; s < S; ++s) {
                g.v[N::W3 / 2 + 2 * q] = fma2(dd[s], a.h2[s][2 * q], g.v[N::W3 / 2 + 2 * q]);
                g.v[N::W3 / 2 + 2 * q + 1] = fma2(dd[s], a.h2[s][2 * q + 1], g.v[N::W3 / 2 + 2 * q + 1]);
                d2[s][2 * q] = mul2(mul2(dd[s], c.a), dlrelu2(a.h2[s][2 * q]));
                d2[s][2 * q + 1] = mul2(mul2(dd[s], c.b), dlrelu2(a.h2[s][2 * q + 1]));
            }
        }
    } else {
        f2 dout[S];
        float d2s[S][kH];
#pragma unroll
        for (int s = 0; s < S; ++s) {
            dout[s] = make_float2(dV[s] * p0[s], dV[s] * p1);
            g.v[N::B3 / 2] = add2(g.v[N::B3 / 2], dout[s]);
        }
#pragma unroll
        for (int q = 0; q < kH / 2; ++q) {
            const f2x2 c = ldw(w, N::W3 / 2 + 2 * q);
#pragma unroll
            for (int s = 0; s < S; ++s) {
                g.v[N::W3 / 2 + 2 * q] = fma2s(elem(a.h2[s], 2 * q), dout[s], g.v[N::W3 / 2 + 2 * q]);
                g.v[N::W3 / 2 + 2 * q + 1] = fma2s(elem(a.h2[s], 2 * q + 1), dout[s], g.v[N::W3 / 2 + 2 * q + 1]);
                const f2 t0 = mul2(dout[s], c.a), t1 = mul2(dout[s], c.b);
                d2s[s][2 * q] = t0.x + t0.y;
                d2s[s][2 * q + 1] = t1.x + t1.y;
            }
        }
#pragma unroll
        for (int s = 0; s < S; ++s)
#pragma unroll
            for (int ip = 0; ip < kH / 2; ++ip)
                d2[s][ip] = mul2(make_float2(d2s[s][2 * ip], d2s[s][2 * ip + 1]), dlrelu2(a.h2[s][ip]));
    }
    f2 d1[S][kH / 2];
#pragma unroll
    for (int s = 0; s < S; ++s)
#pragma unroll
        for (int ip = 0; ip < kH / 2; ++ip) d1[s][ip] = make_float2(0.f, 0.f);
#pragma unroll
    for (int j = 0; j < kH; ++j)                                              // d1[i] = sum_j d2[j] W2[i][j], pairs over i
#pragma unroll
        for (int q = 0; q < kH / 4; ++q) {
            const f2x2 c = ldw(w, (kW2T + j * kH) / 2 + 2 * q);
#pragma unroll
            for (int s = 0; s < S; ++s) {
                const float dj = elem(d2[s], j);
                d1[s][2 * q] = fma2s(dj, c.a, d1[s][2 * q]);
                d1[s][2 * q + 1] = fma2s(dj, c.b, d1[s][2 * q + 1]);
            }
        }
#pragma unroll
    for (int s = 0; s < S; ++s) {
#pragma unroll
        for (int i = 0; i < kH; ++i) {                                        // dW2[i][j] += h1[i] d2[j], pairs over j
            const float hi = elem(a.h1[s], i);
#pragma unroll
            for (int jp = 0; jp < kH / 2; ++jp)
                g.v[(N::W2 + i * kH) / 2 + jp] = fma2s(hi, d2[s][jp], g.v[(N::W2 + i * kH) / 2 + jp]);
        }
#pragma unroll
        for (int jp = 0; jp < kH / 2; ++jp) {
            d1[s][jp] = mul2(d1[s][jp], dlrelu2(a.h1[s][jp]));
            g.v[N::B2 / 2 + jp] = add2(g.v[N::B2 / 2 + jp], d2[s][jp]);
            g.v[N::B1 / 2 + jp] = add2(g.v[N::B1 / 2 + jp], d1[s][jp]);
        }
#pragma unroll
        for (int i = 0; i < N::IN; ++i)
#pragma unroll
            for (int jp = 0; jp < kH / 2; ++jp)
                g.v[(N::W1 + i * kH) / 2 + jp] = fma2s(x[s][i], d1[s][jp], g.v[(N::W1 + i * kH) / 2 + jp]);
    }
}

// per-sample loss and dLoss/dV (without the 1/N of the batch mean)
template <int LOSS>
__device__ __forceinline__ void loss_grad(float y, float v, float& l, float& dv) {
    // products rounded on their own: `loss += l` must not contract into an FMA in one kernel and not in another
    if (LOSS == ORP_LOSS_MSE) { const float e = v - y; l = __fmul_rn(e, e); dv = 2.f * e; }
    else { const float e = y - v; l = fmaxf(__fmul_rn(0.99f, e), __fmul_rn(0.99f - 1.f, e)); dv = e > 0.f ? -0.99f : 0.01f; }   // RP.py:138-142
}

struct FitArgs {
    const void* s_t; const void* s_t1; const float* feat1; const float* feat2; const float* target;
    uint64_t count;
    double inv_scale; float b_t, b_t1;
    int skip;                  // non-zero: samples with a non-finite input or target take no part (ORP_FIT_SKIP_NONFINITE)
    int n_seg;                 // the batch is n_seg segments of `count` samples, segment v starting seg_stride * v samples in
    int g_blocks;              // virtual blocks per segment (fit_blocks(count)): part of the sums' definition, not of the launch
    uint64_t seg_stride;
};

// Segment v of a batch as a batch of its own (orp_fit_batch.n_seg: the virtual shards that make the fit's sums
// independent of the number of GPUs, see reduce_partials).
template <typename T>
__device__ __forceinline__ FitArgs seg_args(const FitArgs& b, int v) {
    FitArgs a = b;
    const uint64_t off = (uint64_t)v * b.seg_stride;
    a.s_t = reinterpret_cast<const char*>(b.s_t) + off * sizeof(T);
    a.s_t1 = reinterpret_cast<const char*>(b.s_t1) + off * sizeof(T);
    a.target = b.target + off;
    if (b.feat1) a.feat1 = b.feat1 + off;
    if (b.feat2) a.feat2 = b.feat2 + off;
    return a;
}

// A price in network units.  fp64 columns: divide in fp64, then Keras' cast.  fp32 columns are in network units already
// (make_args insists on inv_scale == 1 for them; orp_fit_scale_column produces them with the same fp64 division).
template <typename T>
__device__ __forceinline__ float to_net(T v, double inv_scale) {
    if (sizeof(T) == 4) return (float)v;
    return (float)((double)v * inv_scale);
}
template <typename T>
__device__ __forceinline__ float load_price(const void* p, uint64_t i, double inv_scale) {
    return to_net<T>(reinterpret_cast<const T*>(p)[i], inv_scale);
}

template <int NET, typename T>
__device__ __forceinline__ void load_sample(const FitArgs& a, uint64_t i, float (&x)[Net<NET>::IN], float& s_t, float& s_t1) {
    s_t = load_price<T>(a.s_t, i, a.inv_scale);
    s_t1 = load_price<T>(a.s_t1, i, a.inv_scale);
    x[0] = s_t;
    if (NET == ORP_NET_RP) {
        x[1] = a.feat1 ? a.feat1[i] : 1.f;
        x[Net<NET>::IN - 1] = a.feat2 ? a.feat2[i] : 0.f;
    }
}

constexpr int kGsumLen = ORP_FIT_MAX_PARAMS + 2;   // grads, loss sum, count

__device__ __forceinline__ double schedule_lr(int epoch, double lr) {      // Replicating_Portfolio.py:128-136
    if (epoch < 100) return 1e-2;
    if (epoch < 200) return 1e-3;
    if (epoch < 400) return 5e-4;
    return lr;
}

// Adam (Keras: lr_t = lr sqrt(1-b2^t)/(1-b1^t); w -= lr_t m / (sqrt(v) + 1e-7)), fp32 like the Keras variables.
// Every operation is rounded on its own (no FMA contraction), so that the three drivers of a fit (two-stage,
// single-launch step, persistent kernel) and the Keras-faithful mini-batch kernel produce the same bits.
__device__ __forceinline__ void adam_core(float& m, float& v, float& w, float grad, float lr_t) {
    m = __fadd_rn(__fmul_rn(0.9f, m), __fmul_rn(1.f - 0.9f, grad));
    v = __fadd_rn(__fmul_rn(0.999f, v), __fmul_rn(1.f - 0.999f, __fmul_rn(grad, grad)));
    w = __fsub_rn(w, __fdiv_rn(__fmul_rn(lr_t, m), __fadd_rn(__fsqrt_rn(v), 1e-7f)));
}

__device__ __forceinline__ void adam_apply(float* __restrict__ params, double* __restrict__ state, int k, float grad,
                                           float lr_t) {
    float m = (float)state[ST_ADAM_M + k], v = (float)state[ST_ADAM_V + k], w = params[k];
    adam_core(m, v, w, grad, lr_t);
    state[ST_ADAM_M + k] = m;
    state[ST_ADAM_V + k] = v;
    params[k] = w;
}

// End-of-epoch callbacks, run by ONE thread after the epoch's updates are visible in params.
__device__ void epoch_end(int P, const float* __restrict__ params, double* __restrict__ state, double cur) {
    state[ST_LAST_LOSS] = cur;
    int epoch = (int)state[ST_EPOCH];
    double wait = state[ST_WAIT] + 1.0;
    if (cur < state[ST_BEST]) {                       // EarlyStopping(monitor='loss', min_delta=0)
        state[ST_BEST] = cur;
        wait = 0.0;
        for (int k = 0; k < P; ++k) state[ST_BEST_W + k] = params[k];     // restore_best_weights=True
        state[ST_HAVE_BEST] = 1.0;
    }
    state[ST_WAIT] = wait;
    if (wait >= state[ST_PATIENCE] && epoch > 0) { state[ST_STOP] = 1.0; state[ST_STOPPED_EARLY] = 1.0; }
    epoch += 1;
    state[ST_EPOCH] = epoch;
    state[ST_EPOCHS_RUN] = epoch;
    state[ST_TOTAL_EPOCHS] += 1.0;
    if (epoch >= (int)state[ST_EPOCHS]) state[ST_STOP] = 1.0;
}

// One optimiser step, run by one block of >= 128 threads (all of them must call).  flags bit 0: first batch of
// the epoch (LearningRateScheduler.on_epoch_begin), bit 1: last batch (epoch loss = sample-weighted mean of the
// batch losses, then the callbacks).  gsum may live in shared or global memory.
__device__ __forceinline__ void optimiser_step(int P, float* __restrict__ params, double* __restrict__ state,
                                               const double* __restrict__ gsum, int flags) {
    __shared__ float s_lr_t;
    if (threadIdx.x == 0) {
        double lr = state[ST_LR];
        if ((flags & 1) && state[ST_USE_SCHED] != 0.0) lr = schedule_lr((int)state[ST_EPOCH], lr);
        state[ST_LR] = lr;
        const double t = state[ST_ADAM_T] + 1.0;
        state[ST_ADAM_T] = t;
        s_lr_t = (float)(lr * sqrt(1.0 - pow(0.999, t)) / (1.0 - pow(0.9, t)));
    }
    __syncthreads();
    const double inv_n = 1.0 / gsum[ORP_FIT_MAX_PARAMS + 1];
    if (threadIdx.x < P) adam_apply(params, state, threadIdx.x, (float)(gsum[threadIdx.x] * inv_n), s_lr_t);
    __syncthreads();
    if (threadIdx.x == 0) {
        const double loss_acc = ((flags & 1) ? 0.0 : state[ST_LOSS_ACC]) + gsum[ORP_FIT_MAX_PARAMS];
        const double cnt_acc = ((flags & 1) ? 0.0 : state[ST_COUNT_ACC]) + gsum[ORP_FIT_MAX_PARAMS + 1];
        state[ST_LOSS_ACC] = loss_acc;
        state[ST_COUNT_ACC] = cnt_acc;
        if (flags & 2) epoch_end(P, params, state, loss_acc / cnt_acc);
    }
}

__global__ void fit_update_kernel(int P, float* __restrict__ params, double* __restrict__ state,
                                  const double* __restrict__ gsum, int flags) {
    if (state[ST_STOP] != 0.0) return;
    optimiser_step(P, params, state, gsum, flags);
}

// Pairwise (balanced-tree) sum of n values in index order: ((v0 + v1) + (v2 + v3)) + ...  The sums over virtual shards
// and over ranks use it, so that a rank holding the shards [r L, (r + 1) L) of the V = L x world shards adds its own L
// totals first and the ranks' totals afterwards and still reproduces, bit for bit, the one-GPU sum over all V shards
// (L and world powers of two).
template <int MAXN>
__device__ __forceinline__ double tree_sum(double (&v)[MAXN], int n) {
#pragma unroll
    for (int stride = 1; stride < MAXN; stride *= 2)
#pragma unroll
        for (int i = 0; i + stride < MAXN; i += 2 * stride)
            if (i + stride < n) v[i] += v[i + stride];
    return v[0];
}

constexpr int kMaxSeg = ORP_FIT_MAX_SEGMENTS;

// Fixed-order sum of the per-block partials [segment][block][kGsumLen] by ONE block of 256 threads: per segment, lane l
// adds blocks l, l+32, ... in increasing order and a shuffle tree combines the 32 lane sums; the segment totals are
// then added pairwise — the same association on every launch, whichever block finished last, and on any number of
// GPUs that hold whole segments.  Writes gsum and (optionally) a shared-memory copy.
__device__ __forceinline__ void reduce_partials(const double* __restrict__ partials, int n_blocks, int n_seg,
                                                double* __restrict__ gsum, double* s_copy) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int k = warp; k < kGsumLen; k += (int)(blockDim.x >> 5)) {
        double seg[kMaxSeg];
#pragma unroll
        for (int v = 0; v < kMaxSeg; ++v) {
            double t = 0.0;
            if (v < n_seg) {
                for (int b = lane; b < n_blocks; b += 32) t += __ldcg(partials + ((size_t)v * n_blocks + b) * kGsumLen + k);
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) t += __shfl_down_sync(0xffffffffu, t, o);
            }
            seg[v] = t;
        }
        const double t = tree_sum<kMaxSeg>(seg, n_seg);
        if (lane == 0) { gsum[k] = t; if (s_copy) s_copy[k] = t; }
    }
}

#ifndef ORP_FIT_THREADS
#define ORP_FIT_THREADS 256
#endif
constexpr int kFitThreads = ORP_FIT_THREADS;       // scaled drivers, predict and residual kernels
constexpr int kGsum = kGsumLen;

// Sum over the 32 lanes of a warp of N per-lane values, as a reduce-scatter: at every stage a lane keeps one half
// of its values and hands the other half to its partner, so N values cost about N shuffles instead of 5 N.
// Afterwards lane l holds, in a[0 .. N/32), the warp totals of the entries i + (N/32) * bit-weighted(l) — see
// scattered_index().  N must be a multiple of 32; the association is fixed, hence deterministic.
template <int N>
__device__ __forceinline__ void warp_reduce_scatter(float (&a)[N], int lane) {
    static_assert(N % 32 == 0, "pad to a multiple of the warp size");
    int len = N;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        const int half = len / 2;
        const bool up = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < N / 2; ++i) {
            if (i < half) {
                const float lo = a[i], hi = a[i + half];
                const float send = up ? lo : hi, keep = up ? hi : lo;
                a[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
            }
        }
        len = half;
    }
}
template <int N>
__device__ __forceinline__ int scattered_index(int i, int lane) {
    constexpr int R = N / 32;
    return i + R * ((lane & 1) + 2 * ((lane >> 1) & 1) + 4 * ((lane >> 2) & 1) + 8 * ((lane >> 3) & 1) + 16 * ((lane >> 4) & 1));
}

// Block total of the per-thread gradients, loss and sample count as ONE fp64 vector: warp reduce-scatter ->
// shared memory -> sum over the 8 warps in warp order -> dst[index * stride] (index = k for the gradient of
// parameter k, ORP_FIT_MAX_PARAMS for the loss, + 1 for the count).  Shared by every driver of the fit.
// tag != 0: every entry is written as two 8-byte words {low half, tag}, {high half, tag} (dst counts 16-byte
// entries), so that a reader polling for `tag` needs no barrier and no fence (8-byte accesses are single-copy atomic).
__device__ __forceinline__ void st_tagged(double* entry, double val, unsigned int tag) {
    // two 8-byte elements {value half (low 32 bits), tag (high 32 bits)}: each element is ONE 64-bit access in the
    // PTX memory model (a .v4.u32 would be four independent 4-byte accesses), so a half can never be seen with a
    // newer tag than its value
    const unsigned long long t = (unsigned long long)tag << 32;
    const unsigned long long e0 = t | (unsigned int)__double2loint(val), e1 = t | (unsigned int)__double2hiint(val);
    asm volatile("st.volatile.global.v2.b64 [%0], {%1, %2};" ::"l"(entry), "l"(e0), "l"(e1) : "memory");
}
__device__ __forceinline__ uint4 ld_tagged(const double* entry) {
    unsigned long long e0, e1;
    asm volatile("ld.volatile.global.v2.b64 {%0, %1}, [%2];" : "=l"(e0), "=l"(e1) : "l"(entry) : "memory");
    return make_uint4((unsigned int)e0, (unsigned int)(e0 >> 32), (unsigned int)e1, (unsigned int)(e1 >> 32));   // {low, tag, high, tag}
}

// Warp totals of one unit's per-thread gradients, loss and sample count -> red[0 .. P + 2) in shared memory (gradient in
// parameter order, then loss, then count).  No block barrier: the warps of a block drift apart from unit to unit, so one
// warp's reduction (shuffles and selects) overlaps the other warps' arithmetic.
template <int NET>
__device__ __forceinline__ void warp_partial(const Grads<NET>& g, float loss, uint32_t n, float* __restrict__ red) {
    constexpr int P = Net<NET>::P;
    constexpr int N = P + 2;
    constexpr int NB = (N % 32) <= 4 ? (N / 32) * 32 : ((N + 31) / 32) * 32;      // EO: 96 (+ 3 by the classic tree); RP: 128
    constexpr int REST = N > NB ? N - NB : 0;
    const int lane = threadIdx.x & 31;
    float a[NB];
#pragma unroll
    for (int k = 0; k < NB; ++k) a[k] = k < P ? g.get(k) : (k == P ? loss : (k == P + 1 ? (float)n : 0.f));
    warp_reduce_scatter<NB>(a, lane);
#pragma unroll
    for (int i = 0; i < NB / 32; ++i) {
        const int q = scattered_index<NB>(i, lane);
        if (q < N) red[q] = a[i];
    }
#pragma unroll
    for (int k = NB; k < NB + REST; ++k) {
        float v = k < P ? g.get(k) : (k == P ? loss : (float)n);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if (lane == 0) red[k] = v;
    }
}

// Block totals of the units this block worked on in a step, as fp64 vectors.  A UNIT is (segment sv, virtual block vb):
// virtual block vb of g_k takes the tiles vb, vb + g_k, ... of the segment's part, whatever physical block runs it — the
// units of a step are numbered u = sv * g_k + vb and physical block b of G runs u = b, b + G, ... (slot 0, 1, ...), so a
// step of few tiles per segment still spreads over the whole grid, and the sums do not depend on the grid at all.
// For every (slot, column) the warps' totals are added in warp order and written to
// dst[sv * seg_stride + index * col_stride + vb * blk_stride] (index = k for the gradient of parameter k,
// ORP_FIT_MAX_PARAMS for the loss, + 1 for the count).  Call after a block barrier that follows the last warp_partial.
// Shared by every scaled driver of the fit, so they agree bit for bit.
// tag != 0: every entry is written as two 8-byte words {low half, tag}, {high half, tag} (dst counts 16-byte
// entries), so that a reader polling for `tag` needs no barrier and no fence (8-byte accesses are single-copy atomic).
template <int NET>
__device__ __forceinline__ void block_publish(const float (*s_red)[kFitThreads / 32][Net<NET>::P + 2], int n_slots,
                                              uint32_t u0, uint32_t u_stride, uint32_t g_k, double* __restrict__ dst,
                                              size_t seg_stride, size_t col_stride, size_t blk_stride, unsigned int tag = 0u,
                                              int slot0 = 0) {
    constexpr int P = Net<NET>::P;
    constexpr int N = P + 2;
    for (int item = slot0 * N + threadIdx.x; item < n_slots * N; item += kFitThreads) {
        const int slot = item / N, q = item - slot * N;
        const uint32_t u = u0 + (uint32_t)slot * u_stride, sv = u / g_k, vb = u - sv * g_k;
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < kFitThreads / 32; ++w) t += (double)s_red[slot][w][q];
        const size_t at = (size_t)sv * seg_stride + (size_t)(q < P ? q : ORP_FIT_MAX_PARAMS + (q - P)) * col_stride + (size_t)vb * blk_stride;
        if (tag) st_tagged(dst + 2 * at, t, tag); else dst[at] = t;
    }
}

// The same for ONE unit by ONE warp (the persistent kernel publishes a unit as soon as every warp has parked its totals,
// one unit behind the arithmetic, instead of all units in a burst at the end of the step).
template <int NET>
__device__ __forceinline__ void warp_publish(const float (&red)[kFitThreads / 32][Net<NET>::P + 2], uint32_t u, uint32_t g_k,
                                             double* __restrict__ dst, size_t seg_stride, size_t col_stride, unsigned int tag, int lane) {
    constexpr int P = Net<NET>::P;
    constexpr int N = P + 2;
    const uint32_t sv = u / g_k, vb = u - sv * g_k;
    for (int q = lane; q < N; q += 32) {
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < kFitThreads / 32; ++w) t += (double)red[w][q];
        const size_t at = (size_t)sv * seg_stride + (size_t)(q < P ? q : ORP_FIT_MAX_PARAMS + (q - P)) * col_stride + (size_t)vb;
        st_tagged(dst + 2 * at, t, tag);
    }
}

// Split block barrier on a numbered hardware barrier (PTX producer / consumer pattern): the producers' shared-memory
// stores before bar.arrive are visible to the consumer after its bar.sync on the same barrier.
__device__ __forceinline__ void bar_arrive(int id, int n_threads) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n_threads) : "memory"); }
__device__ __forceinline__ void bar_sync(int id, int n_threads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n_threads) : "memory"); }

// Raw inputs of one sample, as loaded (conversion to the network's fp32 happens when the sample is used).
template <int NET, typename T> struct RawSample {
    T s_t, s_t1;
    float y, f1, f2;
};

template <int NET, typename T>
__device__ __forceinline__ void load_raw(const FitArgs& a, uint64_t i, RawSample<NET, T>& r) {
    r.s_t = reinterpret_cast<const T*>(a.s_t)[i];
    r.s_t1 = reinterpret_cast<const T*>(a.s_t1)[i];
    r.y = a.target[i];
    if (NET == ORP_NET_RP) {
        r.f1 = a.feat1 ? a.feat1[i] : 1.f;
        r.f2 = a.feat2 ? a.feat2[i] : 0.f;
    }
}

// the same for the passes that may run without a target (predict)
template <int NET, typename T>
__device__ __forceinline__ void load_raw_any(const FitArgs& a, uint64_t i, RawSample<NET, T>& r) {
    r.s_t = reinterpret_cast<const T*>(a.s_t)[i];
    r.s_t1 = reinterpret_cast<const T*>(a.s_t1)[i];
    r.y = a.target ? a.target[i] : 0.f;
    if (NET == ORP_NET_RP) {
        r.f1 = a.feat1 ? a.feat1[i] : 1.f;
        r.f2 = a.feat2 ? a.feat2[i] : 0.f;
    }
}

template <int NET, typename T, bool SKIP>
__device__ __forceinline__ bool unpack_raw(const FitArgs& a, const RawSample<NET, T>& r, float (&x)[Net<NET>::IN], float& s_t1, float& y) {
    x[0] = to_net<T>(r.s_t, a.inv_scale);
    s_t1 = to_net<T>(r.s_t1, a.inv_scale);
    y = r.y;
    if (NET == ORP_NET_RP) { x[1] = r.f1; x[Net<NET>::IN - 1] = r.f2; }
    if (SKIP) {                                            // opt-in (own instantiation: the default loop carries none of it):
                                                           // a non-finite sample is replaced by zeros and reported
        bool fin = isfinite(s_t1) && isfinite(y);
#pragma unroll
        for (int q = 0; q < Net<NET>::IN; ++q) fin = fin && isfinite(x[q]);
        if (!fin) {
#pragma unroll
            for (int q = 0; q < Net<NET>::IN; ++q) x[q] = 0.f;
            s_t1 = 0.f; y = 0.f;
            return false;
        }
    }
    return true;
}

// A block walks tiles of kFitThreads * U consecutive samples (thread t owns samples t, t + 256, ... of the
// tile: every load is a full-line warp access) and pushes them through the network two at a time.  The loads
// of the NEXT tile are issued before the arithmetic of the current one, so their HBM latency hides behind
// U x ~230 issue slots per warp: at 8 warps per SM (the 97/122 gradient accumulators take the register file)
// there is no other warp to switch to.
constexpr int kFitS = 2;                                   // samples per thread in flight through the network
#ifndef ORP_FIT_UNROLL_EO
#define ORP_FIT_UNROLL_EO 4
#endif
template <int NET> constexpr int kFitUnroll = NET == ORP_NET_EO ? ORP_FIT_UNROLL_EO : 2;

// Issues the loads of this block's FIRST tile of a batch (the persistent kernel does this one optimiser step
// ahead, so that the latency hides behind the grid barrier and the reduction).
template <int NET, typename T>
__device__ __forceinline__ void load_first_tile(const FitArgs& a, uint32_t block, RawSample<NET, T> (&cur)[kFitUnroll<NET>]) {
    constexpr int U = kFitUnroll<NET>;
    const uint64_t base = (uint64_t)block * ((uint64_t)kFitThreads * U) + threadIdx.x;
#pragma unroll
    for (int u = 0; u < U; ++u) {
        cur[u] = RawSample<NET, T>{};
        if (base + (uint64_t)u * kFitThreads < a.count) load_raw<NET, T>(a, base + (uint64_t)u * kFitThreads, cur[u]);
    }
}

// Forward + backward over this block's tiles of one batch: block `block` of `n_blocks` takes tiles block,
// block + n_blocks, ...  `cur` holds the first tile (load_first_tile).  Adds into g / loss / n.
// `nx` (optional): the batch this block works on NEXT, as virtual block `block_nx` — its first tile is requested
// before the arithmetic of THIS batch's last tile and left in `cur`, so that a unit boundary exposes no memory latency
// (requested only after the last tile, every boundary cost ~1.2 us: 8 % of a one-GPU step with eight units).
template <int NET, int LOSS, typename T, bool SKIP>
__device__ __forceinline__ void accumulate_tiles(const FitArgs& a, uint32_t s_wa, uint32_t block, uint32_t n_blocks,
                                                 RawSample<NET, T> (&cur)[kFitUnroll<NET>], Grads<NET>& g, float& loss, uint32_t& n,
                                                 const FitArgs& nx, uint32_t block_nx, bool has_nx) {
    using N = Net<NET>;
    constexpr int U = kFitUnroll<NET>, S = kFitS;
    static_assert(U % S == 0, "tile slots are consumed S at a time");
    const uint64_t tile = (uint64_t)kFitThreads * U, stride = (uint64_t)n_blocks * tile;
    uint64_t base = (uint64_t)block * tile + threadIdx.x;
    RawSample<NET, T> nxt[U];
    for (; base < a.count + threadIdx.x; base += stride) {        // block-uniform trip count (base - threadIdx.x < count)
        const uint64_t nb = base + stride;
        // one load path for both cases (block-uniform choice): the next tile of this batch, or the first tile of the next one
        const bool more = nb < a.count + threadIdx.x;
        FitArgs src = a;
        if (!more) { src.s_t = nx.s_t; src.s_t1 = nx.s_t1; src.target = nx.target; src.feat1 = nx.feat1; src.feat2 = nx.feat2; }
        const uint64_t nidx = more ? nb : (uint64_t)block_nx * tile + threadIdx.x;
        const uint64_t ncount = more ? a.count : (has_nx ? nx.count : 0ull);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            nxt[u] = RawSample<NET, T>{};
            if (nidx + (uint64_t)u * kFitThreads < ncount) load_raw<NET, T>(src, nidx + (uint64_t)u * kFitThreads, nxt[u]);
        }
#pragma unroll
        for (int u = 0; u < U; u += S) {
            if (base + (uint64_t)u * kFitThreads < a.count) {                 // slot u valid; later slots of the group may not be
                float x[S][N::IN], s_t1[S], y[S], phi[S], psi[S], v[S], dv[S];
                bool fin[S];
                Acts<NET, S> act;
#pragma unroll
                for (int s = 0; s < S; ++s) fin[s] = unpack_raw<NET, T, SKIP>(a, cur[u + s], x[s], s_t1[s], y[s]);     // out-of-range slots hold zeros
                net_forward<NET, S>(s_wa, x, s_t1, a.b_t1, act, phi, psi, v);
#pragma unroll
                for (int s = 0; s < S; ++s) {
                    const bool ok = fin[s] && base + (uint64_t)(u + s) * kFitThreads < a.count;
                    float l;
                    loss_grad<LOSS>(y[s], v[s], l, dv[s]);
                    if (!ok) { l = 0.f; dv[s] = 0.f; }
                    loss += l;
                    n += ok ? 1u : 0u;
                }
                net_backward<NET, S>(s_wa, x, s_t1, a.b_t1, dv, act, g);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) cur[u] = nxt[u];
    }
}

// STEP = true: the whole optimiser step in this one launch — the block that arrives last (a ticket counter in
// the state block) adds the partials in the same fixed order as fit_reduce_kernel and applies the update.
template <int NET, int LOSS, typename T, bool STEP, bool SKIP>
__global__ void __launch_bounds__(kFitThreads) fit_grad_kernel(const FitArgs a, float* params, double* state,
                                                               double* __restrict__ partials, double* __restrict__ gsum, int flags) {
    using N = Net<NET>;
    if (state[ST_STOP] != 0.0) return;
    __shared__ __align__(16) float s_w[kSmemW];
    load_weights<NET>(s_w, params, threadIdx.x, kFitThreads);
    __syncthreads();
    const uint32_t s_wa = (uint32_t)__cvta_generic_to_shared(s_w);

    // one UNIT per block: block b works on virtual block b % g_k of segment b / g_k (g_k = a.g_blocks blocks per segment)
    __shared__ float s_red[1][kFitThreads / 32][N::P + 2];
    const uint32_t g_k = (uint32_t)a.g_blocks, sv = blockIdx.x / g_k, vb = blockIdx.x - sv * g_k;
    {
        const FitArgs av = seg_args<T>(a, (int)sv);
        RawSample<NET, T> first[kFitUnroll<NET>];
        load_first_tile<NET, T>(av, vb, first);
        Grads<NET> g;
        g.zero();
        float loss = 0.f;
        uint32_t n = 0;
        accumulate_tiles<NET, LOSS, T, SKIP>(av, s_wa, vb, g_k, first, g, loss, n, av, 0u, false);
        warp_partial<NET>(g, loss, n, s_red[0][threadIdx.x >> 5]);
    }
    __syncthreads();
    block_publish<NET>(s_red, 1, blockIdx.x, 0u, g_k, partials, (size_t)g_k * kGsum, 1, kGsum);
    if (STEP) {
        __shared__ bool s_last;
        __shared__ double s_g[kGsumLen];
        unsigned int* ticket = reinterpret_cast<unsigned int*>(state + ST_TICKET);
        __threadfence();                                   // this block's partial is visible before its ticket
        __syncthreads();
        if (threadIdx.x == 0) s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
        __syncthreads();
        if (!s_last) return;
        __threadfence();
        reduce_partials(partials, a.g_blocks, a.n_seg, gsum, s_g);
        __syncthreads();
        optimiser_step(N::P, params, state, s_g, flags);
        if (threadIdx.x == 0) *ticket = 0u;                // ready for the next launch
    }
}

__global__ void __launch_bounds__(256) fit_reduce_kernel(const double* __restrict__ partials, int n_blocks, int n_seg,
                                                         const double* __restrict__ state, double* __restrict__ gsum) {
    if (state[ST_STOP] != 0.0) return;
    reduce_partials(partials, n_blocks, n_seg, gsum, nullptr);
}

__global__ void fit_init_kernel(double* __restrict__ state) {
    for (int k = threadIdx.x; k < ORP_FIT_STATE_DOUBLES; k += blockDim.x) state[k] = 0.0;
    __syncthreads();
    if (threadIdx.x == 0) { state[ST_LR] = 1e-3; state[ST_STOP] = 1.0; }      // Adam(learning_rate=1e-3), RP.py:175
}

__global__ void fit_begin_kernel(double* __restrict__ state, int epochs, int patience, int use_schedule) {
    if (threadIdx.x != 0) return;
    state[ST_EPOCH] = 0.0; state[ST_EPOCHS] = epochs; state[ST_PATIENCE] = patience; state[ST_USE_SCHED] = use_schedule;
    state[ST_WAIT] = 0.0; state[ST_BEST] = __longlong_as_double(0x7ff0000000000000ll); state[ST_STOP] = epochs > 0 ? 0.0 : 1.0;
    state[ST_STOPPED_EARLY] = 0.0; state[ST_EPOCHS_RUN] = 0.0; state[ST_HAVE_BEST] = 0.0;
}

__global__ void fit_end_kernel(int P, float* __restrict__ params, double* __restrict__ state) {
    // TF 2.x EarlyStopping: best weights come back only when the stop fired before the last epoch
    if (state[ST_STOPPED_EARLY] != 0.0 && state[ST_HAVE_BEST] != 0.0 && threadIdx.x < P)
        params[threadIdx.x] = (float)state[ST_BEST_W + threadIdx.x];
    if (threadIdx.x == 0) state[ST_STOP] = 1.0;
}

// ---- persistent fit: every epoch of a Model.fit call in ONE cooperative launch ----------------------
//
// One block per SM stays resident for the whole fit (cooperative launch: all blocks co-resident).  Per optimiser
// step: each block accumulates its tiles of the batch and reduces them to one fp64 vector, which it publishes as
// tagged entries [parity][column][block]; the owner block of each column adds the blocks' entries in the fixed order
// of fit_reduce_kernel and publishes the column total (to every rank); EVERY block then adds the ranks' totals and
// applies the Adam step to its own shared-memory copy of the weights (moments in registers, callback state in shared
// memory).  All blocks hold bit-identical weights and take identical decisions, so there is no grid-wide barrier and no
// weight round trip through global memory.  Block 0 writes weights and state back when the fit stops or the epochs
// run out.  Per-step fixed cost: ~5 us instead of a kernel boundary (~18 us measured).
__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Hand-over polls give up after a wall-clock budget (not a retry count: host-side skew between ranks — a first-launch
// module load, a slow collective — or a profiler's replay can stretch a healthy wait by orders of magnitude).
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
constexpr unsigned long long kPollBudgetNs = 60ull * 1000ull * 1000ull * 1000ull;      // 60 s
__device__ __forceinline__ bool poll_expired(unsigned long long t_start) { return globaltimer_ns() - t_start > kPollBudgetNs; }

// Optional phase timestamps (build with -DORP_FIT_TRACE; tools/trace_fit.py): blocks 0 and G-1 write %globaltimer at
// eight points of each of the first 512 optimiser steps into the 64 KB tail of the scratch buffer.
#ifdef ORP_FIT_TRACE
#define ORP_TRACE(phase)                                                                                          \
    do {                                                                                                          \
        if (threadIdx.x == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1) && barriers_done_for_trace < 512) { \
            unsigned long long t_;                                                                                \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                                \
            reinterpret_cast<unsigned long long*>(partials)[kScrDoubles +                                          \
                ((blockIdx.x == 0 ? 0 : 512) + barriers_done_for_trace) * 8 + (phase)] = t_;                     \
        }                                                                                                         \
    } while (0)
#else
#define ORP_TRACE(phase) do { } while (0)
#endif

struct RunArgs;
template <typename T> __device__ __forceinline__ FitArgs part_args(const RunArgs& r, int v, int k, uint32_t G, uint32_t& g_k);

struct RunArgs {
    FitArgs a;                 // the whole batch: a.n_seg segments of a.count samples; part k of a segment = its samples
                               // [k * per, min((k + 1) * per, count)); optimiser step k sees part k of EVERY segment
    uint64_t per;
    int n_parts, n_epochs, max_blocks_per_part, unroll;
    int rank, world;           // world = 0: single GPU, no exchange
    double* const* peers;      // device array of `world` exchange-buffer pointers (peers[rank] = this GPU's own)
};

// ---- exchange buffer of the multi-GPU fit (one per rank, cudaMalloc'ed, mapped into every peer through CUDA IPC) ----
// The column totals of every rank are all-gathered by plain stores over NVLink with the flag travelling INSIDE the
// store (the "LL" idea of collective libraries): a double is sent as {low 32 bits, step, high 32 bits, step}.
// 8-byte halves are single-copy atomic, so a receiver that sees the step number it is waiting for in both halves has
// the value — no fence, no separate flag, one NVLink one-way latency.  The owner warp of a column writes its total into
// slot [parity][rank][column] of EVERY rank's buffer; every block of every rank polls the `world` slots in its OWN memory
// and adds them in rank order — the same bits on every GPU.  Two parities suffice: a rank can only send step s + 2 after
// every peer has sent step s + 1, i.e. after all of that peer's blocks have finished reading step s.
constexpr int kPeerMaxWorld = 16;
constexpr int kPeerHdr = 0;                                     // word 0 (uint32): optimiser steps completed so far
constexpr int kPeerSlots = 32;                                  // tagged totals [2][kPeerMaxWorld][kGsumLen][2], in 8-byte words
constexpr int kPeerDoubles = kPeerSlots + 2 * kPeerMaxWorld * kGsumLen * 2;

// Scratch layout of the persistent fit (in doubles from the start of d_scratch; the first region belongs to the
// per-step drivers).  The scratch must be ZERO-FILLED once before its first use: the tags below count up from the
// step counter kept in its header and must never meet stale memory that happens to carry a future tag.
constexpr int kLLBlocks = 256;                                              // grid size limit (fit_blocks caps at it)
constexpr size_t kScrHeader = (size_t)kMaxSeg * kLLBlocks * kGsumLen;       // (per-step drivers' partials in front); word 0 (uint32): optimiser steps done with this scratch
constexpr size_t kScrLLPart = kScrHeader + 16;                              // tagged block totals [2][kMaxSeg][kGsumLen][kLLBlocks][2]
constexpr size_t kScrLLFinal = kScrLLPart + 2ull * kMaxSeg * kGsumLen * kLLBlocks * 2;    // tagged totals [2][kPeerMaxWorld][kGsumLen][2] (one GPU)
constexpr size_t kScrDoubles = kScrLLFinal + 2ull * kPeerMaxWorld * kGsumLen * 2;

// Part k of segment v and the number of blocks that work on it (the grid orp_fit_grad would launch for it).
template <typename T>
__device__ __forceinline__ FitArgs part_args(const RunArgs& r, int v, int k, uint32_t G, uint32_t& g_k) {
    FitArgs a = r.a;
    const uint64_t in_seg = (uint64_t)k * r.per, off = (uint64_t)v * r.a.seg_stride + in_seg;
    a.count = min(r.per, r.a.count - in_seg);
    a.s_t = reinterpret_cast<const char*>(r.a.s_t) + off * sizeof(T);
    a.s_t1 = reinterpret_cast<const char*>(r.a.s_t1) + off * sizeof(T);
    a.target = r.a.target + off;
    if (r.a.feat1) a.feat1 = r.a.feat1 + off;
    if (r.a.feat2) a.feat2 = r.a.feat2 + off;
    const uint64_t tile = (uint64_t)kFitThreads * (r.unroll);
    const uint64_t want = (a.count + tile - 1) / tile;
    g_k = (uint32_t)min(want, (uint64_t)r.max_blocks_per_part);      // = fit_blocks(a.count): no function of the grid
    return a;
}

template <int NET, int LOSS, typename T, bool SKIP>
__global__ void __launch_bounds__(kFitThreads) fit_run_kernel(const RunArgs r, float* params, double* state,
                                                              double* __restrict__ partials, double* __restrict__ gsum) {
    using N = Net<NET>;
    constexpr int P = N::P;
    __shared__ __align__(16) float s_w[kSmemW];
    __shared__ float s_best[ORP_FIT_MAX_PARAMS];
    __shared__ double s_g[kGsumLen];
    __shared__ double s_c[16];                     // head of the control block (ST_LR .. ST_COUNT_ACC)
    __shared__ double s_seg[kGsumLen][kMaxSeg];    // owner stage: segment totals of the columns this block owns
    extern __shared__ __align__(16) float s_dyn[];                 // warp totals of up to kMaxSeg units: [slot][warp][P + 2]
    float (*s_red)[kFitThreads / 32][P + 2] = reinterpret_cast<float (*)[kFitThreads / 32][P + 2]>(s_dyn);
    __shared__ float s_lr_t;
    __shared__ int s_improved;
    constexpr int kLrTable = 256;
    __shared__ double s_num[kLrTable], s_den[kLrTable];            // Adam bias terms of the next 256 steps
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned int* abort_flag = reinterpret_cast<unsigned int*>(state + ST_ABORT);

    load_weights<NET>(s_w, params, tid, kFitThreads);
    float m = 0.f, v = 0.f;
    if (tid < P) { m = (float)state[ST_ADAM_M + tid]; v = (float)state[ST_ADAM_V + tid]; s_best[tid] = (float)state[ST_BEST_W + tid]; }
    if (tid < 16) s_c[tid] = state[tid];
    __syncthreads();
    const uint32_t s_wa = (uint32_t)__cvta_generic_to_shared(s_w);
    const uint32_t G = gridDim.x;
    unsigned int barriers = 0;
    bool alive = true;
    __shared__ int s_peer_ok;
    double* const ll_part = partials + kScrLLPart;                                     // tagged block totals
    double* const ll_final = partials + kScrLLFinal;                                   // tagged totals, single-GPU form
    double* const xbuf = r.world > 0 ? r.peers[r.rank] : nullptr;                      // this GPU's exchange buffer
    const int n_ranks = r.world > 0 ? r.world : 1, my_rank = r.world > 0 ? r.rank : 0;
    const double* const fin_mine = r.world > 0 ? xbuf + kPeerSlots : ll_final;         // where the ranks' totals arrive
    const unsigned int tag0 = *reinterpret_cast<const unsigned int*>(partials + kScrHeader);    // steps done with this scratch
    const unsigned int step0 = r.world > 0 ? *reinterpret_cast<const unsigned int*>(xbuf + kPeerHdr) : tag0;

    RawSample<NET, T> first[kFitUnroll<NET>];
    {
        uint32_t g0, g_same;
        part_args<T>(r, 0, 0, G, g0);
        const uint32_t sg0 = blockIdx.x / g0;
        for (int u = 0; u < kFitUnroll<NET>; ++u) first[u] = RawSample<NET, T>{};
        if (sg0 < (uint32_t)r.a.n_seg) load_first_tile<NET, T>(part_args<T>(r, (int)sg0, 0, G, g_same), blockIdx.x - sg0 * g0, first);
    }
    for (int e = 0; e < r.n_epochs && alive; ++e) {
        if (s_c[ST_STOP] != 0.0) break;                                    // same value in every block
        for (int k = 0; k < r.n_parts && alive; ++k) {
            // ---- this block's share of batch k (its first tile was requested one step ago)
#ifdef ORP_FIT_TRACE
            const unsigned int barriers_done_for_trace = barriers;
#endif
            ORP_TRACE(0);
            // ---- one UNIT per segment: this block's tiles of part k of segment v -> fp32 accumulators -> block total
            //      -> tagged entry [parity][segment][column][block] (blocks without tiles in this part publish nothing)
            ++barriers;
            const unsigned int tag = tag0 + barriers;                          // never 0; grows across launches
            const unsigned int par = tag & 1u;
            const int n_seg = r.a.n_seg;
            double* const part = ll_part + (size_t)par * kMaxSeg * kGsumLen * kLLBlocks * 2;
            uint32_t g_k = 0;
            part_args<T>(r, 0, k, G, g_k);                                     // virtual blocks per segment in this part
            const uint32_t n_units = (uint32_t)n_seg * g_k;
            int n_slots = 0;
            for (uint32_t u = blockIdx.x; u < n_units; u += G, ++n_slots) {
                const uint32_t sg = u / g_k, vb = u - sg * g_k;
                uint32_t g_same;
                const FitArgs a = part_args<T>(r, (int)sg, k, G, g_same);
                Grads<NET> g;
                g.zero();
                float loss = 0.f;
                uint32_t n = 0;
                // this block's NEXT unit (in this step, or its first unit of the next step): accumulate_tiles requests its first
                // tile before the arithmetic of this unit's last tile (wasted only when the fit stops here)
                FitArgs a_nx;
                uint32_t vb_nx = 0u;
                bool has_nx = true;
                if (u + G < n_units) {
                    const uint32_t sg2 = (u + G) / g_k;
                    a_nx = part_args<T>(r, (int)sg2, k, G, g_same);
                    vb_nx = u + G - sg2 * g_k;
                } else {
                    uint32_t g_next;
                    const int k2 = k + 1 < r.n_parts ? k + 1 : 0;
                    part_args<T>(r, 0, k2, G, g_next);
                    const uint32_t sg2 = blockIdx.x / g_next;
                    has_nx = sg2 < (uint32_t)n_seg;
                    a_nx = part_args<T>(r, has_nx ? (int)sg2 : 0, k2, G, g_same);
                    vb_nx = blockIdx.x - sg2 * g_next;
                }
                accumulate_tiles<NET, LOSS, T, SKIP>(a, s_wa, vb, g_k, first, g, loss, n, a_nx, vb_nx, has_nx);
                warp_partial<NET>(g, loss, n, s_red[n_slots][warp]);
                static_assert(kMaxSeg <= 15, "numbered barriers 1 .. kMaxSeg");
                constexpr int NW = kFitThreads / 32;                            // the warps take turns as publishers
                // hand-over inside the block, one unit behind: warp (slot mod 8) publishes unit `slot` once the other seven
                // have parked its totals (numbered barrier slot + 1: they arrive without waiting, the publisher waits) —
                // it does so after its OWN next unit, so the wait is over before it starts and the stores of all but the
                // last unit are in L2 long before the owners look for them
                if (n_slots > 0 && warp == (n_slots - 1) % NW) {
                    bar_sync(n_slots, kFitThreads);
                    warp_publish<NET>(s_red[n_slots - 1], u - G, g_k, part, (size_t)kGsumLen * kLLBlocks, kLLBlocks, tag, lane);
                }
                if (u + G < n_units && warp != n_slots % NW) bar_arrive(n_slots + 1, kFitThreads);
            }
            ORP_TRACE(1);
            // the step's learning rate: Adam bias terms from a table that all threads refill together every 256 steps
            if (((barriers - 1u) & (kLrTable - 1u)) == 0u && tid < kLrTable) {
                const double ti = s_c[ST_ADAM_T] + 1.0 + (double)tid;
                s_num[tid] = sqrt(1.0 - pow(0.999, ti));
                s_den[tid] = 1.0 - pow(0.9, ti);
            }
            if (tid == 0) s_peer_ok = 1;
            __syncthreads();
            // the last unit of the step is on the critical path: all threads publish it (after the step's one block barrier)
            if (n_slots > 0)
                block_publish<NET>(s_red, n_slots, blockIdx.x, G, g_k, part, (size_t)kGsumLen * kLLBlocks, kLLBlocks, 1, tag, n_slots - 1);
            ORP_TRACE(2);
            if (tid == 32) {
                double lr = s_c[ST_LR];
                if (k == 0 && s_c[ST_USE_SCHED] != 0.0) lr = schedule_lr((int)s_c[ST_EPOCH], lr);
                s_c[ST_LR] = lr;
                s_c[ST_ADAM_T] = s_c[ST_ADAM_T] + 1.0;
                const unsigned int i = (barriers - 1u) & (kLrTable - 1u);
                s_lr_t = (float)(lr * s_num[i] / s_den[i]);
            }
            ORP_TRACE(3);

            // ---- column owners: column q belongs to block q mod G (one warp each).  The owner polls the g_k tagged
            //      entries of its column (no barrier: the tag says "this step"), adds them in the fixed order of
            //      fit_reduce_kernel (lane l: blocks l, l + 32, ...; then a shuffle tree) and publishes the total, again
            //      tagged, in slot [parity][rank][column] of every rank — its own memory when there is one GPU,
            //      plain stores over NVLink to the peers' mapped buffers otherwise.
            const unsigned int step_id = step0 + barriers;                    // tag of the published totals: same on every rank
            const unsigned int fpar = step_id & 1u;
            // (owners are taken from the TOP of the grid: when the tiles do not divide evenly the low-numbered blocks carry
            //  the extra tile and finish last — they should not also have a column to add up)
            // A block owns the columns q = G - 1 - b, + G, ...; its (column, segment) pairs are spread over its warps, so the
            // eight segments of a column are polled side by side.
            const int q_first = (int)(G - 1 - blockIdx.x);
            const int n_own = q_first < P + 2 ? (P + 2 - q_first + (int)G - 1) / (int)G : 0;
            for (int item = warp; item < n_own * n_seg; item += kFitThreads / 32) {
                const int j = item / n_seg, sv = item - j * n_seg, q = q_first + j * (int)G;
                const int qq = q < P ? q : ORP_FIT_MAX_PARAMS + (q - P);
                const double* col = part + ((size_t)sv * kGsumLen + qq) * kLLBlocks * 2;
                const unsigned long long t_start = globaltimer_ns();
                double t = 0.0;
                bool ok = true;
                for (uint32_t b0 = 0; b0 < g_k && ok; b0 += 160) {
                    // every retry re-reads ALL of this lane's entries at once (independent loads: one L2 round trip per
                    // retry), so the total is ready at most two round trips after the last block has published
                    uint4 v[5];
                    unsigned int spins = 0;
                    for (;;) {
                        bool ready = true;
#pragma unroll
                        for (int jj = 0; jj < 5; ++jj) v[jj] = ld_tagged(col + 2 * (size_t)min(b0 + 32 * jj + lane, g_k - 1));
#pragma unroll
                        for (int jj = 0; jj < 5; ++jj) ready = ready && v[jj].y == tag && v[jj].w == tag;
                        if (ready) break;
                        if (((++spins & 255u) == 0u) && (poll_expired(t_start) || ld_acquire_u32(abort_flag) != 0u)) { ok = false; break; }
                    }
#pragma unroll
                    for (int jj = 0; jj < 5; ++jj) t += b0 + 32 * jj + lane < g_k ? __hiloint2double((int)v[jj].z, (int)v[jj].x) : 0.0;
                }
                ok = __all_sync(0xffffffffu, ok);
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) t += __shfl_down_sync(0xffffffffu, t, o);
                if (lane == 0) {
                    s_seg[j][sv] = t;
                    if (!ok) { atomicExch(abort_flag, 1u); s_peer_ok = 0; }
                }
            }
            __syncthreads();
            if (s_peer_ok)
                for (int j = warp; j < n_own; j += kFitThreads / 32) {
                    const int q = q_first + j * (int)G;
                    const int qq = q < P ? q : ORP_FIT_MAX_PARAMS + (q - P);
                    double seg[kMaxSeg];
#pragma unroll
                    for (int sv = 0; sv < kMaxSeg; ++sv) seg[sv] = sv < n_seg ? s_seg[j][sv] : 0.0;
                    const double t = tree_sum<kMaxSeg>(seg, n_seg);            // the segments' totals, pairwise in segment order
                    if (lane < n_ranks) {
                        double* dst = (r.world > 0 ? r.peers[lane] + kPeerSlots : ll_final)
                                      + (((size_t)fpar * kPeerMaxWorld + my_rank) * kGsumLen + qq) * 2;
                        st_tagged(dst, t, step_id);
                    }
                }
            ORP_TRACE(4);
            // ---- every block: wait for the ranks' totals in its OWN GPU's memory and add them in rank order (same bits
            //      everywhere).  ONE thread per block watches one entry (a different column per block, so the watchers
            //      spread over the cache lines) with a back-off; a column total can only exist once every block has
            //      published, so when it shows up the other columns are at most a fraction of a microsecond behind and the
            //      per-entry polls below are short.  (148 x 99 threads spinning on 13 cache lines slowed the stragglers down.)
            if (tid == 0) {
                const int qg = (int)(blockIdx.x % (unsigned)(P + 2));
                const int qq = qg < P ? qg : ORP_FIT_MAX_PARAMS + (qg - P);
                const double* w = fin_mine + (((size_t)fpar * kPeerMaxWorld + (n_ranks - 1)) * kGsumLen + qq) * 2;
                const unsigned long long t_start = globaltimer_ns();
                unsigned int spins = 0;
                uint4 v = ld_tagged(w);
                while (v.y != step_id || v.w != step_id) {
                    if (((++spins & 255u) == 0u) && (poll_expired(t_start) || ld_acquire_u32(abort_flag) != 0u)) break;   // the polls below report it
                    v = ld_tagged(w);
                }
            }
            __syncthreads();
            for (int q = tid; q < P + 2; q += kFitThreads) {
                const int qq = q < P ? q : ORP_FIT_MAX_PARAMS + (q - P);
                const double* w0 = fin_mine + ((size_t)fpar * kPeerMaxWorld * kGsumLen + qq) * 2;
                double rk[kPeerMaxWorld];
                const unsigned long long t_start = globaltimer_ns();
#pragma unroll
                for (int s0 = 0; s0 < kPeerMaxWorld; s0 += 8) {                // eight ranks per pass, all their entries per retry
                    uint4 v[8];
                    unsigned int spins = 0;
                    if (s0 < n_ranks)
                        for (;;) {
                            bool ready = true;
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                if (s0 + j < n_ranks) v[j] = ld_tagged(w0 + (size_t)(s0 + j) * kGsumLen * 2);
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                if (s0 + j < n_ranks) ready = ready && v[j].y == step_id && v[j].w == step_id;
                            if (ready) break;
                            if (((++spins & 255u) == 0u) && (poll_expired(t_start) || ld_acquire_u32(abort_flag) != 0u)) {
                                atomicExch(abort_flag, 1u);
                                s_peer_ok = 0;
                                break;
                            }
                        }
#pragma unroll
                    for (int j = 0; j < 8; ++j) rk[s0 + j] = s0 + j < n_ranks ? __hiloint2double((int)v[j].z, (int)v[j].x) : 0.0;
                }
                s_g[qq] = tree_sum<kPeerMaxWorld>(rk, n_ranks);                // pairwise in rank order: same bits on every GPU, and
                                                                               // the continuation of the ranks' own segment trees
            }
            __syncthreads();
            ORP_TRACE(5);
            if (!s_peer_ok) { alive = false; break; }

            ORP_TRACE(6);
            // ---- Adam step on the block's own copy (identical in every block)
            const double inv_n = 1.0 / s_g[ORP_FIT_MAX_PARAMS + 1];
            if (tid < P) {
                const float grad = (float)(s_g[tid] * inv_n);
                float wn = s_w[tid];
                adam_core(m, v, wn, grad, s_lr_t);
                s_w[tid] = wn;
                if (tid >= N::W2 && tid < N::W2 + kH * kH) {
                    const int q = tid - N::W2;
                    s_w[kW2T + (q & 7) * kH + (q >> 3)] = wn;
                }
            }
            if (tid == 0) {
                const double loss_acc = (k == 0 ? 0.0 : s_c[ST_LOSS_ACC]) + s_g[ORP_FIT_MAX_PARAMS];
                const double cnt_acc = (k == 0 ? 0.0 : s_c[ST_COUNT_ACC]) + s_g[ORP_FIT_MAX_PARAMS + 1];
                s_c[ST_LOSS_ACC] = loss_acc;
                s_c[ST_COUNT_ACC] = cnt_acc;
                s_improved = 0;
                if (k == r.n_parts - 1) {                                  // epoch end: EarlyStopping bookkeeping (epoch_end())
                    const double cur = loss_acc / cnt_acc;
                    s_c[ST_LAST_LOSS] = cur;
                    int epoch = (int)s_c[ST_EPOCH];
                    double wait = s_c[ST_WAIT] + 1.0;
                    if (cur < s_c[ST_BEST]) { s_c[ST_BEST] = cur; wait = 0.0; s_improved = 1; s_c[ST_HAVE_BEST] = 1.0; }
                    s_c[ST_WAIT] = wait;
                    if (wait >= s_c[ST_PATIENCE] && epoch > 0) { s_c[ST_STOP] = 1.0; s_c[ST_STOPPED_EARLY] = 1.0; }
                    epoch += 1;
                    s_c[ST_EPOCH] = epoch;
                    s_c[ST_EPOCHS_RUN] = epoch;
                    s_c[ST_TOTAL_EPOCHS] += 1.0;
                    if (epoch >= (int)s_c[ST_EPOCHS]) s_c[ST_STOP] = 1.0;
                }
            }
            __syncthreads();
            if (s_improved && tid < P) s_best[tid] = s_w[tid];             // restore_best_weights=True
            ORP_TRACE(7);
        }
    }

    // ---- write back (block 0)
    __syncthreads();
    if (blockIdx.x == 0) {
        if (tid < P) {
            params[tid] = s_w[tid];
            state[ST_ADAM_M + tid] = m;
            state[ST_ADAM_V + tid] = v;
            state[ST_BEST_W + tid] = s_best[tid];
        }
        if (tid < 16) state[tid] = s_c[tid];
        for (int q = tid; q < kGsumLen; q += kFitThreads) gsum[q] = s_g[q];
        if (tid == 0) {
            *reinterpret_cast<unsigned int*>(partials + kScrHeader) = tag0 + barriers;
            if (r.world > 0) *reinterpret_cast<unsigned int*>(xbuf + kPeerHdr) = step0 + barriers;
        }
    }
}

// a-10 / a-11: roll-back values at prices@t and the hedge-weight sums.
constexpr int kPredictSums = 5;      // sum V, sum phi, sum psi, sum |y - V1|, sum |y - V1| / max(|y|, 1e-7)
constexpr int kPredS = 2;            // samples per thread and iteration (packed arithmetic: one FFMA2 serves both)
constexpr int kPredBlocksPerSM = 3;
template <int NET, typename T>
__global__ void __launch_bounds__(kFitThreads, kPredBlocksPerSM) fit_predict_kernel(const FitArgs a, const float* __restrict__ params,
                                                                  float* __restrict__ values_out, double* __restrict__ partials) {
    using N = Net<NET>;
    __shared__ __align__(16) float s_w[kSmemW];
    load_weights<NET>(s_w, params, threadIdx.x, kFitThreads);
    __syncthreads();
    const uint32_t s_wa = (uint32_t)__cvta_generic_to_shared(s_w);
    double sv = 0.0, sphi = 0.0, spsi = 0.0, sae = 0.0, sape = 0.0;
    // A block walks chunks of kPredS x 256 consecutive samples; a thread owns samples t and t + 256 of the chunk (full-line
    // warp accesses) and pushes both through the network together.  The loads of the thread's NEXT pair are issued before
    // the arithmetic of the current one: with three blocks per SM that is 768 x 2 x 2 samples (37 KB of fp32 columns) in
    // flight per SM, above the HBM latency-bandwidth product per SM (~35 KB); one sample per thread at four blocks per
    // SM (25 KB) left the pass latency-bound (long_scoreboard 72 % of the stall samples, 3.3 TB/s).
    constexpr uint64_t chunk = (uint64_t)kPredS * kFitThreads;
    const uint64_t stride = (uint64_t)gridDim.x * chunk;
    uint64_t i0 = (uint64_t)blockIdx.x * chunk + threadIdx.x;
    RawSample<NET, T> nxt[kPredS];
#pragma unroll
    for (int s = 0; s < kPredS; ++s) {
        nxt[s].s_t = nxt[s].s_t1 = T(0); nxt[s].y = nxt[s].f1 = nxt[s].f2 = 0.f;
        if (i0 + s * kFitThreads < a.count) load_raw_any<NET, T>(a, i0 + s * kFitThreads, nxt[s]);
    }
    for (; i0 < a.count; i0 += stride) {
        RawSample<NET, T> cur[kPredS];
#pragma unroll
        for (int s = 0; s < kPredS; ++s) {
            cur[s] = nxt[s];
            if (i0 + stride + s * kFitThreads < a.count) load_raw_any<NET, T>(a, i0 + stride + s * kFitThreads, nxt[s]);
        }
        float x[kPredS][N::IN], s_t[kPredS], s_t1[kPredS], phi[kPredS], psi[kPredS], v0[kPredS], tgt[kPredS];
        Acts<NET, kPredS> act;
#pragma unroll
        for (int s = 0; s < kPredS; ++s) {
            unpack_raw<NET, T, false>(a, cur[s], x[s], s_t1[s], tgt[s]);
            s_t[s] = x[s][0];
        }
        net_forward<NET, kPredS>(s_wa, x, s_t, a.b_t, act, phi, psi, v0);     // model.predict(X0): prices at t
#pragma unroll
        for (int s = 0; s < kPredS; ++s) {
            const uint64_t i = i0 + s * kFitThreads;
            if (i >= a.count) continue;
            const float v1 = fmaf(phi[s], s_t1[s], psi[s] * a.b_t1);          // model.predict(X1): prices at t+1
            if (a.skip) {                                      // opt-in: a path with a non-finite feature, price or target
                bool fin = isfinite(s_t1[s]) && isfinite(tgt[s]);             // stays marked (NaN) and out of the sums
#pragma unroll
                for (int q = 0; q < N::IN; ++q) fin = fin && isfinite(x[s][q]);
                if (!fin) {
                    if (values_out) values_out[i] = __int_as_float(0x7fc00000);
                    continue;
                }
            }
            if (values_out) values_out[i] = v0[s];
            sv += v0[s]; sphi += phi[s]; spsi += psi[s];
            if (a.target) {                                                   // metrics of model.evaluate(X1, ...), RP.py:215
                const float ae = fabsf(tgt[s] - v1);
                sae += ae;                                                    // 'mae'
                sape += ae / fmaxf(fabsf(tgt[s]), 1e-7f);                     // 'mape' (Keras: |y - v| / max(|y|, epsilon), x 100 on the host)
            }
        }
    }
    __shared__ double s_red[kFitThreads / 32][kPredictSums];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double q[kPredictSums] = {sv, sphi, spsi, sae, sape};
#pragma unroll
    for (int k = 0; k < kPredictSums; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) q[k] += __shfl_down_sync(0xffffffffu, q[k], o);
        if (lane == 0) s_red[warp][k] = q[k];
    }
    __syncthreads();
    if (threadIdx.x < kPredictSums) {
        double t = 0.0;
        for (int w = 0; w < kFitThreads / 32; ++w) t += s_red[w][threadIdx.x];
        partials[(size_t)blockIdx.x * kPredictSums + threadIdx.x] = t;
    }
}

// §8f-1: the per-path replication residual ("VaR" of get_phi_psi_VaR): y - V(prices at t+1), current weights.
template <int NET, typename T>
__global__ void __launch_bounds__(kFitThreads, 4) fit_residual_kernel(const FitArgs a, const float* __restrict__ params,
                                                                     float* __restrict__ resid) {
    using N = Net<NET>;
    __shared__ __align__(16) float s_w[kSmemW];
    load_weights<NET>(s_w, params, threadIdx.x, kFitThreads);
    __syncthreads();
    const uint32_t s_wa = (uint32_t)__cvta_generic_to_shared(s_w);
    for (uint64_t i = (uint64_t)blockIdx.x * kFitThreads + threadIdx.x; i < a.count; i += (uint64_t)gridDim.x * kFitThreads) {
        float x[1][N::IN], s_t, s_t1[1], phi[1], psi[1], v1[1];
        Acts<NET, 1> act;
        load_sample<NET, T>(a, i, x[0], s_t, s_t1[0]);
        net_forward<NET, 1>(s_wa, x, s_t1, a.b_t1, act, phi, psi, v1);          // features at t, prices at t+1 (X1)
        float res = a.target[i] - v1[0];
        if (a.skip) {                                      // opt-in: same rule as the fit — such a path has no residual
            bool fin = isfinite(s_t) && isfinite(s_t1[0]) && isfinite(a.target[i]);      // the rule of unpack_raw<SKIP>
#pragma unroll
            for (int q = 0; q < N::IN; ++q) fin = fin && isfinite(x[0][q]);
            if (!fin) res = __int_as_float(0x7fc00000);
        }
        resid[i] = res;
    }
}

// Fixed-order sum of the per-block partials: warp k owns quantity k, lane l adds blocks l, l + 32, ... in
// increasing order, then a shuffle tree — independent loads instead of one chain of 512 dependent ones.
__global__ void predict_reduce_kernel(const double* __restrict__ partials, int n_blocks, double* __restrict__ sums) {
    const int k = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (k >= kPredictSums) return;
    double t = 0.0;
    for (int b = lane; b < n_blocks; b += 32) t += partials[(size_t)b * kPredictSums + k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_down_sync(0xffffffffu, t, o);
    if (lane == 0) sums[k] = t;
}

// ---- Keras-faithful mini-batch epoch: one block, serial over batches ------------------------------

// Bijection of [0, 2^bits): odd multiply, xorshift, odd multiply — restated in oracle/fit.py::minibatch_perm.
__device__ __forceinline__ uint64_t perm_pow2(uint64_t x, int bits, uint64_t seed) {
    const uint64_t mask = bits >= 64 ? ~0ull : ((1ull << bits) - 1ull);
    const uint64_t a1 = (seed * 0x9E3779B97F4A7C15ull) | 1ull, a2 = ((seed ^ 0xD1B54A32D192ED03ull) * 0xBF58476D1CE4E5B9ull) | 1ull;
    x = (x * a1 + (seed >> 7)) & mask;
    x ^= x >> ((bits + 1) / 2);
    x = (x * a2) & mask;
    x ^= x >> ((bits + 2) / 3);
    return x & mask;
}

constexpr int kMiniThreads = 256;   // 255 registers per thread: the 97/122 gradient accumulators stay in registers

template <int NET, int LOSS, typename T>
__global__ void __launch_bounds__(kMiniThreads, 1) fit_minibatch_kernel(const FitArgs a, float* __restrict__ params,
                                                                        double* __restrict__ state, int batch_size,
                                                                        uint64_t epoch_seed) {
    using N = Net<NET>;
    if (state[ST_STOP] != 0.0) return;
    __shared__ __align__(16) float s_w[kSmemW];
    __shared__ float s_red[kMiniThreads / 32][N::P + 1];
    __shared__ float s_lr, s_t;
    __shared__ double s_loss;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    load_weights<NET>(s_w, params, tid, kMiniThreads);
    const uint32_t s_wa = (uint32_t)__cvta_generic_to_shared(s_w);
    float m = 0.f, v = 0.f;
    if (tid < N::P) { m = (float)state[ST_ADAM_M + tid]; v = (float)state[ST_ADAM_V + tid]; }
    if (tid == 0) {
        double lr = state[ST_LR];
        if (state[ST_USE_SCHED] != 0.0) lr = schedule_lr((int)state[ST_EPOCH], lr);
        state[ST_LR] = lr;
        s_lr = (float)lr; s_t = (float)state[ST_ADAM_T]; s_loss = 0.0;
    }
    __syncthreads();
    int bits = 0;
    while ((1ull << bits) < a.count) ++bits;
    double adam_t = state[ST_ADAM_T];
    const double lr = state[ST_LR];

    for (uint64_t b0 = 0; b0 < a.count; b0 += batch_size) {
        const int nb = (int)min((uint64_t)batch_size, a.count - b0);
        Grads<NET> g;
        g.zero();
        float loss = 0.f;
        for (int r = tid; r < nb; r += kMiniThreads) {
            // position b0 + r of the epoch's permutation (cycle-walk down from the enclosing power of two)
            uint64_t i = perm_pow2(b0 + r, bits, epoch_seed);
            while (i >= a.count) i = perm_pow2(i, bits, epoch_seed);
            float x[1][N::IN], s_t_, s_t1[1], phi[1], psi[1], vv[1], l, dv[1];
            Acts<NET, 1> act;
            load_sample<NET, T>(a, i, x[0], s_t_, s_t1[0]);
            net_forward<NET, 1>(s_wa, x, s_t1, a.b_t1, act, phi, psi, vv);
            loss_grad<LOSS>(a.target[i], vv[0], l, dv[0]);
            loss += l;
            net_backward<NET, 1>(s_wa, x, s_t1, a.b_t1, dv, act, g);
        }
#pragma unroll
        for (int k = 0; k < N::P; ++k) {
            float t = g.get(k);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) t += __shfl_down_sync(0xffffffffu, t, o);
            if (lane == 0) s_red[warp][k] = t;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) loss += __shfl_down_sync(0xffffffffu, loss, o);
        if (lane == 0) s_red[warp][N::P] = loss;
        __syncthreads();
        adam_t += 1.0;
        if (tid <= N::P) {
            float t = 0.f;
#pragma unroll
            for (int w = 0; w < kMiniThreads / 32; ++w) t += s_red[w][tid];
            if (tid < N::P) {
                const float grad = t / (float)nb;
                const float lr_t = (float)(lr * sqrt(1.0 - pow(0.999, adam_t)) / (1.0 - pow(0.9, adam_t)));
                float wn = s_w[tid];
                adam_core(m, v, wn, grad, lr_t);
                s_w[tid] = wn;
                if (tid >= N::W2 && tid < N::W2 + kH * kH) {       // keep the transposed copy in step
                    const int k = tid - N::W2;
                    s_w[kW2T + (k & 7) * kH + (k >> 3)] = wn;
                }
            } else {
                s_loss += (double)t;                  // Keras: epoch loss = sample-weighted mean of batch losses
            }
        }
        __syncthreads();
    }
    if (tid < N::P) { params[tid] = s_w[tid]; state[ST_ADAM_M + tid] = m; state[ST_ADAM_V + tid] = v; }
    __syncthreads();
    __threadfence_block();
    if (tid == 0) {
        state[ST_ADAM_T] = adam_t;
        epoch_end(N::P, s_w, state, s_loss / (double)a.count);
    }
}

static int fit_blocks(uint64_t count, int unroll) {
    const uint64_t tile = (uint64_t)kFitThreads * unroll;
    uint64_t want = (count + tile - 1) / tile;
    if (want < 1) want = 1;
    uint64_t cap = (uint64_t)sm_count();           // 256 threads x ~200 registers: one block per SM
    if (cap > (uint64_t)kLLBlocks) cap = kLLBlocks;
    return (int)(want < cap ? want : cap);
}

static int make_args(const orp_fit_batch* b, FitArgs& a, bool need_target) {
    ORP_REQUIRE(b, "null batch");
    ORP_REQUIRE(b->net == ORP_NET_EO || b->net == ORP_NET_RP, "unknown net %d", b->net);
    ORP_REQUIRE(b->loss == ORP_LOSS_MSE || b->loss == ORP_LOSS_PINBALL, "unknown loss %d", b->loss);
    ORP_REQUIRE(b->paths_dtype == ORP_F64 || b->paths_dtype == ORP_F32, "paths_dtype must be ORP_F64 or ORP_F32");
    ORP_REQUIRE(b->d_s_t && b->d_s_t1, "null price column");
    ORP_REQUIRE(!need_target || b->d_target, "null target");
    ORP_REQUIRE(b->count >= 1, "count must be >= 1");
    a.s_t = b->d_s_t; a.s_t1 = b->d_s_t1; a.feat1 = b->d_feat1; a.feat2 = b->d_feat2; a.target = b->d_target;
    a.count = b->count; a.inv_scale = b->inv_scale;
    a.b_t = (float)(b->bond_t * b->inv_scale); a.b_t1 = (float)(b->bond_t1 * b->inv_scale);
    ORP_REQUIRE(b->paths_dtype == ORP_F64 || b->inv_scale == 1.0,
                "fp32 price columns must be in network units already (inv_scale = 1): convert with orp_fit_scale_column");
    a.skip = (b->reserved & ORP_FIT_SKIP_NONFINITE) ? 1 : 0;
    a.n_seg = b->n_seg ? (int)b->n_seg : 1;
    a.g_blocks = 1;
    a.seg_stride = b->seg_stride;
    ORP_REQUIRE(a.n_seg >= 1 && a.n_seg <= ORP_FIT_MAX_SEGMENTS, "n_seg must be in [1, %d]", ORP_FIT_MAX_SEGMENTS);
    ORP_REQUIRE(a.n_seg == 1 || a.seg_stride >= a.count, "seg_stride must be >= count (segments may not overlap)");
    return ORP_OK;
}

}  // namespace orp

using namespace orp;

extern "C" int orp_fit_n_params(int net) {
    return net == ORP_NET_EO ? Net<ORP_NET_EO>::P : net == ORP_NET_RP ? Net<ORP_NET_RP>::P : -1;
}

extern "C" uint64_t orp_fit_scratch_bytes(void) { return kScrDoubles * sizeof(double) + 65536; }   // + tail for the optional phase trace

#define ORP_FIT_DISPATCH(KERNEL_CALL)                                                              \
    do {                                                                                           \
        const bool f64 = b->paths_dtype == ORP_F64;                                                \
        if (b->net == ORP_NET_EO) {                                                                \
            if (b->loss == ORP_LOSS_MSE) { if (f64) { KERNEL_CALL(ORP_NET_EO, ORP_LOSS_MSE, double); } else { KERNEL_CALL(ORP_NET_EO, ORP_LOSS_MSE, float); } } \
            else { if (f64) { KERNEL_CALL(ORP_NET_EO, ORP_LOSS_PINBALL, double); } else { KERNEL_CALL(ORP_NET_EO, ORP_LOSS_PINBALL, float); } }                 \
        } else {                                                                                   \
            if (b->loss == ORP_LOSS_MSE) { if (f64) { KERNEL_CALL(ORP_NET_RP, ORP_LOSS_MSE, double); } else { KERNEL_CALL(ORP_NET_RP, ORP_LOSS_MSE, float); } } \
            else { if (f64) { KERNEL_CALL(ORP_NET_RP, ORP_LOSS_PINBALL, double); } else { KERNEL_CALL(ORP_NET_RP, ORP_LOSS_PINBALL, float); } }                 \
        }                                                                                          \
    } while (0)

extern "C" int orp_fit_grad(const orp_fit_batch* b, const float* d_params, const double* d_state, double* d_gsum,
                            void* d_scratch, void* stream) {
    FitArgs a;
    int rc = make_args(b, a, true);
    if (rc) return rc;
    ORP_REQUIRE(d_params && d_state && d_gsum && d_scratch, "null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    a.g_blocks = fit_blocks(a.count, b->net == ORP_NET_EO ? kFitUnroll<ORP_NET_EO> : kFitUnroll<ORP_NET_RP>);
    const int grid = a.n_seg * a.g_blocks;                     // one block per unit (segment, virtual block)
    double* partials = reinterpret_cast<double*>(d_scratch);
#define CALL_GRAD(NET, LOSS, T) fit_grad_kernel<NET, LOSS, T, false, false><<<grid, kFitThreads, 0, st>>>(a, const_cast<float*>(d_params), const_cast<double*>(d_state), partials, d_gsum, 0)
#define CALL_GRAD_SKIP(NET, LOSS, T) fit_grad_kernel<NET, LOSS, T, false, true><<<grid, kFitThreads, 0, st>>>(a, const_cast<float*>(d_params), const_cast<double*>(d_state), partials, d_gsum, 0)
    if (a.skip) ORP_FIT_DISPATCH(CALL_GRAD_SKIP); else ORP_FIT_DISPATCH(CALL_GRAD);
#undef CALL_GRAD
#undef CALL_GRAD_SKIP
    ORP_LAUNCH_CHECK("fit_grad_kernel");
    fit_reduce_kernel<<<1, 256, 0, st>>>(partials, a.g_blocks, a.n_seg, d_state, d_gsum);
    ORP_LAUNCH_CHECK("fit_reduce_kernel");
    return ORP_OK;
}

static int fit_step_launch(const orp_fit_batch* b, float* d_params, double* d_state, double* d_gsum, void* d_scratch,
                           int flags, cudaStream_t st) {
    FitArgs a;
    int rc = make_args(b, a, true);
    if (rc) return rc;
    a.g_blocks = fit_blocks(a.count, b->net == ORP_NET_EO ? kFitUnroll<ORP_NET_EO> : kFitUnroll<ORP_NET_RP>);
    const int grid = a.n_seg * a.g_blocks;
    double* partials = reinterpret_cast<double*>(d_scratch);
#define CALL_STEP(NET, LOSS, T) fit_grad_kernel<NET, LOSS, T, true, false><<<grid, kFitThreads, 0, st>>>(a, d_params, d_state, partials, d_gsum, flags)
#define CALL_STEP_SKIP(NET, LOSS, T) fit_grad_kernel<NET, LOSS, T, true, true><<<grid, kFitThreads, 0, st>>>(a, d_params, d_state, partials, d_gsum, flags)
    if (a.skip) ORP_FIT_DISPATCH(CALL_STEP_SKIP); else ORP_FIT_DISPATCH(CALL_STEP);
#undef CALL_STEP
#undef CALL_STEP_SKIP
    ORP_LAUNCH_CHECK("fit_grad_kernel (single-launch step)");
    return ORP_OK;
}

extern "C" int orp_fit_step(const orp_fit_batch* b, float* d_params, double* d_state, double* d_gsum, void* d_scratch,
                            int flags, void* stream) {
    ORP_REQUIRE(d_params && d_state && d_gsum && d_scratch, "null pointer");
    return fit_step_launch(b, d_params, d_state, d_gsum, d_scratch, flags, (cudaStream_t)stream);
}

extern "C" int orp_fit_epochs(const orp_fit_batch* parts, int n_parts, float* d_params, double* d_state, double* d_gsum,
                              void* d_scratch, int n_epochs, void* stream) {
    ORP_REQUIRE(parts && d_params && d_state && d_gsum && d_scratch, "null pointer");
    ORP_REQUIRE(n_parts >= 1 && n_epochs >= 0, "n_parts must be >= 1 and n_epochs >= 0");
    for (int e = 0; e < n_epochs; ++e)
        for (int k = 0; k < n_parts; ++k) {
            const int flags = (k == 0 ? ORP_FIT_EPOCH_FIRST : 0) | (k == n_parts - 1 ? ORP_FIT_EPOCH_LAST : 0);
            int rc = fit_step_launch(parts + k, d_params, d_state, d_gsum, d_scratch, flags, (cudaStream_t)stream);
            if (rc) return rc;
        }
    return ORP_OK;
}

static int fit_run_launch(const orp_fit_batch* b, int n_parts, float* d_params, double* d_state, double* d_gsum,
                          void* d_scratch, int n_epochs, int rank, int world, double* const* d_peers, cudaStream_t st) {
    RunArgs r;
    int rc = make_args(b, r.a, true);
    if (rc) return rc;
    ORP_REQUIRE(d_params && d_state && d_gsum && d_scratch, "null pointer");
    ORP_REQUIRE(n_parts >= 1 && n_epochs >= 0, "n_parts must be >= 1 and n_epochs >= 0");
    if (n_epochs == 0) return ORP_OK;
    // the same split as the host layer's: ceil(count / n_parts) samples per batch, the last one possibly shorter
    uint64_t k = (uint64_t)n_parts < r.a.count ? (uint64_t)n_parts : r.a.count;
    r.per = (r.a.count + k - 1) / k;
    r.n_parts = (int)((r.a.count + r.per - 1) / r.per);
    r.n_epochs = n_epochs;
    r.rank = rank; r.world = world; r.peers = d_peers;
    const int unroll = b->net == ORP_NET_EO ? kFitUnroll<ORP_NET_EO> : kFitUnroll<ORP_NET_RP>;
    r.unroll = unroll;
    r.max_blocks_per_part = fit_blocks(r.per, unroll);
    r.a.g_blocks = r.max_blocks_per_part;
    // the units (segment, virtual block) of a step are dealt round-robin to the grid: at most one block per SM, all co-resident
    int grid = r.a.n_seg * r.max_blocks_per_part;
    if (grid > sm_count()) grid = sm_count();
    double* partials = reinterpret_cast<double*>(d_scratch);
    void* args[] = {(void*)&r, (void*)&d_params, (void*)&d_state, (void*)&partials, (void*)&d_gsum};
    const void* fn = nullptr;
#define PICK_RUN(NET, LOSS, T) fn = (const void*)fit_run_kernel<NET, LOSS, T, false>
#define PICK_RUN_SKIP(NET, LOSS, T) fn = (const void*)fit_run_kernel<NET, LOSS, T, true>
    if (r.a.skip) ORP_FIT_DISPATCH(PICK_RUN_SKIP); else ORP_FIT_DISPATCH(PICK_RUN);
#undef PICK_RUN
#undef PICK_RUN_SKIP
    const size_t dyn = (size_t)kMaxSeg * (kFitThreads / 32) * (size_t)(orp_fit_n_params(b->net) + 2) * sizeof(float);
    ORP_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
    ORP_CUDA(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kFitThreads), args, dyn, st));
    return ORP_OK;
}

extern "C" int orp_fit_run(const orp_fit_batch* b, int n_parts, float* d_params, double* d_state, double* d_gsum,
                           void* d_scratch, int n_epochs, void* stream) {
    return fit_run_launch(b, n_parts, d_params, d_state, d_gsum, d_scratch, n_epochs, 0, 0, nullptr, (cudaStream_t)stream);
}

extern "C" int orp_fit_run_peers(const orp_fit_batch* b, int n_parts, float* d_params, double* d_state, double* d_gsum,
                                 void* d_scratch, int n_epochs, int rank, int world, void* const* d_peer_buffers,
                                 void* stream) {
    ORP_REQUIRE(world >= 1 && world <= kPeerMaxWorld, "world must be in [1, %d]", kPeerMaxWorld);
    ORP_REQUIRE(rank >= 0 && rank < world, "rank must be in [0, world)");
    ORP_REQUIRE(d_peer_buffers, "null peer table");
    return fit_run_launch(b, n_parts, d_params, d_state, d_gsum, d_scratch, n_epochs, rank, world,
                          reinterpret_cast<double* const*>(d_peer_buffers), (cudaStream_t)stream);
}

extern "C" uint64_t orp_peer_buffer_bytes(void) { return (uint64_t)kPeerDoubles * sizeof(double); }

extern "C" int orp_peer_buffer_create(void** d_buf, unsigned char* handle64) {
    ORP_REQUIRE(d_buf && handle64, "null pointer");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handles are 64 bytes");
    void* p = nullptr;
    ORP_CUDA(cudaMalloc(&p, orp_peer_buffer_bytes()));
    ORP_CUDA(cudaMemset(p, 0, orp_peer_buffer_bytes()));
    ORP_CUDA(cudaDeviceSynchronize());
    cudaIpcMemHandle_t h;
    ORP_CUDA(cudaIpcGetMemHandle(&h, p));
    memcpy(handle64, &h, 64);
    *d_buf = p;
    return ORP_OK;
}

extern "C" int orp_peer_buffer_open(const unsigned char* handle64, void** d_ptr) {
    ORP_REQUIRE(handle64 && d_ptr, "null pointer");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    ORP_CUDA(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return ORP_OK;
}

extern "C" int orp_peer_buffer_close(void* d_ptr) {
    if (d_ptr) ORP_CUDA(cudaIpcCloseMemHandle(d_ptr));
    return ORP_OK;
}

extern "C" int orp_peer_buffer_destroy(void* d_buf) {
    if (d_buf) ORP_CUDA(cudaFree(d_buf));
    return ORP_OK;
}

extern "C" int orp_fit_update(int net, float* d_params, double* d_state, const double* d_gsum, int flags, void* stream) {
    const int P = orp_fit_n_params(net);
    ORP_REQUIRE(P > 0, "unknown net %d", net);
    ORP_REQUIRE(d_params && d_state && d_gsum, "null pointer");
    fit_update_kernel<<<1, 128, 0, (cudaStream_t)stream>>>(P, d_params, d_state, d_gsum, flags);
    ORP_LAUNCH_CHECK("fit_update_kernel");
    return ORP_OK;
}

extern "C" int orp_fit_init(double* d_state, void* stream) {
    ORP_REQUIRE(d_state, "null pointer");
    fit_init_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(d_state);
    ORP_LAUNCH_CHECK("fit_init_kernel");
    return ORP_OK;
}

extern "C" int orp_fit_begin(double* d_state, int epochs, int patience, int use_schedule, void* stream) {
    ORP_REQUIRE(d_state, "null pointer");
    ORP_REQUIRE(epochs >= 0 && patience >= 0, "epochs and patience must be >= 0");
    fit_begin_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(d_state, epochs, patience, use_schedule);
    ORP_LAUNCH_CHECK("fit_begin_kernel");
    return ORP_OK;
}

extern "C" int orp_fit_end(int net, float* d_params, double* d_state, void* stream) {
    const int P = orp_fit_n_params(net);
    ORP_REQUIRE(P > 0, "unknown net %d", net);
    ORP_REQUIRE(d_params && d_state, "null pointer");
    fit_end_kernel<<<1, 128, 0, (cudaStream_t)stream>>>(P, d_params, d_state);
    ORP_LAUNCH_CHECK("fit_end_kernel");
    return ORP_OK;
}

extern "C" int orp_fit_predict(const orp_fit_batch* b, const float* d_params, float* d_values_out, double* d_sums,
                               void* d_scratch, void* stream) {
    FitArgs a;
    int rc = make_args(b, a, false);
    if (rc) return rc;
    ORP_REQUIRE(d_params && d_sums && d_scratch, "null pointer");
    ORP_REQUIRE(a.n_seg == 1, "orp_fit_predict takes one segment per call");
    cudaStream_t st = (cudaStream_t)stream;
    uint64_t want = (a.count + kPredS * kFitThreads - 1) / (kPredS * kFitThreads);
    // the resident blocks of every SM (launch bounds), a whole number of waves: 444 blocks on B200;
    // their 5 partial sums each fit the 512 x kGsumLen doubles in front of the scratch header many times over
    uint64_t cap = (uint64_t)kPredBlocksPerSM * (uint64_t)sm_count();
    if (cap * kPredictSums > kScrHeader) cap = kScrHeader / kPredictSums;
    const int grid = (int)(want < cap ? want : cap);
    double* partials = reinterpret_cast<double*>(d_scratch);
    const bool f64 = b->paths_dtype == ORP_F64;
    if (b->net == ORP_NET_EO) {
        if (f64) fit_predict_kernel<ORP_NET_EO, double><<<grid, kFitThreads, 0, st>>>(a, d_params, d_values_out, partials);
        else     fit_predict_kernel<ORP_NET_EO, float><<<grid, kFitThreads, 0, st>>>(a, d_params, d_values_out, partials);
    } else {
        if (f64) fit_predict_kernel<ORP_NET_RP, double><<<grid, kFitThreads, 0, st>>>(a, d_params, d_values_out, partials);
        else     fit_predict_kernel<ORP_NET_RP, float><<<grid, kFitThreads, 0, st>>>(a, d_params, d_values_out, partials);
    }
    ORP_LAUNCH_CHECK("fit_predict_kernel");
    predict_reduce_kernel<<<1, 32 * kPredictSums, 0, st>>>(partials, grid, d_sums);
    ORP_LAUNCH_CHECK("predict_reduce_kernel");
    return ORP_OK;
}

extern "C" int orp_fit_residuals(const orp_fit_batch* b, const float* d_params, float* d_resid_out, void* stream) {
    FitArgs a;
    int rc = make_args(b, a, true);
    if (rc) return rc;
    ORP_REQUIRE(d_params && d_resid_out, "null pointer");
    ORP_REQUIRE(a.n_seg == 1, "orp_fit_residuals takes one segment per call");
    cudaStream_t st = (cudaStream_t)stream;
    uint64_t want = (a.count + kFitThreads - 1) / kFitThreads;
    const uint64_t cap = 4ull * (uint64_t)sm_count();          // four resident blocks per SM, whole waves
    const int grid = (int)(want < cap ? want : cap);
    const bool f64 = b->paths_dtype == ORP_F64;
    if (b->net == ORP_NET_EO) {
        if (f64) fit_residual_kernel<ORP_NET_EO, double><<<grid, kFitThreads, 0, st>>>(a, d_params, d_resid_out);
        else     fit_residual_kernel<ORP_NET_EO, float><<<grid, kFitThreads, 0, st>>>(a, d_params, d_resid_out);
    } else {
        if (f64) fit_residual_kernel<ORP_NET_RP, double><<<grid, kFitThreads, 0, st>>>(a, d_params, d_resid_out);
        else     fit_residual_kernel<ORP_NET_RP, float><<<grid, kFitThreads, 0, st>>>(a, d_params, d_resid_out);
    }
    ORP_LAUNCH_CHECK("fit_residual_kernel");
    return ORP_OK;
}

extern "C" int orp_fit_epoch_minibatch(const orp_fit_batch* b, float* d_params, double* d_state, int batch_size,
                                       uint64_t epoch_seed, void* stream) {
    FitArgs a;
    int rc = make_args(b, a, true);
    if (rc) return rc;
    ORP_REQUIRE(d_params && d_state, "null pointer");
    ORP_REQUIRE(batch_size >= 1, "batch_size must be >= 1");
    ORP_REQUIRE(a.n_seg == 1, "the mini-batch driver takes one segment");
    ORP_REQUIRE(!a.skip, "ORP_FIT_SKIP_NONFINITE is not available in the Keras-faithful mini-batch driver");
    cudaStream_t st = (cudaStream_t)stream;
#define CALL_MINI(NET, LOSS, T) fit_minibatch_kernel<NET, LOSS, T><<<1, kMiniThreads, 0, st>>>(a, d_params, d_state, batch_size, epoch_seed)
    ORP_FIT_DISPATCH(CALL_MINI);
#undef CALL_MINI
    ORP_LAUNCH_CHECK("fit_minibatch_kernel");
    return ORP_OK;
}

// ---- price column in network units: (float)((double)S * inv_scale), the cast Keras applies to its inputs ----------
namespace orp {
template <typename T>
__global__ void __launch_bounds__(256) scale_column_kernel(const T* __restrict__ in, uint64_t n, double inv_scale, float* __restrict__ out) {
    for (uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (uint64_t)gridDim.x * 256)
        out[i] = (float)((double)in[i] * inv_scale);
}
}  // namespace orp

extern "C" int orp_fit_scale_column(const void* d_in, int dtype, uint64_t count, double inv_scale, float* d_out, void* stream) {
    ORP_REQUIRE(d_in && d_out, "null pointer");
    ORP_REQUIRE(dtype == ORP_F64 || dtype == ORP_F32, "dtype must be ORP_F64 or ORP_F32");
    if (count == 0) return ORP_OK;
    uint64_t want = (count + 255) / 256;
    const uint64_t cap = 8ull * (uint64_t)sm_count();
    const int grid = (int)(want < cap ? want : cap);
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == ORP_F64) scale_column_kernel<double><<<grid, 256, 0, st>>>(reinterpret_cast<const double*>(d_in), count, inv_scale, d_out);
    else scale_column_kernel<float><<<grid, 256, 0, st>>>(reinterpret_cast<const float*>(d_in), count, inv_scale, d_out);
    ORP_LAUNCH_CHECK("scale_column_kernel");
    return ORP_OK;
}
