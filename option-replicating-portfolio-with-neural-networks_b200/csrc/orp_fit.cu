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

namespace orp {

// ---- state layout (indices into the double array) ------------------------------------------------
enum {
    ST_LR = 0, ST_ADAM_T, ST_EPOCH, ST_EPOCHS, ST_PATIENCE, ST_USE_SCHED, ST_WAIT, ST_BEST, ST_STOP,
    ST_STOPPED_EARLY, ST_LAST_LOSS, ST_EPOCHS_RUN, ST_HAVE_BEST, ST_TOTAL_EPOCHS, ST_LOSS_ACC, ST_COUNT_ACC,
    ST_ADAM_M = 16, ST_ADAM_V = ST_ADAM_M + ORP_FIT_MAX_PARAMS, ST_BEST_W = ST_ADAM_V + ORP_FIT_MAX_PARAMS,
    ST_END = ST_BEST_W + ORP_FIT_MAX_PARAMS
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

// Forward pass; returns V, optionally phi/psi and the activations needed by the backward pass.
template <int NET>
__device__ __forceinline__ float net_forward(const float* __restrict__ w, const float (&x)[Net<NET>::IN], float p0, float p1,
                                             float (&pre1)[kH], float (&h1)[kH], float (&pre2)[kH], float (&h2)[kH],
                                             float& phi, float& psi) {
    using N = Net<NET>;
#pragma unroll
    for (int j = 0; j < kH; ++j) {
        float acc = w[N::B1 + j];
#pragma unroll
        for (int i = 0; i < N::IN; ++i) acc = fmaf(x[i], w[N::W1 + i * kH + j], acc);
        pre1[j] = acc; h1[j] = lrelu(acc);
    }
#pragma unroll
    for (int j = 0; j < kH; ++j) {
        float acc = w[N::B2 + j];
#pragma unroll
        for (int i = 0; i < kH; ++i) acc = fmaf(h1[i], w[N::W2 + i * kH + j], acc);
        pre2[j] = acc; h2[j] = lrelu(acc);
    }
    float out[N::OUT];
#pragma unroll
    for (int o = 0; o < N::OUT; ++o) {
        float acc = w[N::B3 + o];
#pragma unroll
        for (int i = 0; i < kH; ++i) acc = fmaf(h2[i], w[N::W3 + i * N::OUT + o], acc);
        out[o] = acc;
    }
    phi = out[0];
    psi = NET == ORP_NET_EO ? 1.f - phi : out[N::OUT - 1];
    return fmaf(phi, p0, psi * p1);
}

// Backward pass of V = phi p0 + psi p1 for one sample, accumulating into g[P].
template <int NET>
__device__ __forceinline__ void net_backward(const float* __restrict__ w, const float (&x)[Net<NET>::IN], float p0, float p1,
                                             float dV, const float (&pre1)[kH], const float (&h1)[kH],
                                             const float (&pre2)[kH], const float (&h2)[kH], float (&g)[Net<NET>::P]) {
    using N = Net<NET>;
    float dout[N::OUT];
    if (NET == ORP_NET_EO) dout[0] = dV * (p0 - p1);          // psi = 1 - phi
    else { dout[0] = dV * p0; dout[N::OUT - 1] = dV * p1; }
    float d2[kH];
#pragma unroll
    for (int i = 0; i < kH; ++i) {
        float acc = 0.f;
#pragma unroll
        for (int o = 0; o < N::OUT; ++o) {
            g[N::W3 + i * N::OUT + o] = fmaf(h2[i], dout[o], g[N::W3 + i * N::OUT + o]);
            acc = fmaf(dout[o], w[N::W3 + i * N::OUT + o], acc);
        }
        d2[i] = acc * dlrelu(pre2[i]);
    }
#pragma unroll
    for (int o = 0; o < N::OUT; ++o) g[N::B3 + o] += dout[o];
    float d1[kH];
#pragma unroll
    for (int i = 0; i < kH; ++i) {
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < kH; ++j) {
            g[N::W2 + i * kH + j] = fmaf(h1[i], d2[j], g[N::W2 + i * kH + j]);
            acc = fmaf(d2[j], w[N::W2 + i * kH + j], acc);
        }
        d1[i] = acc * dlrelu(pre1[i]);
    }
#pragma unroll
    for (int j = 0; j < kH; ++j) {
        g[N::B2 + j] += d2[j];
        g[N::B1 + j] += d1[j];
#pragma unroll
        for (int i = 0; i < N::IN; ++i) g[N::W1 + i * kH + j] = fmaf(x[i], d1[j], g[N::W1 + i * kH + j]);
    }
}

// per-sample loss and dLoss/dV (without the 1/N of the batch mean)
template <int LOSS>
__device__ __forceinline__ void loss_grad(float y, float v, float& l, float& dv) {
    if (LOSS == ORP_LOSS_MSE) { const float e = v - y; l = e * e; dv = 2.f * e; }
    else { const float e = y - v; l = fmaxf(0.99f * e, (0.99f - 1.f) * e); dv = e > 0.f ? -0.99f : 0.01f; }   // RP.py:138-142
}

struct FitArgs {
    const void* s_t; const void* s_t1; const float* feat1; const float* feat2; const float* target;
    uint64_t count;
    float inv_scale_f; double inv_scale; float b_t, b_t1;
};

template <typename T>
__device__ __forceinline__ float load_price(const void* p, uint64_t i, double inv_scale) {
    return (float)((double)reinterpret_cast<const T*>(p)[i] * inv_scale);     // divide in fp64, then Keras' cast
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

constexpr int kFitThreads = 256;
constexpr int kGsum = ORP_FIT_MAX_PARAMS + 2;      // grads, loss sum, count

template <int P>
__device__ __forceinline__ void block_reduce_to(float (&g)[P], float loss, uint64_t n, double* __restrict__ dst) {
    __shared__ float s_red[kFitThreads / 32][P + 2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < P; ++k) {
        float v = g[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if (lane == 0) s_red[warp][k] = v;
    }
    float l = loss, c = (float)n;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { l += __shfl_down_sync(0xffffffffu, l, o); c += __shfl_down_sync(0xffffffffu, c, o); }
    if (lane == 0) { s_red[warp][P] = l; s_red[warp][P + 1] = c; }
    __syncthreads();
    for (int k = threadIdx.x; k < P + 2; k += kFitThreads) {
        double t = 0.0;
        for (int w = 0; w < kFitThreads / 32; ++w) t += (double)s_red[w][k];
        dst[k < P ? k : ORP_FIT_MAX_PARAMS + (k - P)] = t;
    }
}

template <int NET, int LOSS, typename T>
__global__ void __launch_bounds__(kFitThreads) fit_grad_kernel(const FitArgs a, const float* __restrict__ params,
                                                               const double* __restrict__ state, double* __restrict__ partials) {
    using N = Net<NET>;
    if (state[ST_STOP] != 0.0) return;
    __shared__ __align__(16) float s_w[ORP_FIT_MAX_PARAMS];
    if (threadIdx.x < N::P) s_w[threadIdx.x] = params[threadIdx.x];
    __syncthreads();

    float g[N::P];
#pragma unroll
    for (int k = 0; k < N::P; ++k) g[k] = 0.f;
    float loss = 0.f;
    uint64_t n = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * kFitThreads + threadIdx.x; i < a.count; i += (uint64_t)gridDim.x * kFitThreads) {
        float x[N::IN], s_t, s_t1, pre1[kH], h1[kH], pre2[kH], h2[kH], phi, psi, l, dv;
        load_sample<NET, T>(a, i, x, s_t, s_t1);
        const float v = net_forward<NET>(s_w, x, s_t1, a.b_t1, pre1, h1, pre2, h2, phi, psi);
        loss_grad<LOSS>(a.target[i], v, l, dv);
        loss += l;
        net_backward<NET>(s_w, x, s_t1, a.b_t1, dv, pre1, h1, pre2, h2, g);
        ++n;
    }
    block_reduce_to<N::P>(g, loss, n, partials + (size_t)blockIdx.x * kGsum);
}

// Fixed-order sum of the per-block partials: lane l adds blocks l, l+32, ... in increasing order, then a
// shuffle tree combines the 32 lane sums — the same association on every launch, whichever block finished last.
__global__ void __launch_bounds__(256) fit_reduce_kernel(const double* __restrict__ partials, int n_blocks,
                                                         const double* __restrict__ state, double* __restrict__ gsum) {
    if (state[ST_STOP] != 0.0) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int k = warp; k < kGsum; k += 8) {
        double t = 0.0;
        for (int b = lane; b < n_blocks; b += 32) t += partials[(size_t)b * kGsum + k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_down_sync(0xffffffffu, t, o);
        if (lane == 0) gsum[k] = t;
    }
}

__device__ __forceinline__ double schedule_lr(int epoch, double lr) {      // Replicating_Portfolio.py:128-136
    if (epoch < 100) return 1e-2;
    if (epoch < 200) return 1e-3;
    if (epoch < 400) return 5e-4;
    return lr;
}

// Adam (Keras: lr_t = lr sqrt(1-b2^t)/(1-b1^t); w -= lr_t m / (sqrt(v) + 1e-7)), fp32 like the Keras variables.
__device__ __forceinline__ void adam_apply(float* __restrict__ params, double* __restrict__ state, int k, float grad,
                                           float lr_t) {
    float m = (float)state[ST_ADAM_M + k], v = (float)state[ST_ADAM_V + k];
    m = 0.9f * m + (1.f - 0.9f) * grad;
    v = 0.999f * v + (1.f - 0.999f) * grad * grad;
    state[ST_ADAM_M + k] = m;
    state[ST_ADAM_V + k] = v;
    params[k] -= lr_t * m / (sqrtf(v) + 1e-7f);
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

// One optimiser step.  flags bit 0: first batch of the epoch (LearningRateScheduler.on_epoch_begin),
// bit 1: last batch (epoch loss = sample-weighted mean of the batch losses, then the callbacks).
__global__ void fit_update_kernel(int P, float* __restrict__ params, double* __restrict__ state,
                                  const double* __restrict__ gsum, int flags) {
    if (state[ST_STOP] != 0.0) return;
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

// a-10 / a-11: roll-back values at prices@t and the hedge-weight sums.
template <int NET, typename T>
__global__ void __launch_bounds__(kFitThreads) fit_predict_kernel(const FitArgs a, const float* __restrict__ params,
                                                                  float* __restrict__ values_out, double* __restrict__ partials) {
    using N = Net<NET>;
    __shared__ __align__(16) float s_w[ORP_FIT_MAX_PARAMS];
    if (threadIdx.x < N::P) s_w[threadIdx.x] = params[threadIdx.x];
    __syncthreads();
    double sv = 0.0, sphi = 0.0, spsi = 0.0, sae = 0.0;
    for (uint64_t i = (uint64_t)blockIdx.x * kFitThreads + threadIdx.x; i < a.count; i += (uint64_t)gridDim.x * kFitThreads) {
        float x[N::IN], s_t, s_t1, pre1[kH], h1[kH], pre2[kH], h2[kH], phi, psi;
        load_sample<NET, T>(a, i, x, s_t, s_t1);
        net_forward<NET>(s_w, x, s_t, a.b_t, pre1, h1, pre2, h2, phi, psi);
        const float v0 = fmaf(phi, s_t, psi * a.b_t);            // model.predict(X0): prices at t
        const float v1 = fmaf(phi, s_t1, psi * a.b_t1);          // model.predict(X1): prices at t+1
        if (values_out) values_out[i] = v0;
        sv += v0; sphi += phi; spsi += psi;
        if (a.target) sae += fabsf(a.target[i] - v1);            // 'mae' metric of model.evaluate(X1, ...)
    }
    __shared__ double s_red[kFitThreads / 32][4];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double q[4] = {sv, sphi, spsi, sae};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) q[k] += __shfl_down_sync(0xffffffffu, q[k], o);
        if (lane == 0) s_red[warp][k] = q[k];
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        double t = 0.0;
        for (int w = 0; w < kFitThreads / 32; ++w) t += s_red[w][threadIdx.x];
        partials[(size_t)blockIdx.x * 4 + threadIdx.x] = t;
    }
}

__global__ void predict_reduce_kernel(const double* __restrict__ partials, int n_blocks, double* __restrict__ sums) {
    if (threadIdx.x >= 4) return;
    double t = 0.0;
    for (int b = 0; b < n_blocks; ++b) t += partials[(size_t)b * 4 + threadIdx.x];
    sums[threadIdx.x] = t;
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
    __shared__ __align__(16) float s_w[ORP_FIT_MAX_PARAMS];
    __shared__ float s_red[kMiniThreads / 32][N::P + 1];
    __shared__ float s_lr, s_t;
    __shared__ double s_loss;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < N::P) s_w[tid] = params[tid];
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
        float g[N::P];
#pragma unroll
        for (int k = 0; k < N::P; ++k) g[k] = 0.f;
        float loss = 0.f;
        for (int r = tid; r < nb; r += kMiniThreads) {
            // position b0 + r of the epoch's permutation (cycle-walk down from the enclosing power of two)
            uint64_t i = perm_pow2(b0 + r, bits, epoch_seed);
            while (i >= a.count) i = perm_pow2(i, bits, epoch_seed);
            float x[N::IN], s_t_, s_t1, pre1[kH], h1[kH], pre2[kH], h2[kH], phi, psi, l, dv;
            load_sample<NET, T>(a, i, x, s_t_, s_t1);
            const float vv = net_forward<NET>(s_w, x, s_t1, a.b_t1, pre1, h1, pre2, h2, phi, psi);
            loss_grad<LOSS>(a.target[i], vv, l, dv);
            loss += l;
            net_backward<NET>(s_w, x, s_t1, a.b_t1, dv, pre1, h1, pre2, h2, g);
        }
#pragma unroll
        for (int k = 0; k < N::P; ++k) {
            float t = g[k];
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
                m = 0.9f * m + (1.f - 0.9f) * grad;
                v = 0.999f * v + (1.f - 0.999f) * grad * grad;
                s_w[tid] -= lr_t * m / (sqrtf(v) + 1e-7f);
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

static int fit_blocks(uint64_t count) {
    uint64_t want = (count + kFitThreads - 1) / kFitThreads;
    if (want < 1) want = 1;
    uint64_t cap = (uint64_t)sm_count();           // 256 threads x ~200 registers: one block per SM
    if (cap > 512) cap = 512;
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
    a.count = b->count; a.inv_scale = b->inv_scale; a.inv_scale_f = (float)b->inv_scale;
    a.b_t = (float)(b->bond_t * b->inv_scale); a.b_t1 = (float)(b->bond_t1 * b->inv_scale);
    return ORP_OK;
}

}  // namespace orp

using namespace orp;

extern "C" int orp_fit_n_params(int net) {
    return net == ORP_NET_EO ? Net<ORP_NET_EO>::P : net == ORP_NET_RP ? Net<ORP_NET_RP>::P : -1;
}

extern "C" uint64_t orp_fit_scratch_bytes(void) { return 512ull * kGsum * sizeof(double); }

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
    const int grid = fit_blocks(a.count);
    double* partials = reinterpret_cast<double*>(d_scratch);
#define CALL_GRAD(NET, LOSS, T) fit_grad_kernel<NET, LOSS, T><<<grid, kFitThreads, 0, st>>>(a, d_params, d_state, partials)
    ORP_FIT_DISPATCH(CALL_GRAD);
#undef CALL_GRAD
    ORP_LAUNCH_CHECK("fit_grad_kernel");
    fit_reduce_kernel<<<1, 256, 0, st>>>(partials, grid, d_state, d_gsum);
    ORP_LAUNCH_CHECK("fit_reduce_kernel");
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
    cudaStream_t st = (cudaStream_t)stream;
    uint64_t want = (a.count + kFitThreads - 1) / kFitThreads;
    const uint64_t cap = (uint64_t)(4 * sm_count() < 512 ? 4 * sm_count() : 512);
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
    predict_reduce_kernel<<<1, 32, 0, st>>>(partials, grid, d_sums);
    ORP_LAUNCH_CHECK("predict_reduce_kernel");
    return ORP_OK;
}

extern "C" int orp_fit_epoch_minibatch(const orp_fit_batch* b, float* d_params, double* d_state, int batch_size,
                                       uint64_t epoch_seed, void* stream) {
    FitArgs a;
    int rc = make_args(b, a, true);
    if (rc) return rc;
    ORP_REQUIRE(d_params && d_state, "null pointer");
    ORP_REQUIRE(batch_size >= 1, "batch_size must be >= 1");
    cudaStream_t st = (cudaStream_t)stream;
#define CALL_MINI(NET, LOSS, T) fit_minibatch_kernel<NET, LOSS, T><<<1, kMiniThreads, 0, st>>>(a, d_params, d_state, batch_size, epoch_seed)
    ORP_FIT_DISPATCH(CALL_MINI);
#undef CALL_MINI
    ORP_LAUNCH_CHECK("fit_minibatch_kernel");
    return ORP_OK;
}
