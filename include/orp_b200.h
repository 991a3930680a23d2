/*
 * orp_b200.h — C ABI of liborp_b200.so: the B200 (sm_100a) hot path of
 * ithakis/Option-Replicating-Portfolio-with-Neural-Networks.
 *
 * Plain pointers and sizes only; no torch / C++ types cross this boundary.  Pointers named d_* are
 * DEVICE pointers on the current CUDA device; everything else is host memory.  `stream` is a
 * cudaStream_t passed as void* (NULL = default stream).  All device work is enqueued on `stream`
 * and NOT synchronised.  The caller owns every buffer; nothing is retained after a call returns.
 *
 * Error convention: every function returns 0 on success and a negative code on failure
 * (ORP_E_*); orp_last_error() returns the message of the calling thread's last failure.  The
 * Python host layer turns ORP_E_VALUE into ValueError (what SciPy raises for a bad d / m,
 * scipy/stats/_qmc.py:1771-1774, 1908-1915) and everything else into RuntimeError.
 *
 * The reference is pure Python, so "the reference interface each entry point replaces" is the
 * Python/SciPy/Keras call named in the comment above each declaration (file:line under
 * /root/reference unless it starts with scipy/).
 */
#ifndef ORP_B200_H
#define ORP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORP_OK 0
#define ORP_E_VALUE (-1)   /* bad argument (-> ValueError) */
#define ORP_E_CUDA (-2)    /* CUDA runtime / launch failure (-> RuntimeError) */
#define ORP_E_STATE (-3)   /* wrong device / missing set-up (-> RuntimeError) */

#define ORP_SOBOL_BITS 30      /* scipy default, scipy/stats/_qmc.py:1780-1781 */
#define ORP_SOBOL_MAXDIM 21201 /* scipy/stats/_sobol.pyi */
#define ORP_SV_WORDS 32        /* packed engine: words 0..29 direction numbers, word 30 = 0, word 31 = shift */
#define ORP_SV_DPAD(d) (((d) + 31) & ~31)   /* dimensions padded to a multiple of 32 */

int orp_version(void);
const char* orp_last_error(void);

/* ---------------------------------------------------------------- a-1  Sobol set-up (host, integer) */

/* Replaces scipy.stats._sobol._initialize_v (called from qmc.Sobol.__init__, scipy/stats/_qmc.py:1793-1794
 * <- Replicating_Portfolio.py:55).  poly[d], vinit[d][18]: Joe-Kuo data (SciPy ships them in
 * _sobol_direction_numbers.npz).  v_out[d][bits]. */
int orp_sobol_direction_numbers(const int64_t* poly, const int64_t* vinit, int d, int bits, uint32_t* v_out);

/* Replaces qmc.Sobol._scramble + _cscramble (scipy/stats/_qmc.py:1812-1828): digital shift from
 * shift_bits[d][bits] (LSB first) and left matrix scramble with ltm_bits[d][bits][bits] (lower triangle
 * used, unit diagonal forced).  The random bits come from numpy.random.default_rng(seed).integers(2, ...)
 * exactly as SciPy draws them.  sv_out[d][bits], shift_out[d]. */
int orp_sobol_scramble(int d, int bits, const uint32_t* v, const uint32_t* shift_bits, const uint32_t* ltm_bits,
                       uint32_t* sv_out, uint32_t* shift_out);

/* Device-friendly packing (no reference counterpart; derived data): WORD-MAJOR sv32_out[32][ORP_SV_DPAD(d)]
 * (word b of dimension k at sv32_out[b * dpad + k], so 32 lanes working on 32 consecutive dimensions read one
 * 128-byte line) and the per-lane Gray-code table tlane_out[d][32][2] used by the warp decomposition of the
 * skip-ahead. */
int orp_sobol_pack(int d, int bits, const uint32_t* sv, const uint32_t* shift, uint32_t* sv32_out, uint32_t* tlane_out);

/* ---------------------------------------------------------------- a-2  Sobol points (device) */

/* Replaces qmc.Sobol.random_base2 / random / fast_forward -> _draw (scipy/stats/_qmc.py:1862-1917,
 * 1932-1957 <- Replicating_Portfolio.py:56): integer points x_i, i in [start, start+count), for all d
 * dimensions; u = x * 2^-30.  layout 0: d_out[count][d] (SciPy's), layout 1: d_out[d][count]. */
int orp_sobol_points(const uint32_t* d_sv32, const uint32_t* d_tlane, int d, uint64_t start, uint64_t count,
                     int layout, uint32_t* d_out, void* stream);

/* a-3: scipy.stats.norm.ppf -> scipy.special.ndtri (Replicating_Portfolio.py:57) on integer points:
 * d_z[i] = ndtri(d_x[i] * 2^-30), ndtri(0) = -inf.  Test/inspection entry; the path kernels inline it. */
int orp_ndtri_u32(const uint32_t* d_x, uint64_t n, double* d_z, void* stream);

/* ---------------------------------------------------------------- a-4/5/8  paths (device) */

enum { ORP_MODEL_GBM_LOG = 0, ORP_MODEL_GBM_ARITH = 1, ORP_MODEL_SV = 2, ORP_MODEL_LINEAR = 3 };
enum { ORP_ORDER_INCREMENTAL = 0, ORP_ORDER_BRIDGE = 1 };
enum { ORP_F64 = 0, ORP_F32 = 1 };

typedef struct orp_path_params {
    int32_t model;       /* ORP_MODEL_* */
    int32_t ordering;    /* ORP_ORDER_* */
    int32_t n_steps;     /* D; the engines have d = D + 1 dimensions, step t uses dimension t, dimension 0 unused */
    int32_t keep_every;  /* the reference's `reduction` (Replicating_Portfolio.py:92-93): column t is stored iff t % keep_every == 0 */
    int32_t out_dtype;   /* ORP_F64 / ORP_F32 (arithmetic is fp64 either way) */
    int32_t sobol_dim;   /* d of the engines (the reference's n_time_steps); 0 means n_steps + 1 */
    double s0;           /* Y / S0 */
    double mu;           /* drift: mu (Replicating_Portfolio.py:65, :287) or r (European Options.ipynb cell 6) */
    double sigma;        /* GBM volatility */
    double dt;           /* fine-grid step */
    double v0, a, b, c;  /* stochastic volatility: s0, a, b, c of Replicating_Portfolio.py:244-249, 285 */
} orp_path_params;

/* Replaces sobol_norm + the Euler loops + np.exp + the column slice:
 *   GBM_ARITH  Replicating_Portfolio.py:60-65, 93
 *   GBM_LOG    European Options.ipynb cell 6 (raw 156-165) and cell 7 (raw 205-207)
 *   SV         Replicating_Portfolio.py:274-289, 93   (second engine d_sv32_b/d_tlane_b drives the volatility)
 *   LINEAR     Replicating_Portfolio.py:72-76 (mortality intensity: mu = c, sigma = ita, s0 = l0)
 * Output is TIME-MAJOR: d_out[col * ld + (i - start)], col = t / keep_every, 0 <= col <= D / keep_every, ld >= count.
 * d_bridge: device schedule from orp_bridge_schedule (ordering = BRIDGE) or NULL. */
int orp_paths(const orp_path_params* p, const uint32_t* d_sv32, const uint32_t* d_tlane,
              const uint32_t* d_sv32_b, const uint32_t* d_tlane_b, const void* d_bridge,
              uint64_t start, uint64_t count, void* d_out, uint64_t ld, void* stream);

/* Brownian-bridge schedule for D steps (NEW feature, no reference counterpart; specification in
 * oracle/bridge.py).  Host function: fills sched_out (orp_bridge_schedule_bytes(D) bytes) which the caller
 * copies to the device. */
uint64_t orp_bridge_schedule_bytes(int n_steps);
int orp_bridge_schedule(int n_steps, void* sched_out);

/* ---------------------------------------------------------------- a-6/7  payoff + reductions (device) */

enum { ORP_PAYOFF_CALL = 0, ORP_PAYOFF_PUT = 1, ORP_PAYOFF_FLOOR = 2 };
#define ORP_MAX_OPTIONS 64

/* Replaces the payoff and summary block: CALL/PUT European Options.ipynb cell 8 (raw 328-329, 344, 347),
 * FLOOR Replicating_Portfolio.py:88-89, means :185.  For option j: sum_i payoff(S[cols[j]][i], strikes[j])
 * and the out-of-the-money count (CALL/PUT: payoff <= 0; FLOOR: S < s0).
 * d_paths time-major [n_cols][ld], dtype as written by orp_paths; every cols[j] must lie in [0, n_cols) (ORP_E_VALUE
 * otherwise: the reference's indexing would raise IndexError).  d_sums[n_opt], d_otm[n_opt] (fp64).
 * d_values_out (optional, fp32 [count]): payoff of option 0 times value_scale — values[:, -1] of
 * Replicating_Portfolio.py:184 / European Options.ipynb cell 13.
 * d_scratch: orp_payoff_scratch_bytes(n_opt) bytes of device scratch (per-block partial sums; the two-stage
 * fixed-order reduction makes the result independent of scheduling).
 * Options are processed in groups of 8 in the order given: list the options of one maturity column next to each other
 * (a strike ladder) and a group reads its column once; when every group of a multi-option call reads ONE column the
 * kernel sorts the group's strikes and does one accumulation per path instead of one per (path, option).  Sums agree
 * between the forms to fp64 rounding (1e-12), out-of-the-money counts exactly.  A NaN strike is ORP_E_VALUE. */
uint64_t orp_payoff_scratch_bytes(int n_opt);
int orp_payoff_reduce(const void* d_paths, int dtype, uint64_t ld, uint64_t count, int kind, int n_opt,
                      const double* strikes, const int32_t* cols, int32_t n_cols, double s0, double value_scale,
                      float* d_values_out, double* d_sums, double* d_otm, void* d_scratch, void* stream);

/* Fused generate + payoff: the paths are NEVER stored (north_star: "keeps them in registers/shared memory when the
 * payoff is fused").  Replaces, in one kernel, sobol_norm + the Euler loop (the model's lines as for orp_paths) AND the
 * payoff block (European Options.ipynb cell 8 raw 328-329, 344, 347; Replicating_Portfolio.py:88-89) for up to
 * ORP_MAX_OPTIONS (strike, maturity) pairs on ONE path set: option j pays off at time step steps[j] in [0, n_steps]
 * (a fine-grid step, not a kept column: keep_every and out_dtype of `p` are ignored, ordering must be incremental).
 * d_sums[j] = sum_i payoff(S_i(steps[j]), strikes[j]), d_otm[j] = out-of-the-money count (FLOOR: S < s0), as
 * orp_payoff_reduce defines them; results agree with orp_paths + orp_payoff_reduce to summation order (1e-12).
 * d_scratch: orp_paths_payoff_scratch_bytes(n_opt) bytes (per-block partial sums, fixed-order second pass). */
uint64_t orp_paths_payoff_scratch_bytes(int n_opt);
int orp_paths_payoff(const orp_path_params* p, const uint32_t* d_sv32, const uint32_t* d_tlane,
                     const uint32_t* d_sv32_b, const uint32_t* d_tlane_b, uint64_t start, uint64_t count,
                     int kind, int n_opt, const double* strikes, const int32_t* steps, double s0,
                     double* d_sums, double* d_otm, void* d_scratch, void* stream);

/* No counterpart in the reference.  Same computation as orp_payoff_reduce (results equal to fp64 rounding of the
 * sums, counts exactly) for callers that cannot
 * allocate device memory when they price — e.g. a service that gives all of the GPU to fixed arenas at start-up
 * and bans the runtime allocator afterwards, or a plug-in inside a framework whose allocator it may not use — and
 * therefore cannot supply d_scratch.  The per-block partial sums go to a fixed 128 KiB array inside the library
 * image instead (resident on every GPU that loads the library, until the process exits).  The caller still
 * passes inputs and outputs.  Limits and rules:
 *   - at most ORP_PAYOFF_NOALLOC_MAX_OPTIONS options per call; more is refused with ORP_E_VALUE (split the
 *     options over several calls, or use orp_payoff_reduce);
 *   - the first call on each device looks up the array's device address (no allocation); make that first call
 *     before starting any stream capture on that device;
 *   - all calls share the one array: two calls whose kernels may run at the same time on a GPU (different
 *     streams, different host threads) race on it — serialise them.  Nothing has to be cleared between calls;
 *   - the entry point itself never calls cudaMalloc / cudaFree or any allocating routine. */
#define ORP_PAYOFF_NOALLOC_MAX_OPTIONS 8
int orp_payoff_reduce_noalloc(const void* d_paths, int dtype, uint64_t ld, uint64_t count, int kind, int n_opt,
                              const double* strikes, const int32_t* cols, int32_t n_cols, double s0, double value_scale,
                              float* d_values_out, double* d_sums, double* d_otm, void* stream);

/* ---------------------------------------------------------------- a-9/10/11  replicating-portfolio fit (device) */

enum { ORP_NET_EO = 0, ORP_NET_RP = 1 };       /* 1->8->8->1 (psi = 1 - phi), 97 params; 3->8->8->2, 122 params */
enum { ORP_LOSS_MSE = 0, ORP_LOSS_PINBALL = 1 }; /* 'mse'; quantile_loss q = .99 (Replicating_Portfolio.py:138-142) */
#define ORP_FIT_MAX_PARAMS 128
#define ORP_FIT_STATE_DOUBLES 1024

int orp_fit_n_params(int net);

/* One time step of the backward loop (Replicating_Portfolio.py:195-201; European Options.ipynb cell 13 raw 3551-3558).
 * Feature/price columns are read in place from the time-major path array and cast to fp32 as Keras does. */
/* orp_fit_batch.reserved flag (no counterpart in the reference, opt-in): a sample whose features, prices or target
 * are not finite takes no part in the fit (no gradient, no loss, not counted), gets NaN as its rolled-back value and
 * is left out of the predict sums.  With the reference's arithmetic Euler step a single exact-zero Sobol coordinate
 * (they exist from 2^22 paths on) otherwise turns the whole fit into NaN. */
#define ORP_FIT_SKIP_NONFINITE 1

/* Segments (no counterpart in the reference, which runs in one process): a batch may consist of n_seg equally long
 * pieces of `count` samples, piece v starting seg_stride * v samples after the base pointers.  Every piece is reduced
 * on its own (fp32 per thread, then fp64 in a fixed order) and the pieces' totals are added pairwise in piece order.
 * With the paths cut into V = 8 virtual shards of the GLOBAL index space, a rank that holds L = V / world of them and
 * passes them as L segments produces — after the pairwise sum over ranks — exactly the bits one GPU produces with 8
 * segments: the fitted weights do not depend on the number of GPUs (1, 2, 4 or 8). */
#define ORP_FIT_MAX_SEGMENTS 8

typedef struct orp_fit_batch {
    int32_t net;            /* ORP_NET_* */
    int32_t loss;           /* ORP_LOSS_* */
    int32_t paths_dtype;    /* dtype of d_s_t / d_s_t1; ORP_F32 columns must already be in network units (inv_scale = 1, see orp_fit_scale_column) */
    int32_t reserved;       /* flags: ORP_FIT_SKIP_NONFINITE (0 = the reference's behaviour: NaN / inf propagate) */
    uint64_t count;         /* local paths */
    const void* d_s_t;      /* S column t   (feature, and price for the roll-back) */
    const void* d_s_t1;     /* S column t+1 (price the fit is scored on) */
    const float* d_feat1;   /* RP net only: N_t / N per path, or NULL (= 1) */
    const float* d_feat2;   /* RP net only: lambda_t per path, or NULL (= 0) */
    const float* d_target;  /* values[:, t+1] */
    double inv_scale;       /* 1 / ADJUSTMENT_FACTOR applied to prices (EO: 1/S0; RP: 1) */
    double bond_t, bond_t1; /* B_t, B_{t+1} (Replicating_Portfolio.py:68) */
    uint32_t n_seg;         /* segments (0 or 1: one piece of `count` samples); at most ORP_FIT_MAX_SEGMENTS */
    uint32_t reserved2;
    uint64_t seg_stride;    /* samples between the starts of consecutive segments (>= count) */
} orp_fit_batch;

uint64_t orp_fit_scratch_bytes(void);

/* Scaled mode, stage 1 of one epoch: fused forward + backward over all local paths; writes the fixed-order
 * reduced vector d_gsum[ORP_FIT_MAX_PARAMS + 2] = (sum of per-sample gradients, sum of per-sample losses, count)
 * in fp64.  Multi-GPU: the caller all-reduces d_gsum (NCCL, sum) between stage 1 and stage 2.
 * No-op when the control block says the fit has stopped.  Replaces the inside of keras Model.fit
 * (Replicating_Portfolio.py:211) for one optimiser step. */
int orp_fit_grad(const orp_fit_batch* b, const float* d_params, const double* d_state, double* d_gsum,
                 void* d_scratch, void* stream);

/* Stage 2: Adam step (Keras formula, eps 1e-7), epoch loss, LearningRateScheduler (Replicating_Portfolio.py:128-136)
 * and EarlyStopping(monitor='loss', restore_best_weights=True) bookkeeping (:174, :206), all on the device so the
 * host can enqueue epochs without synchronising.  d_state: ORP_FIT_STATE_DOUBLES doubles, layout in orp_fit.cu.
 * An epoch may be split into several batches (one grad + update pair each, as Keras does with batch_size):
 * flags bit 0 = first batch of the epoch, bit 1 = last batch (the epoch-end callbacks run then). */
#define ORP_FIT_EPOCH_FIRST 1
#define ORP_FIT_EPOCH_LAST 2
int orp_fit_update(int net, float* d_params, double* d_state, const double* d_gsum, int flags, void* stream);

/* Single-GPU form of one optimiser step: orp_fit_grad + the fixed-order reduction + orp_fit_update in ONE launch
 * (the thread block that finishes last reduces and updates; results are bit-identical to the two-stage form).
 * Same place in Model.fit's inner loop (Replicating_Portfolio.py:211).  d_gsum receives the reduced vector. */
int orp_fit_step(const orp_fit_batch* b, float* d_params, double* d_state, double* d_gsum, void* d_scratch,
                 int flags, void* stream);

/* Enqueues n_epochs epochs of n_parts single-launch steps each (parts[k] = the k-th batch of every epoch, flags
 * set as above) without returning to the caller in between: the epoch loop of Model.fit (Replicating_Portfolio.py:211).
 * Epochs after the stop flag is raised are no-ops on the device. */
int orp_fit_epochs(const orp_fit_batch* parts, int n_parts, float* d_params, double* d_state, double* d_gsum,
                   void* d_scratch, int n_epochs, void* stream);

/* The whole epoch loop of Model.fit (Replicating_Portfolio.py:211) in ONE cooperative launch, single GPU: every segment
 * of the batch is cut into n_parts contiguous ranges of ceil(count / n_parts) samples, and optimiser step k takes range k
 * of every segment; a grid of at most one thread block per SM stays resident; there is NO grid-wide barrier: the blocks
 * hand their totals to per-column owner blocks through tagged entries, and every block applies the same fixed-order
 * sums and Adam step to its own copy of the weights.  Runs at most n_epochs epochs and returns to the host as soon as
 * EarlyStopping fires.  Results are bit-identical to orp_fit_grad -> orp_fit_update (same batch, same segments).
 * d_scratch (orp_fit_scratch_bytes()) must be ZERO-FILLED once before its first use and then left alone between
 * calls: blocks hand their totals to each other through entries tagged with a step counter kept in it (no grid-wide
 * barrier).  Needs a device that supports cooperative launches (every B200 does): all blocks must be co-resident. */
int orp_fit_run(const orp_fit_batch* b, int n_parts, float* d_params, double* d_state, double* d_gsum,
                void* d_scratch, int n_epochs, void* stream);

/* Multi-GPU form of orp_fit_run (one process per GPU, paths sharded by Sobol index range; no counterpart in the
 * reference, which is single-process).  The ranks' [gradient, loss, count] vectors are exchanged INSIDE the
 * persistent kernel: the owner block of each vector column stores its column total into every rank's exchange buffer
 * over NVLink, the step number riding inside every 8-byte element; all blocks poll their own GPU's buffer and add the
 * ranks' totals pairwise in rank order, so every GPU applies the same bits (and, with 8 / world segments per rank, the
 * bits of a one-GPU run).  Every rank must call with the same n_parts / n_epochs / n_seg and an equal local sample
 * count.  All ranks must be on one host with peer access between their GPUs; a hand-over that does not complete within
 * 60 s raises the abort flag (word ST_ABORT of the state block) instead of hanging.
 * d_peer_buffers: DEVICE array of `world` pointers, entry r = rank r's buffer as mapped into this process.
 * Exchange buffers are the one thing this library allocates itself (CUDA IPC needs a cudaMalloc'ed base):
 *   orp_peer_buffer_create  cudaMalloc + zero + IPC handle (64 bytes) to send to the other ranks
 *   orp_peer_buffer_open    maps another rank's buffer (cudaIpcOpenMemHandle); orp_peer_buffer_close unmaps it
 *   orp_peer_buffer_destroy frees this rank's buffer. */
uint64_t orp_peer_buffer_bytes(void);
int orp_peer_buffer_create(void** d_buf, unsigned char* handle64);
int orp_peer_buffer_open(const unsigned char* handle64, void** d_ptr);
int orp_peer_buffer_close(void* d_ptr);
int orp_peer_buffer_destroy(void* d_buf);
int orp_fit_run_peers(const orp_fit_batch* b, int n_parts, float* d_params, double* d_state, double* d_gsum,
                      void* d_scratch, int n_epochs, int rank, int world, void* const* d_peer_buffers, void* stream);

/* A price column in network units: d_out[i] = (float)((double)d_in[i] * inv_scale) — the division by ADJUSTMENT_FACTOR
 * (European Options.ipynb cell 13 raw 3551-3554) followed by the cast Keras applies to float64 inputs.  The fit kernels
 * do this themselves on the fly; a caller that runs many epochs on the same two columns converts each column once,
 * passes the fp32 copies with paths_dtype = ORP_F32, inv_scale = 1 (and bond_t, bond_t1 already multiplied by the
 * scale) and gets the same bits with 12 instead of 20 bytes per sample and no fp64 arithmetic in the epoch loop. */
int orp_fit_scale_column(const void* d_in, int dtype, uint64_t count, double inv_scale, float* d_out, void* stream);

/* Starts a Model.fit call: resets the early-stopping state, keeps weights and Adam moments (they persist across
 * time steps, Replicating_Portfolio.py:203-211).  use_schedule = 1 for the first fit. */
int orp_fit_begin(double* d_state, int epochs, int patience, int use_schedule, void* stream);
/* Initialises weights-independent optimiser state (Adam moments, step, lr = 1e-3). */
int orp_fit_init(double* d_state, void* stream);
/* Ends a fit: restores the best weights when the stop fired early (TF 2.x semantics). */
int orp_fit_end(int net, float* d_params, double* d_state, void* stream);

/* Replaces model.predict(X0) / get_phi_psi_VaR means (Replicating_Portfolio.py:212, 103-115, 229-235):
 * d_values_out[i] = phi_i * S_t * inv_scale + psi_i * B_t * inv_scale (fp32), and
 * d_sums[5] = (sum V, sum phi, sum psi, sum |y - V1|, sum |y - V1| / max(|y|, 1e-7)) with y = target and
 * V1 = V(prices at t+1): the last two are the 'mae' and 'mape' / 100 metrics of model1.evaluate (Replicating_Portfolio.py:215)
 * before the division by the path count. */
int orp_fit_predict(const orp_fit_batch* b, const float* d_params, float* d_values_out, double* d_sums,
                    void* d_scratch, void* stream);

/* Keras-faithful mode: ONE epoch of mini-batch training (batch_size samples per Adam step, visiting samples in
 * the order perm(epoch_seed)), serial by construction, run by a single thread block.  Replaces Model.fit's inner
 * loop with batch_size=512, shuffle=True (Replicating_Portfolio.py:211). */
int orp_fit_epoch_minibatch(const orp_fit_batch* b, float* d_params, double* d_state, int batch_size,
                            uint64_t epoch_seed, void* stream);

/* ---------------------------------------------------------------- §8f-1  per-step risk metrics (device) */

/* d_resid_out[i] = target_i - (phi_i * S_{t+1} + psi_i * B_{t+1}) in network units: the array the reference calls VaR
 * (get_phi_psi_VaR, Replicating_Portfolio.py:120), computed with the weights as they stand after the step's fits. */
int orp_fit_residuals(const orp_fit_batch* b, const float* d_params, float* d_resid_out, void* stream);

/* Radix select for np.quantile(VaR, q) (Replicating_Portfolio.py:122) without sorting or host copies of the array.
 * orp_select_hist ADDS to d_hist[ORP_SELECT_BINS] (caller zeroes it) the histogram of the 8-bit digit at `shift` of the
 * order-preserving key, over the elements whose key agrees with `prefix` on the bits of `mask`; bin 256 counts NaNs.
 * Four passes (shift 24, 16, 8, 0) pin the k-th smallest element; the caller picks the digit between passes (after
 * summing the ranks' histograms on several GPUs).  key -> float: key & 0x80000000 ? key ^ 0x80000000 : ~key.
 * orp_select_next: d_out[0] += #{key <= `key`}, d_out[1] = min(d_out[1], smallest key above `key`) — the successor
 * needed for numpy's linear interpolation (caller initialises d_out = {0, 0xffffffff}). */
#define ORP_SELECT_BINS 257
int orp_select_hist(const float* d_x, uint64_t n, int shift, uint32_t prefix, uint32_t mask,
                    unsigned long long* d_hist, void* stream);
int orp_select_next(const float* d_x, uint64_t n, uint32_t key, unsigned long long* d_out, void* stream);

/* The same selection with the digit picked ON THE DEVICE, for n_q <= ORP_SELECT_MAX_Q order statistics at once (the two
 * quantiles of Replicating_Portfolio.py:122 share every pass over the array; no host round trip between passes).
 * d_state[q]: the caller sets k = rank of the wanted element (0-based, over all ranks' elements) and zeroes the rest.
 *   for shift in 24, 16, 8, 0:
 *     orp_select_hist_multi  ADDS the n_q histograms (d_hist[n_q][ORP_SELECT_BINS], zero before the first pass);
 *     [several GPUs: sum d_hist over the ranks];
 *     orp_select_pick        descends into the digit that holds rank k (updates k, prefix, mask), sets `nan` when the
 *                            array holds a NaN and `bad` when k is beyond the element count, and zeroes d_hist again;
 *   orp_select_next_multi    d_out[q][0] += #{key <= prefix_q}, d_out[q][1] = min(., smallest key above prefix_q)
 *                            (caller initialises d_out[q] = {0, 0xffffffff}; several GPUs: sum / min over the ranks).
 * Afterwards d_state[q].prefix is the order-preserving key of the k-th smallest element. */
#define ORP_SELECT_MAX_Q 4
typedef struct orp_select_state {
    unsigned long long k;      /* in: rank wanted; during the passes: rank inside the digits chosen so far */
    uint32_t prefix, mask;     /* digits chosen so far and the bits they cover */
    uint32_t nan, bad;         /* out: a NaN was seen / k was beyond the number of elements */
    unsigned long long reserved;
} orp_select_state;
int orp_select_hist_multi(const float* d_x, uint64_t n, int shift, int n_q, const orp_select_state* d_state,
                          unsigned long long* d_hist, void* stream);
int orp_select_pick(unsigned long long* d_hist, int shift, int n_q, orp_select_state* d_state, void* stream);
int orp_select_next_multi(const float* d_x, uint64_t n, int n_q, const orp_select_state* d_state,
                          unsigned long long* d_out, void* stream);

/* ---------------------------------------------------------------- §8f-2  survivors (device) */

/* Replaces the binomial thinning loop of Replicating_Portfolio.py:78-84 for path counts the host loop cannot reach:
 * N_t ~ Binomial(N_{t-1}, exp(-lambda_t dt)) per path, lambda time-major as written by orp_paths (ORP_MODEL_LINEAR),
 * d_lambda[t * ld + j], t = 0..n_steps.  The reference's NumPy legacy stream is sequential and irreproducible in
 * parallel; the draws here come from Philox4x32-10 keyed by `seed` and indexed by (start + j, t), so they do not
 * depend on the sharding.  Parity with the reference is distributional.  d_out[col * ld_out + j] = N_t / n0 (fp32)
 * at t = col * keep_every, col = 0 .. n_steps / keep_every. */
int orp_survivors(const void* d_lambda, int dtype, uint64_t ld, uint64_t start, uint64_t count, int n_steps,
                  int keep_every, double n0, double dt, uint64_t seed, float* d_out, uint64_t ld_out, void* stream);

/* ---------------------------------------------------------------- measurement support */

/* No counterpart in the reference.  Launches a register-only loop of one instruction kind on every SM (2 blocks of 512
 * threads per SM, 8 independent chains per thread, `iters` x 64 instructions per chain; 8 per chain for the MMA probe)
 * and returns the number of WARP instructions issued; the caller times the launch with CUDA events.  bench.py uses
 * ORP_PROBE_FFMA2 (two fp32 FMAs per lane and instruction) as the measured denominator of the fit kernel's roofline:
 * MEASURED_PEAKS.json has no fp32 figure. */
enum { ORP_PROBE_FFMA2 = 0, ORP_PROBE_FFMA = 1, ORP_PROBE_FADD = 2, ORP_PROBE_LOP3 = 3, ORP_PROBE_MMA_TF32 = 4 };
int orp_probe_throughput(int kind, int iters, float* d_sink, double* warp_instructions_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ORP_B200_H */
