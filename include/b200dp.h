/*
 * b200dp — hand-written sm_100a kernels for sharded data-parallel training.
 *
 * Plain C ABI: every entry point takes raw device pointers and the CUDA stream to
 * launch on (as void*), returns 0 on success (otherwise a cudaError_t value or a
 * negative B200DP_ERR_* code; b200dp_last_error() has the text) and never
 * synchronises the host.  Unless an entry point says otherwise it allocates nothing:
 * scratch memory is passed in by the caller.
 *
 * "Reference" citations name the call site in the reference scripts whose library
 * call the kernel replaces (the reference itself ships no native code).
 */
#ifndef B200DP_H
#define B200DP_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200DP_ERR_INVALID (-1)
#define B200DP_ERR_UNSUPPORTED (-2)
#define B200DP_ERR_TIMEOUT (-3)

typedef enum { B200DP_F32 = 0, B200DP_F16 = 1, B200DP_BF16 = 2 } b200dp_dtype;

const char* b200dp_last_error(void);
int b200dp_version(void);

/* ------------------------------------------------------------------ multi-tensor ops
 * A "tensor list" is described on the device by an array of b200dp_tensor_desc and a
 * chunk table: chunk c covers elements [chunks[c].y * B200DP_CHUNK, +B200DP_CHUNK) of
 * tensor chunks[c].x.  One CTA per chunk.  The caller builds both tables once and
 * reuses them while the pointers stay the same. */
#define B200DP_CHUNK 8192

typedef struct {
    void* p;      /* fp32 parameter (master copy)            */
    void* g;      /* gradient, dtype given per call          */
    void* m;      /* fp32 first moment / momentum buffer     */
    void* v;      /* fp32 second moment (AdamW only)         */
    void* p_lp;   /* optional 16-bit copy of the parameter   */
    int64_t n;    /* elements                                */
} b200dp_tensor_desc;

typedef struct { int32_t x, y; } b200dp_chunk;

typedef struct {
    float lr, beta1, beta2, eps, weight_decay;
    int adamw;            /* 1: decoupled weight decay (AdamW); 0: L2 (Adam)   */
    int maximize;
    const float* lr_dev;  /* optional device scalar overriding lr (LR schedules under CUDA graphs) */
} b200dp_adam_hparams;

typedef struct {
    float lr, momentum, dampening, weight_decay;
    int nesterov;
    int maximize;
    const float* lr_dev;  /* optional device scalar overriding lr */
} b200dp_sgd_hparams;

/* Fused AdamW over a tensor list (reference: AdamW built at Stoke-DDP.py:226-235 /
 * Fairscale-DDP.py:78-86, stepped at Stoke-DDP.py:82 / Fairscale-DDP.py:101).
 *   step       device float: number of steps INCLUDING this one (see b200dp_step_advance)
 *   grad_mult  device float or NULL: every gradient is multiplied by it first
 *              (1/loss_scale x clip coefficient — AMP unscale and clipping fused in)
 *   found_inf  device float or NULL: if non-zero the kernel returns without touching
 *              anything (the GradScaler "skip this step" decision, made on the device)
 *   lp_dtype   dtype of desc.p_lp copies (ignored where p_lp is NULL)               */
int b200dp_multi_adam(const b200dp_tensor_desc* descs, const b200dp_chunk* chunks, int n_chunks,
                      b200dp_dtype grad_dtype, b200dp_dtype lp_dtype, b200dp_adam_hparams hp,
                      const float* step, const float* grad_mult, const float* found_inf, void* stream);

/* Fused SGD (momentum / nesterov / weight decay).  `first_step` non-zero initialises the
 * momentum buffer with the gradient (torch semantics). */
int b200dp_multi_sgd(const b200dp_tensor_desc* descs, const b200dp_chunk* chunks, int n_chunks,
                     b200dp_dtype grad_dtype, b200dp_dtype lp_dtype, b200dp_sgd_hparams hp, int first_step,
                     const float* grad_mult, const float* found_inf, void* stream);

/* step += (found_inf == 0) ? 1 : 0   (one thread; keeps the step count on the device) */
int b200dp_step_advance(float* step, const float* found_inf, void* stream);

/* In-place g *= inv_scale over a tensor list, found_inf = 1 if any element is non-finite
 * (reference: GradScaler configured at Stoke-DDP.py:182-184; replaces
 * _amp_foreach_non_finite_check_and_unscale_). */
int b200dp_multi_unscale_check(const b200dp_tensor_desc* descs, const b200dp_chunk* chunks, int n_chunks,
                               b200dp_dtype grad_dtype, const float* inv_scale, float* found_inf, void* stream);

/* Sum of squares of every gradient in a tensor list plus a non-finite check, in one
 * read of the gradients (reference: ClipGradNormConfig at Stoke-DDP.py:253).
 *   scratch   n_chunks floats of caller-provided scratch (per-CTA partial sums)
 *   counter   one uint32 of caller-provided scratch, must be zero on entry; the kernel
 *             leaves it zero again
 *   out[0]    sum of squares (deterministic: fixed-order final reduction by the last CTA)
 *   out[1]    1.0 if any element is non-finite, else left unchanged                  */
int b200dp_multi_sumsq_check(const b200dp_tensor_desc* descs, const b200dp_chunk* chunks, int n_chunks,
                             b200dp_dtype grad_dtype, float* scratch, unsigned int* counter, float* out,
                             void* stream);

/* grad_mult = inv_scale * min(1, max_norm / (sqrt(sumsq) * inv_scale + 1e-6));
 * total_norm = sqrt(sumsq) * inv_scale.  inv_scale may be NULL (= 1).  One thread. */
int b200dp_clip_coef(const float* sumsq, const float* inv_scale, float max_norm, float* grad_mult,
                     float* total_norm, void* stream);

/* Dynamic loss-scale update (replaces _amp_update_scale_); also clears found_inf. */
int b200dp_scaler_update(float* scale, int* growth_tracker, float* found_inf, float growth_factor,
                         float backoff_factor, int growth_interval, void* stream);

/* Flat sum of squares of n floats (two-stage, deterministic).
 *   scratch: at least b200dp_sumsq_scratch_floats(n) floats; counter: one zeroed uint32. */
int64_t b200dp_sumsq_scratch_floats(int64_t n);
int b200dp_sumsq_f32(const float* x, int64_t n, float* scratch, unsigned int* counter, float* out, void* stream);

/* Sum of squares for callers that cannot allocate device memory (host callbacks and
 * serving loops where a cudaMalloc would stall the GPU, processes whose memory is fully
 * pre-partitioned, C hosts with no allocator to borrow a workspace from).  Same result as
 * b200dp_sumsq_f32, but the partial sums and the ticket live in a fixed device array
 * inside the library (65,792 bytes held on every GPU that loads it, until exit), so the
 * caller passes only input and output.  No counterpart in the reference: the reference
 * has no native code and no such constraint.
 *   - n may be at most B200DP_NOALLOC_MAX_ELEMS; a larger n returns B200DP_ERR_INVALID
 *     and launches nothing.
 *   - All calls on a GPU share the one array: calls that can overlap on the GPU (different
 *     streams) race; keep them on one stream.  The ticket word is cleared per call.
 *   - The first call on each GPU resolves the array's device address
 *     (cudaGetSymbolAddress, that GPU current); make it before any stream capture.
 *   - Calls no cudaMalloc / cudaFree / allocating routine. */
#define B200DP_NOALLOC_MAX_ELEMS (1ll << 28)
int b200dp_sumsq_f32_noalloc(const float* x, int64_t n, float* out, void* stream);

/* ------------------------------------------------------------------ losses
 * Softmax cross-entropy with mean reduction, forward and backward in one pass over the
 * logits (reference: classification loss implied by BASELINE configs; nn.MSELoss at
 * Fairscale-DDP.py:76 is b200dp_mse below).
 *   logits [rows, classes] (dtype), labels int64 [rows]
 *   loss_rows  float [rows]   per-row loss
 *   dlogits    [rows, classes] (dtype) = (softmax - onehot) / rows   (may be NULL)
 *   loss_mean  device float, overwritten (uses loss_rows as its partials: no other scratch) */
int b200dp_cross_entropy(const void* logits, const int64_t* labels, int rows, int classes, b200dp_dtype dtype,
                         float* loss_rows, void* dlogits, float* loss_mean, float label_smoothing, void* stream);
/* x *= *scale in place (backward of the above: dlogits *= upstream gradient) */
int b200dp_scale_inplace(void* x, int64_t n, b200dp_dtype dtype, const float* scale, void* stream);

/* Mean squared error: out = mean((a-b)^2); optionally grad = 2 (a-b) / n.
 *   scratch: b200dp_sumsq_scratch_floats(n) floats; counter: one zeroed uint32. */
int b200dp_mse(const void* a, const void* b, int64_t n, b200dp_dtype dtype, void* grad, float* scratch,
               unsigned int* counter, float* out, void* stream);

/* ------------------------------------------------------------------ normalisation (NHWC)
 * BatchNorm over channels-last activations x[rows, C] (rows = N*H*W), 16-bit storage,
 * fp32 statistics (reference: BatchNorm inside the CNN of Fairscale-DDP.py:74 and the
 * SyncBatchNorm conversion at Stoke-DDP.py:192).
 *
 * forward, training:
 *   1. b200dp_bn_reduce        sums[0:C] = sum x, sums[C:2C] = sum x^2      (one read of x)
 *      (multi-GPU: all-reduce `sums` and the row count between 1 and 2)
 *   2. b200dp_bn_apply         y = act(gamma*(x-mean)*invstd + beta + residual), also writes
 *                              mean/invstd and updates the running statistics
 * backward:
 *   3. b200dp_bn_bwd_reduce    dsums[0:C] = sum dz, dsums[C:2C] = sum dz*xhat, dz = dy*(y>0)
 *      (multi-GPU: all-reduce `dsums`)
 *   4. b200dp_bn_bwd_apply     dx = gamma*invstd*(dz - mean(dz) - xhat*mean(dz*xhat)); dres = dz
 *
 * relu_mask (optional, rows*C/8 bytes): bn_apply records y > 0 as one bit per element and the
 * two backward kernels read that instead of y (1/16 of the bytes); pass NULL to use y itself.
 * scratch for the two reductions: b200dp_bn_scratch_floats(rows, C) floats + 64 zeroed uint32
 * completion counters (one per 256-channel column). */
int64_t b200dp_bn_scratch_floats(int64_t rows, int C);
int b200dp_bn_reduce(const void* x, int64_t rows, int C, b200dp_dtype dtype, float* scratch,
                     unsigned int* counter, float* sums, void* stream);
int b200dp_bn_apply(const void* x, const void* residual, void* y, int64_t rows, int C, b200dp_dtype dtype,
                    const float* sums, float count, const float* gamma, const float* beta, float eps,
                    float momentum, float* running_mean, float* running_var, float* mean_out,
                    float* invstd_out, int relu, unsigned char* relu_mask, void* stream);
/* Same, and *num_batches_tracked (int64, may be NULL) += 1 from inside the kernel: nn.BatchNorm's step counter without a
 * launch of its own per layer and step. */
int b200dp_bn_apply_tracked(const void* x, const void* residual, void* y, int64_t rows, int C, b200dp_dtype dtype,
                            const float* sums, float count, const float* gamma, const float* beta, float eps,
                            float momentum, float* running_mean, float* running_var, float* mean_out,
                            float* invstd_out, int relu, unsigned char* relu_mask, long long* num_batches_tracked,
                            void* stream);
/* inference: y = act(gamma*(x-running_mean)*rsqrt(running_var+eps) + beta + residual) */
int b200dp_bn_infer(const void* x, const void* residual, void* y, int64_t rows, int C, b200dp_dtype dtype,
                    const float* gamma, const float* beta, const float* running_mean,
                    const float* running_var, float eps, int relu, void* stream);
int b200dp_bn_bwd_reduce(const void* dy, const void* y, const void* x, int64_t rows, int C,
                         b200dp_dtype dtype, const float* mean, const float* invstd, int relu,
                         float* scratch, unsigned int* counter, float* dsums, const unsigned char* relu_mask,
                         void* stream);
int b200dp_bn_bwd_apply(const void* dy, const void* y, const void* x, void* dx, void* dres, int64_t rows,
                        int C, b200dp_dtype dtype, const float* mean, const float* invstd,
                        const float* gamma, const float* dsums, float count, int relu,
                        const unsigned char* relu_mask, void* stream);

/* NHWC max pooling (reference: the stem pooling of the CNN models).  The forward also writes one
 * byte per output element — the winning tap's position in the window — which turns the backward
 * into a gather with no index tensor, zero-fill or atomics.  C must be a multiple of 8. */
int b200dp_maxpool_fwd(const void* x, void* y, unsigned char* idx, int N, int H, int W, int C, int K, int stride, int pad,
                       b200dp_dtype dtype, void* stream);
int b200dp_maxpool_bwd(const void* dy, const unsigned char* idx, void* dx, int N, int H, int W, int C, int K, int stride,
                       int pad, b200dp_dtype dtype, void* stream);

/* LayerNorm over the last dimension, x[rows, D] (reference: SwinIR / ViT blocks,
 * Stoke-DDP.py:206-208).  mean/rstd are float[rows].  The backward is ONE pass over dy and x: dx plus
 * one [dbeta | dgamma] partial per CTA in `scratch` (b200dp_ln_scratch_floats(rows, D) floats), summed
 * in a fixed order by a second small kernel; `counter` is unused (kept for ABI stability). */
int64_t b200dp_ln_scratch_floats(int64_t rows, int D);
int b200dp_layernorm_fwd(const void* x, void* y, int64_t rows, int D, b200dp_dtype dtype, const float* gamma,
                         const float* beta, float eps, float* mean, float* rstd, void* stream);
int b200dp_layernorm_bwd(const void* dy, const void* x, void* dx, int64_t rows, int D, b200dp_dtype dtype,
                         const float* gamma, const float* mean, const float* rstd, float* scratch,
                         unsigned int* counter, float* dgamma, float* dbeta, void* stream);
/* dx += add_in (gradient of a branch that bypassed the layer, e.g. the residual stream) in the same pass */
int b200dp_layernorm_bwd_add(const void* dy, const void* x, void* dx, int64_t rows, int D, b200dp_dtype dtype,
                             const float* gamma, const float* mean, const float* rstd, float* scratch,
                             unsigned int* counter, float* dgamma, float* dbeta, const void* add_in, void* stream);
/* out[C] (fp32) = column sums of x[rows][C] (C % 8 == 0): the bias gradient of a linear layer.  Two small launches
 * (per-CTA partials, then a multi-CTA second stage); scratch holds b200dp_column_sums_scratch_floats(rows, C) floats. */
int64_t b200dp_column_sums_scratch_floats(int64_t rows, int C);
int b200dp_column_sums(const void* x, int64_t rows, int C, b200dp_dtype dtype, float* scratch, float* out, void* stream);

/* Zero-padded rows: x / y (dy / dx / add_in) have rows `pitch` elements apart, the first D of which are normalised;
 * columns D .. pitch of y and dx are written as zeros, so a width padded for the tensor cores (SwinIR-S: 60 -> 64)
 * stays padded with zeros through the layer.  D % 4 == 0, pitch % 4 == 0, D <= pitch <= 256.  add_in may be NULL. */
int b200dp_layernorm_fwd_pitched(const void* x, void* y, int64_t rows, int D, int pitch, b200dp_dtype dtype,
                                 const float* gamma, const float* beta, float eps, float* mean, float* rstd, void* stream);
int b200dp_layernorm_bwd_pitched(const void* dy, const void* x, void* dx, int64_t rows, int D, int pitch, b200dp_dtype dtype,
                                 const float* gamma, const float* mean, const float* rstd, float* scratch,
                                 unsigned int* counter, float* dgamma, float* dbeta, const void* add_in, void* stream);

/* ------------------------------------------------------------------ tensor-core GEMM (tcgen05 / TMEM / TMA)
 * C[M,N] (+)= A(MxK) * B(NxK)^T (+ bias[N]), 16-bit inputs, fp32 accumulation in TMEM.
 *   a_mn / b_mn = 0: operand stored [rows][K] (K contiguous);  = 1: stored [K][rows] (rows
 *   contiguous) — so forward, dgrad and wgrad of a linear layer need no transposes:
 *     forward  Y  = X W^T      gemm(X,  W, M=rows, N=out, K=in,   a_mn=0, b_mn=0)
 *     dgrad    dX = dY W       gemm(dY, W, M=rows, N=in,  K=out,  a_mn=0, b_mn=1)
 *     wgrad    dW = dY^T X     gemm(dY, X, M=out,  N=in,  K=rows, a_mn=1, b_mn=1)
 *   lda / ldb: row pitch in elements of A / B as stored (multiples of 8); ldc likewise for C.
 *   out_dtype: the input dtype or B200DP_F32.  accumulate_c: C += instead of C =.
 * Reference: the linear layers of the transformer models (Stoke-DDP.py:206-208) and the 1x1
 * convolutions of the CNNs, which the reference leaves to cuBLAS / cuDNN. */
int b200dp_gemm(const void* A, const void* B, void* C, const float* bias, int M, int N, int K, int lda, int ldb,
                int ldc, int a_mn, int b_mn, b200dp_dtype in_dtype, b200dp_dtype out_dtype, int accumulate_c,
                void* stream);

/* Implicit-GEMM NHWC convolution on tcgen05, operands fetched by TMA straight from the
 * activation tensor (no im2col buffer).  mode 0: forward, y[N][P][Q][Cout] = conv(x[N][H][W][Cin], w);
 * mode 1: input gradient, y = dX[N][H][W][Cin] from x = dY[N][P][Q][Cout] (stride 2 runs as four stride-1
 * problems, one per parity class of the input pixel).
 * w is [Cout][R][S][Cin] (channels-last weight), same 16-bit dtype as x.  The reduction-side channel
 * count must be a multiple of 64, the output-side one a multiple of 8; stride 1 or 2; square padding.
 * Reference: the convolutions of the CNN models (Fairscale-DDP.py:74), left to cuDNN there. */
int b200dp_conv2d(int mode, const void* x, const void* w, void* y, const float* bias, int N, int H, int W, int Cin,
                  int Cout, int R, int S, int stride, int pad, b200dp_dtype dtype, void* stream);

/* Weight gradient of the same convolution: dw[Cout][R][S][Cin] (fp32) = sum over pixels of
 * dy[N][P][Q][Cout] x x[N][H][W][Cin] shifted by the tap; both operands fetched by 4-D TMA as
 * MN-major tiles (pixels are the reduction), split over pixel boxes and combined with TMA reduce-add
 * (so the summation order is not fixed).  dw is zeroed first unless accumulate != 0. */
int b200dp_conv2d_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout, int R, int S,
                        int stride, int pad, b200dp_dtype dtype, int accumulate, void* stream);

/* ------------------------------------------------------------------ peer-memory collectives
 * One kernel per rank; ranks meet at flag barriers in peer-mapped memory (csrc/collectives.cu
 * explains the protocol).  Every buffer argument is a table of B200DP_MAX_RANKS pointers:
 * entry p is rank p's buffer as mapped into THIS process (entry `rank` is the local one).
 * Tables of rank-local objects (scalars, optimizer state) only need entry `rank`.
 * `mc` arguments are the NVLS multicast address of the same buffer, or NULL to use plain
 * peer loads/stores.  With ctx.emulate = 1 all ranks run on one GPU (tests).
 *
 * Reference: these replace the NCCL collectives the reference's libraries issue —
 * DDP bucket all-reduce (Stoke-DDP.py:248), ShardedDDP reduce-to-owner (Fairscale-DDP.py:89,
 * Stoke-DDP.py:250), OSS post-step broadcast (Fairscale-DDP.py:101, Stoke-DDP.py:197-199),
 * FSDP all-gather / reduce-scatter (Stoke fairscale_fsdp flag), SyncBN / loss / norm scalar
 * all-reduces (Stoke-DDP.py:192,86,253). */
#define B200DP_MAX_RANKS 8

typedef struct {
    int rank;            /* this process's rank within the group (ignored when emulating)        */
    int world;           /* number of ranks, <= B200DP_MAX_RANKS                                  */
    int emulate;         /* 1: single-GPU emulation, grid.y = world, cooperative launch           */
    int max_blocks;      /* rows in each flag array                                               */
    uint32_t* flags[B200DP_MAX_RANKS];   /* [2*max_blocks][B200DP_MAX_RANKS] words per rank, zeroed once: barrier rows, then the
                                            non-finite notice rows of the NVLS all-reduce */
    uint32_t* epochs[B200DP_MAX_RANKS];  /* [max_blocks] per-CTA barrier counters (local entry used)      */
    uint32_t* status[B200DP_MAX_RANKS];  /* one word per rank: non-zero after a barrier timed out         */
    uint64_t timeout_ns; /* spin limit per barrier; 0 = wait forever                              */
    uint64_t skew_ns;    /* stress knob, 0 in production: before every barrier each (rank, CTA) first
                            idles a pseudo-random time below this, so arrival orders are shuffled  */
} b200dp_peer_ctx;

/* `rounds` barriers back to back (protocol self-test / latency probe). */
int b200dp_peer_barrier(const b200dp_peer_ctx* ctx, int blocks, int rounds, void* stream);

/* In-place all-reduce of n elements: buf = pre_mult * (*mult_dev) * sum over ranks, and
 * found_inf = 1 if any result is non-finite (W1: all-reduce + AMP unscale + inf check).
 * algo 0 = one-shot (small), 1 = two-shot; with mc != NULL the NVLS two-shot is used.
 * n must be a multiple of 16 bytes worth of elements. */
int b200dp_peer_allreduce(const b200dp_peer_ctx* ctx, void* const* bufs, void* mc, int64_t n, b200dp_dtype dtype,
                          float pre_mult, const void* const* mult_dev, void* const* found_inf, int algo, int blocks,
                          void* stream);

/* outs[rank][i] (= or +=) pre_mult * (*mult_dev) * sum_p in[p][rank*n_per_rank + i], fp32 out
 * (W5: reduce-scatter + scale + cast to the fp32 shard gradient). */
int b200dp_peer_reduce_scatter(const b200dp_peer_ctx* ctx, void* const* in_bufs, void* mc, int64_t n_per_rank,
                               b200dp_dtype dtype, void* const* outs, float pre_mult, const void* const* mult_dev,
                               void* const* found_inf, int accumulate, int blocks, void* stream);

/* Every rank pushes cast(shard) into slot `rank` of every rank's output buffer (all-gather
 * fused with the fp32 -> compute-dtype cast; one multicast store per 16 bytes with NVLS). */
int b200dp_peer_all_gather(const b200dp_peer_ctx* ctx, void* const* outs, void* mc, const void* const* shards,
                           int64_t n_per_rank, b200dp_dtype shard_dtype, b200dp_dtype out_dtype, int blocks,
                           void* stream);

/* ZeRO step on owner-ordered flat buffers (W2 + W3): rank r owns elements [seg[r], seg[r+1])
 * (multiples of 8).  Reduce the owned gradient segment across ranks (or read it pre-reduced
 * from local_grad), apply AdamW to the rank-local fp32 master/m/v, and publish the new
 * parameter values into every rank's parameter buffer (fp32 or 16-bit wire). */
/* hp.adamw selects the update rule: 1 = AdamW, 0 = Adam with L2, 2 = SGD — then beta1 is the momentum, beta2 the
 * dampening, eps != 0 selects Nesterov, weight_decay is L2, `m` is the momentum buffer and `v` is not touched. */
int b200dp_peer_zero_adam(const b200dp_peer_ctx* ctx, void* const* grad_bufs, void* grad_mc, void* const* param_bufs,
                          void* param_mc, const int64_t* seg, b200dp_dtype wire_dtype, void* const* master,
                          void* const* m, void* const* v, b200dp_adam_hparams hp, const void* const* step,
                          float pre_mult, const void* const* mult_dev, const void* const* found_inf, int prereduced,
                          const void* const* local_grad, int blocks, void* stream);

/* Reduce-to-owner only: local_out[rank] = scaled sum of the owned segment; norm_out[rank] =
 * {sum of squares of it, non-finite flag}.  partials: `blocks` floats; counters: zeroed uint32. */
int b200dp_peer_reduce_to_owner(const b200dp_peer_ctx* ctx, void* const* grad_bufs, const int64_t* seg,
                                void* const* local_out, float pre_mult, const void* const* mult_dev,
                                void* const* norm_out, void* const* partials, void* const* counters, int blocks,
                                void* stream);

/* Root's buffer (raw bytes, multiple of 16) replicated into every rank's buffer: module
 * buffers such as BatchNorm running statistics, every forward (replaces the per-forward
 * _broadcast_coalesced of the data-parallel wrappers). */
int b200dp_peer_broadcast(const b200dp_peer_ctx* ctx, void* const* bufs, void* mc, int64_t nbytes, int root, int blocks,
                          void* stream);

/* In-place sum across ranks of n <= slot_floats fp32 values (SyncBN moments, loss, norm, inf
 * flag: W6 / W11).  staging[p]: 2*slot_floats floats of peer memory; call_counter: local uint32. */
int b200dp_peer_small_allreduce(const b200dp_peer_ctx* ctx, void* const* staging, void* const* io, int n,
                                int slot_floats, void* const* call_counter, void* stream);

/* First-layer convolutions (Cin <= 4: RGB stems) on the tcgen05 tensor cores, without im2col: the input rows of a tile are
 * staged in shared memory with channels padded so that consecutive output pixels start 16 bytes apart, and the un-swizzled
 * UMMA layout (LBO 16, SBO 128) reads the overlapping windows straight out of them.
 *   fwd    y[N][P][Q][Cout] = conv(x[N][H][W][Cin], w[Cout][R][S][Cin]) (+ bias[Cout]), 16-bit NHWC
 *   wgrad  dw[Cout][R][S][Cin] (fp32) += dy^T im2col(x); ATOMIC accumulation, the caller zeroes dw
 * Envelope: Cin <= 4, stride 1 or 2, R <= 7, Cout a multiple of 8 up to 64 (b200dp_conv2d_smallc_supported).
 * Reference: the first convolution of the CNN models (ResNet stem 7x7/2; the SR net of Fairscale-DDP.py:74;
 * SwinIR conv_first and the VGG layer of feat_loss, Stoke-DDP.py:206-224) — cuDNN first_layer kernels there. */
int b200dp_conv2d_smallc_supported(int Cin, int Cout, int R, int S, int stride);
int b200dp_conv2d_smallc_fwd(const void* x, const void* w, void* y, const float* bias, int N, int H, int W, int Cin, int Cout,
                             int R, int S, int stride, int pad, b200dp_dtype dtype, void* stream);
int b200dp_conv2d_smallc_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout, int R, int S,
                               int stride, int pad, b200dp_dtype dtype, void* stream);

/* Attention core on the tcgen05 tensor cores: out = softmax(q k^T * scale + bias + mask) v, forward and backward,
 * over the PACKED projections (no per-head copies):
 *   qkv / dqkv [B][S][3][H][d],  out / dout [B][S][H][d],  lse [B][H][S] fp32 (row log-sum-exp, written by the forward),
 *   bias [H][S][S] fp32 or NULL (window attention: relative-position bias),
 *   mask [nW][S][S] fp32 or NULL, applied to batch entry b as mask[b % nW] (shifted windows),
 *   dbias [H][S][S] fp32 or NULL: gradient of bias, ACCUMULATED with atomic adds (caller zeroes it); S <= 64 only.
 * Envelope: S <= 256 tokens, even d <= 64 (b200dp_attention_supported).  Sequences of <= 64 tokens are stacked two
 * to a 128-row tile.  Operands arrive by TMA when d == 64 and S > 64, else by a zero-padding copy in the kernel.
 * Reference: the attention of the transformer models (SwinIR built at Stoke-DDP.py:206-208, window_size 8;
 * ViT-B/16 of BASELINE config 4); the library path there is scaled_dot_product_attention. */
int b200dp_attention_supported(int S, int d);
int b200dp_attention_fwd(const void* qkv, void* out, float* lse, const float* bias, const float* mask, int nW, int B, int H,
                         int S, int d, float scale, b200dp_dtype dtype, void* stream);
int b200dp_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, const float* bias,
                         const float* mask, int nW, float* dbias, int B, int H, int S, int d, float scale,
                         b200dp_dtype dtype, void* stream);

/* (Shifted-)window attention addressed on the IMAGE-ordered token grid: qkv [images][Hi][Wi][qkv_pitch] (the first 3*H*d
 * elements of a token row are q|k|v in [3][H][d] order), out [images][Hi][Wi][out_pitch].  The cyclic shift by `shift`
 * tokens and the cut into ws x ws windows are part of the kernels' addressing, so the roll / window_partition /
 * window_reverse copies of a Swin block do not exist.  lse [windows][H][ws*ws], bias [H][S][S], mask
 * [windows per image][S][S] and dbias keep the layouts of the plain entry points.  Padding columns (past 3*H*d / H*d)
 * are never read; the backward writes only the first 3*H*d elements of each dqkv row (the caller zeroes the padding).
 * Envelope: ws*ws <= 64, Hi % ws == Wi % ws == 0, even d <= 64, even pitches.
 * Reference: SwinIR's SwinTransformerBlock (built at Stoke-DDP.py:206-208, window_size 8). */
int b200dp_window_attention_fwd(const void* qkv, void* out, float* lse, const float* bias, const float* mask, int images,
                                int Hi, int Wi, int ws, int shift, int H, int d, int qkv_pitch, int out_pitch, float scale,
                                b200dp_dtype dtype, void* stream);
int b200dp_window_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                                const float* bias, const float* mask, float* dbias, int images, int Hi, int Wi, int ws,
                                int shift, int H, int d, int qkv_pitch, int out_pitch, int dout_pitch, float scale,
                                b200dp_dtype dtype, void* stream);

/* Attention layout moves (ViT / SwinIR): W per-head tensors [B][H][S][d] <-> one token-major tensor
 * [B][S][W][H][d] in one streaming pass (W = 3: packed q/k/v gradient; W = 1: merge heads).  Rows of
 * d elements must be multiples of 16 bytes.  Reference: the reshape / permute / stack chains around
 * nn.MultiheadAttention in the transformer models (ATen copies there). */
int b200dp_heads_to_tokens(const void* const* srcs, void* out, int B, int H, int S, int d, int W, int elem_bytes,
                           long long stride_b, long long stride_h, long long stride_s /* source strides, elements */, void* stream);
int b200dp_tokens_to_heads(const void* in, void* const* dsts, int B, int H, int S, int d, int W, int elem_bytes, void* stream);

/* GEMM / forward convolution whose epilogue also accumulates, per output column (channel), the sum
 * and the sum of squares of the values it stores: colstats[0:N] += sum, colstats[N:2N] += sum of
 * squares (fp32 atomics from per-lane register partials; the caller zeroes colstats).  These are the
 * batch statistics of the BatchNorm that follows the layer, which then skips its own reduction pass
 * over the activation (b200dp_bn_reduce).  K-major 16-bit operands, 16-bit output, no bias for the
 * GEMM form.  Reference: every conv -> BatchNorm pair of the CNN models (cuDNN conv + ATen batch_norm
 * there, two passes over the activation). */
int b200dp_gemm_colstats(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                         b200dp_dtype dtype, float* colstats, void* stream);
int b200dp_conv2d_colstats(const void* x, const void* w, void* y, const float* bias, int N, int H, int W, int Cin, int Cout,
                           int R, int S, int stride, int pad, b200dp_dtype dtype, float* colstats, void* stream);

/* Large-product GEMM with a fused epilogue on the CTA-pair (cta_group::2) kernel; 16-bit in / out, A K-major:
 *   mode 1  C = A B^T (+bias), C2 = gelu(C)      pre-activation and activation from one epilogue (b_mn = 0)
 *   mode 2  C = (A B^T) * gelu'(aux)              input gradient through a GELU (dgrad form, b_mn = 1)
 *   mode 3  C = A B^T (+bias) + aux               residual add (b_mn = 0)
 * b200dp_gemm_fused_supported(M, N) says whether the shape is inside the kernel's envelope.
 * Reference: nn.Linear + nn.GELU + the residual adds of the transformer blocks (separate ATen kernels there). */
int b200dp_gemm_fused_supported(int M, int N);
int b200dp_gemm_fused(const void* A, const void* B, void* C, void* C2, const float* bias, const void* aux, int M, int N, int K,
                      int lda, int ldb, int ldc, int ldc2, int ld_aux, int b_mn, b200dp_dtype dtype, int mode, void* stream);

/* All-gather fused into the consuming GEMM (W4): Y[M,N] = X[M,K] W^T (+bias) with W[N,K] row-sharded
 * over the ranks (w_shards[p] = rank p's [rows_per_rank][K] shard in peer-mapped memory).  One kernel:
 * the otherwise idle warp of every CTA pulls chunks of 1-8 weight rows over NVLink into the local copy
 * w_full and bumps the arrival counter of the chunk's weight tile; the TMA producer waits until the
 * tile it is about to consume is complete.
 * Locally owned tiles are computed first, so remote transfers overlap the math; each remote byte is
 * read once per rank.  tile_flags: ceil(N/128)+1 words, zero before the
 * first call (the kernel leaves them zero again; one buffer per stream).  ctx == NULL: shards are
 * local (tests).  Reference: FSDP's parameter all-gather / OSS's broadcast followed by the first
 * consuming linear (Stoke fairscale_fsdp / fairscale_oss flags), two NCCL + cuBLAS calls there. */
int b200dp_gemm_gather(const void* X, const void* const* w_shards, void* w_full, void* Y, const float* bias, int M, int N,
                       int K, int ldx, int ldy, int world, int rank, int rows_per_rank, b200dp_dtype dtype,
                       unsigned int* tile_flags, const b200dp_peer_ctx* ctx, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200DP_H */
