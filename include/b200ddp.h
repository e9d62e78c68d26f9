/* libb200ddp — C ABI of the sm_100a kernels and the native runtime.
 *
 * Every function returns 0 on success, a cudaError_t (> 0) when a CUDA call
 * failed, or B200_ERR_INVALID (< 0) when an argument was rejected;
 * b200_last_error() then holds a one-line description (thread-local).
 * Pointers are device pointers unless the name says host.  `stream` is a
 * cudaStream_t passed as void*.  Nothing here synchronises the stream, and
 * nothing allocates unless the comment says so, so calls can be captured in
 * CUDA graphs.
 *
 * Citations "ref:" point at the reference (northflank-examples/
 * distributed-pytorch-example) call site whose work the entry point replaces;
 * "torch:" at the PyTorch component the reference delegated to.
 */
#ifndef B200DDP_H
#define B200DDP_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_ERR_INVALID (-1)

/* ---------------------------------------------------------------- runtime */
const char* b200_last_error(void);
/* Number of kernel launches issued through this library by the calling
 * process (bench.py reports it as gpu_launches). */
uint64_t b200_launch_count(void);
/* SM count etc. of the current device: out[0]=sm_count, out[1]=cc_major,
 * out[2]=cc_minor, out[3]=max dynamic smem per block (opt-in). */
int b200_device_info(int* out4);
/* Loads all kernels and sets their attributes; call once per process before the
 * first CUDA-graph capture. */
int b200_init(void);

/* ---------------------------------------------------------------- optimizer
 * One pass over a contiguous arena slice: p/m/v fp32 (master), g fp32 or bf16,
 * optional bf16 shadow copy of the updated parameters for the compute kernels.
 * ref: train.py:139 optimizer.step(), :249 optim.Adam; torch:optim/adam.py
 * (_multi_tensor_adam: ~8 foreach launches + a host step tensor per parameter).
 * kind: 0 Adam (L2 weight decay), 1 AdamW (decoupled), 2 SGD with momentum.
 * The step number t used for bias correction is step_dev[0] + 1 when step_dev
 * is non-null (device counter, graph-replay safe), else step_host.  When
 * `advance` is non-zero the last block to finish publishes step_dev[0] = t.
 * step_dev points at two int32: {t, ticket}. */
typedef struct {
    int kind;
    float lr, beta1, beta2, eps, weight_decay, momentum;
    float grad_scale; /* 1/world_size on the single-GPU + NCCL path */
} b200_opt_hparams;

int b200_optimizer_step(const b200_opt_hparams* hp, float* master, const void* grad,
                        int grad_is_bf16, float* exp_avg, float* exp_avg_sq, void* shadow_bf16,
                        int64_t n, int* step_dev, int step_host, int advance, void* stream);

/* ---------------------------------------------------------------- loss
 * Fused softmax cross-entropy, forward + backward in one pass.
 * ref: train.py:137,250 (CrossEntropyLoss mean) and the autograd backward of
 * it at train.py:138; torch: log_softmax + nll_loss + their two backward ops.
 * logits [rows, ld] (bf16 when logits_bf16 else fp32), targets int64.
 * dlogits (same dtype/ld as logits, may alias logits, may be null) receives
 * (softmax - onehot) * grad_scale.  A target outside [0, classes) — e.g. an
 * ignore index of -100 — is never dereferenced: its row adds no loss and gets
 * plain softmax * grad_scale; the mean still divides by `rows`.  loss_out[0] = mean loss, loss_out[1] =
 * number of rows whose argmax equals the target (as float).  `scratch` holds
 * per-block partials and a ticket: b200_ce_scratch_bytes(rows) bytes, zeroed
 * once at allocation (the kernel leaves the ticket at zero). */
int64_t b200_ce_scratch_bytes(int64_t rows);
int b200_cross_entropy(const void* logits, int logits_bf16, int64_t ld, const int64_t* targets,
                       int64_t rows, int classes, void* dlogits, float grad_scale, float* loss_out,
                       void* scratch, void* stream);

/* Same inputs, outputs and results as b200_cross_entropy, for callers that
 * must not allocate.  (The numbers agree within the tests' tolerance, the code
 * that produces them can differ: for bf16 rows with more than 2048 classes
 * b200_cross_entropy runs its vectorised row kernel, this entry always runs the
 * generic team-per-row kernel.)  (No counterpart in the reference: it has no native code
 * and no allocation-free path; this exists for embedders of this library.)
 *
 * Meant for code that lives inside another component's memory budget — a plugin
 * in a serving process whose device memory is pre-partitioned, a latency-bound
 * loop that cannot afford the device-wide stall of cudaMalloc/cudaFree, a plain
 * C caller that holds only the tensors it was handed.  Those callers have nowhere
 * to get the `scratch` buffer from, so the library carries one itself: a
 * file-scope __device__ array of B200_CE_STATIC_BYTES that every GPU loading
 * the library holds until the process exits.  The caller still passes all
 * inputs and outputs.  The call performs no cudaMalloc/cudaFree and calls no
 * allocating library routine.
 *
 * Limit: the array has B200_CE_STATIC_MAX_BLOCKS partial-sum slots, i.e. at
 * most 8 * B200_CE_STATIC_MAX_BLOCKS rows when classes <= 2048 and at most
 * B200_CE_STATIC_MAX_BLOCKS rows above that.  Larger problems are rejected
 * with B200_ERR_INVALID and a message; use b200_cross_entropy for them.
 *
 * The first call on each device looks the array's device address up
 * (cudaGetSymbolAddress, with that device current); make that first call
 * before starting any stream capture.  The address is cached per device
 * ordinal for the life of the process: after a cudaDeviceReset the cached
 * value is stale, so a process that resets a device must not call this again
 * on it.
 *
 * All calls on a device share the one array, and it keeps the previous call's
 * contents; each call re-zeroes the word that has to start at zero (an async
 * memset on `stream`) and overwrites the rest.  Two calls whose kernels can
 * overlap on the GPU (different streams, no dependency) race on the array:
 * serialise them. */
#define B200_CE_STATIC_MAX_BLOCKS 1024
#define B200_CE_STATIC_BYTES (16 + 8 * B200_CE_STATIC_MAX_BLOCKS)
int b200_cross_entropy_static(const void* logits, int logits_bf16, int64_t ld,
                              const int64_t* targets, int64_t rows, int classes, void* dlogits,
                              float grad_scale, float* loss_out, void* stream);

/* ---------------------------------------------------------------- GEMM
 * D[M,N] = epilogue(A[M,K] * B[N,K]^T) on tcgen05 tensor cores: operands staged
 * by TMA into 128B-swizzled shared memory, fp32 accumulators in TMEM (double
 * buffered), persistent warp-specialised CTAs (TMA warp, MMA warp, 4 epilogue
 * warps).  ref: train.py:39-45 nn.Linear forward (G1/G4/G5) and the five
 * backward GEMMs of train.py:138 (G8/G10/G11); torch: cuBLAS via ATen.
 *
 * Operand storage: "K-major" means the reduction index is contiguous
 * (A stored [M,K], B stored [N,K] — the nn.Linear weight layout).  "MN-major"
 * means the other index is contiguous (A stored [K,M], B stored [K,N]); dgrad
 * and wgrad use it so no transposed copies of weights/activations exist.
 * Leading dimensions are in elements and must be multiples of 8 (16 bytes);
 * base pointers must be 16-byte aligned. */
enum {
    B200_EPI_NONE = 0,        /* D = acc (+bias) */
    B200_EPI_RELU_DROPOUT,    /* D = dropout(relu(acc + bias), p); mask from Philox(seed, step) */
    B200_EPI_GATE_POSITIVE,   /* D = acc * (aux > 0 ? gate_scale : 0)   (ReLU+dropout backward) */
    B200_EPI_GELU,            /* D = gelu_tanh(acc + bias); aux_out = acc + bias */
    B200_EPI_GELU_BWD,        /* D = acc * gelu_tanh'(aux) */
    B200_EPI_SOFTMAX_CE       /* fused cross-entropy on the tile row (N <= tile N) */
};

typedef struct {
    int M, N, K;
    const void* A; int64_t lda; int a_mn_major; /* bf16 */
    const void* B; int64_t ldb; int b_mn_major; /* bf16 */
    void* D; int64_t ldd; int d_fp32;           /* bf16 or fp32 output */
    const float* bias;                          /* fp32 [N] or null */
    int epilogue;
    const void* aux; int64_t ldaux;             /* bf16 gate / pre-activation input */
    void* aux_out;                              /* bf16 [M, ldd] pre-activation output */
    float* colsum;                              /* fp32 [N] += column sums of D (bias grad) or null */
    float dropout_p;                            /* RELU_DROPOUT: drop prob; GATE_POSITIVE: 1/(1-p) is gate_scale */
    float gate_scale;
    uint64_t seed;                              /* Philox key */
    const int* step_dev;                        /* Philox stream = step_dev[0] (null: 0) */
    uint32_t layer_id;                          /* decorrelates masks between layers */
    /* SOFTMAX_CE only */
    const int64_t* targets;                     /* [M] */
    float ce_grad_scale;                        /* dlogits = (softmax - onehot) * this */
    float* loss_out;                            /* [2]: += sum of row losses * loss_scale, += correct */
    float loss_scale;                           /* 1/M for a mean */
    int ce_write_grad;                          /* 1: D receives dlogits (training); 0: D receives logits */
    int tile_n;                                 /* 0 = choose; else one of 16,32,64,128,256 */
    int max_ctas;                               /* 0 = one per SM */
    /* Batched GEMM (attention): batch index = outer * batch_inner + inner (e.g. sequence, head);
     * per-operand element strides; strides multiples of 8.  Plain epilogue only. */
    int batch_inner, batch_outer;               /* 0/1 = not batched */
    int64_t a_stride_inner, a_stride_outer, b_stride_inner, b_stride_outer, d_stride_inner, d_stride_outer;
    int causal;                                 /* 0 none; 1 output lower-triangular: tiles above the diagonal
                                                 * are skipped (never written); 2 reduce k < row tile end (A is
                                                 * lower-triangular); 3 reduce k >= row tile start (A^T is) */
    int tile_m;                                 /* 0 = choose; 128; 256 = dual-accumulator 256x256 tiles
                                                 * (needs tile_n 256, plain non-batched problem) */
    int split_k;                                /* 0 = off; 1 = choose; >1 = that many K slices.  Plain
                                                 * fp32 D only, zeroed by the caller: slices are combined
                                                 * with 16-byte vector atomics (wgrad: long K, few tiles) */
    int cta_pair;                               /* 0 = choose; 1 = force the two-SM kernel (tcgen05
                                                 * cta_group::2, one 256x256 tile per SM pair; needs a
                                                 * non-batched problem, a light epilogue and bf16 D);
                                                 * -1 = never */
    /* Implicit-GEMM convolution: one operand is the im2col VIEW of an NHWC bf16 activation
     * [conv_n, conv_h, conv_w, conv_c] (conv_c a multiple of 64), read through a TMA im2col-mode tensor map
     * — nothing is gathered in memory.  Row m of the view is output pixel (n, p, q), column
     * k = (r * conv_s + s) * conv_c + c.
     *   a_conv: A (pointer to the activation) is the view: M = N*P*Q, K = R*S*C  (forward; B = weights
     *           laid out [Cout, R*S*C])
     *   b_conv: B is the view read MN-major: N = R*S*C, K = N*P*Q  (weight gradient with A = dY^T)
     * ref: BASELINE.json config 3 (ResNet-50); the reference's baseline runs cuDNN convolutions. */
    int a_conv, b_conv;
    int conv_n, conv_h, conv_w, conv_c, conv_r, conv_s, conv_stride, conv_pad;
} b200_gemm_desc;

int b200_gemm(const b200_gemm_desc* d, void* stream);
/* The kernel b200_gemm would run for `d`, without launching anything (pure host logic; usable
 * without a GPU): out[0] = family (0 general 128 x N, 1 lean-epilogue 128 x N, 2 256 x 256 dual
 * accumulator, 3 SM pair / cta_group::2), out[1] = N tile.  sms <= 0: the current device (148). */
int b200_gemm_plan(const b200_gemm_desc* d, int sms, int* out2);

/* ---------------------------------------------------------------- elementwise
 * ref: train.py:133 `.to(device)` + the implicit fp32 input; the bf16 compute
 * path needs the batch in bf16. */
int b200_cast_f32_to_bf16(const float* src, void* dst, int64_t n, void* stream);
int b200_cast_bf16_to_f32(const void* src, float* dst, int64_t n, void* stream);
/* Stage one batch into the captured step program's fixed input buffers: features
 * fp32 -> bf16 and labels copied, one launch.  ref: train.py:133. */
#define B200_STAGE_MAX_ZERO 8
/* zero_ptrs/zero_counts (host arrays, nzero <= B200_STAGE_MAX_ZERO): fp32 regions the same
 * launch clears — the bias-gradient slots and the loss accumulator that the step's GEMM
 * epilogues add into with atomics — so the captured step graph carries no memset nodes. */
int b200_stage_batch(const float* x, const int64_t* y, void* x_bf16, int64_t* y_out, int64_t nx,
                     int64_t ny, float* const* zero_ptrs, const int* zero_counts, int nzero,
                     void* stream);

/* ---------------------------------------------------------------- normalisation / reductions
 * For the transformer configuration (BASELINE.json config 4).  ref: not in the
 * reference's own model; its GPT-2 baseline uses ATen native_layer_norm(+backward)
 * and sum().  All tensors bf16 unless noted; cols must be a multiple of 256 in
 * [256, 2048] for LayerNorm.  mean/rstd are fp32 [rows], saved for backward.
 * layernorm_bwd: dx = LN'(dy) (+ dx_add when non-null: the residual-stream
 * gradient joins in the same pass); dgamma/dbeta [cols] bf16 (grads_bf16) or
 * fp32, reduced deterministically through `scratch`
 * (b200_layernorm_bwd_scratch_bytes).  b200_colsum: out[c] = sum_r x[r, c]
 * (bias gradient), scratch of b200_colsum_scratch_bytes. */
int b200_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean,
                       float* rstd, int64_t rows, int cols, float eps, void* stream);
int64_t b200_layernorm_bwd_scratch_bytes(int64_t rows, int cols);
int b200_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* mean,
                       const float* rstd, void* dx, const void* dx_add, void* dgamma, void* dbeta,
                       int grads_bf16, int64_t rows, int cols, void* scratch, void* stream);
int64_t b200_colsum_scratch_bytes(int64_t rows, int cols);
int b200_colsum(const void* x, int64_t ld, int64_t rows, int cols, void* out, int out_bf16,
                void* scratch, void* stream);

/* ---------------------------------------------------------------- attention (GPT-2 config)
 * Row kernels of the causal attention core; the matrix products around them are batched
 * causal b200_gemm calls (see batch_* / causal in b200_gemm_desc).  ref: none in the
 * reference; its GPT-2 baseline uses a library flash-attention kernel via torch SDPA.
 * scores: [batch_heads, T, T] bf16 row-major, in place.  fwd: P = softmax(scale * S) over
 * columns <= row; bwd: dS = P * (dP - rowsum(dP * P)) * scale written over dP.  Both zero the
 * columns between the row index and the end of its 128-row block (the consuming GEMMs read
 * whole tiles up to that boundary); columns beyond are never read. */
int b200_causal_softmax_fwd(void* scores, int64_t batch_heads, int T, float scale, void* stream);
int b200_causal_softmax_bwd(const void* probs, void* dprobs, int64_t batch_heads, int T, float scale,
                            void* stream);

/* Fused causal attention (flash-style): scores never leave the chip.  qkv: packed
 * [B, T, 3, H, D] bf16 (the qkv linear's output), D = 64, T a multiple of 128.
 * fwd: out [B, T, H*D] bf16, lse [B, H, T] fp32 (natural-log sum-exp of the scaled scores).
 * bwd: dqkv [B, T, 3, H, D] bf16 from dout [B, T, H*D]; `dq_acc` is an fp32 [B, H, T, D]
 * workspace (zeroed by the call), `delta` an fp32 [B, H, T] workspace. */
int b200_flash_attn_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, int D, void* stream);
int b200_flash_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                        float* dq_acc, float* delta, int B, int T, int H, int D, void* stream);

/* ---------------------------------------------------------------- convolution as GEMM
 * For BASELINE.json config 3 (ResNet-50).  ref: none in the reference itself; its
 * ResNet-50 baseline calls cuDNN.  NHWC bf16 tensors, C a multiple of 8.
 * im2col: col[(n,p,q), (r,s,c)] = x[n, p*stride+r-pad, q*stride+s-pad, c] (zero outside);
 * the forward/wgrad GEMMs of b200_gemm consume it.  col2im: the adjoint gather (no
 * atomics) that turns the dgrad GEMM's output back into dx[n,h,w,c].  1x1 stride-1
 * convolutions need neither. */
int b200_im2col(const void* x, void* col, int N, int H, int W, int C, int R, int S, int stride,
                int pad, void* stream);
int b200_col2im(const void* dcol, void* dx, int N, int H, int W, int C, int R, int S, int stride,
                int pad, void* stream);

/* ---------------------------------------------------------------- batch normalisation
 * For BASELINE.json config 3 (ResNet-50).  ref: none in the reference; its ResNet-50
 * baseline uses ATen batch_norm (+ separate ReLU / add kernels).  x, y, residual, dy, dx,
 * dres: [rows, C] bf16 (NHWC activations flattened), C a power of two in [8, 2048];
 * gamma, beta, mean, rstd, running_*: fp32 [C].
 * fwd: training != 0 computes batch statistics (and updates running_* when non-null, unbiased
 * variance, `momentum`), otherwise mean/rstd must already hold the inference statistics
 * (b200_batchnorm_eval_rstd).  y = relu?((x - mean) * rstd * gamma + beta (+ residual)).
 * relu_mask (optional, uint8 [rows * C/8]): with relu the forward writes bit k of byte v = [y[8v+k] > 0];
 * the backward reads it instead of y (1/16 of the bytes) when it is passed, else y itself.
 * bwd: g = dy * [y > 0] when relu; dgamma/dbeta [C] bf16 or fp32; dx; dres = g when non-null
 * (gradient of the fused residual branch).  scratch: b200_batchnorm_scratch_bytes(C). */
int64_t b200_batchnorm_scratch_bytes(int C);
int b200_batchnorm_fwd(const void* x, const void* residual, void* y, void* relu_mask, const float* gamma,
                       const float* beta,
                       float* mean, float* rstd, float* running_mean, float* running_var, int64_t rows,
                       int C, float eps, float momentum, int training, int relu, void* scratch,
                       void* stream);
int b200_batchnorm_bwd(const void* dy, const void* x, const void* y, const void* relu_mask,
                       const float* mean, const float* rstd,
                       const float* gamma, void* dx, void* dres, void* dgamma, void* dbeta, int grads_bf16,
                       int64_t rows, int C, int relu, void* scratch, void* stream);
int b200_batchnorm_eval_rstd(const float* running_var, float* rstd, int C, float eps, void* stream);

/* ---------------------------------------------------------------- peer-memory collectives
 * Symmetric allocation: every rank allocates the same size and exchanges the
 * 64-byte CUDA IPC handle through the process group (the only host-side
 * collective involved); b200_symm_open maps a peer's allocation into this
 * process.  b200_symm_alloc allocates (cudaMalloc) and zero-fills.
 * ref: train.py:233 DDP(model) — torch's Reducer stages gradients through
 * bucket buffers and calls ProcessGroup::allreduce; here peers read each other's
 * gradient arenas directly over NVLink. */
#define B200_MAX_RANKS 8
int b200_symm_alloc(int64_t bytes, void** ptr, void* handle64);
int b200_symm_open(const void* handle64, void** ptr);
int b200_symm_close(void* ptr);
int b200_symm_free(void* ptr);

typedef struct {
    int rank, world;
    void* grad[B200_MAX_RANKS];  /* rank j's gradient arena as mapped in this process */
    void* push[B200_MAX_RANKS];  /* rank j's parameter arena (two-shot writes it) or null */
    void* flags[B200_MAX_RANKS]; /* rank j's flag block: (2*B200_MAX_RANKS+2) uint32, zero-initialised */
    void* grad_mc;               /* NVLS multicast address of all gradient arenas, or null */
    void* push_mc;               /* NVLS multicast address of all parameter arenas, or null */
    int timeout_ms;              /* spin bound before the kernel traps; 0 = 30 s */
    int max_blocks;              /* 0 = default */
    /* Sharded algorithms (1, 2) with a bf16 push keep the fp32 master current only on the shard
     * owner.  f32_mask (device, one byte per 8 arena elements) marks the parameters whose fp32
     * value forward reads (biases, norm scales/shifts): for those the kernel also stores the fp32
     * master into every rank's master_peer[] arena.  Null = no such parameters. */
    void* master_peer[B200_MAX_RANKS];
    const void* f32_mask;
    /* Double-buffered gradients: the producer of step t (1-based) writes the arena at byte offset
     * ((t-1)&1) * parity_stride and the kernel reads there.  0 = single buffer.  With two buffers a
     * one-shot kernel (algo 0) needs no exit barrier: set no_exit_wait. */
    int64_t parity_stride;
    int no_exit_wait;
} b200_comm;
#define B200_COMM_FLAG_WORDS (2 * B200_MAX_RANKS + 2)

/* One bucket [start, start+n) (elements, multiples of 8) of the gradient arena:
 * all-reduce over peer memory, scale by hp->grad_scale, optimizer update, all in
 * one kernel.  ref: train.py:138 (DDP hook all-reduce + divide) and :139
 * (optimizer.step).  algo 0 = one-shot (every rank updates its full replica),
 * algo 1 = two-shot (rank r updates shard r of master/exp_avg/exp_avg_sq and
 * writes the new parameters into every rank's `push` arena: bf16 when
 * push_is_bf16, else fp32).  master/exp_avg/exp_avg_sq/shadow are arena BASE
 * pointers (the kernel indexes them with the same element offsets as grad).
 * algo 2 = NVLS: two-shot sharding with multimem.ld_reduce (in-switch reduction)
 * and multimem.st (in-switch broadcast); needs grad_mc/push_mc.
 * Must be launched by every rank, in the same order, or peers time out. */
int b200_allreduce_optimizer(const b200_comm* c, int algo, const b200_opt_hparams* hp,
                             int64_t start, int64_t n, int grad_is_bf16, float* master,
                             float* exp_avg, float* exp_avg_sq, void* shadow_bf16,
                             int push_is_bf16, int* step_dev, int step_host, int advance,
                             void* stream);

/* Plain all-reduce of c->grad[*][start, start+n) * scale.  algo 0 writes the
 * result to `out` (local, indexed like the arena); algo 1 works in place on
 * every rank's arena.  ref: train.py:276 (metric all-reduce) and the NCCL
 * all-reduce sweep of BASELINE.json. */
int b200_allreduce(const b200_comm* c, int algo, int64_t start, int64_t n, int is_bf16, void* out,
                   float scale, void* stream);

/* ---------------------------------------------------------------- embedding (GPT-2, BASELINE.json config 4)
 * forward: out[b, t, :] = wte[idx[b, t], :] + wpe[t, :] (bf16, C % 8 == 0; ids are clamped to [0, V)).
 * backward: g_wte[v, :] += sum of dx over the tokens with id v — on top of what is already there (the
 * tied LM-head gradient) — and g_wpe[t, :] = sum_b dx[b, t, :] (rows T..wpe_rows-1 zeroed).  Either
 * gradient pointer may be null; each is bf16 or fp32 per its flag.  Deterministic: ids are sorted
 * on the device and every row is summed in position order, no atomics.  `scratch`:
 * b200_embedding_scratch_bytes(B*T) bytes, any contents.
 * ref: train.py:136,138 at GPT-2 scale; torch: aten::embedding / embedding_dense_backward. */
int b200_embedding_fwd(const int64_t* idx, const void* wte, const void* wpe, void* out, int64_t tokens,
                       int T, int C, int V, void* stream);
int64_t b200_embedding_scratch_bytes(int64_t tokens);
int b200_embedding_bwd(const int64_t* idx, const void* dx, void* g_wte, int g_wte_bf16, void* g_wpe,
                       int g_wpe_bf16, int Bsz, int T, int wpe_rows, int C, int V, void* scratch,
                       void* stream);

/* ---------------------------------------------------------------- fused MLP training step
 * The reference's model (ref train.py:32-50: 784-256-256-10 with ReLU + dropout) and its step
 * (ref train.py:135-139: zero_grad, forward, criterion, backward, optimizer.step) as TWO launches:
 * one 8-CTA-cluster kernel for forward + cross-entropy + backward (csrc/mlp_fused.cu: tcgen05 MMAs
 * on TMA-loaded weight tiles, activations kept transposed in shared memory, split-K partials through
 * an L2-resident scratch ordered by hardware cluster barriers) and the fused all-reduce + optimizer
 * kernel.  Shapes: hidden width 256 (two hidden layers), in_features a multiple of 8 up to 1024,
 * up to 16 classes, batch up to 64 per call; other MLPs run the GEMM-per-layer program.
 * The session object owns the TMA descriptors and, for the end-to-end path, a device input ring and
 * a mapped pinned loss ring (allocated by b200_mlp_create / b200_mlp_host_init, nothing afterwards). */
#define B200_MLP_MAX_BUCKETS 4
typedef struct {
    int in_features, hidden, classes;
    const void *W1, *W2, *W3;          /* bf16 [hidden,in], [hidden,hidden], [classes,hidden] */
    const float *b1, *b2, *b3;         /* fp32 */
    float *gW1, *gb1, *gW2, *gb2, *gW3, *gb3;   /* fp32 gradient arena slices (buffer 0) */
    int64_t grad_parity_stride;        /* bytes to gradient buffer 1; step t (1-based) writes buffer (t-1)&1 */
    void* scratch;                     /* b200_mlp_scratch_bytes() bytes, any contents */
    float* loss_out;                   /* [2]: mean loss, number of correct rows */
    int* step_dev;                     /* optimizer step counter {t, ticket} (dropout stream, parity) */
    float dropout_p1, dropout_p2;
    uint64_t seed;
    /* the update chained after the backward by b200_mlp_train / b200_mlp_host_step */
    b200_opt_hparams hp;
    float *master, *grad, *exp_avg, *exp_avg_sq;   /* arena bases */
    void* shadow;                      /* bf16 parameter arena base */
    int n_buckets;
    int64_t bucket_start[B200_MLP_MAX_BUCKETS], bucket_n[B200_MLP_MAX_BUCKETS];
    int bucket_algo[B200_MLP_MAX_BUCKETS];
    const b200_comm* comm;             /* null: single GPU (flat optimizer kernel) */
    b200_comm comm_copy;               /* filled by b200_mlp_create */
} b200_mlp_desc;
int64_t b200_mlp_scratch_bytes(void);
int b200_mlp_create(void** session, const b200_mlp_desc* d);
int b200_mlp_destroy(void* session);
/* Number of extra 8-CTA clusters that run the update (all-reduce + optimizer) inside the step launch;
 * 0 = the update is launched separately after the step kernel (more than one bucket, or the launch
 * would not be fully resident).  Env B200DDP_MLP_UPDATE_CLUSTERS overrides the default (0 disables). */
int b200_mlp_update_clusters(void* session);
/* forward (+ backward when train) on device inputs x fp32 [bsz,in], y int64 [bsz]; no update.
 * logits_out: optional fp32 [bsz, classes]. */
int b200_mlp_forward_backward(void* session, const float* x, const int64_t* y, int bsz, int train,
                              float* logits_out, void* stream);
/* `count` full training steps over a device-resident pool of batches (x_base [pool,bsz,in],
 * y_base [pool,bsz]); step i uses batch (first + i) % pool. */
int b200_mlp_train(void* session, const float* x_base, const int64_t* y_base, int bsz, int64_t pool,
                   int64_t first, int64_t count, void* stream);
/* end-to-end path (ref train.py:133,141): pinned host batch in, previous step's loss out */
int b200_mlp_host_init(void* session, int bsz, int slots);
int b200_mlp_host_step(void* session, const void* host_x, const void* host_y, int bsz, float* prev_loss,
                       int* prev_valid, void* stream);
int b200_mlp_host_finish(void* session, float* last_loss, int* valid, void* stream);
/* `n` steps on `n` pinned host batches (host_x[i] fp32 [bsz,in], host_y[i] int64 [bsz]) in one call:
 * every step's inputs are copied host -> device and its loss comes back into losses[i], but a launch
 * runs up to 256 steps — the kernel itself waits for each step's input copy.  Needs the update CTAs
 * (b200_mlp_update_clusters > 0) and b200_mlp_host_init; may be mixed with b200_mlp_host_step calls. */
int b200_mlp_host_run(void* session, const void* const* host_x, const void* const* host_y, int64_t n, int bsz,
                      float* losses, void* stream);

/* Tuning aids (tools/prof_mlp.py, tools/fa_check.py); they synchronise the device.
 * b200_mlp_debug_trace: clock64 of thread 0 of the 8 compute CTAs at 24 phase boundaries of the last
 * fused-step launch made with B200DDP_MLP_TRACE set (out: long long[8 * 24]).
 * b200_mlp_debug_utrace: globaltimer (ns) marks of the compute cluster and of the first / last update
 * CTA for the last two single-launch steps (out: unsigned long long[34]).
 * b200_flash_debug_trace: clock64 of one softmax thread per step with B200DDP_FA_TRACE set (long long[128]). */
int b200_mlp_debug_trace(long long* out);
int b200_mlp_debug_utrace(unsigned long long* out);
int b200_flash_debug_trace(long long* out);

/* Broadcast of [offset_bytes, offset_bytes + nbytes) of rank `root`'s c->push arena into every
 * rank's (multiples of 16 bytes).  One kernel on every rank with an entry and an exit flag barrier;
 * capturable like the other collectives.  ref: train.py:136 (DDP forward -> _sync_buffers, the
 * per-forward broadcast of module buffers). */
int b200_broadcast(const b200_comm* c, int root, int64_t offset_bytes, int64_t nbytes, void* stream);

/* ---------------------------------------------------------------- symmetric memory, VMM flavour
 * cuMemCreate allocations shared as POSIX file descriptors and the NVLS multicast object bound over
 * them (csrc/vmm.cu).  The caller moves the descriptors between processes (SCM_RIGHTS).  All sizes
 * are the padded size from b200_vmm_padded_size, identical on every rank.  These calls allocate.
 * ref: what c10d provides behind train.py:71,233; torch: CUDASymmetricMemory. */
int b200_vmm_support(int* out2);                       /* out[0]: fd-shareable VMM, out[1]: multicast */
int b200_vmm_padded_size(int64_t bytes, int ranks, int64_t* out);
int b200_vmm_create(int64_t bytes, void** mapping, void** ptr);
int b200_vmm_export(void* mapping, int* fd);
int b200_vmm_import(int fd, int64_t bytes, void** mapping, void** ptr);
int b200_vmm_release(void* mapping);
int b200_mc_create(int64_t bytes, int ranks, void** mapping);
int b200_mc_import(int fd, int64_t bytes, void** mapping);
int b200_mc_add_device(void* mc);
int b200_mc_bind_map(void* mc, void* local_mapping, void** mc_ptr);
int b200_close_fd(int fd);

/* ---------------------------------------------------------------- pooling (NHWC bf16, C % 8 == 0)
 * BASELINE.json config 3 (ResNet-50); no counterpart in the reference (its baseline runs ATen's
 * pooling kernels).  idx: one byte per output element, the winning tap r*S + s (first maximum in
 * scan order, as torch); the backward is a deterministic gather. */
int b200_maxpool_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, int R, int S, int stride,
                     int pad, void* stream);
int b200_maxpool_bwd(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, int R, int S,
                     int stride, int pad, void* stream);
/* global average pool: x [N, HW, C] -> y [N, C]; backward dx = dy / HW */
int b200_avgpool_fwd(const void* x, void* y, int N, int HW, int C, void* stream);
int b200_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream);

/* ---------------------------------------------------------------- host step runner
 * One call per training step on the end-to-end path (ref train.py:133,141: `.to(device)` x2, the
 * step, `loss.item()`): H2D of the pinned host inputs on a copy stream, staging, cudaGraphLaunch of
 * a captured step, D2H of the loss, and a wait for the PREVIOUS step's loss.  csrc/runner.cu. */
typedef struct {
    void* graph_exec;          /* cudaGraphExec_t of the captured step */
    void* stream;              /* compute stream the graph is launched on (0 = default stream) */
    int mode;                  /* 0: D2D copy of the inputs into dst[]; 1: MLP staging kernel
                                * (dst[0] = bf16 x, dst[1] = int64 labels, clears the zero regions) */
    int slots;                 /* device input ring, 2..4 */
    int64_t in_bytes[2];       /* bytes of input 0 (fp32 x / tokens) and input 1 (labels; 0 = none) */
    void* dst[2];              /* static device buffers the captured step reads */
    const float* loss_dev;     /* device scalar the captured step writes */
    float* zero_ptrs[B200_STAGE_MAX_ZERO];   /* mode 1: fp32 regions cleared by the staging kernel */
    int zero_counts[B200_STAGE_MAX_ZERO];
    int nzero;
} b200_runner_desc;
int b200_runner_create(void** out, const b200_runner_desc* d);
/* *prev_valid = 1 and *prev_loss = loss of the step before this one (0 on the first call) */
int b200_runner_step(void* runner, const void* host_in0, const void* host_in1, float* prev_loss, int* prev_valid);
int b200_runner_finish(void* runner, float* last_loss, int* valid);
int b200_runner_destroy(void* runner);

#ifdef __cplusplus
}
#endif
#endif /* B200DDP_H */
