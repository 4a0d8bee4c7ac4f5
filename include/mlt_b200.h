/* libmlt_b200: hand-written sm_100a kernels + native runtime of the ml-trainer B200 framework.
 *
 * Plain C ABI.  Conventions:
 *   - every function returns 0 on success; on failure a non-zero code, with a message available from
 *     mlt_last_error() (thread-local);
 *   - pointers are device pointers unless the name says otherwise; `stream` is a cudaStream_t;
 *   - nothing here allocates device memory unless its comment says so: callers pass outputs and scratch.
 * The comments cite the reference call site (file:line under /root/reference) each entry point replaces.
 */
#ifndef MLT_B200_H_
#define MLT_B200_H_

#ifdef __cplusplus
extern "C" {
#endif

#define MLT_B200_VERSION 100

/* activation ids for fused epilogues */
#define MLT_ACT_NONE 0
#define MLT_ACT_RELU 1
#define MLT_ACT_GELU 2
#define MLT_ACT_TANH 3

/* optimizer rule ids: reference optimizer table, src/trainer.py:123-138 */
#define MLT_RULE_SGD 0
#define MLT_RULE_ADAM 1
#define MLT_RULE_ADAMW 2
#define MLT_RULE_ADAGRAD 3
#define MLT_RULE_ADAMAX 4
#define MLT_RULE_MEAN_ONLY 5 /* no update: the fused all-reduce kernels just return the mean gradient */

/* ---- library ------------------------------------------------------------------------------------ */
const char* mlt_last_error(void);
int mlt_version(void);
int mlt_device_info(int device, int* sm_count, int* cc_major, int* cc_minor, long long* total_mem);

/* Host-side gradient bucket planner (replaces DDP's bucket assignment behind src/trainer.py:98).
 * out_bounds gets nb+1 entries; returns nb. */
int mlt_plan_buckets(const long long* sizes_bytes, int n, long long cap_bytes, long long first_cap_bytes,
                     int* out_bounds);

/* ---- optimizer: src/trainer.py:188 optimizer.step() --------------------------------------------- */
/* One fused launch over n elements (n % 4 == 0): grad scale + cast + rule + master write + bf16 shadow write.
 * hyper_dev points at 8 device floats {lr, wd, b1|momentum, b2, eps, bias_corr1, sqrt(bias_corr2), grad_scale}. */
int mlt_fused_update(int rule, void* master_f32, const void* grad, int grad_is_bf16, void* shadow_bf16, void* s1,
                     void* s2, const void* hyper_dev, long long n, void* stream);

/* ---- loss / metric: src/trainer.py:185-186 (criterion, loss.item()) and :165-166 (accuracy) ------ */
/* Mean softmax cross-entropy + dlogits (=(softmax-onehot)/B) + optional accuracy, and accumulation into
 * acc3 = {loss_sum, metric_sum, steps}.  workspace16: 16 zero-initialised bytes, self-resetting. */
int mlt_softmax_xent(const void* logits, int is_bf16, const void* targets_i64, void* dlogits, void* loss_out,
                     void* acc3, void* workspace16, int B, int C, int want_acc, void* stream);
int mlt_argmax_accuracy(const void* logits, int is_bf16, const void* targets_i64, void* out_scalar, int B, int C,
                        void* stream);
/* The reference's other criteria and metric, fused like the cross-entropy (src/trainer.py:140-150, 160-163).
 * mlt_regression_loss: kind 0 = L1Loss, 1 = MSELoss (also custom_loss_function, src/utils/functions.py:15-17); loss, its
 * gradient w.r.t. pred (NULL: forward only) and, for rows of C, the mcrmse metric; acc3 = {loss_sum, metric_sum, steps}.
 * workspace: mlt_regression_loss_ws_bytes(C) bytes, zero before the first call, left zero by every call.
 * mlt_nll_loss: NLLLoss on log-probabilities + gradient + arg-max accuracy; 16-byte workspace as mlt_softmax_xent. */
long long mlt_regression_loss_ws_bytes(int C);
int mlt_regression_loss(int kind, const void* pred, int pred_bf16, const void* target, int target_bf16, void* dpred,
                        void* loss_out, void* acc3, void* workspace, long long ws_bytes, long long n, int C, int want_mcrmse,
                        void* stream);
int mlt_nll_loss(const void* logp, int is_bf16, const void* targets_i64, void* dlogp, void* loss_out, void* acc3,
                 void* workspace16, int B, int C, int want_acc, void* stream);

/* ---- small-shape CNN path: src/model.py:18-23 (conv/relu/pool/fc of MLModel) and their backward ---- */
int mlt_conv2d_direct_fwd(const void* x, const void* w, const void* bias, void* y, void* pool_idx, int B, int Cin,
                          int H, int W, int Cout, int KH, int KW, int relu, int pool, void* stream);
long long mlt_conv2d_direct_bwd_scratch_bytes(int B, int Cin, int Cout, int KH, int KW);
/* scratch: mlt_conv2d_direct_bwd_scratch_bytes() bytes, zero-initialised once (tickets self-reset). */
int mlt_conv2d_direct_bwd(const void* x, const void* w, const void* dy, const void* y, const void* pool_idx, void* dx,
                          void* dw, void* db, void* scratch, long long scratch_bytes, int B, int Cin, int H, int W,
                          int Cout, int KH, int KW, int relu, int pool, int accumulate, void* stream);
/* Same computation as mlt_conv2d_direct_bwd for callers that cannot allocate device memory (a plug-in inside an
 * application that already gave the whole GPU to its own pool; a loop that cannot afford cudaMalloc's implicit
 * synchronisation).  The reduction scratch is a MLT_CONV2D_BWD_NOALLOC_SCRATCH_BYTES array built into the library
 * (resident on every GPU that loads it, until exit); larger problems are rejected with an error.  The first call on a
 * device resolves the array's address: make it before any stream capture.  Calls share the array, so calls whose GPU
 * work overlaps race.  This entry point has no counterpart in the reference. */
#define MLT_CONV2D_BWD_NOALLOC_SCRATCH_BYTES (256 * 1024)
int mlt_conv2d_direct_bwd_noalloc(const void* x, const void* w, const void* dy, const void* y, const void* pool_idx,
                                  void* dx, void* dw, void* db, int B, int Cin, int H, int W, int Cout, int KH, int KW,
                                  int relu, int pool, int accumulate, void* stream);
/* is_bf16: dtype of x / w / bias / y (0 fp32, 1 bf16); grads_bf16: dtype of the dw / db sinks. */
int mlt_linear_small_fwd(const void* x, const void* w, const void* bias, void* y, int B, int K, int N, int act,
                         int is_bf16, void* stream);
int mlt_linear_small_bwd(const void* dy, const void* y, const void* x, const void* w, void* dx, void* dw, void* db,
                         int B, int K, int N, int act, int accumulate, int is_bf16, int grads_bf16, void* stream);

/* ---- tensor-core GEMM: the cuBLAS calls behind nn.Linear (src/model.py:21-23) and every BERT / ResNet-FC GEMM ------ */
/* D[M,N] = A * B^T on tcgen05 (bf16 in, fp32 accumulate in TMEM, TMA-fed), fused epilogue:
 *   + bias[N] (fp32 or bf16), activation, aux_mode 1: also store the pre-activation (bf16) to aux,
 *   aux_mode 2: multiply by act'(aux) instead (the dgrad-through-activation epilogue).
 * a_mn = 0: A stored [M][K];  a_mn = 1: A stored [K][M].  b_mn = 0: B stored [N][K];  b_mn = 1: B stored [K][N].
 * Row strides (elements) must be multiples of 8.  block_n: 0 = auto, or 64 / 128 / 256.
 * splitk_ws (optional fp32 scratch): when the output has few tiles and K is long (weight gradients) the reduction is
 * split over up to 16 K-slices whose partials land in the scratch and are summed by a second kernel. */
int mlt_gemm_bf16(const void* A, long long lda, int a_mn, const void* B, long long ldb, int b_mn, void* out,
                  long long ldo, int out_f32, int M, int N, int K, const void* bias, int bias_f32, int act, int aux_mode,
                  void* aux, long long ld_aux, int block_n, void* splitk_ws, long long splitk_ws_bytes, void* stream);

/* nb1*nb2 independent GEMMs sharing M,N,K; operand (z1,z2) starts at base + z1*s1 + z2*s2 elements (attention heads). */
int mlt_gemm_bf16_batched(const void* A, long long lda, long long a_s1, long long a_s2, int a_mn, const void* B,
                          long long ldb, long long b_s1, long long b_s2, int b_mn, void* out, long long ldo, long long o_s1,
                          long long o_s2, int out_f32, int M, int N, int K, int nb1, int nb2, int block_n, void* stream);

/* ---- bandwidth-bound helpers around the GEMMs (ATen elementwise / sum kernels in the reference's backward) ------- */
int mlt_act_backward(const void* dy, const void* saved, void* dz, long long n, int act, int from_pre, int is_bf16,
                     void* stream);
long long mlt_colsum_scratch_bytes(int M, int N);
int mlt_colsum(const void* x, int x_bf16, void* out, int out_bf16, void* scratch, long long scratch_bytes, int M, int N,
               long long ld, int accumulate, void* stream);
int mlt_cast(const void* src, int src_bf16, void* dst, int dst_bf16, long long n, void* stream);
/* fp32 [R, C] (row pitch ldx) -> the tripled bf16 operand of the three-term split GEMM (x = hi + lo):
 * mode 0: out [R, 3C] = [t0 | t1 | t2] along the columns, mode 1: out [3R, C] stacked along the rows;
 * pattern 0: (hi, hi, lo), pattern 1: (hi, lo, hi).  One operand takes pattern 0, the other pattern 1. */
int mlt_split3_bf16(const void* x_f32, long long ldx, void* out_bf16, long long R, long long C, int mode, int pattern, void* stream);

/* ---- communication runtime: replaces init_process_group + DistributedDataParallel (src/trainer.py:59,98) --------- */
/* Symmetric arena: one VMM allocation per rank, exported as a POSIX fd, passed to the peers over unix sockets
 * (SCM_RIGHTS) and mapped by every rank at its own address; optionally bound to an NVSwitch multicast object.
 * Protocol (the caller supplies a barrier between phases, e.g. a gloo barrier):
 *   h = mlt_symm_create(...)  | barrier | mlt_symm_exchange(h) | barrier |
 *   mlt_symm_multicast_setup(h, phase 0) | barrier | (h, phase 1) | barrier | (h, phase 2) | barrier
 * These calls allocate device memory (they are set-up, not hot path). */
void* mlt_symm_create(int rank, int world, int device, long long bytes, const char* job_tag);
int mlt_symm_exchange(void* handle);
int mlt_symm_multicast_setup(void* handle, int phase);   /* returns 0 ok, >0 error, -1 multicast unsupported */
void* mlt_symm_peer_ptr(void* handle, int peer);           /* arena base of `peer` as mapped in this process */
void* mlt_symm_multicast_ptr(void* handle);                /* NULL when NVLS is unavailable */
long long mlt_symm_bytes(void* handle);
int mlt_symm_destroy(void* handle);

typedef struct {
  void* base[8];              /* arena base of every rank, as mapped in this process */
  void* mc_base;              /* multicast mapping of the arena, or NULL */
  int world, rank;
  long long flags_off;        /* byte offset of the flag block (mlt_p2p_flag_block_bytes()) */
  long long recv_off;         /* byte offset of the one-shot receive slots: [2][world][recv_slot_bytes] */
  long long recv_slot_bytes;
  int mc_flags;               /* != 0 (needs mc_base): flags and the in-step gradient push go out as ONE multimem.st each */
  int reserved;
} mlt_p2p_ctx;

long long mlt_p2p_flag_block_bytes(void);
/* Fused all-reduce (mean via hyper.grad_scale) + optimizer update of one gradient bucket.
 * algo 0 one-shot push | 1 two-shot P2P | 2 NVLS multimem.  grad/model offsets are byte offsets in the arena;
 * master/s1/s2/hyper are arrays with one local pointer per rank handled by this launch (1 in production;
 * `ranks_here` > 1 emulates several ranks on one GPU with a cooperative launch, for tests). */
int mlt_p2p_allreduce_update(const mlt_p2p_ctx* ctx, int algo, int rule, int grad_is_bf16, long long grad_off,
                             long long model_off, int model_is_master, void* const* master, void* const* s1,
                             void* const* s2, void* const* hyper, long long n, int nblocks, int ranks_here, void* stream);
/* The same, with the bucket's own byte offset inside the one-shot receive slots (buckets of one step must not share slot
 * bytes: their CTAs' double-buffer parities are independent).  recv_bucket_off + bucket bytes <= recv_slot_bytes. */
int mlt_p2p_allreduce_update_at(const mlt_p2p_ctx* ctx, int algo, int rule, int grad_is_bf16, long long grad_off,
                                long long model_off, int model_is_master, void* const* master, void* const* s1,
                                void* const* s2, void* const* hyper, long long n, int nblocks, int ranks_here,
                                long long recv_bucket_off, void* stream);
int mlt_p2p_broadcast(const mlt_p2p_ctx* ctx, long long off, long long nbytes, int root, int ranks_here, void* stream);
int mlt_p2p_reduce_scalars(const mlt_p2p_ctx* ctx, long long scalar_off, void* vals_dev, int k, int ranks_here, void* stream);
int mlt_p2p_check_error(const mlt_p2p_ctx* ctx, unsigned int* out6_host);
/* Epoch-end form: k floats in pinned host memory `pin` (>= 160 bytes) are summed over all ranks in place and this rank's timeout
 * diagnostic is read in the same round trip (one stream synchronisation).  Returns like mlt_p2p_check_error. */
int mlt_p2p_reduce_scalars_host(const mlt_p2p_ctx* ctx, long long scalar_off, void* vals_dev_table, void* vals_dev, void* pin, int k,
                                void* stream);
/* wait limit of every flag wait of this rank's collectives in seconds (<= 0: the default, ~20 s).  When a wait runs past it the
 * kernel records a diagnostic, poisons this rank and its peers (later collectives return without touching weights) and
 * mlt_p2p_check_error reports it. */
int mlt_p2p_set_timeout(const mlt_p2p_ctx* ctx, double seconds);

/* ---- the reference model's whole train step as one kernel: src/model.py:7-24 + src/trainer.py:181-188 ------------- */
long long mlt_lenet_scratch_bytes(int B);
/* forward + cross-entropy + accuracy + backward + (multi-GPU) in-kernel one-shot all-reduce + optimizer rule.
 * param_off: arena element offsets of conv1.w, conv1.b, conv2.w, conv2.b, fc1.w, fc1.b, fc2.w, fc2.b, fc3.w, fc3.b.
 * scratch: mlt_lenet_scratch_bytes(B), zero-initialised once.  ctx NULL (or world 1) = single GPU. */
int mlt_lenet_step(void* master, void* s1, void* s2, const void* hyper, long long numel, const long long* param_off,
                   int rule, const void* x, const void* y_i64, int B, int train, int want_acc, void* acc3,
                   void* loss_out, void* scratch, long long scratch_bytes, const mlt_p2p_ctx* ctx, void* stream);

/* ---- image augmentation on the device: the reference's CIFAR pipeline (src/utils/functions.py:5-12) as one kernel ---------- */
typedef struct {
  int C, H, W;                /* source image (uint8) */
  int out_h, out_w;           /* output window (== H, W for random crops) */
  int pad;                    /* zero padding before the crop (RandomCrop(size, padding)) */
  int flip;                   /* != 0: horizontal flip with probability 1/2 */
  int hwc;                    /* != 0: source is [N, H, W, C] (CIFAR's own layout), else [N, C, H, W] */
  float mean[4], std[4];      /* Normalize(mean, std) on the [0, 1] image */
  unsigned long long seed;
} mlt_augment_cfg;
/* uint8 batch on the device -> normalised [B, C, out_h, out_w] fp32 / bf16; sample b draws from stream position pos0 + b;
 * train == 0: centre window, no flip. */
int mlt_augment_u8(const void* src_u8, void* dst, int dst_bf16, int B, const mlt_augment_cfg* cfg, unsigned long long pos0, int train,
                   void* stream);
/* the random draw of one stream position: out3 = {top, left, flip} */
int mlt_augment_params(const mlt_augment_cfg* cfg, unsigned long long pos, int* out3);

/* ---- native epoch executor + data feeder: the Python hot loop of src/trainer.py:180-197, in C++ ------------------ */
typedef struct {            /* fixed arguments of mlt_lenet_step; x / y / B come from the executor's ring slots */
  void *master, *s1, *s2;
  const void* hyper;
  long long numel;
  const long long* param_off;
  int rule, train, want_acc;
  void *acc3, *loss_out, *scratch;
  long long scratch_bytes;
  const mlt_p2p_ctx* ctx;
} mlt_lenet_call;

/* Allocates (set-up only): `slots` pinned staging buffers + device buffers for max_batch samples, a copy stream, events. */
void* mlt_loop_create(int device, int slots, long long max_batch, long long x_row_bytes, long long y_row_bytes);
void* mlt_loop_slot_x(void* loop, int slot);
void* mlt_loop_slot_y(void* loop, int slot);
int mlt_loop_num_slots(void* loop);
int mlt_loop_set_direct_rows(void* loop, int on);
/* uint8 image datasets: every step first runs mlt_augment_u8 from the slot's uint8 rows into `out_dev` ([max_batch, C, out_h, out_w],
 * owned by the caller) and the step reads `out_dev`.  cfg NULL switches it off.  pos0: stream position of the next run's first
 * sample; train as in mlt_augment_u8. */
int mlt_loop_set_augment(void* loop, const mlt_augment_cfg* cfg, void* out_dev, int out_bf16, unsigned long long pos0, int train);   /* x rows are page-locked: gather by DMA (one copy per sample) */
int mlt_loop_destroy(void* loop);
/* Runs `steps` steps: host gather -> pinned -> H2D (copy stream) -> hyper row H2D -> step -> loss D2H, ring-buffered.
 * step_kind 0: single-kernel LeNet step (`lenet`);  1: one captured cudaGraphExec_t per slot (`graph_execs`). */
int mlt_loop_run(void* loop, const void* host_x, const void* host_y, const long long* indices, long long steps, int batch,
                 const int* batch_sizes, const float* hyper_rows, void* hyper_dev, int step_kind, const mlt_lenet_call* lenet,
                 void* const* graph_execs, const void* loss_dev, float* loss_host, void* compute_stream);
int mlt_loop_stats(void* loop, long long* h2d_bytes, long long* d2h_bytes, long long* launches);
/* host-side timing of the last run, microseconds since entry: {first step enqueued, all steps enqueued, device drained} */
int mlt_loop_timing(void* loop, double* out3);
/* ... and how long the enqueueing thread waited for the gather threads (MLT_GATHER_THREADS, default 2) in that run */
double mlt_loop_gather_wait_us(void* loop);

/* ---- transformer pieces: ATen layer_norm / softmax / embedding kernels of the BERT workload --------------------- */
/* y = LN(x [+ res]) * gamma + beta (bf16 activations; gamma/beta fp32 or bf16); hsum (optional) receives x + res. */
int mlt_layernorm_fwd(const void* x, const void* res, const void* gamma, const void* beta, int params_bf16, void* y,
                      void* hsum, void* mean, void* rstd, int T, int H, float eps, void* stream);
long long mlt_layernorm_bwd_scratch_bytes(int T, int H);
int mlt_layernorm_bwd(const void* dy, const void* h, const void* mean, const void* rstd, const void* gamma, int params_bf16,
                      void* dx, void* dgamma, void* dbeta, int grads_bf16, void* scratch, long long scratch_bytes, int T,
                      int H, void* stream);
/* RMSNorm (+ residual add) forward / one-pass backward (csrc/rmsnorm.cu): bf16 [T, H] activations, H a multiple of 8 up to 8192,
 * gamma fp32 or bf16, rstd fp32 [T].  hsum (optional, with res) receives the rounded pre-norm sum h = x + res.
 * backward: dx and dgamma from one read of dy and h; scratch = mlt_rmsnorm_bwd_scratch_bytes(), zeroed once, left zero. */
int mlt_rmsnorm_fwd(const void* x, const void* res, const void* gamma, int params_bf16, void* y, void* hsum, void* rstd, int T, int H,
                    float eps, void* stream);
long long mlt_rmsnorm_bwd_scratch_bytes(void);
int mlt_rmsnorm_bwd(const void* dy, const void* h, const void* rstd, const void* gamma, int params_bf16, void* dx, void* dgamma,
                    int grads_bf16, void* scratch, long long scratch_bytes, int T, int H, void* stream);
int mlt_softmax_rows(void* scores_bf16, long long rows, int L, float scale, void* stream);               /* in place */
int mlt_softmax_rows_bwd(void* dp_bf16, const void* p_bf16, long long rows, int L, float scale, void* stream);
/* Fused attention on tcgen05 (scores never leave the SM).  q, k, v: [B, S, H*64] bf16 with row stride ld_qkv elements
 * (0 = H*64; 3*H*64 when they are column slices of a packed QKV projection); out [B, S, H*64] bf16; lse [B, H, S] fp32 or
 * NULL.  Head dim 64, S a multiple of 64 up to 512, no mask. */
int mlt_attention_fwd(const void* q, const void* k, const void* v, long long ld_qkv, void* out, void* lse, int B, int S, int H,
                      int D, void* stream);
/* Its backward: out / lse from the forward, dout the gradient of out, delta a [B, H, S] fp32 scratch; dq, dk, dv bf16 with
 * row stride ld_dqkv (0 = H*64). */
int mlt_attention_bwd(const void* q, const void* k, const void* v, long long ld_qkv, const void* out, const void* dout,
                      const void* lse, void* delta, void* dq, void* dk, void* dv, long long ld_dqkv, int B, int S, int H, int D,
                      void* stream);
/* The same with a key-padding mask: key_bias [B, S] fp32 (NULL = none) is added to every score of that key after the 1/sqrt(d)
 * scaling - HuggingFace's extended attention mask (0 for tokens, a large negative number for padding).  S: multiple of 64, <= 512. */
int mlt_attention_fwd_masked(const void* q, const void* k, const void* v, long long ld_qkv, const void* key_bias, void* out, void* lse,
                             int B, int S, int H, int D, void* stream);
int mlt_attention_bwd_masked(const void* q, const void* k, const void* v, long long ld_qkv, const void* key_bias, const void* out,
                             const void* dout, const void* lse, void* delta, void* dq, void* dk, void* dv, long long ld_dqkv, int B,
                             int S, int H, int D, void* stream);
/* The same with dropout on the attention probabilities drawn inside the kernels (p_drop in [0, 1), resolution 1/256; 0 = none).
 * The mask is a counter hash of (seed, *step_ptr, call_id, sample, head, query, key): nothing is stored, and the backward call
 * regenerates it from the same three values.  step_ptr: device int64 (NULL = 0), read when the kernel starts. */
int mlt_attention_fwd_ex(const void* q, const void* k, const void* v, long long ld_qkv, const void* key_bias, void* out, void* lse,
                         int B, int S, int H, int D, float p_drop, unsigned long long seed, const void* step_ptr,
                         unsigned int call_id, void* stream);
int mlt_attention_bwd_ex(const void* q, const void* k, const void* v, long long ld_qkv, const void* key_bias, const void* out,
                         const void* dout, const void* lse, void* delta, void* dq, void* dk, void* dv, long long ld_dqkv, int B,
                         int S, int H, int D, float p_drop, unsigned long long seed, const void* step_ptr, unsigned int call_id,
                         void* stream);
/* Analysis hook: device buffer of 8 int64 per CTA that the dQ kernel fills with %globaltimer stamps (NULL clears it). */
int mlt_attention_set_debug(void* buf);
/* Implicit-GEMM convolution on tcgen05 for NHWC bf16 tensors with Cin a multiple of 64: the patch matrix is never written,
 * operand tiles are gathered by TMA im2col loads.  w2 / dw: [Co, KH*KW*Cin]. */
int mlt_conv_fwd_implicit(const void* x, int B, int H, int W, int C, const void* w2, long long ldw, int Co, int KH, int KW, int stride,
                          int pad, void* y, long long ldy, void* stream);
int mlt_conv_wgrad_implicit(const void* dy, long long lddy, const void* x, int B, int H, int W, int C, int Co, int KH, int KW,
                            int stride, int pad, void* dw, long long lddw, int out_f32, void* splitk_ws, long long splitk_ws_bytes,
                            void* stream);
/* Data gradient of a stride-1 convolution as an implicit GEMM over the patches of dy with the filter read flipped;
 * dy [B, OH, OW, Co] (Co a multiple of 64), w2 [Co, KH*KW, C] contiguous, dx [B*H*W, C]. */
int mlt_conv_dgrad_implicit(const void* dy, int B, int OH, int OW, int Co, const void* w2, int C, int KH, int KW, int pad, void* dx,
                            long long lddx, void* stream);
/* Test hook for the TMA im2col path: the [pixels x 64] patch tile of output pixels m0.., filter tap (r, s), channels c0..c0+63 of
 * an NHWC bf16 tensor, loaded with one cp.async.bulk.tensor...im2col and written back un-swizzled. */
int mlt_im2col_probe(const void* x, int B, int H, int W, int C, int KH, int KW, int stride, int pad, long long m0, int r, int s, int c0,
                     int pixels, void* out, void* stream);
/* Column sums of a contiguous bf16 [M, N] matrix, N a multiple of 8 (bias gradients of the GEMM layers): flat 16-byte streams,
 * fp32 atomics fold.  scratch: mlt_colsum_flat_scratch_bytes(N), zeroed once, left zero by every call. */
long long mlt_colsum_flat_scratch_bytes(int N);
int mlt_colsum_flat(const void* x, void* out, int out_bf16, void* scratch, long long scratch_bytes, int M, int N, int accumulate,
                    void* stream);
/* Dropout with a Philox4x32-10 mask that is a function of (seed, *step_ptr, call_id, element): the same call on the upstream
 * gradient is the backward pass; *step_ptr lives in device memory so captured graphs replay with fresh masks. */
int mlt_dropout(const void* x, void* y, long long n, float p, unsigned long long seed, const void* step_ptr, unsigned int call_id,
                int is_bf16, void* stream);
/* `nsteps` LeNet steps in ONE launch (see csrc/lenet_step.cu): ring of nslots device buffers gated by *ready, or a device pool
 * of pool_n batches; per-step batch sizes / hyper rows / losses from device tables; *progress (pinned host) = steps finished. */
int mlt_lenet_steps(void* master, void* s1, void* s2, const void* hyper, long long numel, const long long* param_off, int rule,
                    const void* const* xs, const void* const* ys, int nslots, long long pool_n, int B, const int* batch_sizes,
                    const void* hyper_table, void* loss_base, const void* ready, void* progress, int nsteps, int train,
                    int want_acc, void* acc3, void* scratch, long long scratch_bytes, const mlt_p2p_ctx* ctx, void* stream);
int mlt_lenet_set_debug(void* buf);
/* Analysis hook: {start, end} %globaltimer stamps of every training step (CTA 0), `pairs` x 2 int64 on the device; NULL clears. */
int mlt_lenet_set_stamps(void* buf, int pairs);
int mlt_embedding_sum_fwd(const void* ids_i64, const void* word, const void* pos, const void* type, int params_bf16,
                          void* out_bf16, int T, int S, int H, void* stream);
/* the same, with the row count V of the word table: ids outside [0, V) are clamped instead of read / written out of bounds */
int mlt_embedding_sum_fwd_v(const void* ids_i64, const void* word, const void* pos, const void* type, int params_bf16,
                            void* out_bf16, int T, int S, int H, long long V, void* stream);
int mlt_embedding_bwd(const void* ids_i64, const void* dh_bf16, void* dword, long long V, void* dpos, int grads_bf16, int B,
                      int S, int H, void* stream);

/* ---- ResNet pieces (NHWC bf16): cuDNN convolution / batch-norm / pooling calls of the ResNet-50 workload ---------- */
/* conv = im2col (skipped for 1x1 stride 1) + mlt_gemm_bf16; dgrad = GEMM + col2im gather; wgrad = split-K GEMM. */
int mlt_im2col_nhwc(const void* x, void* col, int B, int H, int W, int C, int KH, int KW, int stride, int pad, int OH, int OW,
                    void* stream);
int mlt_col2im_nhwc(const void* dcol, void* dx, int B, int H, int W, int C, int KH, int KW, int stride, int pad, int OH, int OW,
                    void* stream);
int mlt_nchw_to_nhwc_bf16(const void* x_f32, void* y_bf16, int B, int C, int H, int W, int Cp, void* stream);
int mlt_maxpool_nhwc_fwd(const void* x, void* y, void* idx_u8, int B, int H, int W, int C, int K, int stride, int pad, int OH,
                         int OW, void* stream);
int mlt_maxpool_nhwc_bwd(const void* dy, const void* idx_u8, void* dx, int B, int H, int W, int C, int K, int stride, int pad,
                         int OH, int OW, void* stream);
int mlt_avgpool_nhwc_fwd(const void* x, void* y, int B, int HW, int C, void* stream);
int mlt_avgpool_nhwc_bwd(const void* dy, void* dx, int B, int HW, int C, void* stream);
/* Training BatchNorm over [M = B*H*W, C]: stats (+ running-stat update) -> apply (+ residual, + ReLU); backward in two passes.
 * scratch: mlt_bn_scratch_bytes(M, C), zero-initialised once. */
long long mlt_bn_scratch_bytes(int M, int C);
int mlt_bn_stats(const void* x, void* mean, void* rstd, void* run_mean, void* run_var, float momentum, float eps, void* scratch,
                 long long scratch_bytes, int M, int C, void* stream);
/* relu_mask (optional, M*C/8 bytes): the forward writes one bit per element (output > 0); the backward then reads that byte
 * per 8 elements instead of 16 bytes of y (pass y = NULL there). */
int mlt_bn_apply(const void* x, const void* mean, const void* rstd, const void* gamma, const void* beta, int params_bf16,
                 const void* residual, void* y, void* relu_mask, long long M, int C, int relu, void* stream);
int mlt_bn_bwd(const void* dy, const void* y, const void* relu_mask, const void* x, const void* mean, const void* rstd, const void* gamma, int params_bf16,
               void* dx, void* dres, void* dgamma, void* dbeta, int grads_bf16, void* scratch, long long scratch_bytes, int M, int C,
               int relu, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MLT_B200_H_ */
