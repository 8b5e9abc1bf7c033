/* C ABI of libttd_b200.so — the sm_100a kernel library of tensorflow_train_distributed_b200.
 *
 * Every function returns 0 on success; on failure it returns a non-zero code and
 * ttd_last_error() describes what went wrong. Pointers are device pointers unless a
 * parameter says "host". `stream` is a cudaStream_t passed as void*; all work is
 * enqueued on it and nothing synchronises unless stated.
 *
 * Reference citations are to /root/reference/distribute_training.py (the TF1 script
 * whose behaviour each kernel reproduces).
 */
#ifndef TTD_B200_H_
#define TTD_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TTD_MAX_PEERS 8 /* replicas of one NVLink domain */

const char* ttd_last_error(void);
/* library self-description: "ttd_b200 sm_100a <build stamp>" */
const char* ttd_version(void);
/* number of kernels launched by this library in this process since load (bench "gpu_launches") */
uint64_t ttd_launch_count(void);

/* ---- epilogue modes of the grouped tcgen05 GEMM ------------------------------------------ */
enum {
  TTD_EPI_FWD = 0,      /* out = dropout(act(acc + bias))            (reference :54-58) */
  TTD_EPI_DGRAD = 1,    /* out = acc * dropout_mask/keep * act'(aux) (autodiff of :54-58) */
  TTD_EPI_WGRAD = 2,    /* outf = scale * acc, fp32                  (autodiff of :54-55) */
  TTD_EPI_FWD_XENT = 3, /* logits = acc + bias; mean sparse softmax CE + top-1 + dlogits (reference :77-82, :96, :125) */
  TTD_EPI_WGRAD_UPDATE = 4 /* single replica: weight gradient AND the optimiser step of that variable group in one
                              epilogue — SGD + L1/L2 + EMA on the fp32 master/shadow, bf16 compute copies (`out` = TF
                              layout, `outT` = transposed) and the fp32 bias copy (reference :150-152, :153-161). Net
                              kernel only (ttd_gemm_net_launch_ex carries the hyper-parameters). */
};
enum { TTD_ACT_NONE = 0, TTD_ACT_RELU = 1, TTD_ACT_ELU = 2 };

/* One GEMM problem D[M,N] = A[M,K] * B[N,K]^T. A and B are bf16, K contiguous ("K-major"),
 * leading dimensions in elements and multiples of 8 (TMA needs 16-byte row pitch). */
typedef struct TtdGemmDesc {
  const void* a; int64_t lda;   /* bf16 [M, lda] */
  const void* b; int64_t ldb;   /* bf16 [N, ldb] */
  int32_t M, N, K;
  int32_t block_n;              /* CTA tile width: multiple of 16 in [16, 256]; 0 = choose */
  int32_t mode;                 /* TTD_EPI_* */
  int32_t act;                  /* TTD_ACT_* (FWD: applied; DGRAD: derivative taken w.r.t. aux) */
  float drop_rate;              /* dropout rate of the produced tensor (FWD) / of aux (DGRAD); 0 = off */
  uint32_t layer_id;            /* stream id for the dropout RNG; FWD and its DGRAD must agree */
  const float* bias;            /* fp32 [N] or NULL */
  void* out; int64_t ldo;       /* bf16 [M, ldo] row-major result or NULL */
  void* outT; int64_t ldt;      /* bf16 [N, ldt] transposed result or NULL */
  float* outf; int64_t ldf;     /* fp32 [M, ldf] result (WGRAD: required; FWD*: optional copy of pre-dropout output/logits) */
  const void* aux; int64_t ldaux; /* DGRAD: bf16 [M, ldaux] saved (post-dropout) activation */
  float scale;                  /* WGRAD / XENT gradient scale */
  const int64_t* labels;        /* XENT: int64 [M] */
  float* metrics;               /* XENT: float[2] accumulators: += mean loss, += top-1 accuracy */
  /* Pull-fused B operand (first problem of a group only): K is cut into owner ranges
   * [k_bound[o], k_bound[o+1]) (multiples of 64); the tiles of range o are fetched by TMA from b_peer[o],
   * replica o's peer-mapped copy of `b` (same shape and ld). This is the parameter pull of the first layer:
   * an all-gather of W1 fused into the 784->hidden GEMM (reference M1 traffic, :186-190 placement). */
  const void* b_peer[8];
  int32_t k_bound[9];
  int32_t n_b_peers;            /* 0 = B is read locally */
  /* TTD_EPI_WGRAD_UPDATE: the problem's [M, N] result is the gradient of the variable group [rows+1, cols] */
  int32_t upd_regularize;       /* 1: kernel rows take the L1/L2 term */
  float* upd_w;                 /* fp32 master [M, ldf] (ldf = row pitch of the group in the flat buffer = N) */
  float* upd_ema;               /* fp32 shadow [M, ldf] or NULL */
  float* upd_bias;              /* fp32 [N] compute copy of the bias row or NULL */
  /* Net kernel: A taken from a pool of a_slots matrices of a_slot_rows rows each, stacked along the rows (the
   * transposed batches of a device-resident data set): slot (*step_ptr % a_slots) is read. 0 / 1 = plain A. */
  int32_t a_slot_rows, a_slots;
} TtdGemmDesc;

/* Launch up to 4 independent GEMM problems in ONE kernel (grid = sum of their tiles).
 * tcgen05.mma (cta_group::1, M=128) with TMEM accumulators, TMA-fed, warp-specialised.
 * step_ptr: device uint64 local-step counter keyed into the dropout RNG (may be NULL if no dropout).
 * flags bit0: programmatic dependent launch. */
int ttd_gemm_group(const TtdGemmDesc* descs_host, int count, uint64_t seed, const uint64_t* step_ptr,
                   uint32_t flags, void* stream);

/* Chain of small DEPENDENT problems (K <= 256, N <= 256 each) run back to back by one CTA with the same
 * tcgen05 pipeline and epilogues: the narrow layers behind the first one cost a __syncthreads per GEMM instead
 * of a kernel launch. Build once (synchronous; encodes tensor maps into `device_buf`, which must hold
 * count * ttd_gemm_chain_problem_bytes() bytes, 64-byte aligned), launch every step. flags bit0: PDL. */
uint64_t ttd_gemm_chain_problem_bytes(void);
int ttd_gemm_chain_build(const TtdGemmDesc* descs_host, int count, void* device_buf, void* stream);
int ttd_gemm_chain_launch(const void* device_buf, int count, uint64_t seed, const uint64_t* step_ptr,
                          uint32_t flags, void* stream);

/* ---- stand-alone softmax cross-entropy (any class count) --------------------------------- */
/* logits fp32 [B, ldl]; labels int64 [B]; writes dlogits bf16 [B, ldd] (and transposed [C, ldt] if
 * non-NULL) = scale*(softmax - onehot)/B, and accumulates mean loss / accuracy into metrics[0..1].
 * Reference :77-82 (loss), :96 and :125 (in_top_k accuracy). */
int ttd_softmax_xent(const float* logits, int64_t ldl, const int64_t* labels, int32_t B, int32_t C,
                     void* dlogits, int64_t ldd, void* dlogitsT, int64_t ldt, float scale, float* metrics,
                     void* stream);

/* ---- input staging ------------------------------------------------------------------------ */
/* x fp32 [B, F] -> xb bf16 [B, ldx] and xT bf16 [F, ldt] (row F of xT is the caller's ones row).
 * Device-resident dataset: with pool_batches > 1, x is [pool_batches, B, F] and labels_pool is
 * [pool_batches, B]; batch (*step_ptr % pool_batches) is staged and its labels copied to labels_out.
 * Reference :191-192, :224-226 (the fed batch). */
int ttd_prep_input(const float* x, int32_t B, int32_t F, void* xb, int64_t ldx, void* xT, int64_t ldt,
                   const int64_t* labels_pool, int64_t* labels_out, int32_t pool_batches,
                   const uint64_t* step_ptr, float* xf /* optional fp32 copy [B, F] of the staged batch */, void* stream);
/* Same with launch flags: bit0 = programmatic dependent launch (overlaps the previous kernel's tail). */
int ttd_prep_input_ex(const float* x, int32_t B, int32_t F, void* xb, int64_t ldx, void* xT, int64_t ldt,
                      const int64_t* labels_pool, int64_t* labels_out, int32_t pool_batches,
                      const uint64_t* step_ptr, float* xf, uint32_t flags, void* stream);

/* ---- parameter store: variable table ------------------------------------------------------ */
/* One entry per dense layer group: kernel [rows, cols] immediately followed by bias [cols] in the
 * flat fp32 buffer (so the group is a [rows+1, cols] matrix). */
typedef struct TtdVarDesc {
  int64_t offset;       /* element offset of the group in the flat buffers */
  int32_t rows, cols;   /* kernel shape [in, out] (TF layout, reference A.1) */
  int32_t regularize;   /* 1: the kernel rows take the L1/L2 term (reference :55 — hidden layers only) */
  int32_t reserved;     /* flags; bit0: w_t is owner-local (read remotely by the pull-fused first GEMM), do not broadcast it */
  void* w_rm; int64_t ld_rm;  /* bf16 [rows+1, ld_rm] compute copy, TF layout (dgrad B operand) or NULL */
  void* w_t;  int64_t ld_t;   /* bf16 [cols, ld_t]   transposed compute copy (fwd B operand) or NULL */
  float* bias_c;              /* fp32 [cols] compute copy of the bias (what the forward epilogue reads) or NULL */
  float* w_f32; int64_t ld_f32;   /* fp32 [rows, ld_f32] compute copy, TF layout (fused MLP tail) or NULL */
  float* wt_f32; int64_t ld_tf32; /* fp32 [cols, ld_tf32] transposed compute copy (fused MLP tail) or NULL */
} TtdVarDesc;

typedef struct TtdHyper {
  float lr0;            /* INITIAL_LEARNING_RATE (:14) */
  float decay_rate;     /* LEARNING_RATE_DECAY_FACTOR (:13) */
  int64_t decay_steps;  /* int(60000 / BATCH_SIZE) (:120-122) */
  int32_t staircase;    /* :140 */
  int32_t reg_kind;     /* 0 none, 1 L1, 2 L2 */
  float reg_scale;      /* :39 regularizer_scale */
  float ema_decay;      /* MOVING_AVERAGE_DECAY (:12); <=0 disables the shadow update */
  int32_t ema_warmup;   /* 1: min(decay, (1+n)/(10+n)) because num_updates=global_step (:156) */
  float grad_scale;     /* multiplies the summed gradient (1/replicas_to_aggregate for the sync mean) */
  void* host_mirror;    /* pinned host ring of TtdStepState (device-visible address) or NULL: the kernel that ends a
                           step writes the step's state into slot (local_step - 1) % mirror_slots */
  int32_t mirror_slots;
  int32_t reserved;
} TtdHyper;

/* He / variance-scaling init (reference :49, A.5): truncated normal, sigma = sqrt(1.3*2/fan_in),
 * bias row zero; also fills ema (if non-NULL) with the same values and refreshes the bf16 copies. */
int ttd_init_params(float* w, float* ema, const TtdVarDesc* vars_host, int nvars, uint64_t seed, void* stream);

/* Device-resident step state (one per process, 64 bytes). The host reads it only at log cadence. */
typedef struct TtdStepState {
  int64_t global_step;     /* shared step counter (reference :116); every sync replica keeps an identical copy */
  uint64_t local_step;     /* this worker's step count (reference :217, :236); keys dropout RNG and barrier epochs */
  float metrics_acc[2];    /* running (mean loss, accuracy) of the step in flight */
  float metrics_out[2];    /* (loss, accuracy) of the last completed step (reference :225-226 fetches) */
  uint32_t done_counter;   /* scratch: blocks finished in the current update launch */
  float lr_last;           /* learning rate applied by the last update */
  uint32_t reserved[6];
} TtdStepState;

/* Who contributes gradients and who receives the refreshed bf16 copies. Single GPU: n_grads =
 * n_copies = 1, grads[0] = local gradient buffer, copy_delta[0] = 0, barrier = 0. */
typedef struct TtdPeers {
  const float* grads[TTD_MAX_PEERS];   /* flat gradient buffer of replica r (peer-mapped over NVLink) */
  int64_t copy_delta[TTD_MAX_PEERS];   /* byte distance from a local bf16-copy address to replica r's */
  uint32_t* flags[TTD_MAX_PEERS];      /* replica r's signal pad: u32[grid][TTD_MAX_PEERS], zero-initialised */
  const float* grads_mc;               /* multicast (NVLS) mapping of the gradient buffers, or NULL */
  int64_t mc_delta;                    /* byte distance from a local arena address to its multicast alias; 0 = none */
  int32_t n_grads, n_copies, rank;
  int32_t barrier;                     /* 1: bracket the update with cross-replica barriers */
  int32_t grid;                        /* barrier mode: CTA count, identical on all replicas, <= 592 (all co-resident) */
  int32_t reserved;                    /* bit0: diagnostic phase clocks as the max over all CTAs (host zeroes TtdStepState.reserved) */
} TtdPeers;

/* Fused gradient push over the flat range [lo, hi) owned by the caller:
 *   g = grad_scale * sum_r grads[r][i]  (+ L1/L2 term)   lr from state->global_step (staircase decay)
 *   w -= lr*g ; ema -= (1-d)(ema-w) ; bf16 copies refreshed on every replica in `peers`
 *   last CTA: ++global_step (if bump_global_step bit0), ++local_step, metrics_out <- metrics_acc <- 0.
 *   bump_global_step bit1: launch with programmatic dependent launch.
 * hi < 0: up to the end of the last group. lo/hi need no alignment (a boundary inside a 4-element group of a
 * variable whose row pitch is a multiple of 4 selects the element-wise path for the whole launch).
 * Reference :136-140 (lr), :150-152 (apply + global_step), :153-161 (EMA), :50-55 (regulariser),
 * :144-148 (SyncReplicasOptimizer mean-of-N aggregation when n_grads > 1). */
int ttd_apply_update(float* w, float* ema, const TtdVarDesc* vars_host, int nvars, int64_t lo, int64_t hi,
                     const TtdHyper* hyper_host, const TtdPeers* peers_host, TtdStepState* state,
                     int32_t bump_global_step, void* stream);

/* ---- asynchronous replicas (reference default: plain GradientDescentOptimizer, :149-152) ------------- */
/* Hogwild on a GPU-sharded parameter server: no barrier, no aggregation. Each worker step is
 *   ttd_async_pull   read every owner's fp32 master over NVLink -> local bf16/fp32 compute copies
 *                    (+ the caller's own range: EMA shadow update)
 *   forward / backward
 *   ttd_async_push   w_owner[i] += -lr * (g[i] + reg)  as system-scope NVLink atomics (use_locking=False),
 *                    then ONE atomic increment of the shared global_step held by owner 0. */
typedef struct TtdAsyncPeers {
  float* w_peer[TTD_MAX_PEERS];       /* fp32 master of replica r (peer-mapped) */
  int64_t bounds[TTD_MAX_PEERS + 1];  /* owner k holds [bounds[k], bounds[k+1]) */
  TtdStepState* state0;               /* owner 0's step state (peer-mapped): the shared global_step */
  int32_t n_owners, rank;
} TtdAsyncPeers;
int ttd_async_pull(float* w, float* ema, const TtdVarDesc* vars_host, int nvars, int64_t own_lo, int64_t own_hi,
                   const TtdHyper* hyper_host, const TtdAsyncPeers* peers_host, TtdStepState* state, void* stream);
int ttd_async_push(float* w, const float* grads, const TtdVarDesc* vars_host, int nvars, const TtdHyper* hyper_host,
                   const TtdAsyncPeers* peers_host, TtdStepState* state, void* stream);

/* Diagnostic: two GPUs bounce a flag `iters` times through each other's memory (`mine` is polled,
 * `theirs` is the peer mapping of the other side's flag); *out_ns = total device time. variant selects the
 * ordering recipe: 0 volatile, 1 fence+relaxed, 2 release/acquire, 3 atomicExch, 4 red.release. */
int ttd_diag_pingpong(uint32_t* mine, uint32_t* theirs, int rank, int iters, int variant,
                      unsigned long long* out_ns, void* stream);

/* ---- fused MLP tail ---------------------------------------------------------------------------------- */
/* Every layer behind the first — forward, softmax cross-entropy, dgrad chain and all weight gradients — in ONE
 * launch on a thread-block cluster (fp32 CUDA-core register tiles; phase boundaries are cluster barriers).
 * Reference semantics: :54-61 (layers), :80-82 (loss), :96/:125 (accuracy), :152 (gradients).
 * Leading dimensions are multiples of 4 floats; buffers are zero-initialised and padded as stated. */
#define TTD_TAIL_MAX_LAYERS 6
typedef struct TtdTailLayer {
  int32_t in, out;
  int32_t act_kind;          /* TTD_ACT_* applied to this layer's output (ignored for the last layer) */
  float drop_rate;           /* dropout on this layer's output (ignored for the last layer) */
  uint32_t layer_id;         /* dropout RNG stream */
  int32_t reserved;
  const float* w;  int64_t ld_w;    /* fp32 [pad4(in)][ld_w], TF layout */
  const float* wt; int64_t ld_wt;   /* fp32 [pad4(out)][ld_wt], transposed */
  const float* bias;                /* fp32 [out] */
  float* dw;                        /* fp32 [in+1][out] -> flat gradient buffer (row `in` = bias gradient) */
  float* act;  int64_t ld_act;      /* fp32 [B][ld_act] post-dropout activation; logits for the last layer */
  float* actT; int64_t ld_actT;     /* fp32 [pad4(out+1)][ld_actT] transposed, row `out` all ones (hidden layers) */
  float* dz;   int64_t ld_dz;       /* fp32 [pad4(B)][ld_dz] gradient w.r.t. this layer's pre-activation */
} TtdTailLayer;
typedef struct TtdTailDesc {
  int32_t B, n_layers;
  int32_t cluster;           /* CTAs in the cluster: 0 = 8 (portable maximum), up to 16 */
  int32_t in_act_kind;       /* activation / dropout of the tensor feeding the tail (the first layer's output) */
  float in_drop_rate;
  uint32_t in_layer_id;
  const void* a_in;  int64_t ld_a_in;   /* bf16 [B][ld_a_in] input activations (post-dropout) */
  const void* a_inT; int64_t ld_a_inT;  /* bf16 [in0+1][ld_a_inT] transposed, last row all ones */
  void* dz_in;  int64_t ld_dz_in;       /* out: bf16 [B][ld_dz_in] gradient w.r.t. the first layer's pre-activation */
  void* dz_inT; int64_t ld_dz_inT;      /* out: bf16 [in0][ld_dz_inT] the same, transposed */
  const int64_t* labels;                /* int64 [B] */
  float* metrics;                       /* float[2]: += mean loss, += accuracy */
  const uint64_t* step_ptr;             /* local step (dropout RNG) */
  TtdTailLayer layer[TTD_TAIL_MAX_LAYERS];
} TtdTailDesc;
/* flags bit0: programmatic dependent launch */
int ttd_mlp_tail(const TtdTailDesc* desc_host, uint64_t seed, uint32_t flags, void* stream);

/* ---- column kernel: forward + loss + data gradients of a narrow MLP in ONE launch ------------------------- */
/* The batch is cut into slices of 16 examples; one CTA owns a slice through every layer. Weights are the
 * 128-row tcgen05 operand (layers behind the first stay resident in shared memory, the first layer's are streamed
 * once per cluster by TMA multicast), the slice's activations are the 16-column operand and never leave shared
 * memory between layers: an epilogue writes the next layer's swizzled K-major operand in place. Outputs are the
 * transposed activations / pre-activation gradients the weight-gradient GEMMs read (row `out` of aT is the caller's
 * ones row). Dropout masks use the same Philox stream as the per-layer GEMM epilogues (seed, step, layer, row, col).
 * Limits: <= TTD_COL_MAX_LAYERS layers, hidden widths <= 256, <= 32 classes, operands must fit 227 KB.
 * Reference semantics: :54-61 (layers), :80-82 (loss), :96/:125 (accuracy), :152 (gradients). */
#define TTD_COL_MAX_LAYERS 6
typedef struct TtdColumnLayer {
  int32_t in, out;
  uint32_t layer_id;            /* dropout RNG stream of this layer's output */
  int32_t reserved;
  const void* w_t;  int64_t ld_t;   /* bf16 [out, ld_t]: transposed compute copy (forward operand) */
  const void* w_rm; int64_t ld_rm;  /* bf16 [in, ld_rm]: TF-layout compute copy (data-gradient operand); unused for layer 0 */
  const float* bias;                /* fp32 [out] */
  void* aT;  int64_t ld_aT;         /* out: bf16 [out(+1), ld_aT] post-dropout activation, transposed (hidden layers) */
  void* dzT; int64_t ld_dzT;        /* out: bf16 [out, ld_dzT] gradient w.r.t. the pre-activation, transposed */
} TtdColumnLayer;
typedef struct TtdColumnDesc {
  int32_t B, n_layers;
  int32_t act;                  /* TTD_ACT_* of the hidden layers */
  float drop_rate;              /* dropout on every hidden layer's output */
  int32_t cluster;              /* CTAs sharing the first layer's weight stream: 0 = largest of 8/4/2/1 that divides ceil(B/16) */
  int32_t pool_batches;         /* batches in x_pool / labels_pool; batch (*step_ptr % pool_batches) is used */
  const void* x_pool; int64_t ld_x;   /* bf16 [pool_batches * B, ld_x] input rows */
  const int64_t* labels_pool;         /* int64 [pool_batches * B] */
  float* metrics;                     /* float[2]: += mean loss, += accuracy */
  const uint64_t* step_ptr;           /* device: local step (dropout RNG, pool slot) */
  const int64_t* global_step_ptr;     /* device: global step for the learning-rate schedule, or NULL */
  float* lr_out;                      /* device float[2]: this step's learning rate and EMA decay (NULL: not computed) */
  TtdHyper hyper;                     /* schedule constants for lr_out */
  TtdColumnLayer layer[TTD_COL_MAX_LAYERS];
} TtdColumnDesc;
/* Plans shared memory / TMEM and encodes the tensor maps once; returns an opaque handle (NULL + error on failure). */
void* ttd_mlp_column_create(const TtdColumnDesc* desc_host);
void ttd_mlp_column_destroy(void* handle);
/* flags bit0: programmatic dependent launch. */
int ttd_mlp_column_launch(void* handle, uint64_t seed, uint32_t flags, void* stream);
/* Diagnostic: device buffer of 64 uint64 receiving CTA 0's stage timestamps (globaltimer ns); NULL = off. */
void ttd_mlp_column_set_trace(void* device_buf_64_u64);
/* 1 if a model of these sizes fits the kernel (hidden widths, class count, shared memory), else 0. */
int ttd_mlp_column_fits(int32_t input_dim, const int32_t* layer_sizes, int32_t n_layers);

/* ---- LeNet-5 convolution stages (BASELINE config 5; the reference itself has no CNN) --------------------- */
/* One stage = conv KHxKW stride 1 (SAME or VALID) + bias + activation + 2x2 max-pool, NHWC fp32 activations.
 * Weights: bf16 [KH*KW*Cin, ldw] (the store's TF-layout compute copy of the HWIO kernel), bias fp32 [Cout]. */
typedef struct TtdConvDesc {
  int32_t B, H, W, Cin, Cout, KH, KW;
  int32_t same_pad; /* 1: SAME (pad (K-1)/2), 0: VALID */
} TtdConvDesc;
/* pooled fp32 [B,PH,PW,Cout] (post-activation), argmax u8 (winning tap 0..3 of every pooled cell); optional
 * bf16 flatten [B, ldflat] (+ transposed [PH*PW*Cout, ldflatT]) feeding the first dense layer's GEMM. */
int ttd_conv_pool_fwd(const TtdConvDesc* d, const float* x, const void* w, int64_t ldw, const float* bias,
                      int32_t act, float* pooled, uint8_t* argmax, void* flat, int64_t ldflat, void* flatT,
                      int64_t ldflatT, void* stream);
/* dzp = dL/d(pooled) * act'(pooled), fp32 or bf16 [B, ld_dzp]. Writes dw fp32 [KH*KW*Cin + 1, Cout] (last row =
 * bias gradient) into the flat gradient buffer and, if dx != NULL, dx = dL/dx * act_below'(x_act) (fp32, shape of x). */
int ttd_conv_pool_bwd(const TtdConvDesc* d, const void* dzp, int64_t ld_dzp, int32_t dzp_is_bf16,
                      const uint8_t* argmax, const float* x, const void* w, int64_t ldw, float* dw, float* dx,
                      const float* x_act, int32_t act_below, float* scratch, int64_t scratch_floats, void* stream);
/* scratch (optional, NULL: one CTA per kernel row): zero-initialised once by the caller, then owned by this stage;
 * holds ceil64(K+1) arrival counters followed by slices x (K+1) x 16 partial sums. */

/* cudaMemcpyAsync(device -> device) on `stream`; lets Python snapshot peer-mapped shards for checkpoints. */
int ttd_copy_d2d(void* dst, const void* src, uint64_t nbytes, void* stream);

/* ---- net kernel: every GEMM of a step (forward, data and weight gradients) in ONE launch of co-resident CTAs;
 * a tile is a task, dependent problems hand over through arrival counters in global memory. dep_mask[i] bit d:
 * problem d (< i) must be complete before problem i starts. The buffer (ttd_gemm_net_bytes, 64-byte aligned)
 * is encoded once by _build (synchronous; returns the CTA count to launch with, <= grid <= SM count). */
uint64_t ttd_gemm_net_bytes(int count, int grid, int max_tasks);
int ttd_gemm_net_build(const TtdGemmDesc* descs, int count, const uint32_t* dep_mask, int grid, int max_tasks,
                       void* device_buf, void* stream);
int ttd_gemm_net_launch(const void* device_buf, int count, int grid, int max_tasks, int ctas, uint64_t seed,
                        const uint64_t* step_ptr, uint32_t flags, void* stream);
/* Same launch with the optimiser state for TTD_EPI_WGRAD_UPDATE problems: hyper (regulariser, grad_scale, EMA on/off,
 * host mirror), lr_slot = device float[2] {learning rate, EMA decay} of this step (written by the column kernel),
 * and — finalize bit0 — the last CTA to leave ends the step: ++local_step, metrics_out <- metrics_acc <- 0, host
 * mirror; bit1: ++global_step as well (reference :152). */
int ttd_gemm_net_launch_ex(const void* device_buf, int count, int grid, int max_tasks, int ctas, uint64_t seed,
                           const uint64_t* step_ptr, uint32_t flags, const TtdHyper* hyper_host, const float* lr_slot,
                           TtdStepState* state, int32_t finalize, void* stream);
/* trace buffer: 4 uint64 per problem (24 problems max) + [96] kernel start after the dependency, [97] last CTA done */
void ttd_gemm_net_set_trace(void* device_buf_4_u64_per_problem);

/* ---- step feeder: the host side of a training step in one native call (reference: the C++ session run loop
 * behind mon_sess.run, distribute_training.py:225). Host batches are copied on `copy_stream` into a two-slot
 * device pool (slot = step & 1; the captured step picks local_step % 2 on the device); events order the copy
 * behind the slot's last reader and the step behind the copy. All handles are raw CUDA handles. */
void* ttd_feeder_create(void* compute_stream, void* copy_stream, void* x_slot0, void* x_slot1, void* y_slot0,
                        void* y_slot1, size_t x_bytes, size_t y_bytes);
void ttd_feeder_destroy(void* feeder);
/* Convert every fed slot on the copy stream (fp32 [B,F] -> bf16 rows at xb + slot*B*ldx and transposed bf16 rows at
 * xT + slot*(F+1)*ldt): the input staging leaves the step's critical path. xb = NULL switches it off. */
int ttd_feeder_set_prep(void* feeder, int32_t B, int32_t F, void* xb, int64_t ldx, void* xT, int64_t ldt);
int ttd_feeder_feed(void* feeder, const void* host_x, const void* host_y, uint64_t step);
int ttd_feeder_before_step(void* feeder); /* returns the fed slot or -1; < -1 never; -1 with an error set on failure */
int ttd_feeder_after_step(void* feeder);
/* feed (host_x != NULL) + cudaGraphLaunch(graph_exec) + optional 64-byte TtdStepState read-back to pinned memory */
int ttd_feeder_step(void* feeder, void* graph_exec, const void* host_x, const void* host_y, uint64_t step,
                    const void* state_dev, void* state_host);

/* The same update for hosts that may not allocate device memory (no counterpart in the reference).
 * Meant for callers embedded in a process that has given all of HBM to its own arena, or that cannot
 * afford the device-wide synchronisation of cudaMalloc/cudaFree inside their loop. One gradient source,
 * one GPU, no replica barriers. The kernel's completion counter and the variable table live in a 8 KiB
 * array inside the library (resident on every GPU that loads it, until exit), so at most
 * TTD_STATIC_MAX_VARS groups per call; more is refused with an error and nothing is launched.
 * state->done_counter is not used. All calls on a GPU share that array: calls whose kernels could overlap
 * on the GPU race and must be ordered by the caller. The first call on each GPU looks up the array's
 * device address (cudaGetSymbolAddress); make it before any stream capture begins. The call itself is refused
 * on a capturing stream (it copies the table from host memory every time). hi < 0 means "to the end of the last
 * group", as in ttd_apply_update; hyper->host_mirror is honoured the same way. lo/hi need no alignment. */
#define TTD_STATIC_MAX_VARS 64
int ttd_apply_update_static(float* w, float* ema, const float* grad, const TtdVarDesc* vars_host, int nvars,
                            int64_t lo, int64_t hi, const TtdHyper* hyper_host, TtdStepState* state,
                            int32_t bump_global_step, void* stream);

/* Cross-GPU barrier watchdog: a replica that waits longer than this inside the update kernel's barriers traps
 * (which destroys its CUDA context) instead of hanging the GPU. Default 600 s, or TTD_BARRIER_TIMEOUT_SECS; it must
 * outlast the chief's checkpoint writes, during which the other replicas already wait in their next update. */
void ttd_set_barrier_timeout_ms(uint64_t ms);

/* Rewrite only the bf16 compute copies from the fp32 master (after a restore). */
int ttd_refresh_copies(const float* w, const TtdVarDesc* vars_host, int nvars, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TTD_B200_H_ */
