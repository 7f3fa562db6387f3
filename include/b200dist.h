/* b200dist — C ABI of the sm_100a kernel library (libb200dist.so).
 *
 * Every function returns 0 on success; on failure it returns non-zero and the text is
 * available from b200_last_error() (per calling thread). All pointers are device pointers
 * unless stated otherwise; `stream` is a cudaStream_t passed as void*.
 *
 * The Python package loads this library with ctypes (jax_distributed_tuts_b200/ops/_lib.py);
 * nothing here depends on PyTorch.
 */
#ifndef B200DIST_H_
#define B200DIST_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_ABI_VERSION 17

const char* b200_last_error(void);
int b200_abi_version(void);
/* Digest of the sources this library was built from (csrc/, include/), or "unstamped". The Python
 * loader compares it with the tree it runs from and rebuilds a library that does not match. */
const char* b200_source_hash(void);
/* Number of kernels launched by this library since the last reset (host-side counter). */
long long b200_launch_count(void);
void b200_reset_launch_count(void);
/* Programmatic dependent launch for the launches that follow (it is a launch attribute, so inside a
 * stream capture it is baked into the graph): 1 = on, 0 = off, -1 = default (on, unless the
 * environment has B200DIST_NO_PDL=1, which always wins). Back-to-back short kernels gain from the
 * overlapped prologue; long persistent kernels do not, and early-launched CTAs only tie up SMs. */
void b200_set_pdl(int mode);

/* ------------------------------------------------------------------------------------------
 * GEMM on tcgen05 tensor cores with TMEM accumulators and TMA-fed operand tiles.
 *
 *   D[M,N] (+)= epilogue( A[M,K] * B[K,N] ),  bf16 inputs, fp32 accumulation.
 *
 * Operand storage (all row-major bf16, `ld*` = row pitch in elements, multiple of 8):
 *   a_mn = 0: A stored [M][K] (K contiguous)      a_mn = 1: A stored [K][M] (M contiguous)
 *   b_mn = 0: B stored [N][K] (K contiguous)      b_mn = 1: B stored [K][N] (N contiguous)
 * This covers the three products of a Dense layer whose kernel is kept in the reference's
 * (in, out) layout (reference: data_paral.py:86-98):
 *   forward  y  = x  * W      a_mn=0, b_mn=1
 *   dgrad    dx = dy * W^T    a_mn=0, b_mn=0
 *   wgrad    dW = x^T * dy    a_mn=1, b_mn=1   (accumulate=1 sums minibatches in place,
 *                                               reference: util.py:69-77)
 *
 * Sharded B (FSDP all-gather fused into the consuming GEMM; reference: param_sharding.py:129-175
 * gathers the weight first and multiplies afterwards): B may be given as `b_nsrc` equally sized
 * shards that live on different GPUs (peer-mapped pointers). The TMA producer reads each shard in
 * place, so no gathered copy of the weight is ever materialised.
 *   b_shard_mode = 0  single B (b_src[0])
 *   b_shard_mode = 1  shards split K: shard s holds K rows/cols [s*b_shard_len, (s+1)*b_shard_len)
 *   b_shard_mode = 2  shards split N: shard s holds N rows/cols [s*b_shard_len, (s+1)*b_shard_len)
 *   b_rot             shard visited first (callers pass their own rank so that peers are pulled
 *                     evenly).
 * ------------------------------------------------------------------------------------------ */
enum { B200_ACT_NONE = 0, B200_ACT_SILU = 1, B200_ACT_GELU = 2 };

typedef struct b200_gemm_args {
  const void* a;
  const void* b_src[8];
  void* d;
  int32_t m, n, k;
  int32_t lda, ldb, ldd;
  int32_t a_mn, b_mn;
  int32_t b_nsrc, b_shard_mode, b_shard_len, b_rot;
  int32_t d_fp32;     /* 1: D is fp32, 0: D is bf16 */
  int32_t accumulate; /* 1: D += result (fp32 D only) */
  float alpha;        /* result scale applied to the accumulator first */
  /* forward epilogue: v = alpha*acc + bias; pre = v; v = act(v); v = dropout(v); v += resid */
  const void* bias; /* bf16 [N] or NULL */
  void* pre;        /* bf16 [M][ldpre] or NULL: saved pre-activation */
  int32_t ldpre;
  int32_t act; /* B200_ACT_* */
  const void* resid; /* bf16 [M][ldres] or NULL */
  int32_t ldres;
  /* backward epilogue (bwd = 1): v = alpha*acc * dropout_mask_scale * act'(aux) */
  int32_t bwd;
  const void* aux; /* bf16 [M][ldaux]: the forward's saved pre-activation */
  int32_t ldaux;
  /* dropout (either direction); disabled when drop_key == NULL */
  const uint64_t* drop_key; /* device pointer to the 64-bit key state */
  uint64_t drop_salt;       /* host constant: seed = mix64(*drop_key ^ drop_salt) */
  float drop_rate;
  int32_t block_n; /* 0 = choose, else 64/128/256 */
  /* Gradient accumulation folded into one GEMM: rows [i*drop_rows_per_salt, (i+1)*...) form
   * minibatch i with its own RNG stream drop_salts[i] and element index restarting at 0, so the
   * masks equal those of a loop over minibatches (reference: util.py:51,58-66).
   * drop_nsalt = 0 uses drop_salt for all rows. */
  int32_t drop_nsalt, drop_rows_per_salt;
  uint64_t drop_salts[16];
  /* 0 = choose, 1 = single-CTA 128-row tiles, 2 = CTA pairs (cta_group::2, 256-row tiles: the two
   * SMs of a TPC share the B operand, one third less L2->SM traffic per flop) */
  int32_t pair_mode;
  /* split-K for fp32 outputs: 0 = choose (when the M x N tile grid cannot fill the chip),
   * 1 = never, n > 1 = that many K ranges. Partials are summed with red.global.add. */
  int32_t split_k;
  /* Column sums of the stored output, added atomically: dbias[n] += sum_m D[m][n] (fp32 [N], 16-byte
   * aligned, caller zeroes it). Lets a dgrad GEMM emit the bias gradient of the layer it feeds
   * without a separate pass over D. NULL = off. */
  float* dbias;
  /* Softmax cross-entropy fused into the epilogue of a classifier head with n <= 32 classes (the
   * reference's Dense(num_classes) -> optax.softmax_cross_entropy_with_integer_labels,
   * data_paral.py:150-167): D receives dlogits = (softmax(A*B + bias) - onehot(label)) * xent_scale
   * as bf16 instead of the logits, xent_metrics[0] += sum of the row losses, xent_metrics[1] +=
   * number of rows whose argmax is the label, and dbias (if given) the column sums of the stored
   * dlogits. Forward only, no activation / dropout / residual. NULL labels = off. */
  const int64_t* xent_labels;
  float* xent_metrics;
  float xent_scale;
} b200_gemm_args;

int b200_gemm_bf16(const b200_gemm_args* args, void* stream);

/* ------------------------------------------------------------------------------------------
 * Softmax cross-entropy with integer labels, forward + backward in one pass
 * (reference: data_paral.py:181-187).
 *   logits fp32 [rows][ld_logits]; labels int64 [rows]
 *   dlogits (bf16, pitch ld_dlogits, may be NULL) = (softmax - onehot) * grad_scale
 *   metrics (fp32[2], accumulated atomically): [0] += sum of per-row losses, [1] += #correct
 *   row_loss (fp32 [rows], may be NULL)
 * ------------------------------------------------------------------------------------------ */
int b200_softmax_xent(const float* logits, int ld_logits, const int64_t* labels, int rows,
                      int classes, float grad_scale, void* dlogits, int ld_dlogits,
                      float* metrics, float* row_loss, void* stream);
/* Same, and dbias[c] += sum_rows dlogits[row][c] (the head's bias gradient; caller zeroes it). */
int b200_softmax_xent_dbias(const float* logits, int ld_logits, const int64_t* labels, int rows,
                            int classes, float grad_scale, void* dlogits, int ld_dlogits,
                            float* metrics, float* row_loss, float* dbias, void* stream);
/* bf16-logit variant for large vocabularies (GPT-2 head): logits are overwritten in place by
 * dlogits. */
int b200_softmax_xent_bf16_inplace(void* logits_bf16, int ld, const int64_t* labels, int rows,
                                   int classes, float grad_scale, float* metrics, float* row_loss,
                                   void* stream);

/* Column sums: out[N] (+)= sum_rows x[rows][ld] (bias gradients). x bf16, out fp32. */
int b200_colsum_bf16(const void* x, int ld, int rows, int cols, float* out, int accumulate,
                     void* stream);

/* ------------------------------------------------------------------------------------------
 * AdamW (optax.adamw semantics; reference: data_paral.py:106-108,214-217)
 * over a flat fp32 buffer, fused with the gradient mean-scale and the bf16 working-copy cast.
 *   step_count: device int32 holding the number of updates done so far (t = count + 1);
 *   incremented by the kernel when bump_step != 0.
 * ------------------------------------------------------------------------------------------ */
typedef struct b200_adamw_hparams {
  float lr, beta1, beta2, eps, weight_decay, grad_scale;
} b200_adamw_hparams;

int b200_adamw(float* p, float* m, float* v, const float* g, void* p_bf16, int64_t n,
               const b200_adamw_hparams* hp, int32_t* step_count, int bump_step, int zero_grad,
               void* stream);

/* fp32 -> bf16 cast of a flat buffer. */
int b200_cast_f32_bf16(const float* src, void* dst, int64_t n, void* stream);
int b200_cast_bf16_f32(const void* src, float* dst, int64_t n, void* stream);
/* x[i] = scale * x[i] (fp32) / fill with zero */
int b200_scale_f32(float* x, int64_t n, float scale, void* stream);

/* RNG key evolution on the device: key[0] <- mix64(key[0] ^ salt_next); out_step[0] <-
 * mix64(old ^ salt_step). Mirrors `rng, step_rng = split(state.rng)` (data_paral.py:199). */
int b200_rng_split(uint64_t* key, uint64_t* out_step, uint64_t salt_next, uint64_t salt_step,
                   void* stream);

/* The same plus `*step_count += 1` (end-of-step bookkeeping in one launch). */
int b200_rng_split_bump(uint64_t* key, uint64_t* out_step, uint64_t salt_next, uint64_t salt_step,
                        int32_t* step_count, void* stream);
/* Start-of-step housekeeping in one launch: dst = bf16(src) and two small fp32 regions zeroed
 * (either may be NULL / 0). */
int b200_cast_zero(const float* src, void* dst, int64_t n, float* z0, int64_t n0, float* z1, int64_t n1,
                   void* stream);

/* Activation / dropout as a stand-alone op (used when it cannot ride in a GEMM epilogue).
 * y = dropout(act(x)) ; bwd: dx = dy * mask_scale * act'(x). bf16 in/out. */
int b200_act_dropout_fwd(const void* x, void* y, int64_t rows, int cols, int ld, int act,
                         const uint64_t* drop_key, uint64_t drop_salt, float drop_rate,
                         void* stream);
int b200_act_dropout_bwd(const void* x, const void* dy, void* dx, int64_t rows, int cols, int ld,
                         int act, const uint64_t* drop_key, uint64_t drop_salt, float drop_rate,
                         void* stream);

/* ------------------------------------------------------------------------------------------
 * Transformer ops (not in the reference; required by BASELINE.json configs 2-5).
 * ------------------------------------------------------------------------------------------ */
/* LayerNorm over the last dim. x,y bf16 [rows][cols]; gamma/beta bf16 [cols];
 * mean/rstd fp32 [rows] saved for backward. */
int b200_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean,
                       float* rstd, int rows, int cols, float eps, void* stream);
/* dx bf16 = LN-backward(dy) (+ dresid if not NULL: the gradient of the residual branch; must not
 * alias dx); dgamma/dbeta fp32 [cols] accumulated (+=). */
int b200_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* mean,
                       const float* rstd, const void* dresid, void* dx, float* dgamma, float* dbeta,
                       int rows, int cols, void* stream);
/* Same, and dxsum[c] += sum_rows dx[row][c] (fp32 [cols], accumulated): dx is the gradient of the
 * residual stream, whose column sums are the bias gradient of the projection that wrote into it. */
int b200_layernorm_bwd_dxsum(const void* dy, const void* x, const void* gamma, const float* mean,
                             const float* rstd, const void* dresid, void* dx, float* dgamma, float* dbeta,
                             float* dxsum, int rows, int cols, void* stream);
/* out[b,t,:] = wte[tok[b,t],:] + wpe[t,:]   (bf16 tables, int64 tokens) */
int b200_embedding_fwd(const int64_t* tokens, const void* wte, const void* wpe, void* out,
                       int batch, int seq, int dim, void* stream);
/* The same lookup over row-sharded tables that live on several GPUs (FSDP): shard s of a table
 * holds rows [s*rows_per_shard, (s+1)*rows_per_shard) and is read through its peer mapping. */
int b200_embedding_fwd_sharded(const int64_t* tokens, void* const* wte_peers, int n_wte,
                               int64_t wte_rows_per_shard, void* const* wpe_peers, int n_wpe,
                               int64_t wpe_rows_per_shard, void* out, int batch, int seq, int dim,
                               void* stream);
/* dwte[tok] += dout ; dwpe[t] += dout   (fp32 grads, atomics) */
int b200_embedding_bwd(const int64_t* tokens, const void* dout, float* dwte, float* dwpe, int batch,
                       int seq, int dim, void* stream);
/* Causal self-attention on tcgen05 (S = QK^T and O = PV on the tensor cores with TMEM
 * accumulators, K/V tiles by TMA, online softmax in registers on rows read with tcgen05.ld).
 * qkv bf16 [batch][seq][3][heads][64]; out bf16 [batch][seq][heads][64];
 * lse fp32 [batch][heads][seq] (log2 units of the scaled scores). head_dim is 64; seq must be a
 * multiple of 128. */
int b200_attention_fwd(const void* qkv, void* out, float* lse, int batch, int seq, int heads,
                       void* stream);
/* Backward: dqkv bf16, same layout as qkv. `scratch` is caller-provided device memory of
 * b200_attention_bwd_scratch_bytes() bytes, 16-byte aligned: the fp32 dQ accumulator that the
 * kv-tile CTAs add into, and the row dot products delta = sum(dO*O). It is overwritten. */
long long b200_attention_bwd_scratch_bytes(int batch, int seq, int heads);
int b200_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse,
                       void* dqkv, float* scratch, int batch, int seq, int heads, void* stream);

/* Attention backward for callers that cannot allocate (this entry point has no counterpart in
 * the reference). Same maths and outputs as b200_attention_bwd, but the scratch sits in a fixed
 * device array inside the library, so the caller passes only inputs and outputs. Meant for hosts
 * without a device allocator of their own (a bare C program, a plugin) and for runtimes that
 * plan device memory once and forbid cudaMalloc afterwards.
 *   limit   batch*seq*heads <= B200_ATTN_BWD_NOALLOC_MAX_ROWS, else an error is returned;
 *   cost    ~4.1 MiB of device memory per GPU for the life of the process;
 *   sharing one array per GPU, reused by every call: calls that overlap on the GPU race, so keep
 *           them on one stream (or order them with events); nothing needs pre-zeroing, each call
 *           clears what it accumulates into;
 *   set-up  the first call on a device resolves the array's address; do it before any stream
 *           capture begins.
 * It performs no cudaMalloc/cudaFree and calls no allocating library routine. */
#define B200_ATTN_BWD_NOALLOC_MAX_ROWS 16384
int b200_attention_bwd_noalloc(const void* qkv, const void* out, const void* dout, const float* lse,
                               void* dqkv, int batch, int seq, int heads, void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused collective + optimizer kernel over NVSwitch (csrc/comm.cu).
 *
 * All ranks launch the same kernel on buffers that live in symmetric memory (same size on every
 * rank, every rank's copy mapped into every process; a multicast/NVLS address when the fabric
 * has one). One launch does, for a list of segments of the flat gradient buffer:
 *
 *   "gradients ready" exchange  every block stores its epoch into the signal buffer of every peer
 *                  (one-way, st.relaxed.sys) and polls its own buffer for theirs (ld.acquire.sys)
 *   g  = sum over ranks of grad[seg]            multimem.ld_reduce (in-switch) or P2P loads
 *   g *= grad_scale                             (mean over devices and minibatches)
 *   AdamW on this rank's fp32 master / m / v for the segment
 *   publish: new bf16 (and, on request, fp32) parameters   multimem.st / P2P stores to every rank  (DP)
 *            or bf16 shard written locally, to be pulled by the all-gather GEMM                    (FSDP)
 *   metrics psum (optional, epoch-tagged 8-byte stores into the peers' signal buffers)
 *   fence.sys + "published" counter increment in every peer's signal buffer. The kernel does NOT wait
 *   for the peers' increments: that half of the end barrier belongs to whoever reads the published
 *   parameters (or overwrites the gradients) next, i.e. the first kernel of the next step:
 *   b200_cast_zero_wait or b200_comm_wait_published.
 *
 * Reference semantics: DP  = pmean(grads) then replicated apply_gradients
 *                            (data_paral.py:210-217);
 *                      FSDP = psum_scatter(g)/axis_size into the shard, sharded apply_gradients
 *                            (param_sharding.py:134-138,353).
 * ------------------------------------------------------------------------------------------ */
typedef struct b200_seg {
  int64_t src_off;   /* element offset of the segment in every rank's gradient buffer        */
  int64_t src_pitch; /* elements between consecutive rows (== row_len when contiguous)       */
  int64_t dst_off;   /* broadcast mode: element offset in every rank's param buffers         */
  int64_t loc_off;   /* element offset in this rank's local m / v (/ p_shard / bf16_shard)   */
  int32_t rows, row_len;
  int32_t flags;     /* B200_SEG_*                                                           */
  int32_t reserved;
} b200_seg;
/* A segment must span fewer than 2^31 elements, pitch included (rows * max(row_len, src_pitch)): the
 * kernel keeps 32-bit offsets relative to the segment. The engines assert it when they build the lists. */
/* The segment is reduced by several ranks redundantly (a replicated leaf of a sharded run): sum it
 * with peer loads in rank order instead of the in-switch reduction, so all replicas get the same bits. */
#define B200_SEG_ORDERED 1

typedef struct b200_comm_ctx {
  int32_t rank, world;
  void* const* peer_grad;      /* host array[world]: fp32 gradient buffers                   */
  void* const* peer_param;     /* host array[world]: fp32 master buffers (broadcast mode)    */
  void* const* peer_bf16;      /* host array[world]: bf16 working copies (broadcast mode)    */
  void* const* peer_signal;    /* host array[world]: uint32 signal buffers (B200_SIGNAL_WORDS)*/
  void* mc_grad;               /* multicast addresses of the same buffers, or NULL           */
  void* mc_param;
  void* mc_bf16;
} b200_comm_ctx;

#define B200_SIGNAL_WORDS 8192
#define B200_COMM_MAX_BLOCKS 128

/* broadcast = 1 (DP):   p/bf16 written to all ranks at dst_off; m,v local at loc_off.
 * broadcast = 0 (FSDP): p_local/bf16_local/m/v are this rank's shard buffers at loc_off.
 * segs: device array of nseg segments. step_metrics / metrics_acc: optional fp32[nmetrics]
 * (nmetrics <= 8): metrics_acc += sum over ranks of step_metrics. */
int b200_fused_reduce_adamw(const b200_comm_ctx* ctx, const b200_seg* segs_dev, int nseg,
                            int broadcast, float* p_local, void* bf16_local, float* m, float* v,
                            const b200_adamw_hparams* hp, int32_t* step_count, int bump_step,
                            const float* step_metrics, float* metrics_acc, int nmetrics,
                            int blocks, void* stream);
/* Same, with the end-of-step bookkeeping folded in (one launch less per step): the thread block
 * that finishes last sets *rng_key = mix64(*rng_key ^ salt_next) (what b200_rng_split_bump does)
 * and increments *step_count. done_word: a device uint32 that is 0 before the first call; the
 * kernel leaves it at 0. end == NULL behaves like b200_fused_reduce_adamw. */
typedef struct b200_step_end {
  uint64_t* rng_key;
  uint64_t salt_next;
  int32_t* step_count;
  uint32_t* done_word;
} b200_step_end;
int b200_fused_reduce_adamw_end(const b200_comm_ctx* ctx, const b200_seg* segs_dev, int nseg,
                                int broadcast, float* p_local, void* bf16_local, float* m, float* v,
                                const b200_adamw_hparams* hp, int32_t* step_count, int bump_step,
                                const float* step_metrics, float* metrics_acc, int nmetrics,
                                int blocks, const b200_step_end* end, void* stream);
/* The same launch described by one struct (what the engines call). publish_fp32 = 0 keeps the fp32
 * master of a DP slice on its owner only (the bf16 working copy is still stored to every rank): one
 * third of the publish traffic; readers of the full fp32 tree gather the owners' slices first. */
typedef struct b200_fused_update_args {
  const b200_comm_ctx* ctx;
  const b200_seg* segs_dev;
  int32_t nseg, broadcast, publish_fp32, nmetrics, blocks;
  float* p_local;
  void* bf16_local;
  float* m;
  float* v;
  const b200_adamw_hparams* hp;
  const int32_t* step_count;
  const float* step_metrics;
  float* metrics_acc;
  const b200_step_end* end; /* or NULL */
} b200_fused_update_args;
int b200_fused_update(const b200_fused_update_args* args, void* stream);
/* `nranks` ranks emulated on ONE GPU: a single cooperative launch whose blockIdx.y selects the rank,
 * every rank's "peer" pointers being ordinary local buffers (no multicast). It runs the same device
 * code as the real launch, flag protocol included, so the cross-rank logic is testable on a one-GPU
 * box. scratch_dev: b200_fused_update_emulated_scratch_bytes() bytes of device memory. */
int b200_fused_update_emulated(const b200_fused_update_args* ranks, int nranks, void* scratch_dev, void* stream);
long long b200_fused_update_emulated_scratch_bytes(void);

/* Second half of the split-phase end barrier as a kernel of its own (the engines fold it into
 * b200_cast_zero_wait instead): returns when every publish this rank has issued so far has landed
 * from every peer too. signal_local: this rank's own signal buffer. */
int b200_comm_wait_published(const void* signal_local, int rank, int world, void* stream);
/* b200_cast_zero preceded by that wait inside the same kernel. signal_local = NULL or world = 1: no wait.
 * peer_signal (host array[world]) is needed only in the deferred-publish mode below. */
int b200_cast_zero_wait(const float* src, void* dst, int64_t n, float* z0, int64_t n0, float* z1, int64_t n1,
                        const void* signal_local, void* const* peer_signal, int rank, int world, void* stream);
/* How the update kernel orders its parameter stores before the "published" increments
 * (B200DIST_DEFER_PUBLISH in the environment, or this call; -1 = back to the environment):
 *   2 (default)  one thread per block executes fence.sys behind a block barrier (the fence is cumulative);
 *   0            every thread executes fence.sys (measured ~5 us slower on a 70 us step);
 *   1            deferred: the update kernel ends without fence or increments; the next step's
 *                b200_cast_zero_wait sends the signal (the update's stores are complete at its kernel
 *                boundary) and then waits. Relies on full kernel-boundary ordering between the two
 *                kernels (CUDA graph replays, plain stream order), not on programmatic dependent launch. */
void b200_comm_set_defer_publish(int on);
int b200_comm_get_defer_publish(void);
/* How long a rank polls for a peer before the kernel traps (default 30 s, or the environment's
 * B200DIST_COMM_TIMEOUT_MS): a dead peer makes the survivors' launches fail instead of hang. */
void b200_comm_set_timeout_ms(long long ms);
/* Profiling aid: device buffer of uint64 [blocks][8]; thread 0 of every block of the fused update
 * stores %globaltimer at kernel start, after the exchange, after the update loop, after the fence +
 * publish, at the end. NULL switches it off. Read at launch (capture) time. */
void b200_comm_set_stamps(void* dev_u64);

/* Collectives without an optimizer on symmetric-memory buffers (comm/collectives.py backend "fused";
 * reference: psum/pmean data_paral.py:212,223, psum_scatter param_sharding.py:136, all_gather :140).
 *   mode 0  all-reduce in place:  buf[i] = scale * sum_p buf_p[i] on every rank       (fp32, n % (4*world) == 0)
 *   mode 1  reduce-scatter:       out[i] = scale * sum_p buf_p[rank*n/world + i]      (fp32, out local, n/world elements)
 *   mode 2  all-gather:           out[p*n + i] = buf_p[i] in 16-byte units            (n units per rank, out local)
 * Complete when the kernel ends (flag exchanges on both sides). */
typedef struct b200_collective_args {
  int32_t rank, world, mode, blocks;
  void* const* peer_buf;     /* host array[world] */
  void* const* peer_signal;  /* host array[world] */
  void* mc_buf;              /* multicast address of buf or NULL */
  void* out;
  int64_t n;
  float scale;
} b200_collective_args;
int b200_fused_collective(const b200_collective_args* args, void* stream);
int b200_fused_collective_emulated(const b200_collective_args* ranks, int nranks, void* scratch_dev, void* stream);

/* All-gather through peer loads (no collective launch): for every segment and every rank p, `rows`
 * rows of `row_bytes` bytes are read from peer_base[p] + src_off (row pitch src_pitch) and written
 * to dst + dst_off + p*peer_stride (row pitch dst_pitch). Row shards: rows = 1 and peer_stride =
 * row_bytes; column shards: peer_stride = row_bytes and dst_pitch = the full row. The FSDP engines
 * use it for sharded biases and, when the consuming GEMM has many row tiles, to assemble a layer's
 * weights once per use, prefetched one layer ahead (peer memory is not cached in the local L2, so
 * the in-GEMM gather would otherwise re-pull each weight tile for every row tile).
 * Reference: all_gather at param_sharding.py:140. */
typedef struct b200_gather_seg {
  int64_t src_off, dst_off, row_bytes, rows, src_pitch, dst_pitch, peer_stride;
} b200_gather_seg;
int b200_peer_gather(void* const* peer_base, int world, const b200_gather_seg* segs_dev, int nseg,
                     void* dst, int blocks, void* stream);
/* Cross-GPU barrier on the signal buffers (every rank launches it): one flag exchange. */
int b200_signal_barrier(const b200_comm_ctx* ctx, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200DIST_H_ */
