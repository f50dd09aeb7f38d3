/* libb200gate -- C ABI of the B200-native worker engine behind V-Gate's
 * `model.engine_type: b200` backend.
 *
 * This is the boundary a replacement engine must export.  It stands where the
 * reference's vLLM adapter reaches into its engine:
 *
 *   b200_engine_create / _load_tensor / _finalize   <- VLLMBackend.load_model
 *                                                      (vgate/backends/vllm_backend.py:61-86)
 *   b200_submit + b200_wait                         <- VLLMBackend.generate / _agenerate_one
 *                                                      (vllm_backend.py:97-132)
 *   b200_submit + b200_poll (+ b200_abort)          <- VLLMBackend.stream_generate
 *                                                      (vllm_backend.py:134-146; generator close main.py:372-376)
 *   ttft_s / gen_time_s out-params                  <- _extract_metrics (vllm_backend.py:22-32)
 *   b200_engine_destroy                             <- VLLMBackend.shutdown (vllm_backend.py:148-149)
 *   temperature / top_p / max_tokens arguments      <- create_sampling_params (vllm_backend.py:88-95),
 *                                                      wire form vgate/worker_api.py:70-73
 *
 * Conventions: every call returns 0 (B200_OK) or a negative B200_ERR_* code; no
 * exception crosses the boundary; b200_last_error() gives the text of the last
 * failure on the calling thread.  Plain pointers and sizes only.  The engine owns
 * all of its device memory, one scheduler thread and its CUDA stream.  submit /
 * wait / poll / abort / stats are thread-safe and may be called concurrently from
 * any number of threads (V-Gate calls generate() from executor threads,
 * vgate/worker_api.py:112, vgate/batcher.py:374).
 *
 * The b200_op_* entry points expose each kernel of the path on raw device
 * pointers for parity tests and profilers; they launch on `stream` (a
 * cudaStream_t passed as void*, NULL = legacy default stream) and do not
 * synchronise.
 */
#ifndef B200GATE_H
#define B200GATE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK 0
#define B200_ERR_INVALID (-1)   /* bad argument / unsupported shape */
#define B200_ERR_CUDA (-2)      /* a CUDA call failed; engine is poisoned */
#define B200_ERR_CAPACITY (-3)  /* request can never fit (prompt >= max_model_len, KV pool) */
#define B200_ERR_STATE (-4)     /* call made in the wrong engine state */
#define B200_ERR_NOT_FOUND (-5) /* unknown request id */
#define B200_ERR_ABORTED (-6)   /* request was aborted or engine shut down */

#define B200_DTYPE_F16 0
#define B200_DTYPE_I32 1
#define B200_DTYPE_F32 2
#define B200_DTYPE_BF16 3

#define B200_ABI_VERSION 2

typedef struct b200_engine b200_engine;

/* Model + runtime shape.  Zero-initialise, set struct_size = sizeof(b200_config). */
typedef struct b200_config {
  int32_t struct_size;
  int32_t device;          /* CUDA ordinal inside this process (CUDA_VISIBLE_DEVICES picks the GPU) */
  int32_t hidden;          /* 1536 */
  int32_t intermediate;    /* 8960 */
  int32_t n_layers;        /* 28 */
  int32_t n_heads;         /* 12 */
  int32_t n_kv_heads;      /* 2 */
  int32_t head_dim;        /* 128 (only value supported) */
  int32_t vocab;           /* 151936 */
  int32_t group_size;      /* AWQ group, 128 */
  int32_t tie_embeddings;  /* 1: lm_head = embed_tokens */
  float rms_eps;           /* 1e-6 */
  float rope_theta;        /* 1e6 */
  int32_t max_seqs;        /* concurrent sequences (decode batch ceiling), e.g. 256 */
  int32_t max_model_len;   /* prompt + generated tokens per sequence, e.g. 2048 */
  int32_t max_batch_tokens;/* prefill token budget per step, e.g. 8192 */
  int64_t kv_bytes;        /* KV pool size; 0 = max_seqs * max_model_len tokens */
  int32_t eos_token_id;    /* -1 disables EOS (synthetic checkpoints) */
  int32_t use_cuda_graphs; /* 1: capture the decode step per batch bucket */
  int32_t gemm_path;       /* 0 auto, 1 force CUDA-core GEMV, 2 force tcgen05 GEMM */
  int32_t act_dtype;       /* 16-bit storage of activations, KV cache, norm weights, biases, embeddings: B200_DTYPE_F16 (default)
                              or B200_DTYPE_BF16 */
  int32_t dense_weights;   /* 0: AWQ int4 linears (qweight / qzeros / scales, fp16 activations only);
                              1: dense 16-bit linears ("...proj.weight" [out, in] in act_dtype) -- Qwen2.5-7B/14B bf16 */
  int32_t tp_size;         /* tensor-parallel group size (0 or 1: none).  One engine per rank, each created with the
                              dimensions of ITS shard: n_heads / tp, n_kv_heads / tp, intermediate / tp; hidden and vocab whole */
  int32_t tp_rank;         /* 0 = the leader (owns the scheduler, takes the requests) */
  int32_t reserved[4];
} b200_config;

typedef struct b200_stats {
  int64_t steps_prefill, steps_decode;
  int64_t tokens_prefill, tokens_generated;
  double gpu_seconds_prefill, gpu_seconds_decode; /* CUDA-event time of the forward passes */
  int64_t decode_batch_sum;        /* sum over decode steps of the batch size */
  double hbm_bytes_decode;         /* algorithmic bytes (weights + KV read + KV write) over decode steps */
  int64_t kernel_launches;         /* kernels this library launched (graph nodes counted per replay) */
  int64_t requests_finished, requests_aborted;
  int32_t running, waiting;
  int32_t kv_blocks_total, kv_blocks_free;
  int64_t prefix_hit_tokens;       /* prompt tokens served from the prefix cache (their K/V blocks were shared) */
  int64_t prefix_query_tokens;     /* prompt tokens of all admitted requests */
} b200_stats;

int32_t b200_abi_version(void);
const char* b200_last_error(void);

/* ---- engine lifecycle -------------------------------------------------------------- */
int32_t b200_engine_create(const b200_config* cfg, b200_engine** out);
/* HF tensor names ("model.layers.3.self_attn.q_proj.qweight", ...).  `data` is a host
 * pointer; the engine copies and repacks, the caller keeps ownership. */
int32_t b200_engine_load_tensor(b200_engine* e, const char* name, const void* data, int32_t dtype,
                                int32_t ndim, const int64_t* shape);
/* All tensors loaded: build kernel-side layouts, allocate KV pool, warm up, capture graphs,
 * start the scheduler thread. */
int32_t b200_engine_finalize(b200_engine* e);
void b200_engine_destroy(b200_engine* e);

/* ---- requests ---------------------------------------------------------------------- */
/* Returns a request id >= 0, or a negative error code.  Greedy iff temperature < 1e-5.  As in vLLM, a
 * generation stops when the sequence reaches max_model_len: max_tokens is clamped to
 * max_model_len - n_prompt, and only a prompt of max_model_len tokens or more is refused. */
int64_t b200_submit(b200_engine* e, const int32_t* prompt_ids, int32_t n_prompt, float temperature,
                    float top_p, int32_t max_tokens, uint64_t seed);
/* Blocks until the request finishes; copies up to `cap` generated ids; releases the request. */
int32_t b200_wait(b200_engine* e, int64_t rid, int32_t* out_ids, int32_t cap, int32_t* n_out,
                  double* ttft_s, double* gen_time_s);
/* Non-blocking (timeout_ms = 0) or bounded wait for tokens not yet returned by poll.
 * When *finished becomes 1 the request is released. */
int32_t b200_poll(b200_engine* e, int64_t rid, int32_t* out_ids, int32_t cap, int32_t* n_new,
                  int32_t* finished, int32_t timeout_ms);
/* Stops the request and releases it: the caller must not wait on or poll `rid` afterwards (a thread
 * already blocked in b200_wait / b200_poll on it gets B200_ERR_ABORTED). */
int32_t b200_abort(b200_engine* e, int64_t rid);
int32_t b200_stats_get(b200_engine* e, b200_stats* out);
/* Tensor parallelism (north_star: "used only for the larger-model config"; the reference has no counterpart --
 * vgate/backends/sglang_backend.py:40 hard-codes tp_size = 1 -- so SURVEY.md section 8e is the specification):
 * q / k / v and gate / up are column-parallel (each rank holds its heads / intermediate columns), o_proj and
 * down_proj row-parallel, and their partial outputs meet in an NCCL all-reduce (sum) inside the step; residual
 * stream, norms, embeddings and LM head are replicated.  Bring-up, per rank r (threads may run concurrently):
 *   b200_engine_create(cfg with tp_size, tp_rank = r, shard dimensions) -> load that rank's tensors ->
 *   b200_engine_tp_init(e_r, id)   [id from ONE b200_tp_unique_id call; blocks until all ranks arrive: one thread per rank] ->
 *   b200_engine_finalize(e_r);   then b200_engine_tp_attach(e_0, e_r) for r > 0.  Requests go to e_0 only.
 * NCCL is loaded on first use (libnccl.so.2, or the path in VGATE_B200_NCCL_LIB). */
int32_t b200_tp_unique_id(void* id128);                       /* 128 bytes out */
int32_t b200_engine_tp_init(b200_engine* e, const void* id128);
int32_t b200_engine_tp_attach(b200_engine* leader, b200_engine* follower);
/* Liveness for health probes: B200_OK while the engine serves; B200_ERR_CUDA (b200_last_error() = the first
 * failure) once a CUDA error has poisoned it and every request was failed.  What /health should report and the
 * reference's unconditional route cannot (main.py:288-295; probed by vgate/health_checker.py:307-312). */
int32_t b200_engine_status(b200_engine* e);

/* Synchronous batch generate on the engine's stream for benchmarks and parity tests:
 * all prompts have the same length, greedy, exactly max_tokens each, device-resident
 * loop without per-request bookkeeping.  `prompt_ids` is a HOST pointer [n_seqs, n_prompt];
 * `out_ids` a HOST pointer [n_seqs, max_tokens].  If `logits_out` (HOST, fp32
 * [n_seqs, n_logit_steps, vocab]) is non-NULL the logits of the first n_logit_steps
 * generated positions are returned as well.  gpu_seconds gets the CUDA-event time
 * [0] prefill, [1] decode.  Must not overlap with submit()ed requests. */
int32_t b200_generate_batch(b200_engine* e, const int32_t* prompt_ids, int32_t n_seqs, int32_t n_prompt,
                            int32_t max_tokens, int32_t* out_ids, float* logits_out,
                            int32_t n_logit_steps, double* gpu_seconds);

/* ---- kernels on raw device pointers (tests, ncu) -------------------------------------- */
/* AWQ checkpoint tensors -> kernel layout.  qweight int32 [K, N/8], qzeros int32 [K/G, N/8],
 * scales fp16 [K/G, N]  ->  wq uint32 [N, K/8] (k-packed, AWQ nibble order along K),
 * sc fp16 [N, K/G], zp fp16 [N, K/G] holding 1024 + zero. */
int32_t b200_op_awq_repack(const int32_t* qweight, const int32_t* qzeros, const void* scales, int32_t K,
                           int32_t N, int32_t G, uint32_t* wq, void* sc, void* zp, void* stream);
/* kernel layout -> dense fp16 W[N, K] = fp16((q - z) * s): the bit-exact unpack check. */
int32_t b200_op_awq_dequant(const uint32_t* wq, const void* sc, const void* zp, int32_t K, int32_t N,
                            int32_t G, void* w_out, void* stream);
/* y[b, n] (fp32, row stride ldy) = (accumulate ? y : 0) + sum_k x[b, k] * W[n, k] + bias[n].
 * x fp16 [B, K] row stride ldx.  CUDA-core W4A16 GEMV (128-bit loads, warp-shuffle reduce). */
int32_t b200_op_w4_gemv(const uint32_t* wq, const void* sc, const void* zp, const void* x, int32_t ldx,
                        float* y, int32_t ldy, const void* bias, int32_t B, int32_t K, int32_t N, int32_t G,
                        int32_t accumulate, void* stream);
/* Kernel layout -> the chunk layout the tcgen05 GEMM streams with TMA bulk copies: for every
 * (128-row tile, quantisation group) one contiguous chunk [scales fp16 x128][1024+zero fp16 x128]
 * [packed k tiles].  `chunks` must hold b200_op_w4_tile_bytes(K, N, G) bytes.  N % 128 == 0,
 * G in {64, 128}.  pair_rows = N / 2 (a multiple of 64) lays a fused [gate; up] matrix out so that
 * every tile holds 64 gate rows and the 64 matching up rows (needed by the SwiGLU epilogue); 0 = plain. */
int32_t b200_op_w4_tile(const uint32_t* wq, const void* sc, const void* zp, int32_t K, int32_t N, int32_t G,
                        int32_t pair_rows, void* chunks, void* stream);
int64_t b200_op_w4_tile_bytes(int32_t K, int32_t N, int32_t G);
/* tcgen05/TMEM W4A16 GEMM, same result contract as b200_op_w4_gemv for any T (token tiles of up to
 * 128).  epilogue: 0 store (+bias), 1 accumulate into y, 2 fused SwiGLU -- chunks tiled with
 * pair_rows = N / 2, y is then an fp16 matrix act[T, N / 2] = silu(gate) * up with row stride ldy halves.
 * n_splits > 1: split-K over whole quantisation groups; y then receives n_splits partial tensors,
 * partial s at y + s * part_stride (bias NULL, epilogue 0), to be summed by the consumer
 * (b200_op_rmsnorm and b200_op_rope_kv do).  n_splits must equal
 * b200_op_w4_effective_splits(K, G, n_splits). */
int32_t b200_op_w4_gemm_umma(const void* chunks, const void* x, int32_t ldx, float* y, int32_t ldy,
                             const void* bias, int32_t T, int32_t K, int32_t N, int32_t G, int32_t epilogue,
                             int32_t n_splits, int64_t part_stride, void* stream);
int32_t b200_op_w4_effective_splits(int32_t K, int32_t G, int32_t n_splits);
/* tcgen05/TMEM GEMM with dense fp16 weights W[N, K] fed by TMA (lm_head); N % 128 == 0, K % 64 == 0. */
int32_t b200_op_f16_gemm_umma(const void* w, const void* x, int32_t ldx, float* y, int32_t ldy, int32_t T,
                              int32_t K, int32_t N, int32_t accumulate, void* stream);
/* The same kernel with every epilogue of the W4 GEMM, for dense 16-bit checkpoints (Qwen2.5-7B / 14B bf16: vLLM runs
 * their linears as plain GEMMs, model_executor/layers/linear.py UnquantizedLinearMethod.apply): W[N, K], x[T, K] and bias
 * in `dtype` (B200_DTYPE_F16 or B200_DTYPE_BF16), fp32 accumulation in tensor memory.  epilogue 0 store (+bias),
 * 1 accumulate into y, 2 fused SwiGLU -- W then holds gate and up rows interleaved in blocks of 64 (rows 128 t .. 128 t + 63
 * gate features 64 t .., rows 128 t + 64 .. the matching up features) and y is a 16-bit matrix act[T, N / 2] with row
 * stride ldy.  n_splits > 1: split-K over 64-wide k tiles into n_splits partial tensors (epilogue 0, no bias), n_splits ==
 * b200_op_dense_effective_splits(K, n_splits). */
int32_t b200_op_dense_gemm_umma(const void* w, const void* x, int32_t ldx, float* y, int32_t ldy, const void* bias, int32_t T,
                                int32_t K, int32_t N, int32_t dtype, int32_t epilogue, int32_t n_splits, int64_t part_stride,
                                void* stream);
int32_t b200_op_dense_effective_splits(int32_t K, int32_t n_splits);
/* Greedy lm_head (vLLM sampler.py:223-236 greedy branch over LogitsProcessor output): out[t] =
 * argmax_n x[t] . W[n]  (highest logit, lowest index on ties) without writing the [T, N] logits: the
 * GEMM epilogue reduces every 128-row tile to {max, argmax}, a second kernel finishes per token.
 * scratch: device, (N / 128) * T * 8 bytes.  N % 128 == 0, K % 64 == 0. */
int32_t b200_op_lm_head_argmax(const void* w, const void* x, int32_t ldx, int32_t T, int32_t K, int32_t N,
                               float* scratch, int32_t* out, void* stream);
/* Same contract with dense fp16 weights W[N, K] (lm_head). */
int32_t b200_op_f16_gemv(const void* w, const void* x, int32_t ldx, float* y, int32_t ldy, int32_t B,
                         int32_t K, int32_t N, int32_t accumulate, void* stream);
/* h[t] (+= sum of n_parts partial tensors parts[p, t, :]) ; xn[r] = fp16(rmsnorm(h[rows ? rows[r] : r]) * g).
 * h fp32 [T, H] is updated in place when n_parts > 0. */
int32_t b200_op_rmsnorm(float* h, const float* parts, int32_t n_parts, int64_t part_stride,
                        const void* gamma, float eps, const int32_t* rows, int32_t n_rows, int32_t H,
                        void* xn, void* stream);
/* qkv fp32 [T, (nh + 2 nkv) * 128], given as the sum of n_parts >= 1 tensors (part p at
 * qkv + p * part_stride) plus an optional fp16 bias -> q fp16 [T, nh * 128] (rotated), K/V written
 * (K rotated) into the paged cache at slot[t] (= block * 16 + offset, -1 = skip).
 * cos/sin fp32 [max_pos, 64]. */
int32_t b200_op_rope_kv(const float* qkv, int32_t n_parts, int64_t part_stride, const void* bias,
                        const int32_t* pos, const int32_t* slot, const float* cos_t, const float* sin_t,
                        int32_t T, int32_t n_heads, int32_t n_kv, void* q_out, void* k_cache, void* v_cache,
                        void* stream);
/* Paged GQA attention, one query row per t: out fp16 [T, nh * 128].  Row t attends to the first
 * ctx_len[t] tokens of block table row bt_row[t].  K/V cache fp16 [n_blocks, n_kv, 16, 128].
 * n_splits > 1 spreads the 64-token chunks of every row round-robin over n_splits CTAs
 * (flash-decoding); the partial results meet in `scratch` (b200_op_attn_scratch_bytes) and are
 * merged in split order by the last CTA to finish, tracked by `counters` (int32 [T, n_kv], all zero
 * before the call and again after it).  n_splits == 1 needs neither. */
int32_t b200_op_paged_attn(const void* q, const void* k_cache, const void* v_cache,
                           const int32_t* block_tables, int32_t bt_stride, const int32_t* bt_row,
                           const int32_t* ctx_len, int32_t T, int32_t n_heads, int32_t n_kv, int32_t n_splits,
                           float* scratch, int32_t* counters, void* out, void* stream);
int64_t b200_op_attn_scratch_bytes(int32_t T, int32_t n_heads, int32_t n_kv, int32_t n_splits);
/* b200_op_paged_attn without the scratch and counters arguments, for callers that cannot allocate
 * device memory (a serving loop that must never hit a device-wide sync, a host that exposes no
 * allocator, a frozen memory pool).  It has no counterpart in the reference.  The split-KV partials
 * and the arrival counters live in arrays built into the library (about 4.3 MB on every GPU that
 * loads it, for the life of the process), so the sizes are capped:
 *     T * n_heads * n_splits <= 8192   and   T * n_kv <= 2048
 * (e.g. 64 rows x 12 heads x 8 splits); larger calls fail with B200_ERR_CAPACITY and should use
 * b200_op_paged_attn.  The call itself allocates nothing and frees nothing.  The first call on each
 * device looks up the arrays' device addresses with that device current -- make that call before any
 * stream capture.  All calls on a device share the arrays, which keep whatever the previous call left
 * behind; the counters are cleared on the stream before each launch, the partials need no clearing.
 * Two calls that overlap on the GPU (different streams) race on the shared arrays: serialise them. */
int32_t b200_op_paged_attn_noalloc(const void* q, const void* k_cache, const void* v_cache,
                                   const int32_t* block_tables, int32_t bt_stride, const int32_t* bt_row,
                                   const int32_t* ctx_len, int32_t T, int32_t n_heads, int32_t n_kv,
                                   int32_t n_splits, void* out, void* stream);
/* Decode step: b200_op_rope_kv + b200_op_paged_attn in one kernel.  Every row is the single new
 * token of a distinct sequence at position pos[t] (ctx_len[t] = pos[t] + 1): its q is rotated in the
 * kernel, its k (rotated) and v are written to the paged cache at slot[t] and take part in the
 * attention.  qkv is given like in b200_op_rope_kv (n_parts partial tensors + optional bias). */
int32_t b200_op_attn_decode_fused(const float* qkv, int32_t n_parts, int64_t part_stride, const void* bias,
                                  const int32_t* pos, const int32_t* slot, const float* cos_t, const float* sin_t,
                                  void* k_cache, void* v_cache, const int32_t* block_tables, int32_t bt_stride,
                                  const int32_t* bt_row, const int32_t* ctx_len, int32_t T, int32_t n_heads,
                                  int32_t n_kv, int32_t n_splits, float* scratch, int32_t* counters, void* out,
                                  void* stream);
/* gu fp32 [T, 2 I] (gate | up) -> act fp16 [T, I] = silu(gate) * up. */
int32_t b200_op_swiglu(const float* gu, int32_t T, int32_t I, void* act, void* stream);
/* h fp32 [T, H] = embed fp16 [V, H] rows ids[t]. */
int32_t b200_op_embed(const void* table, const int32_t* ids, int32_t T, int32_t H, float* h, void* stream);
/* Per row: greedy argmax when temperature[r] < 1e-5, else temperature + top-p + exponential-race
 * draw with the counter hash (seed[r], step[r]).  logits fp32 [B, V] -> out int32 [B]. */
int32_t b200_op_sample(const float* logits, int32_t B, int32_t V, const float* temperature,
                       const float* top_p, const uint32_t* seed, const int32_t* step, int32_t* out,
                       void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200GATE_H */
