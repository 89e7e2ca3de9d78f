/*
 * b200_decode.h — C-ABI of the B200-native decode path.
 *
 * The reference (psurabh/Enterprise-Inference v1.3.0) has NO in-process FFI for this path:
 * its boundary is the container contract of core/helm-charts/vllm (image + args + env + HTTP,
 * /root/reference/core/helm-charts/vllm/templates/deployment.yaml:57,66-87 and
 * /root/reference/docs/api-spec.yaml:6-49).  The arithmetic runs inside the third-party
 * image vllm-cpu-release-repo:v0.10.2 (values.yaml:23-28).  Each entry point below therefore
 * cites the step of that image's decode loop it replaces (the
 * [V22] citations are the vLLM 0.22 wheel used as the readable statement of the math).
 *
 * Conventions
 *  - every function returns 0 on success, a negative b200_status otherwise; the text of the
 *    last error of the calling thread is b200_last_error(); nothing throws across the ABI;
 *  - every tensor is a CALLER-OWNED DEVICE pointer (e.g. torch tensor.data_ptr()) unless the
 *    parameter name starts with host_; bf16 unless stated; row-major; explicit leading dims;
 *  - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, nothing syncs
 *    unless documented;
 *  - no hidden allocation after b200_ctx_create(): scratch is either inside the ctx or passed
 *    by the caller as (workspace, workspace_bytes) with a *_workspace_bytes() query;
 *  - a ctx is not re-entrant: one engine thread per rank drives it; TP ranks are processes.
 */
#ifndef B200_DECODE_H
#define B200_DECODE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B200_API __attribute__((visibility("default")))
#else
#define B200_API
#endif

typedef struct b200_ctx b200_ctx;

typedef enum b200_status {
  B200_OK = 0,
  B200_ERR_INVALID = -1,      /* bad argument / shape */
  B200_ERR_CUDA = -2,         /* a CUDA runtime/driver call failed */
  B200_ERR_UNSUPPORTED = -3,  /* shape outside what the kernels were built for */
  B200_ERR_NCCL = -4,
  B200_ERR_STATE = -5         /* call order (weights not set, ...) */
} b200_status;

B200_API const char* b200_last_error(void);
B200_API int b200_abi_version(void);
/* 1 when a CUDA device of compute capability 10.x is current, else 0 (no error text). */
B200_API int b200_device_ok(void);

/* ---- model / context ------------------------------------------------------------------ */
typedef struct b200_model_cfg {
  int32_t hidden, n_layers, n_q_heads, n_kv_heads, head_dim; /* GLOBAL sizes; head_dim == 128 */
  int32_t intermediate, vocab;
  int32_t qkv_bias;            /* Qwen2: 1 */
  float rms_eps;
  int32_t block_size;          /* KV block, tokens: 128 (xeon-values.yaml:72-73 --block-size) */
  int32_t max_tokens;          /* max token rows in one step (--max-num-batched-tokens 2048) */
  int32_t max_blocks_per_seq;  /* ceil(max-model-len / block_size) */
  int32_t tp_rank, tp_world;   /* tensor parallel (deployment.yaml:72-78 --tensor-parallel-size) */
} b200_model_cfg;

typedef enum b200_weight_kind {
  B200_W_EMBED = 0,     /* [vocab, hidden]                      (a1)  replicated on every rank */
  B200_W_LN1 = 1,       /* [hidden]                             (a2)  */
  B200_W_QKV = 2,       /* [(hq+2hkv)_local*128, hidden]        (a3)  q|k|v rows, column-parallel */
  B200_W_QKV_BIAS = 3,  /* [(hq+2hkv)_local*128]                      */
  B200_W_O = 4,         /* [hidden, hq_local*128]               (a7)  row-parallel */
  B200_W_LN2 = 5,       /* [hidden]                                   */
  B200_W_GATE_UP = 6,   /* [2*I_local, hidden] gate rows then up rows (a8) */
  B200_W_DOWN = 7,      /* [hidden, I_local]                    (a8)  row-parallel */
  B200_W_FINAL_NORM = 8,/* [hidden]                             (a9)  */
  B200_W_LM_HEAD = 9,   /* [vocab_local, hidden]                (a9)  vocab-parallel */
  B200_W_GATE_UP_INTERLEAVED = 10 /* [2*I_local, hidden] in 128-row slabs: 64 gate rows, then the 64 matching
                                     up rows — lets the gate_up linear apply SiLU*mul in its epilogue (set
                                     this OR B200_W_GATE_UP; needs I_local % 64 == 0) */
} b200_weight_kind;

/* Replaces: model construction inside the image's worker (one per TP rank,
 * xeon-values.yaml:78-79 --distributed_executor_backend mp).  Allocates the ctx's private
 * activation/scratch buffers (the only cudaMalloc calls in the library). */
B200_API int b200_ctx_create(const b200_model_cfg* cfg, b200_ctx** out);
B200_API int b200_ctx_destroy(b200_ctx* ctx);
/* Bytes of device memory the ctx allocated in b200_ctx_create(). */
B200_API int64_t b200_ctx_workspace_bytes(const b200_ctx* ctx);

/* Replaces: weight loading ([V22] vllm/model_executor/models/llama.py:437-444 packing map).
 * The pointer is borrowed for the life of the ctx; `layer` is ignored for EMBED/FINAL_NORM/
 * LM_HEAD.  rows/cols are checked against the cfg (local TP shard sizes). */
B200_API int b200_set_weight(b200_ctx* ctx, int32_t layer, int32_t kind, const void* w, int64_t rows, int64_t cols);

/* KV cache, caller-owned, one allocation per K and V:
 *   [n_layers][n_blocks][n_kv_heads_local][block_size][head_dim] bf16
 * so one (block, kv head) is a contiguous 32 KiB tile — one bulk-async copy per tile.
 * Replaces the image-internal cache ([V22] vllm/v1/attention/backends/cpu_attn.py:88-100). */
B200_API int b200_set_kv_cache(b200_ctx* ctx, void* k_cache, void* v_cache, int64_t n_blocks);

/* cos|sin cache [max_pos][head_dim] bf16 ([V22] rotary_embedding/base.py:58-63,82-92). */
B200_API int b200_set_rope_cache(b200_ctx* ctx, const void* cos_sin, int64_t max_pos);

/* NCCL communicator for TP (replaces the image's shm all-reduce,
 * [V22] vllm/distributed/device_communicators/cpu_communicator.py:229-276).
 * b200_nccl_unique_id writes 128 bytes to host_id (rank 0 calls it and shares the bytes);
 * b200_nccl_init is collective over the tp_world ranks of the ctx. */
B200_API int b200_nccl_unique_id(void* host_id128);
B200_API int b200_nccl_init(b200_ctx* ctx, const void* host_id128);

/* Peer-memory all-reduce (optional, after b200_nccl_init): every rank exports the CUDA-IPC handle of its
 * all-reduce buffers (64 bytes), the ranks exchange them, and each imports all `world` handles (rank
 * order).  From then on the two all-reduces per layer run as ONE kernel per rank over NVLink peer loads
 * (flag barrier + rank-ordered fp32 sum) instead of NCCL; NCCL stays the fallback when this is skipped. */
B200_API int b200_tp_ipc_export(b200_ctx* ctx, void* host_handle64);
B200_API int b200_tp_ipc_import(b200_ctx* ctx, const void* host_handles, int32_t world);

/* ---- the hot path: one scheduler step ----------------------------
 * Replaces ModelRunner.execute_model of the image for one batch of token rows
 * ([V22] vllm/model_executor/models/llama.py:395-434 + sampler.py:235-236).
 * Row t is one new token: ids[t], positions[t], slot_mapping[t] = block_id*block_size+off,
 * ctx_lens[t] = tokens it attends to (position+1), row_to_seq[t] = its row of block_table
 * (NULL: row t).  Decode rows and (chunked) prefill rows have the same form, a prefill chunk
 * being T rows of one sequence with ctx_lens 1..T (+offset).
 * logit_rows[n_logit_rows] selects the rows that get lm_head + greedy sampling (NULL: all).
 * out_token_ids[n_logit_rows] int32; out_logits (nullable) bf16 [n_logit_rows][vocab].
 * max_ctx_len is a HOST hint >= max(ctx_lens) that sizes the split-KV grid.
 * With tp_world > 1 it is collective over the ranks. */
B200_API int b200_decode_step(b200_ctx* ctx, const int32_t* ids, const int32_t* positions,
                     const int32_t* slot_mapping, const int32_t* block_table, int32_t bt_stride,
                     const int32_t* row_to_seq, const int32_t* ctx_lens, int32_t n_tokens,
                     int32_t max_ctx_len, const int32_t* logit_rows, int32_t n_logit_rows,
                     int32_t* out_token_ids, void* out_logits, void* stream);

/* Same step, driven from PINNED HOST memory: one packed int32 buffer
 *   [ids | positions | slot_mapping | ctx_lens | row_to_seq | logit_rows](each n_tokens, the
 *   last n_logit_rows) followed by block_table [n_seqs][bt_stride],
 * copied H2D, the step, and the sampled ids copied D2H into host_out_token_ids.  The copies
 * are enqueued on `stream`; the caller synchronises.  This is the call the serving engine
 * makes every iteration (the `e2e` figure of bench.py).
 * An id of -(j+1) stands for the token the PREVIOUS b200_decode_step_host call on this ctx chose for
 * its logit row j (it is still on the device): an engine can enqueue step n+1 while step n runs and
 * read step n's tokens afterwards (vLLM's async scheduling).  host_packed and host_out_token_ids must
 * stay untouched until their copies have completed, so such an engine alternates two buffers of each. */
B200_API int b200_decode_step_host(b200_ctx* ctx, const int32_t* host_packed, int32_t n_tokens,
                          int32_t n_seqs, int32_t bt_stride, int32_t max_ctx_len,
                          int32_t n_logit_rows, int32_t* host_out_token_ids, void* stream);

/* Graph replay: 1 (default) captures each distinct step shape into a CUDA graph; 0 = eager
 * launches (the chart's --enforce-eager is accepted by the server but does not force this). */
B200_API int b200_ctx_set_graphs(b200_ctx* ctx, int32_t enabled);
/* Number of kernels one step of the given shape launches (bench.py's gpu_launches). */
B200_API int b200_ctx_last_step_launches(const b200_ctx* ctx);
/* Per-layer hidden/residual taps for parity tests: after a step, copies the layer's MLP
 * output and residual rows ([n_tokens][hidden] bf16 each) — only valid with graphs off. */
B200_API int b200_ctx_set_tap(b200_ctx* ctx, int32_t layer, void* out_hidden, void* out_residual);

/* ---- per-kernel entry points (unit tests, ncu) ------------------------------------------ */

/* Launch attribute for the per-kernel entry points called from THIS thread: 1 = programmatic dependent
 * launch (the kernels of b200_decode_step always use it unless B200_PDL=0): a kernel's prologue and first
 * weight tiles overlap the previous kernel's tail.  Default 0. */
B200_API int b200_set_launch_pdl(int32_t enabled);

/* y[T][n_out] = x[T][k] . w[n_out][k]^T (+bias), fp32 accumulate, one rounding to bf16.
 * Replaces sgl-kernel weight_packed_linear / oneDNN matmul of the image
 * ([V22] vllm/model_executor/layers/utils.py:277-299).  impl: 0 auto, 1 CUDA-core GEMV
 * (HBM-bound, T small), 2 tcgen05/TMEM weight-streaming GEMM.  Scratch (split-K partials
 * and slab tickets) comes from the caller: the first 64 KiB of the workspace must be ZERO before
 * its first use (cudaMemset once); every launch leaves them zero again. */
B200_API int64_t b200_linear_workspace_bytes(int32_t n_tokens, int32_t n_out, int32_t k);
B200_API int b200_linear_bf16(const void* x, int64_t ldx, const void* w, const void* bias, void* y,
                     int64_t ldy, int32_t n_tokens, int32_t n_out, int32_t k, int32_t impl,
                     void* workspace, int64_t workspace_bytes, void* stream);

/* Fused epilogues of the tcgen05 linear (what b200_decode_step uses): mode 1 = a8's SiLU*mul over
 * interleaved gate|up slabs (w as B200_W_GATE_UP_INTERLEAVED; y is [T][n_out/2]); mode 2 = a4+a5:
 * w is the fused qkv projection, RoPE is applied to q/k heads, q goes to y[T][ldy], k/v rows go to
 * the paged cache at slot_mapping[t] (pointers unused by mode 1 may be NULL).  Results are
 * bit-identical to b200_linear_bf16(impl=2) followed by b200_silu_mul / b200_rope_kv_append. */
B200_API int b200_linear_fused_bf16(const void* x, int64_t ldx, const void* w, const void* bias, void* y,
                           int64_t ldy, int32_t n_tokens, int32_t n_out, int32_t k, int32_t mode,
                           const int32_t* positions, const int32_t* slot_mapping, const void* cos_sin,
                           void* k_cache_layer, void* v_cache_layer, int32_t n_q_heads, int32_t n_kv_heads,
                           int32_t block_size, void* workspace, int64_t workspace_bytes, void* stream);

/* (a2 folded into a3/a7/a8/a9) RMSNorm without a kernel of its own, for steps of at most 64 rows.  The residual
 * stream lives in fp32: sres[T][hidden] holds s = x + res UNROUNDED (what [V22] vllm/ir/ops/layernorm.py:40-58
 * takes the variance of and normalises; the bf16 residual of the next layer is bf16(s), recomputed on read) and
 * ssq[T][hidden/128] the sums of squares of s per 128-wide slab.
 *  - b200_embed_stream: (a1) sres[t] = float(table[ids[t]]) and its ssq;
 *  - b200_linear_residual_bf16: (a7, a8-down) x = bf16(x_in . w^T + bias); sres = x + bf16(sres); ssq rewritten
 *    (n_out = hidden, a multiple of 128; no y is written);
 *  - b200_linear_normed_bf16: (a3, a8-gate_up, a9) the linear over y_in[t] = bf16(bf16(s * rinv) * norm_w),
 *    rinv = 1/sqrt(sum(ssq[row]) / k + eps), s = sres[rows ? rows[t] : t] (k = hidden); the rows are normalised
 *    inside the kernel while the weights stream.  mode / trailing arguments as b200_linear_fused_bf16 (0 = plain);
 *  - b200_rmsnorm_stream: the same normalised rows written out as bf16 [n_rows][hidden].
 * The summation order of the variance is the one b200_rmsnorm uses, so a step through these entry points is
 * bit-identical to embed -> rmsnorm -> linear -> ... with bf16 activations. */
B200_API int b200_embed_stream(const int32_t* ids, const void* table, float* sres, float* ssq, int32_t n_tokens,
                      int32_t hidden, int32_t vocab, void* stream);
B200_API int b200_linear_residual_bf16(const void* x, int64_t ldx, const void* w, const void* bias, float* sres,
                              float* ssq, int32_t n_tokens, int32_t n_out, int32_t k, void* workspace,
                              int64_t workspace_bytes, void* stream);
B200_API int b200_linear_normed_bf16(const float* sres, const float* ssq, const void* norm_w, float eps,
                            const int32_t* rows, const void* w, const void* bias, void* y, int64_t ldy,
                            int32_t n_tokens, int32_t n_out, int32_t k, int32_t mode, const int32_t* positions,
                            const int32_t* slot_mapping, const void* cos_sin, void* k_cache_layer,
                            void* v_cache_layer, int32_t n_q_heads, int32_t n_kv_heads, int32_t block_size,
                            void* workspace, int64_t workspace_bytes, void* stream);
B200_API int b200_rmsnorm_stream(const float* sres, const float* ssq, const int32_t* rows, const void* w, void* y,
                        int32_t n_rows, int32_t hidden, float eps, void* stream);

/* (a2) y = rmsnorm(x)*w, or with res != NULL: s=x+res; res=bf16(s); y=rmsnorm(s)*w
 * ([V22] vllm/ir/ops/layernorm.py:10-21,40-58). y may alias x. */
B200_API int b200_rmsnorm(const void* x, void* res, const void* w, void* y, int32_t n_tokens,
                 int32_t hidden, float eps, void* stream);

/* (a1) x[t] = table[ids[t]] ([V22] llama.py:371-375,392-393). */
B200_API int b200_embed(const int32_t* ids, const void* table, void* out, int32_t n_tokens,
               int32_t hidden, int32_t vocab, void* stream);

/* (a4+a5) in-place NeoX RoPE on the q and k heads of qkv[T][(hq+2hkv)*128] in bf16
 * arithmetic, then k,v rows written to the paged cache at slot_mapping[t]
 * ([V22] rotary_embedding/common.py:160-183; cpu_attn.py:342-362 reshape_and_cache). */
B200_API int b200_rope_kv_append(void* qkv, const int32_t* positions, const int32_t* slot_mapping,
                        const void* cos_sin, void* k_cache_layer, void* v_cache_layer,
                        int32_t n_tokens, int32_t n_q_heads, int32_t n_kv_heads,
                        int32_t block_size, void* stream);

/* (a6) paged attention, one query row per token, split-KV with LSE merge
 * ([V22] cpu_attn.py:377-396).  q rows are the first hq*128 columns of a [T][ldq] buffer. */
B200_API int64_t b200_paged_attn_workspace_bytes(int32_t n_tokens, int32_t n_q_heads, int32_t max_ctx_len,
                                        int32_t block_size);
B200_API int b200_paged_attn_decode(const void* q, int64_t ldq, const void* k_cache_layer,
                           const void* v_cache_layer, const int32_t* block_table, int32_t bt_stride,
                           const int32_t* row_to_seq, const int32_t* ctx_lens, void* out,
                           int64_t ldo, int32_t n_tokens, int32_t n_q_heads, int32_t n_kv_heads,
                           int32_t block_size, int32_t max_ctx_len, float scale, void* workspace,
                           int64_t workspace_bytes, void* stream);

/* (a6, prefill rows) Same contract and results as b200_paged_attn_decode, for steps dominated by the rows of
 * a prefill chunk: up to 4 consecutive rows that belong to one sequence (row_to_seq) share ONE pass over that
 * sequence's K/V cache with a per-row causal mask (their query heads fill the 16 MMA columns); rows of
 * different sequences inside a group fall back to one pass each.  Chunked prefill in the reference:
 * xeon-values.yaml:80-83 (--enable_chunked_prefill, --max-num-batched-tokens 2048). */
B200_API int b200_paged_attn_prefill(const void* q, int64_t ldq, const void* k_cache_layer,
                            const void* v_cache_layer, const int32_t* block_table, int32_t bt_stride,
                            const int32_t* row_to_seq, const int32_t* ctx_lens, void* out, int64_t ldo,
                            int32_t n_tokens, int32_t n_q_heads, int32_t n_kv_heads, int32_t block_size,
                            int32_t max_ctx_len, float scale, void* workspace, int64_t workspace_bytes,
                            void* stream);

/* No counterpart in the reference: a variant of b200_paged_attn_decode for callers that cannot
 * allocate device memory after start-up (engines with a fixed memory plan, bindings that cannot call
 * cudaMalloc).  Same inputs, outputs and results, but the split-KV partials live in a 2 MiB array
 * built into the library (b200_paged_attn_static_scratch_bytes(); held by every GPU that loads the
 * library until exit) and the call allocates nothing.  Shapes whose partials exceed the array return
 * B200_ERR_UNSUPPORTED with the sizes in b200_last_error().  The first call on each device looks up
 * the array's device address: make it before any stream capture.  All calls on a device share the
 * array — keep them on one stream; calls that overlap on the GPU race. */
B200_API int64_t b200_paged_attn_static_scratch_bytes(void);
B200_API int b200_paged_attn_decode_static(const void* q, int64_t ldq, const void* k_cache_layer,
                                  const void* v_cache_layer, const int32_t* block_table, int32_t bt_stride,
                                  const int32_t* row_to_seq, const int32_t* ctx_lens, void* out,
                                  int64_t ldo, int32_t n_tokens, int32_t n_q_heads, int32_t n_kv_heads,
                                  int32_t block_size, int32_t max_ctx_len, float scale, void* stream);

/* (a8) out[T][I] = bf16(bf16(silu(g))*u), gu = [T][2I] gate|up ([V22] activation.py:137-141). */
B200_API int b200_silu_mul(const void* gu, void* out, int32_t n_tokens, int32_t intermediate, void* stream);

/* (a10) greedy: fp32 argmax of bf16 logits, ties -> lowest index ([V22] sampler.py:90-91,235-236).
 * Also writes the winning value (fp32) when out_max != NULL (used by the TP (max,idx) reduce). */
B200_API int b200_argmax(const void* logits, int64_t ld, int32_t n_rows, int32_t vocab, int32_t index_offset,
                int32_t* out_ids, float* out_max, void* stream);

/* (a10, widened) stochastic sampling for requests that do not ask for greedy decoding: per row
 * logits / temperature, top-k mask (values below the k-th largest dropped, ties at it kept), top-p on the
 * renormalised rest (the smallest-probability tokens with cumulative mass <= 1 - p dropped), one draw
 * ([V22] vllm/v1/sample/sampler.py:90-91,235-256, ops/topk_topp_sampler.py; fields of docs/api-spec.yaml).
 * temperature[r] <= 0 -> the argmax of b200_argmax.  top_k[r] <= 0 and top_p[r] >= 1 disable the filters.
 * The draw is a function of (seeds[r], draws[r], logits row): same inputs, same token.  All arrays are device
 * pointers of n_rows elements; temperature/top_k/top_p/min_p/seeds/draws may be NULL (0, off, off, off, 0, 0).
 * min_p[r] in (0, 1]: tokens whose probability is below min_p x the largest one are dropped before top-k/top-p
 * ([V22] vllm/v1/sample/logits_processor/builtin.py:101).  vLLM compares in probability space; this kernel compares
 * the logit with max + T log(min_p), the same cut up to fp32 rounding of a token that sits exactly on it. */
B200_API int b200_sample(const void* logits, int64_t ld, int32_t n_rows, int32_t vocab, const float* temperature,
                const int32_t* top_k, const float* top_p, const float* min_p, const uint64_t* seeds,
                const uint64_t* draws, int32_t* out_ids, void* stream);

/* Sampling over the logit rows of the LAST b200_decode_step[_host] of this ctx (they are still in the ctx's
 * buffer; with tensor parallelism the vocab shards are all-gathered first: a collective, call it on every rank
 * with the same arguments).  host_params = five host arrays back to back,
 *   float temperature[n] | int32 top_k[n] | float top_p[n] | float min_p[n] (0: off) | uint64 seed[n] | uint64 draw[n],
 * 32 n bytes; host_out_ids receives n token ids (asynchronously: synchronise the stream before reading). */
B200_API int b200_ctx_sample_last(b200_ctx* ctx, int32_t n_rows, const void* host_params, int32_t* host_out_ids,
                         void* stream);

/* (a10, widened) OpenAI `logit_bias`, `frequency_penalty`, `presence_penalty` and vLLM's `repetition_penalty` on logit
 * rows [n_rows][ld] (bf16, modified in place), in the image's order and arithmetic ([V22] vllm/v1/sample/sampler.py:366-416,
 * vllm/model_executor/layers/utils.py:51-89, vllm/_custom_ops.py:360-373), fp32 per touched token:
 *   x += bias;  if the token is in the prompt or the output: x *= (x > 0 ? 1/rep : rep);
 *   x -= freq * (occurrences in the output);  x -= pres * (occurs in the output ? 1 : 0)
 * then ONE rounding back to bf16 (the image keeps the processed logits in fp32: the only deviation).  Sparse: row r owns
 * the entries offsets[r] .. offsets[r+1]-1 of (ids, counts, seen, bias), distinct tokens per row; row_params = (rep, freq,
 * pres) per row.  Device pointers.  b200_ctx_adjust_last applies the same to the logit rows of the LAST step of a ctx from
 * ONE packed host buffer (float row_params[3n] | int32 offsets[n+1] | int32 ids[E] | int32 counts[E] | int32 seen[E] |
 * float bias[E]); call it before b200_ctx_sample_last, on every tensor-parallel rank (each adjusts its vocabulary shard). */
B200_API int b200_logit_adjust(void* logits, int64_t ld, int32_t n_rows, int32_t vocab, const float* row_params,
                      const int32_t* offsets, const int32_t* ids, const int32_t* counts, const int32_t* seen,
                      const float* bias, void* stream);
B200_API int b200_ctx_adjust_last(b200_ctx* ctx, int32_t n_rows, const void* host_packed, int32_t n_entries, void* stream);

/* (a10, widened) OpenAI `logprobs` / `top_logprobs`: log softmax over the raw logits (vLLM's default logprobs mode,
 * [V22] vllm/v1/sample/sampler.py:90-118).  out_logprob[r] = l[token_ids[r]] - logsumexp(l); out_top_ids /
 * out_top_logprobs [r][n_top] = the n_top (<= 32) most likely tokens, value descending, ties by lowest index
 * (may be NULL when n_top == 0).  Device pointers. */
B200_API int b200_logprobs(const void* logits, int64_t ld, int32_t n_rows, int32_t vocab, const int32_t* token_ids,
                  int32_t n_top, float* out_logprob, int32_t* out_top_ids, float* out_top_logprobs, void* stream);

/* The same over the logit rows of the LAST step of this ctx (a collective under tensor parallelism, like
 * b200_ctx_sample_last).  host_out receives float logprob[n] | int32 top_ids[n][n_top] | float top_logprobs[n][n_top]
 * (4 n (1 + 2 n_top) bytes), asynchronously. */
B200_API int b200_ctx_logprobs_last(b200_ctx* ctx, int32_t n_rows, const int32_t* host_token_ids, int32_t n_top,
                           void* host_out, void* stream);

/* (a0) device-side scheduler advance for pure-decode rows: ids[t]=sampled[t], positions[t]+=1,
 * ctx_lens[t]=positions[t]+1, slot_mapping[t] from the block table.  The image's scheduler does this
 * on the host every iteration ([V22] vllm/v1/attention/backends/cpu_attn.py:107-117 metadata build);
 * on the device it lets several decode steps run without a host round trip.  The caller guarantees
 * the block table already covers the next position. */
B200_API int b200_advance_decode_rows(int32_t* ids, int32_t* positions, int32_t* slot_mapping,
                                      int32_t* ctx_lens, const int32_t* block_table, int32_t bt_stride,
                                      const int32_t* row_to_seq, const int32_t* sampled, int32_t n_rows,
                                      int32_t block_size, void* stream);

/* (a11) sum all-reduce of a bf16 buffer over the ctx's TP communicator. */
B200_API int b200_allreduce_bf16(b200_ctx* ctx, void* buf, int64_t count, void* stream);

/* (a11) The collective of one half-layer exactly as b200_decode_step runs it, as an entry point of its own (tests,
 * timing): `partial` [T][hidden] bf16 is this rank's row-parallel partial sum (NULL: reuse what the exchange buffer
 * holds); buf_idx 0 = the O projection's buffer, 1 = the down projection's (they alternate inside a step, and so
 * must consecutive calls).  norm_w == NULL: y = all-reduce(partial).  norm_w != NULL: the fused form the step uses,
 * s = bf16(all-reduce(partial)) + res; res = bf16(s); y = rmsnorm(s) * norm_w (above 16 rows the peer-memory path deals
 * the rows to the ranks: res is then current only for the rows t with t mod tp_world == tp_rank; res == NULL and
 * y == NULL: run on the ctx's own residual / normed rows without any copy, for timing).  Every rank must make the same call.
 * Replaces the shm all-reduce of the image's CPU communicator followed by its fused_add_rms_norm
 * ([V22] vllm/distributed/device_communicators/cpu_communicator.py:229-276; vllm/ir/ops/layernorm.py:40-58). */
B200_API int b200_tp_allreduce(b200_ctx* ctx, int32_t buf_idx, const void* partial, int32_t n_tokens, void* res,
                      const void* norm_w, void* y, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_DECODE_H */
