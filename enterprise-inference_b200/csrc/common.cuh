// Shared device/host helpers for the sm_100a decode kernels.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "b200_decode.h"

namespace b200 {

// ---- error plumbing (never throw across the C ABI) --------------------------------------
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define B200_CUDA(expr)                                                        \
  do {                                                                         \
    cudaError_t _e = (expr);                                                   \
    if (_e != cudaSuccess) return ::b200::cuda_fail(_e, #expr, __FILE__, __LINE__); \
  } while (0)

#define B200_REQUIRE(cond, code, ...)          \
  do {                                         \
    if (!(cond)) {                             \
      ::b200::set_error(__VA_ARGS__);          \
      return (code);                           \
    }                                          \
  } while (0)

#define B200_LAUNCH_CHECK() B200_CUDA(cudaGetLastError())

constexpr int kHeadDim = 128;
constexpr int kNumSMs = 148;

// ---- bf16 helpers -----------------------------------------------------------------------
// A bf16 is the top half of an fp32: widening is a shift / mask (exact).
__device__ __forceinline__ float bf16lo(uint32_t p) { return __uint_as_float(p << 16); }
__device__ __forceinline__ float bf16hi(uint32_t p) { return __uint_as_float(p & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float round_bf16(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

// Streaming 128-bit load that does not allocate in L1 (weights / KV are read once).
__device__ __forceinline__ uint4 ld_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
// Cached 128-bit read-only load (activations re-read by every warp).
__device__ __forceinline__ uint4 ld_ro(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// (a8, R5) bf16(silu(g)) for a bf16-valued g: accurate expf, the division as a correctly rounded reciprocal
// times g (<= 1 fp32 ulp from g / d, i.e. the bf16 result differs once in ~2^15 elements; an IEEE division
// costs three times as much and this sits in the gate_up epilogue).  One definition for the fused epilogue
// and the stand-alone kernel: they stay bit-identical.
__device__ __forceinline__ float silu_to_bf16(float g) { return round_bf16(g * __frcp_rn(1.0f + expf(-g))); }

// Warp max of floats in ONE instruction (redux.sync on an order-preserving integer image of the float)
// instead of five shuffle + max rounds.  -inf maps through unchanged; NaNs do not occur where it is used
// (masked scores are -inf before the reduction).
__device__ __forceinline__ float warp_max_redux(float v) {
  int i = __float_as_int(v);
  i = i >= 0 ? i : (i ^ 0x7fffffff);
  i = __reduce_max_sync(0xffffffffu, i);
  return __int_as_float(i >= 0 ? i : (i ^ 0x7fffffff));
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---- Programmatic dependent launch -------------------------------------------------------
// Every kernel of the step waits for its predecessor's memory here and lets its successor
// start its prologue early; with no PDL attribute on the launch both are no-ops.
// One lane of a converged warp (elect.sync): the predicate for single-thread instructions (tcgen05.mma, TMA) inside
// warp-uniform code, which keeps their operands in uniform registers.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "elect.sync _|p, 0xffffffff;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;"); }

// ---- launch helper -----------------------------------------------------------------------
struct LaunchOpts {
  bool pdl = false;  // programmatic stream serialization (set by the ctx inside a step)
};
LaunchOpts& launch_opts();
int& launch_counter();

template <typename... KArgs, typename... Args>
inline cudaError_t launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                          Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  int na = 0;
  if (launch_opts().pdl) {
    at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = at;
  cfg.numAttrs = na;
  ++launch_counter();
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// Same, with a thread-block cluster shape (1,1,1 = no cluster).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                                  dim3 cluster, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[2];
  int na = 0;
  if (launch_opts().pdl) {
    at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (cluster.x * cluster.y * cluster.z > 1) {
    at[na].id = cudaLaunchAttributeClusterDimension;
    at[na].val.clusterDim.x = cluster.x;
    at[na].val.clusterDim.y = cluster.y;
    at[na].val.clusterDim.z = cluster.z;
    ++na;
  }
  cfg.attrs = at;
  cfg.numAttrs = na;
  ++launch_counter();
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE property of a kernel: set it once per (instantiation,
// device) — `done` is a function-local static of the calling launcher — so that a process driving several GPUs works.
template <typename K>
inline cudaError_t ensure_dynamic_smem(K kernel, int bytes, bool (&done)[64]) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64 && done[dev]) return cudaSuccess;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess && dev >= 0 && dev < 64) done[dev] = true;
  return e;
}

// ---- kernels' host entry points (internal, C++ linkage) ----------------------------------
int linear_gemv(const __nv_bfloat16* x, int64_t ldx, const __nv_bfloat16* w, const __nv_bfloat16* bias,
                __nv_bfloat16* y, int64_t ldy, int T, int n_out, int k, cudaStream_t s);
bool linear_tc_supported(int T, int n_out, int k);
int64_t linear_tc_workspace_bytes(int T, int n_out, int k);
// Optional fusions of the tcgen05 linear.  Epilogue modes: 0 plain, 1 silu*mul over an interleaved gate|up slab,
// 2 RoPE + KV append over a fused qkv projection, 3 add into the fp32 residual stream (sres/ssq, no y).
// norm_in: the input rows are not read from x but normalised on the fly from the fp32 residual stream
// (RMSNorm folded into the consumer: sres, ssq, norm_w, eps, optional row indirection `rows`).
struct LinearExtra {
  int mode = 0;
  bool norm_in = false;
  float* sres = nullptr;         // [T][hidden] fp32: x + residual, unrounded
  float* ssq = nullptr;          // [T][hidden/128] fp32: per-slab sums of squares of sres
  const __nv_bfloat16* norm_w = nullptr;
  const int32_t* rows = nullptr;
  float eps = 0.f;
  const int32_t* positions = nullptr;
  const int32_t* slots = nullptr;
  const __nv_bfloat16* cos_sin = nullptr;
  __nv_bfloat16* kc = nullptr;
  __nv_bfloat16* vc = nullptr;
  int hq = 0, hkv = 0, block_size = 0;
};
int linear_tc(const __nv_bfloat16* x, int64_t ldx, const __nv_bfloat16* w, const __nv_bfloat16* bias,
              __nv_bfloat16* y, int64_t ldy, int T, int n_out, int k, void* ws, int64_t ws_bytes,
              const LinearExtra* ex, cudaStream_t s);
int rmsnorm(const __nv_bfloat16* x, __nv_bfloat16* res, const __nv_bfloat16* w, __nv_bfloat16* y, int T,
            int hidden, float eps, cudaStream_t s);
// fp32 residual stream (RMSNorm folded into the linears): embedding rows -> sres/ssq; stream rows -> normed bf16
// (prev, n_prev: tokens a negative id -(j+1) refers to, see elementwise.cu; nullptr = no such ids)
int embed_stream(const int32_t* ids, const __nv_bfloat16* table, float* sres, float* ssq, int T, int hidden, int vocab,
                 cudaStream_t s, const int32_t* prev = nullptr, int n_prev = 0);
int rmsnorm_stream(const float* sres, const float* ssq, const int32_t* rows, const __nv_bfloat16* w, __nv_bfloat16* y,
                   int R, int hidden, float eps, cudaStream_t s);
int embed(const int32_t* ids, const __nv_bfloat16* table, __nv_bfloat16* out, int T, int hidden, int vocab,
          cudaStream_t s, const int32_t* prev = nullptr, int n_prev = 0);
int rope_kv_append(__nv_bfloat16* qkv, const int32_t* positions, const int32_t* slots,
                   const __nv_bfloat16* cos_sin, __nv_bfloat16* kc, __nv_bfloat16* vc, int T, int hq,
                   int hkv, int block_size, cudaStream_t s);
int64_t paged_attn_workspace_bytes(int T, int hq, int max_ctx, int block_size);
int paged_attn(const __nv_bfloat16* q, int64_t ldq, const __nv_bfloat16* kc, const __nv_bfloat16* vc,
               const int32_t* block_table, int bt_stride, const int32_t* row_to_seq, const int32_t* ctx_lens,
               __nv_bfloat16* out, int64_t ldo, int T, int hq, int hkv, int block_size, int max_ctx,
               float scale, void* ws, int64_t ws_bytes, cudaStream_t s);
int paged_attn_ex(const __nv_bfloat16* q, int64_t ldq, const __nv_bfloat16* kc, const __nv_bfloat16* vc,
                  const int32_t* block_table, int bt_stride, const int32_t* row_to_seq, const int32_t* ctx_lens,
                  __nv_bfloat16* out, int64_t ldo, int T, int hq, int hkv, int block_size, int max_ctx,
                  float scale, void* ws, int64_t ws_bytes, int pack, cudaStream_t s);
int sample_rows(const __nv_bfloat16* logits, int64_t ld, int n_rows, int vocab, int seg_len, int64_t seg_stride,
                const float* temperature, const int32_t* top_k, const float* top_p, const float* min_p,
                const uint64_t* seeds, const uint64_t* draws, int32_t* out, cudaStream_t s);
int logit_adjust_rows(__nv_bfloat16* logits, int64_t ld, int n_rows, int vocab_local, int index_offset, const float* row_params,
                      const int32_t* offsets, const int32_t* ids, const int32_t* counts, const int32_t* seen, const float* bias,
                      cudaStream_t s);
int logprob_rows(const __nv_bfloat16* logits, int64_t ld, int n_rows, int vocab, int seg_len, int64_t seg_stride,
                 const int32_t* tok, int n_top, float* out_lp, int32_t* top_ids, float* top_lp, cudaStream_t s);
// once per step when pack > 0: builds the prefill attention's row plan in ws; 1 = built (use pack = 2), 0 = not used
int paged_attn_prefill_plan_step(const int32_t* row_to_seq, const int32_t* ctx_lens, int T, int hq, int hkv, int block_size,
                                 void* ws, int64_t ws_bytes, cudaStream_t s);
int advance_rows(int32_t* ids, int32_t* positions, int32_t* slots, int32_t* ctx_lens, const int32_t* block_table,
                 int bt_stride, const int32_t* row_to_seq, const int32_t* sampled, int n, int block_size,
                 cudaStream_t s);
int silu_mul(const __nv_bfloat16* gu, __nv_bfloat16* out, int T, int inter, cudaStream_t s);

// Peer-memory tensor-parallel all-reduce (allreduce.cu).  One IPC-shareable allocation per rank:
//   [ partial 0 | partial 1 | gather 0 | gather 1 | flags: 2 kinds x 2 buffers x 8 ranks x 128 B | state ]
struct ArParams {
  const __nv_bfloat16* bufs[8];   // every rank's partial for this buffer index (peer-mapped)
  unsigned int* flags[8];         // every rank's flag block
  unsigned int* state;            // local: counter[2] | ticket[2]
  __nv_bfloat16* out;
  long long n_chunks;             // 16-byte chunks
  int rank, world, buf_idx;
  __nv_bfloat16* gather[8];       // two-shot: every rank's normed-row buffer for this buffer index (peer-mapped)
};
int allreduce_oneshot(const ArParams& p, cudaStream_t s);
int allreduce_add_rmsnorm(const ArParams& p, int T, __nv_bfloat16* res, const __nv_bfloat16* w, __nv_bfloat16* y, int hidden,
                          float eps, cudaStream_t s);
int allreduce_twoshot_rmsnorm(const ArParams& p, int T, __nv_bfloat16* res, const __nv_bfloat16* w, int hidden, float eps,
                              cudaStream_t s);
int argmax_rows(const __nv_bfloat16* logits, int64_t ld, int rows, int vocab, int index_offset,
                int32_t* out_ids, float* out_max, cudaStream_t s);

}  // namespace b200
