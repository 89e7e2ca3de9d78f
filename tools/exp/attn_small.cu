// (a6) Paged attention for SMALL decode problems: a handful of rows over a few hundred cached tokens.
//
// The tensor-core kernel (attn_tc.cu) is built for throughput: TMEM allocation, a 192 KiB TMA ring, a dozen barriers
// and a four-role pipeline cost ~9 us per launch even when the whole problem is 1 MB (8B batch 1 at context 256:
// profiles/r02_launches_bs1_summary.txt), and at batch 1 that latency is paid 32 times per step while HBM idles.
// Here one CTA of 128 threads serves one (row, kv head): no TMEM, no TMA, no split, no workspace.
//   1. thread = cached token: the G query heads' scores against that token's K row (fp32, q.k then * scale: R4);
//   2. block-wide max and sum per head, p = exp(s - m) kept in fp32 in shared memory (R4 keeps P in fp32);
//   3. thread = head dimension: out[h][d] = sum_t p[h][t] V[t][d] / l[h], one rounding to bf16.
// Bound: latency (the problem is far below the HBM roofline by construction); algorithmic bytes ctx * 128 * 2 * 2.
// Serves ctx <= kSmallMaxCtx; anything larger goes to the streaming kernels.
//
// Replaces: paged_attention_v1 of the vLLM-CPU image ([V22] vllm/v1/attention/backends/cpu_attn.py:377-396).
#include <stdlib.h>

#include "common.cuh"

namespace b200 {

namespace {

constexpr int kSmallThreads = 256;
constexpr int kSmallWarps = kSmallThreads / 32;
constexpr int kSmallMaxCtx = 512;

// Every phase is a handful of L2 round trips, so each one issues ALL of a thread's loads before the first use:
//   scores: two threads per token (64 dims = 8 x 16 B each), 128 tokens per pass;
//   output: 16 threads per token row (16 B = 8 dims each), 16 token groups; the groups are summed through shuffles
//           and one shared-memory pass per head.
template <int G>
__global__ void __launch_bounds__(kSmallThreads)
paged_attn_small_kernel(const __nv_bfloat16* __restrict__ q, int64_t ldq, const __nv_bfloat16* kc, const __nv_bfloat16* vc,
                        const int32_t* __restrict__ block_table, int bt_stride, const int32_t* __restrict__ row_to_seq,
                        const int32_t* __restrict__ ctx_lens, __nv_bfloat16* __restrict__ out, int64_t ldo, int hkv,
                        int block_size, float scale) {
  __shared__ float qs[G][kHeadDim];
  __shared__ float ps[G][kSmallMaxCtx];
  __shared__ float red[G][kSmallWarps];
  __shared__ float osum[kSmallWarps][kHeadDim];
  __shared__ int blk_s[kSmallMaxCtx / 64];
  const int kvh = blockIdx.x, row = blockIdx.y, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // context length and block ids are step inputs (written at the start of the step): safe to read before the wait
  const int ctx = min(ctx_lens[row], kSmallMaxCtx);
  const int seq = row_to_seq ? row_to_seq[row] : row;
  const int nblk = (ctx + block_size - 1) / block_size;
  if (tid < nblk) blk_s[tid] = block_table[(int64_t)seq * bt_stride + tid];
  pdl_wait();                 // q and the newest K/V row come from the previous kernel
  pdl_launch_dependents();
  for (int i = tid; i < G * kHeadDim; i += kSmallThreads)
    qs[i / kHeadDim][i % kHeadDim] = __bfloat162float(q[(int64_t)row * ldq + kvh * G * kHeadDim + i]);
  __syncthreads();
  // ---- scores ----
  float mx[G];
#pragma unroll
  for (int h = 0; h < G; ++h) mx[h] = -INFINITY;
  const int half = tid & 1;
  for (int t0 = 0; t0 < ctx; t0 += kSmallThreads / 2) {   // uniform trip count: the shuffle below needs the whole warp
    const int t = t0 + (tid >> 1);
    const bool valid = t < ctx;
    const int tc = valid ? t : 0;
    const int64_t base = (((int64_t)blk_s[tc / block_size] * hkv + kvh) * block_size + tc % block_size) * kHeadDim + half * 64;
    const uint4* kr = reinterpret_cast<const uint4*>(kc + base);
    uint4 kk[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) kk[c] = valid ? __ldcg(kr + c) : make_uint4(0u, 0u, 0u, 0u);
    float acc[G];
#pragma unroll
    for (int h = 0; h < G; ++h) acc[h] = 0.f;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const float kf[8] = {bf16lo(kk[c].x), bf16hi(kk[c].x), bf16lo(kk[c].y), bf16hi(kk[c].y),
                           bf16lo(kk[c].z), bf16hi(kk[c].z), bf16lo(kk[c].w), bf16hi(kk[c].w)};
#pragma unroll
      for (int h = 0; h < G; ++h) {
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[h] = fmaf(qs[h][half * 64 + c * 8 + j], kf[j], acc[h]);
      }
    }
#pragma unroll
    for (int h = 0; h < G; ++h) {
      const float sc = (acc[h] + __shfl_xor_sync(0xffffffffu, acc[h], 1)) * scale;
      if (valid) {
        if (half == 0) ps[h][t] = sc;
        mx[h] = fmaxf(mx[h], sc);
      }
    }
  }
#pragma unroll
  for (int h = 0; h < G; ++h) {
    const float wm = warp_max(mx[h]);
    if (lane == 0) red[h][warp] = wm;
  }
  __syncthreads();
  float m[G], sum[G];
#pragma unroll
  for (int h = 0; h < G; ++h) {
    float v = red[h][0];
#pragma unroll
    for (int w = 1; w < kSmallWarps; ++w) v = fmaxf(v, red[h][w]);
    m[h] = v;
    sum[h] = 0.f;
  }
  __syncthreads();   // red is reused for the sums
  for (int t = tid; t < ctx; t += kSmallThreads) {
#pragma unroll
    for (int h = 0; h < G; ++h) {
      const float pv = expf(ps[h][t] - m[h]);
      ps[h][t] = pv;
      sum[h] += pv;
    }
  }
#pragma unroll
  for (int h = 0; h < G; ++h) {
    const float wsum = warp_sum(sum[h]);
    if (lane == 0) red[h][warp] = wsum;
  }
  __syncthreads();
  // ---- output ----
  const int ch = tid & 15, tg = tid >> 4;   // 16-byte chunk of the row, token group (16 groups)
  float o[G][8];
#pragma unroll
  for (int h = 0; h < G; ++h)
#pragma unroll
    for (int j = 0; j < 8; ++j) o[h][j] = 0.f;
  for (int t0 = tg; t0 < ctx; t0 += 16 * 8) {
    uint4 vv[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int t = t0 + 16 * i;
      if (t < ctx)
        vv[i] = __ldcg(reinterpret_cast<const uint4*>(
            vc + (((int64_t)blk_s[t / block_size] * hkv + kvh) * block_size + t % block_size) * kHeadDim + ch * 8));
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int t = t0 + 16 * i;
      if (t < ctx) {
        const float vf[8] = {bf16lo(vv[i].x), bf16hi(vv[i].x), bf16lo(vv[i].y), bf16hi(vv[i].y),
                             bf16lo(vv[i].z), bf16hi(vv[i].z), bf16lo(vv[i].w), bf16hi(vv[i].w)};
#pragma unroll
        for (int h = 0; h < G; ++h) {
          const float pv = ps[h][t];
#pragma unroll
          for (int j = 0; j < 8; ++j) o[h][j] = fmaf(pv, vf[j], o[h][j]);
        }
      }
    }
  }
  // sum the 16 token groups: lanes l and l+16 of a warp hold two groups of the same chunk, then 8 warps through smem
#pragma unroll
  for (int h = 0; h < G; ++h) {
#pragma unroll
    for (int j = 0; j < 8; ++j) o[h][j] += __shfl_xor_sync(0xffffffffu, o[h][j], 16);
    if (lane < 16) {
#pragma unroll
      for (int j = 0; j < 8; ++j) osum[warp][ch * 8 + j] = o[h][j];
    }
    __syncthreads();
    if (tid < kHeadDim) {
      float tot = 0.f;
#pragma unroll
      for (int w = 0; w < kSmallWarps; ++w) tot += osum[w][tid];
      float l = 0.f;
#pragma unroll
      for (int w = 0; w < kSmallWarps; ++w) l += red[h][w];
      out[(int64_t)row * ldo + (kvh * G + h) * kHeadDim + tid] = __float2bfloat16_rn(l > 0.f ? tot / l : 0.f);
    }
    __syncthreads();
  }
}

}  // namespace

// decode rows only (one row per work item); the caller has checked paged_attn_small_ok
bool paged_attn_small_ok(int T, int hq, int hkv, int block_size, int max_ctx) {
  if (hkv <= 0 || hq % hkv) return false;
  const int G = hq / hkv;
  const char* env = getenv("B200_ATTN_SMALL");   // read per call (host side, outside graph replay): tests switch it
  const int on = env ? atoi(env) : 1;
  // few work items and short contexts: beyond that the streaming kernels win (their fixed cost amortises)
  return on && (G == 1 || G == 2 || G == 4 || G == 5 || G == 8) && max_ctx <= kSmallMaxCtx && block_size >= 64 &&
         (int64_t)T * hkv <= 2 * kNumSMs;
}

int paged_attn_small(const __nv_bfloat16* q, int64_t ldq, const __nv_bfloat16* kc, const __nv_bfloat16* vc,
                     const int32_t* block_table, int bt_stride, const int32_t* row_to_seq, const int32_t* ctx_lens,
                     __nv_bfloat16* out, int64_t ldo, int T, int hq, int hkv, int block_size, float scale, cudaStream_t s) {
  if (T == 0) return B200_OK;
  const int G = hq / hkv;
#define B200_SMALL_CASE(g)                                                                                                   \
  B200_CUDA(launch(paged_attn_small_kernel<g>, dim3(hkv, T), dim3(kSmallThreads), 0, s, q, ldq, kc, vc, block_table, bt_stride, \
                   row_to_seq, ctx_lens, out, ldo, hkv, block_size, scale))
  switch (G) {
    case 1: B200_SMALL_CASE(1); break;
    case 2: B200_SMALL_CASE(2); break;
    case 4: B200_SMALL_CASE(4); break;
    case 5: B200_SMALL_CASE(5); break;
    case 8: B200_SMALL_CASE(8); break;
    default: B200_REQUIRE(false, B200_ERR_UNSUPPORTED, "attn_small: G=%d", G);
  }
#undef B200_SMALL_CASE
  return B200_OK;
}

}  // namespace b200
