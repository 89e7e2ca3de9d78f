// (a6) Paged attention for single-query rows (decode, and prefill rows seen as independent
// queries with their own context length), split-KV with log-sum-exp merge.
//
// Roofline: HBM.  Algorithmic bytes per launch = sum over rows of ctx_len * n_kv_heads * 128 * 2 * 2
// (every cached K and V element of the row's context read once; q/out are noise).
//
// Data path: the cache layout [block][kv head][block_size][128] makes a 64-token K (or V) tile of
// one kv head one contiguous 16 KiB run, so a tile is ONE bulk-async copy (cp.async.bulk ->
// SASS UBLKCP) completing on an mbarrier; a 3-stage ring keeps 64 KiB of KV in flight per CTA and
// two CTAs fit per SM.  One CTA = (row, kv head, split); the G = hq/hkv query heads of the group
// share every K/V byte fetched (GQA).  Scores, softmax and the P.V accumulation are fp32 (R4 of
// oracle/decode_oracle.py); the online-softmax running max/sum are kept per head per CTA.
//
// Replaces: paged_attention_v1/v2 CPU kernels / IPEX single_query_cached_kv_attention of the
// vLLM-CPU image ([V22] vllm/v1/attention/backends/cpu_attn.py:377-396).
#include <stdlib.h>

#include "common.cuh"

namespace b200 {

namespace {

constexpr int kTile = 64;                        // tokens per smem tile
constexpr int kStages = 3;
constexpr int kAttnThreads = 128;
constexpr int kTileBytes = kTile * kHeadDim * 2;  // 16 KiB per K or V tile

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra.uni WAIT_DONE;\n"
      "bra.uni WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

struct AttnSmem {
  alignas(128) unsigned char k[kStages][kTileBytes];
  alignas(128) unsigned char v[kStages][kTileBytes];
  alignas(8) uint64_t full[kStages];
  float p[8][kTile];  // scores, then probabilities, [head][token]
  float alpha[8];
  float m[8];
  float l[8];
};

template <int G>
__global__ void __launch_bounds__(kAttnThreads)
paged_attn_kernel(const __nv_bfloat16* __restrict__ q, int64_t ldq, const __nv_bfloat16* __restrict__ kc,
                  const __nv_bfloat16* __restrict__ vc, const int32_t* __restrict__ block_table, int bt_stride,
                  const int32_t* __restrict__ row_to_seq, const int32_t* __restrict__ ctx_lens,
                  __nv_bfloat16* __restrict__ out, int64_t ldo, int hkv, int block_size, int tiles_per_split,
                  int n_splits, float scale, float* __restrict__ part_o, float* __restrict__ part_ml) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  AttnSmem& sm = *reinterpret_cast<AttnSmem*>(smem_raw);
  const int split = blockIdx.x, kvh = blockIdx.y, t = blockIdx.z;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int hq = hkv * G;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < kStages; ++s) mbar_init(&sm.full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (tid < 8) {
    sm.m[tid] = -INFINITY;
    sm.l[tid] = 0.f;
  }
  __syncthreads();
  pdl_wait();
  pdl_launch_dependents();

  const int ctx = ctx_lens[t];
  const int seq = row_to_seq ? row_to_seq[t] : t;
  const int n_tiles_total = (ctx + kTile - 1) / kTile;
  const int tile_begin = split * tiles_per_split;
  const int tile_end = min(tile_begin + tiles_per_split, n_tiles_total);
  const int n_tiles = max(tile_end - tile_begin, 0);
  const int32_t* bt = block_table + (int64_t)seq * bt_stride;

  auto issue = [&](int i) {  // called by thread 0: fetch tile (tile_begin + i) into stage i % kStages
    const int st = i % kStages;
    const int tok0 = (tile_begin + i) * kTile;
    const int64_t blk = bt[tok0 / block_size];
    const int off = tok0 % block_size;
    const int64_t e = ((blk * hkv + kvh) * block_size + off) * kHeadDim;
    mbar_expect_tx(&sm.full[st], 2 * kTileBytes);
    bulk_g2s(sm.k[st], kc + e, kTileBytes, &sm.full[st]);
    bulk_g2s(sm.v[st], vc + e, kTileBytes, &sm.full[st]);
  };
  if (tid == 0)
    for (int i = 0; i < min(kStages, n_tiles); ++i) issue(i);

  // QK mapping: 8 lanes per token, lane j owns dims [8j,8j+8) and [64+8j,64+8j+8): each 16-byte
  // shared load of a quarter warp is one contiguous 128 bytes (conflict free).
  const int j8 = lane & 7, tsub = lane >> 3;
  float qr[G][16];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const __nv_bfloat16* qp = q + (int64_t)t * ldq + (kvh * G + g) * kHeadDim;
    const uint4 a = *reinterpret_cast<const uint4*>(qp + 8 * j8);
    const uint4 b = *reinterpret_cast<const uint4*>(qp + 64 + 8 * j8);
    const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      qr[g][2 * i] = bf16lo(w[i]);
      qr[g][2 * i + 1] = bf16hi(w[i]);
    }
  }
  // PV mapping: 16 threads per token (8 dims each), 8 token streams.
  const int dg = tid & 15, ts = tid >> 4;
  float acc[G][8];
#pragma unroll
  for (int g = 0; g < G; ++g)
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[g][i] = 0.f;

  for (int it = 0; it < n_tiles; ++it) {
    const int st = it % kStages;
    const int tok0 = (tile_begin + it) * kTile;
    const int n_valid = min(kTile, ctx - tok0);
    mbar_wait(&sm.full[st], (it / kStages) & 1);

    // ---- scores ----
    const unsigned char* kt = sm.k[st];
#pragma unroll
    for (int i = 0; i < kTile / 16; ++i) {
      const int tok = warp * (kTile / 4) + i * 4 + tsub;
      const uint4 a = *reinterpret_cast<const uint4*>(kt + tok * 256 + 16 * j8);
      const uint4 b = *reinterpret_cast<const uint4*>(kt + tok * 256 + 128 + 16 * j8);
      const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
      float kf[16];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        kf[2 * e] = bf16lo(w[e]);
        kf[2 * e + 1] = bf16hi(w[e]);
      }
      float sc[G];
#pragma unroll
      for (int g = 0; g < G; ++g) {
        float d = 0.f;
#pragma unroll
        for (int e = 0; e < 16; ++e) d = fmaf(qr[g][e], kf[e], d);
        d += __shfl_xor_sync(0xffffffffu, d, 1);
        d += __shfl_xor_sync(0xffffffffu, d, 2);
        d += __shfl_xor_sync(0xffffffffu, d, 4);
        sc[g] = d;
      }
#pragma unroll
      for (int g = 0; g < G; ++g)
        if (j8 == g) sm.p[g][tok] = tok < n_valid ? sc[g] * scale : -INFINITY;
    }
    __syncthreads();

    // ---- online softmax: warp w owns heads w, w+4 ----
    for (int g = warp; g < G; g += 4) {
      const float s0 = sm.p[g][lane], s1 = sm.p[g][lane + 32];
      const float m_old = sm.m[g];
      const float m_new = fmaxf(m_old, warp_max(fmaxf(s0, s1)));  // finite: token tok0 is always valid
      const float p0 = __expf(s0 - m_new), p1 = __expf(s1 - m_new);
      sm.p[g][lane] = p0;
      sm.p[g][lane + 32] = p1;
      const float a = __expf(m_old - m_new);  // exp(-inf) = 0 on the first tile
      const float lsum = warp_sum(p0 + p1);
      if (lane == 0) {
        sm.alpha[g] = a;
        sm.m[g] = m_new;
        sm.l[g] = sm.l[g] * a + lsum;
      }
    }
    __syncthreads();

    // ---- P.V ----
    const unsigned char* vt = sm.v[st];
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const float a = sm.alpha[g];
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[g][i] *= a;
    }
#pragma unroll 2
    for (int tok = ts; tok < n_valid; tok += 8) {
      const uint4 a = *reinterpret_cast<const uint4*>(vt + tok * 256 + 16 * dg);
      const uint32_t w[4] = {a.x, a.y, a.z, a.w};
      float vf[8];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        vf[2 * e] = bf16lo(w[e]);
        vf[2 * e + 1] = bf16hi(w[e]);
      }
#pragma unroll
      for (int g = 0; g < G; ++g) {
        const float p = sm.p[g][tok];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[g][e] = fmaf(p, vf[e], acc[g][e]);
      }
    }
    __syncthreads();  // everyone is done with stage st and with sm.p
    if (tid == 0 && it + kStages < n_tiles) issue(it + kStages);
  }

  // ---- reduce the 8 token streams through shared memory (reuse K stage 0: 8*G*128*4 <= 32 KiB) ----
  float* red = reinterpret_cast<float*>(sm.k[0]);  // [8][G][128]; G=8 spills into k[1], also idle now
#pragma unroll
  for (int g = 0; g < G; ++g)
#pragma unroll
    for (int e = 0; e < 8; ++e) red[(ts * G + g) * kHeadDim + dg * 8 + e] = acc[g][e];
  __syncthreads();
  const int d = tid;  // 128 threads = 128 dims
#pragma unroll
  for (int g = 0; g < G; ++g) {
    float o = 0.f;
#pragma unroll
    for (int s = 0; s < 8; ++s) o += red[(s * G + g) * kHeadDim + d];
    const int head = kvh * G + g;
    if (n_splits == 1) {
      const float l = sm.l[g];
      out[(int64_t)t * ldo + head * kHeadDim + d] = __float2bfloat16_rn(l > 0.f ? o / l : 0.f);
    } else {
      const int64_t pi = ((int64_t)t * hq + head) * n_splits + split;
      part_o[pi * kHeadDim + d] = o;
      if (d == 0) {
        part_ml[pi * 2] = sm.m[g];
        part_ml[pi * 2 + 1] = sm.l[g];
      }
    }
  }
}

// Merge the splits of one (row, head): o = sum_s o_s * exp(m_s - m) / sum_s l_s * exp(m_s - m).
// The (m, l) pairs are loaded by one thread per split (parallel, not a dependent chain), the weights are
// shared through smem, and each thread's partial-output loads are issued in batches of 8 before the
// first dependent FMA — in-order issue would otherwise pay one L2 round trip per split.
__global__ void __launch_bounds__(kHeadDim)
attn_merge_kernel(const float* __restrict__ part_o, const float* __restrict__ part_ml, __nv_bfloat16* __restrict__ out,
                  int64_t ldo, int hq, int n_splits) {
  __shared__ float wgt[128];
  __shared__ float red_m[4], red_l[4];
  pdl_wait();
  pdl_launch_dependents();
  const int head = blockIdx.x, t = blockIdx.y, d = threadIdx.x;
  const int64_t base = ((int64_t)t * hq + head) * n_splits;
  float ms = -INFINITY, ls = 0.f;
  if (d < n_splits) {
    ms = part_ml[(base + d) * 2];
    ls = part_ml[(base + d) * 2 + 1];
    if (!(ls > 0.f)) ms = -INFINITY;
  }
  const float wm = warp_max(ms);
  if ((d & 31) == 0) red_m[d >> 5] = wm;
  __syncthreads();
  const float m = fmaxf(fmaxf(red_m[0], red_m[1]), fmaxf(red_m[2], red_m[3]));
  const float w = (d < n_splits && ls > 0.f) ? __expf(ms - m) : 0.f;
  wgt[d] = w;
  const float wl = warp_sum(ls * w);
  if ((d & 31) == 0) red_l[d >> 5] = wl;
  __syncthreads();
  const float l = red_l[0] + red_l[1] + red_l[2] + red_l[3];
  float o = 0.f;
  for (int s0 = 0; s0 < n_splits; s0 += 8) {
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = (s0 + j < n_splits) ? part_o[(base + s0 + j) * kHeadDim + d] : 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (s0 + j < n_splits) o = fmaf(v[j], wgt[s0 + j], o);
  }
  out[(int64_t)t * ldo + head * kHeadDim + d] = __float2bfloat16_rn(l > 0.f ? o / l : 0.f);
}

inline int pow2_at_least(int v, int lo) {
  int p = lo;
  while (p < v) p <<= 1;
  return p;
}

// Split count is a pure function of (T, hkv, bucketed max_ctx) so that a captured graph stays valid
// for every step in the same bucket.
void split_plan(int T, int hkv, int max_ctx, int* n_splits, int* tiles_per_split) {
  const int tiles = pow2_at_least(max(max_ctx, 1), 128) / kTile;
  const int ctas = max(T * hkv, 1);
  int s = (2 * kNumSMs + ctas - 1) / ctas;
  s = max(1, min(s, tiles / 2));  // at least 2 tiles (128 tokens) per split
  s = min(s, 64);
  *tiles_per_split = (tiles + s - 1) / s;
  *n_splits = (tiles + *tiles_per_split - 1) / *tiles_per_split;
}

template <int G>
int launch_attn(const __nv_bfloat16* q, int64_t ldq, const __nv_bfloat16* kc, const __nv_bfloat16* vc,
                const int32_t* block_table, int bt_stride, const int32_t* row_to_seq, const int32_t* ctx_lens,
                __nv_bfloat16* out, int64_t ldo, int T, int hkv, int block_size, int n_splits, int tiles_per_split,
                float scale, float* part_o, float* part_ml, cudaStream_t s) {
  static bool attr_set[64] = {};
  const size_t smem = sizeof(AttnSmem);
  B200_CUDA(ensure_dynamic_smem(paged_attn_kernel<G>, (int)smem, attr_set));
  B200_CUDA(launch(paged_attn_kernel<G>, dim3(n_splits, hkv, T), dim3(kAttnThreads), smem, s, q, ldq, kc, vc,
                   block_table, bt_stride, row_to_seq, ctx_lens, out, ldo, hkv, block_size, tiles_per_split, n_splits,
                   scale, part_o, part_ml));
  return B200_OK;
}

}  // namespace

// ---- tcgen05 variant (attn_tc.cu) -----------------------------------------------------------
bool paged_attn_tc_supported(int hq, int hkv, int block_size);
int64_t paged_attn_prefill_plan_bytes(int T);
bool paged_attn_prefill_fa_supported(int hq, int hkv, int block_size);
int paged_attn_prefill_plan(const int32_t* row_to_seq, const int32_t* ctx_lens, int T, int hq, int hkv, void* ws,
                            int64_t ws_bytes, cudaStream_t s);
int paged_attn_prefill_fa(const __nv_bfloat16* q, int64_t ldq, const __nv_bfloat16* kc, const __nv_bfloat16* vc,
                          const int32_t* block_table, int bt_stride, const int32_t* row_to_seq, const int32_t* ctx_lens,
                          __nv_bfloat16* out, int64_t ldo, int T, int hq, int hkv, int block_size, float scale, void* ws,
                          int64_t ws_bytes, bool plan_ready, cudaStream_t s);
int paged_attn_tc(const __nv_bfloat16* q, int64_t ldq, const __nv_bfloat16* kc, const __nv_bfloat16* vc,
                  const int32_t* block_table, int bt_stride, const int32_t* row_to_seq, const int32_t* ctx_lens,
                  __nv_bfloat16* out, int64_t ldo, int T, int hq, int hkv, int block_size, int max_ctx, float scale,
                  void* ws, int64_t ws_bytes, int pack, cudaStream_t s);
int attn_tc_pack_rows(int G);

// 128-token tiles, persistent CTAs (one per SM) that take work items round-robin in split-major
// order.  Rows are cut into chunks of C tiles so that there are several items per CTA (about 8 per
// SM): chunks cost the same, later splits of shorter rows are empty and skipped, so the static
// round-robin is balanced without a device-side queue.  Pure function of (T, hkv, bucketed max_ctx).
void attn_split_plan_tc(int T, int hkv, int max_ctx, int* n_splits, int* tiles_per_split) {
  const int tiles = pow2_at_least(max(max_ctx, 1), 128) / 128;
  const int64_t base = max((int64_t)T * hkv, (int64_t)1);
  int s = 1;
  if (base < 2 * kNumSMs) {   // enough rows x kv heads already: splitting only adds partials and a merge launch
    // >= 1024 tokens per item: with 512, 64 rows x 1 kv head at a context of 530 (one TP8 rank of 70B; the plan sees the
    // power-of-two bucket 1024) were cut into a 4-tile and a 1-tile item plus a merge launch
    const int64_t chunk = max((int64_t)8, (int64_t)tiles * base / (8 * kNumSMs));
    s = (int)min((int64_t)64, (tiles + chunk - 1) / chunk);
  }
  s = max(1, min(s, tiles));
  *tiles_per_split = (tiles + s - 1) / s;
  *n_splits = (tiles + *tiles_per_split - 1) / *tiles_per_split;
}

int launch_attn_merge(const float* part_o, const float* part_ml, __nv_bfloat16* out, int64_t ldo, int T, int hq,
                      int n_splits, cudaStream_t s) {
  B200_CUDA(launch(attn_merge_kernel, dim3(hq, T), dim3(kHeadDim), 0, s, part_o, part_ml, out, ldo, hq, n_splits));
  return B200_OK;
}

static bool use_tc_attn(int hq, int hkv, int block_size) {
  static const int env = getenv("B200_ATTN_TC") ? atoi(getenv("B200_ATTN_TC")) : 1;
  return env != 0 && paged_attn_tc_supported(hq, hkv, block_size);
}

// Bytes of split-KV scratch one launch of this shape uses (0: a single split writes the output).
int64_t paged_attn_scratch_need(int T, int hq, int hkv, int block_size, int max_ctx) {
  int ns, tps;
  if (use_tc_attn(hq, hkv, block_size))
    attn_split_plan_tc(max(T, 1), hkv, max_ctx, &ns, &tps);
  else
    split_plan(max(T, 1), hkv, max_ctx, &ns, &tps);
  return ns > 1 ? (int64_t)T * hq * ns * (kHeadDim + 2) * (int64_t)sizeof(float) : 0;
}

int64_t paged_attn_workspace_bytes(int T, int hq, int max_ctx, int block_size) {
  (void)block_size;
  // upper bound over every hkv >= 1 and both kernels' split plans
  int ns, tps, ns2;
  split_plan(T, 1, max_ctx, &ns, &tps);
  for (int r = 1; r <= 4; ++r) {   // row-at-a-time and every packed group size
    attn_split_plan_tc((T + r - 1) / r, 1, max_ctx, &ns2, &tps);
    ns = max(ns, ns2);
  }
  // + the work plan of the prefill kernel (attn_prefill.cu), which uses no split partials
  return (int64_t)T * hq * ns * (kHeadDim + 2) * sizeof(float) + paged_attn_prefill_plan_bytes(T) + 256;
}

static int paged_attn_cc(const __nv_bfloat16* q, int64_t ldq, const __nv_bfloat16* kc, const __nv_bfloat16* vc,
               const int32_t* block_table, int bt_stride, const int32_t* row_to_seq, const int32_t* ctx_lens,
               __nv_bfloat16* out, int64_t ldo, int T, int hq, int hkv, int block_size, int max_ctx, float scale,
               void* ws, int64_t ws_bytes, cudaStream_t s) {
  B200_REQUIRE(hkv > 0 && hq % hkv == 0, B200_ERR_INVALID, "attn: hq=%d not a multiple of hkv=%d", hq, hkv);
  B200_REQUIRE(block_size % kTile == 0, B200_ERR_UNSUPPORTED, "attn: block_size=%d must be a multiple of %d",
               block_size, kTile);
  B200_REQUIRE(ldq % 8 == 0 && ldo % 8 == 0, B200_ERR_INVALID, "attn: ldq/ldo must be multiples of 8");
  B200_REQUIRE(T <= 65535, B200_ERR_UNSUPPORTED, "attn: more than 65535 rows in one launch");
  if (T == 0) return B200_OK;
  int n_splits, tps;
  split_plan(T, hkv, max_ctx, &n_splits, &tps);
  float* part_o = nullptr;
  float* part_ml = nullptr;
  if (n_splits > 1) {
    const int64_t need = (int64_t)T * hq * n_splits * (kHeadDim + 2) * sizeof(float);
    B200_REQUIRE(ws && ws_bytes >= need, B200_ERR_INVALID, "attn: workspace too small (%lld < %lld)",
                 (long long)ws_bytes, (long long)need);
    part_o = static_cast<float*>(ws);
    part_ml = part_o + (int64_t)T * hq * n_splits * kHeadDim;
  }
  const int G = hq / hkv;
  int rc;
#define B200_ATTN_CASE(g)                                                                                          \
  case g:                                                                                                          \
    rc = launch_attn<g>(q, ldq, kc, vc, block_table, bt_stride, row_to_seq, ctx_lens, out, ldo, T, hkv, block_size, \
                        n_splits, tps, scale, part_o, part_ml, s);                                                 \
    break;
  switch (G) {
    B200_ATTN_CASE(1)
    B200_ATTN_CASE(2)
    B200_ATTN_CASE(4)
    B200_ATTN_CASE(5)
    B200_ATTN_CASE(8)
    default:
      B200_REQUIRE(false, B200_ERR_UNSUPPORTED, "attn: GQA group %d not built (1,2,4,5,8)", G);
  }
#undef B200_ATTN_CASE
  if (rc) return rc;
  if (n_splits > 1)
    B200_CUDA(launch(attn_merge_kernel, dim3(hq, T), dim3(kHeadDim), 0, s, (const float*)part_o,
                     (const float*)part_ml, out, ldo, hq, n_splits));
  return B200_OK;
}



// pack > 0: consecutive query rows of one sequence (a prefill chunk) share one pass over the K/V cache
//           (pack == 2: the caller already built the row plan in `ws` with paged_attn_prefill_plan_step);
// pack < 0: decode-only step (each row its own sequence, only position ctx-1 is new): early cache requests
int paged_attn_ex(const __nv_bfloat16* q, int64_t ldq, const __nv_bfloat16* kc, const __nv_bfloat16* vc,
                  const int32_t* block_table, int bt_stride, const int32_t* row_to_seq, const int32_t* ctx_lens,
                  __nv_bfloat16* out, int64_t ldo, int T, int hq, int hkv, int block_size, int max_ctx, float scale,
                  void* ws, int64_t ws_bytes, int pack, cudaStream_t s);

// Once per step, before the layer loop: builds the prefill kernel's row plan in `ws` when that kernel will
// serve the step (pack > 0).  Returns < 0 on error, 1 if the plan was built (pass pack = 2 to the layers), 0 if not.
int paged_attn_prefill_plan_step(const int32_t* row_to_seq, const int32_t* ctx_lens, int T, int hq, int hkv, int block_size,
                                 void* ws, int64_t ws_bytes, cudaStream_t s) {
  static const int fa_env = getenv("B200_ATTN_FA") ? atoi(getenv("B200_ATTN_FA")) : 1;
  if (!(fa_env && use_tc_attn(hq, hkv, block_size) && paged_attn_prefill_fa_supported(hq, hkv, block_size) && ws &&
        ws_bytes >= paged_attn_prefill_plan_bytes(T)))
    return 0;
  const int rc = paged_attn_prefill_plan(row_to_seq, ctx_lens, T, hq, hkv, ws, ws_bytes, s);
  return rc ? rc : 1;
}

int paged_attn(const __nv_bfloat16* q, int64_t ldq, const __nv_bfloat16* kc, const __nv_bfloat16* vc,
               const int32_t* block_table, int bt_stride, const int32_t* row_to_seq, const int32_t* ctx_lens,
               __nv_bfloat16* out, int64_t ldo, int T, int hq, int hkv, int block_size, int max_ctx, float scale,
               void* ws, int64_t ws_bytes, cudaStream_t s) {
  return paged_attn_ex(q, ldq, kc, vc, block_table, bt_stride, row_to_seq, ctx_lens, out, ldo, T, hq, hkv, block_size,
                       max_ctx, scale, ws, ws_bytes, 0, s);
}

int paged_attn_ex(const __nv_bfloat16* q, int64_t ldq, const __nv_bfloat16* kc, const __nv_bfloat16* vc,
                  const int32_t* block_table, int bt_stride, const int32_t* row_to_seq, const int32_t* ctx_lens,
                  __nv_bfloat16* out, int64_t ldo, int T, int hq, int hkv, int block_size, int max_ctx, float scale,
                  void* ws, int64_t ws_bytes, int pack, cudaStream_t s) {
  // prefill-dominated steps: query rows on the MMA M dimension (attn_prefill.cu); B200_ATTN_FA=0: the row
  // groups of attn_tc.cu instead
  static const int fa_env = getenv("B200_ATTN_FA") ? atoi(getenv("B200_ATTN_FA")) : 1;
  if (pack > 0 && fa_env && use_tc_attn(hq, hkv, block_size) && paged_attn_prefill_fa_supported(hq, hkv, block_size) &&
      ws && ws_bytes >= paged_attn_prefill_plan_bytes(T))
    return paged_attn_prefill_fa(q, ldq, kc, vc, block_table, bt_stride, row_to_seq, ctx_lens, out, ldo, T, hq, hkv,
                                 block_size, scale, ws, ws_bytes, /*plan_ready=*/pack == 2, s);
  if (use_tc_attn(hq, hkv, block_size))
    return paged_attn_tc(q, ldq, kc, vc, block_table, bt_stride, row_to_seq, ctx_lens, out, ldo, T, hq, hkv, block_size,
                         max_ctx, scale, ws, ws_bytes, pack, s);
  return paged_attn_cc(q, ldq, kc, vc, block_table, bt_stride, row_to_seq, ctx_lens, out, ldo, T, hq, hkv, block_size,
                       max_ctx, scale, ws, ws_bytes, s);
}

}  // namespace b200

// =================================================================================================
// b200_paged_attn_decode_static — the same split-KV attention for callers that may not allocate.
//
// Who needs it: an engine that sizes all of its device memory once at start-up and forbids any
// later allocation (a fixed memory plan under a static partition of the GPU, or a host runtime whose
// device-memory calls go through an arena the library cannot reach), and bindings written in
// languages/FFIs where handing the library a scratch pointer of the right size is awkward.  Such a
// caller cannot satisfy b200_paged_attn_decode's (workspace, workspace_bytes) contract by calling
// cudaMalloc, so the split partials live in the array below instead: it is part of the library's
// image, every GPU that loads the library carries its kStaticScratchBytes until the process exits,
// and nothing is allocated or freed by this entry point (no cudaMalloc/cudaFree, no torch).
//
// `g_attn_scratch` names DEVICE memory.  In host code that name is only a handle for the runtime —
// not an address a kernel can use — so the address is looked up with cudaGetSymbolAddress once per
// device (with that device current) and cached; the FIRST call on a device performs this look-up,
// which is why that first call should be made before any stream capture begins.
//
// Sharing: every call on a device uses the same array and finds whatever the previous call left in
// it.  Nothing in it has to start at zero — each launch writes every partial (o, m, l) it later
// merges, including those of empty splits, before the merge kernel reads them — so no clearing is
// done.  Two calls whose kernels overlap on the GPU (different streams, or different threads
// driving one device) would write the same bytes and race: issue them on one stream, or serialise.
// =================================================================================================
namespace b200 {
namespace {
constexpr int64_t kStaticScratchBytes = 2 << 20;  // 2 MiB: every decode-sized batch; the largest split plans (about 2.4 MiB) are refused
static __device__ __align__(256) unsigned char g_attn_scratch[kStaticScratchBytes];
constexpr int kMaxDevices = 64;
void* g_scratch_addr[kMaxDevices] = {};
}  // namespace
}  // namespace b200

extern "C" int64_t b200_paged_attn_static_scratch_bytes(void) { return b200::kStaticScratchBytes; }

extern "C" int b200_paged_attn_decode_static(const void* q, int64_t ldq, const void* k_cache_layer,
                                             const void* v_cache_layer, const int32_t* block_table, int32_t bt_stride,
                                             const int32_t* row_to_seq, const int32_t* ctx_lens, void* out, int64_t ldo,
                                             int32_t n_tokens, int32_t n_q_heads, int32_t n_kv_heads,
                                             int32_t block_size, int32_t max_ctx_len, float scale, void* stream) {
  using namespace b200;
  B200_REQUIRE(q && k_cache_layer && v_cache_layer && block_table && ctx_lens && out, B200_ERR_INVALID,
               "paged_attn_static: null argument");
  B200_REQUIRE(n_tokens >= 0 && n_q_heads > 0 && n_kv_heads > 0 && max_ctx_len > 0, B200_ERR_INVALID,
               "paged_attn_static: bad sizes");
  // exact need of this launch (0 when a single split writes the output directly)
  const int64_t need = paged_attn_scratch_need(n_tokens, n_q_heads, n_kv_heads, block_size, max_ctx_len);
  B200_REQUIRE(need <= kStaticScratchBytes, B200_ERR_UNSUPPORTED,
               "paged_attn_static: this shape needs %lld bytes of split scratch, the built-in array holds %lld "
               "(rows=%d q_heads=%d); use b200_paged_attn_decode with a caller workspace",
               (long long)need, (long long)kStaticScratchBytes, n_tokens, n_q_heads);
  int dev = -1;
  B200_CUDA(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < kMaxDevices, B200_ERR_UNSUPPORTED, "paged_attn_static: device ordinal %d >= %d", dev,
               kMaxDevices);
  if (!g_scratch_addr[dev]) {  // set-up, once per device, that device current
    void* p = nullptr;
    B200_CUDA(cudaGetSymbolAddress(&p, g_attn_scratch));
    B200_REQUIRE(p != nullptr, B200_ERR_CUDA, "paged_attn_static: cudaGetSymbolAddress returned a null address");
    g_scratch_addr[dev] = p;
  }
  return paged_attn((const __nv_bfloat16*)q, ldq, (const __nv_bfloat16*)k_cache_layer,
                    (const __nv_bfloat16*)v_cache_layer, block_table, bt_stride, row_to_seq, ctx_lens,
                    (__nv_bfloat16*)out, ldo, n_tokens, n_q_heads, n_kv_heads, block_size, max_ctx_len, scale,
                    g_scratch_addr[dev], kStaticScratchBytes, (cudaStream_t)stream);
}
