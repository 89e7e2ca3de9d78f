// The small ops of the decode step: embedding gather (a1), (add+)RMSNorm (a2), RoPE + KV append
// (a4+a5), SiLU*mul (a8), greedy argmax (a10).  All are latency/launch bound at decode sizes;
// their bytes are noise next to the weight stream, so the design goal is: vectorised 16-byte
// accesses, one pass over the data, rounding points identical to oracle/decode_oracle.py R1/R3/R5/R6.
#include "common.cuh"

namespace b200 {

namespace {

// ---------------------------------------------------------------------------------------
// (a1) embedding
// ---------------------------------------------------------------------------------------
// A negative id -(j+1) stands for "the token the previous step chose for its logit row j", still on the device
// (prev): a host that schedules step n+1 while step n runs does not know that token yet.
__device__ __forceinline__ int resolve_id(int id, int vocab, const int32_t* prev, int n_prev) {
  if (id < 0 && prev != nullptr && -id - 1 < n_prev) id = prev[-id - 1];
  return id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);  // ids are validated on the host; clamp, never fault
}

__global__ void __launch_bounds__(128)
embed_kernel(const int32_t* __restrict__ ids, const __nv_bfloat16* __restrict__ table,
             __nv_bfloat16* __restrict__ out, int hidden, int vocab, const int32_t* prev, int n_prev) {
  pdl_wait();
  pdl_launch_dependents();
  const int t = blockIdx.x;
  const int id = resolve_id(ids[t], vocab, prev, n_prev);
  const uint4* src = reinterpret_cast<const uint4*>(table + (int64_t)id * hidden);
  uint4* dst = reinterpret_cast<uint4*>(out + (int64_t)t * hidden);
  for (int c = threadIdx.x; c < hidden / 8; c += blockDim.x) dst[c] = __ldg(src + c);
}

// ---------------------------------------------------------------------------------------
// (a2) RMSNorm / fused add + RMSNorm.  One CTA per token row, the row lives in registers.
//   s = x (+ res) in fp32; res' = bf16(s); var over UNROUNDED s; n = bf16(s / sqrt(var+eps));
//   y = bf16(n * w)                                   [V22] vllm/ir/ops/layernorm.py:10-21,40-58
// ---------------------------------------------------------------------------------------
// The sum of squares is taken in a fixed order that the linears' fused-norm path reproduces exactly (gemm_tc.cu,
// finalize4<kResidual> / norm_b_tiles): per 128-wide slab, lane l squares its 4 values 4l..4l+3 in order (fma chain),
// the warp tree (xor 16, 8, 4, 2, 1) adds the lanes, and the slab sums are added in ascending slab order.
constexpr int kNormThreads = 256;
constexpr int kNormWarps = kNormThreads / 32;
constexpr int kNormMaxSlabs = 8;  // per warp: hidden <= 8 * 8 * 128 = 8192

template <bool HAS_RES>
__global__ void __launch_bounds__(kNormThreads)
rmsnorm_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* res, const __nv_bfloat16* __restrict__ w,
               __nv_bfloat16* y, int hidden, float eps) {
  __shared__ float part[kNormWarps * kNormMaxSlabs];
  pdl_wait();
  pdl_launch_dependents();
  const int t = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nslab = hidden >> 7;
  const uint2* xr = reinterpret_cast<const uint2*>(x + (int64_t)t * hidden);
  uint2* rr = HAS_RES ? reinterpret_cast<uint2*>(res + (int64_t)t * hidden) : nullptr;
  float v[kNormMaxSlabs][4];
#pragma unroll
  for (int i = 0; i < kNormMaxSlabs; ++i) {
    const int slab = warp + i * kNormWarps;
    if (slab < nslab) {
      const int c = slab * 32 + lane;   // 8-byte chunk index
      const uint2 a = xr[c];
      if (HAS_RES) {
        const uint2 b = rr[c];
        v[i][0] = bf16lo(a.x) + bf16lo(b.x);
        v[i][1] = bf16hi(a.x) + bf16hi(b.x);
        v[i][2] = bf16lo(a.y) + bf16lo(b.y);
        v[i][3] = bf16hi(a.y) + bf16hi(b.y);
        rr[c] = make_uint2(pack_bf16(v[i][0], v[i][1]), pack_bf16(v[i][2], v[i][3]));
      } else {
        v[i][0] = bf16lo(a.x);
        v[i][1] = bf16hi(a.x);
        v[i][2] = bf16lo(a.y);
        v[i][3] = bf16hi(a.y);
      }
      float ss = fmaf(v[i][0], v[i][0], 0.f);
      ss = fmaf(v[i][1], v[i][1], ss);
      ss = fmaf(v[i][2], v[i][2], ss);
      ss = fmaf(v[i][3], v[i][3], ss);
      ss = warp_sum(ss);
      if (lane == 0) part[slab] = ss;
    }
  }
  __syncthreads();
  float tot = 0.f;
  for (int i = 0; i < nslab; ++i) tot += part[i];
  const float r = 1.0f / sqrtf(tot / (float)hidden + eps);
  const uint2* wr = reinterpret_cast<const uint2*>(w);
  uint2* yr = reinterpret_cast<uint2*>(y + (int64_t)t * hidden);
#pragma unroll
  for (int i = 0; i < kNormMaxSlabs; ++i) {
    const int slab = warp + i * kNormWarps;
    if (slab < nslab) {
      const int c = slab * 32 + lane;
      const uint2 wq = __ldg(wr + c);
      const float n0 = round_bf16(v[i][0] * r), n1 = round_bf16(v[i][1] * r);
      const float n2 = round_bf16(v[i][2] * r), n3 = round_bf16(v[i][3] * r);
      yr[c] = make_uint2(pack_bf16(n0 * bf16lo(wq.x), n1 * bf16hi(wq.x)), pack_bf16(n2 * bf16lo(wq.y), n3 * bf16hi(wq.y)));
    }
  }
}

// Embedding rows into the fp32 residual stream: sres[t] = float(table[ids[t]]), ssq[t][slab] in the order above.
__global__ void __launch_bounds__(kNormThreads)
embed_stream_kernel(const int32_t* __restrict__ ids, const __nv_bfloat16* __restrict__ table, float* __restrict__ sres,
                    float* __restrict__ ssq, int hidden, int vocab, const int32_t* prev, int n_prev) {
  pdl_wait();
  pdl_launch_dependents();
  const int t = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nslab = hidden >> 7;
  const int id = resolve_id(ids[t], vocab, prev, n_prev);
  const uint2* src = reinterpret_cast<const uint2*>(table + (int64_t)id * hidden);
  float4* dst = reinterpret_cast<float4*>(sres + (int64_t)t * hidden);
  for (int slab = warp; slab < nslab; slab += kNormWarps) {
    const uint2 a = __ldg(src + slab * 32 + lane);
    const float4 f = make_float4(bf16lo(a.x), bf16hi(a.x), bf16lo(a.y), bf16hi(a.y));
    dst[slab * 32 + lane] = f;
    float ss = fmaf(f.x, f.x, 0.f);
    ss = fmaf(f.y, f.y, ss);
    ss = fmaf(f.z, f.z, ss);
    ss = fmaf(f.w, f.w, ss);
    ss = warp_sum(ss);
    if (lane == 0) ssq[(int64_t)t * nslab + slab] = ss;
  }
}

// Stream rows -> normalised bf16 rows (the consumer side of the folded norm as a kernel of its own, for consumers
// that would re-normalise the same rows in too many CTAs: lm_head over many logit rows).
__global__ void __launch_bounds__(kNormThreads)
rmsnorm_stream_kernel(const float* __restrict__ sres, const float* __restrict__ ssq, const int32_t* __restrict__ rows,
                      const __nv_bfloat16* __restrict__ w, __nv_bfloat16* __restrict__ y, int hidden, float eps) {
  pdl_wait();
  pdl_launch_dependents();
  const int t = blockIdx.x;
  const int row = rows ? rows[t] : t;
  const int nslab = hidden >> 7;
  float tot = 0.f;
  for (int i = 0; i < nslab; ++i) tot += __ldcg(ssq + (int64_t)row * nslab + i);
  const float r = 1.0f / sqrtf(tot / (float)hidden + eps);
  const float4* sr = reinterpret_cast<const float4*>(sres + (int64_t)row * hidden);
  const uint2* wr = reinterpret_cast<const uint2*>(w);
  uint2* yr = reinterpret_cast<uint2*>(y + (int64_t)t * hidden);
  for (int c = threadIdx.x; c < hidden / 4; c += kNormThreads) {
    const float4 s4 = __ldcg(sr + c);
    const uint2 wq = __ldg(wr + c);
    const float n0 = round_bf16(s4.x * r), n1 = round_bf16(s4.y * r), n2 = round_bf16(s4.z * r), n3 = round_bf16(s4.w * r);
    yr[c] = make_uint2(pack_bf16(n0 * bf16lo(wq.x), n1 * bf16hi(wq.x)), pack_bf16(n2 * bf16lo(wq.y), n3 * bf16hi(wq.y)));
  }
}

// ---------------------------------------------------------------------------------------
// (a4+a5) NeoX RoPE in bf16 arithmetic on q and k heads, then K/V rows into the paged cache.
// grid (token, head), 64 threads: thread i owns the pair (i, i+64).
// cache layout per layer: [block][kv head][block_size][128].
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(64)
rope_kv_kernel(__nv_bfloat16* qkv, const int32_t* __restrict__ positions, const int32_t* __restrict__ slots,
               const __nv_bfloat16* __restrict__ cos_sin, __nv_bfloat16* kc, __nv_bfloat16* vc, int hq, int hkv,
               int block_size) {
  pdl_wait();
  pdl_launch_dependents();
  const int t = blockIdx.x, head = blockIdx.y, i = threadIdx.x;
  const int row_w = (hq + 2 * hkv) * kHeadDim;
  __nv_bfloat16* hp = qkv + (int64_t)t * row_w + head * kHeadDim;
  const int slot = slots[t];
  if (head < hq + hkv) {
    const int pos = positions[t];
    const __nv_bfloat16* cs = cos_sin + (int64_t)pos * kHeadDim;
    const float c = __bfloat162float(cs[i]), s = __bfloat162float(cs[i + 64]);
    const float x1 = __bfloat162float(hp[i]), x2 = __bfloat162float(hp[i + 64]);
    // every elementwise op rounds to bf16 (R3)
    const float o1 = round_bf16(round_bf16(x1 * c) - round_bf16(x2 * s));
    const float o2 = round_bf16(round_bf16(x2 * c) + round_bf16(x1 * s));
    if (head < hq) {
      hp[i] = __float2bfloat16_rn(o1);
      hp[i + 64] = __float2bfloat16_rn(o2);
    } else if (slot >= 0) {
      const int kvh = head - hq;
      const int64_t blk = slot / block_size, off = slot % block_size;
      __nv_bfloat16* dst = kc + ((blk * hkv + kvh) * block_size + off) * kHeadDim;
      dst[i] = __float2bfloat16_rn(o1);
      dst[i + 64] = __float2bfloat16_rn(o2);
    }
  } else if (slot >= 0) {
    const int kvh = head - hq - hkv;
    const int64_t blk = slot / block_size, off = slot % block_size;
    __nv_bfloat16* dst = vc + ((blk * hkv + kvh) * block_size + off) * kHeadDim;
    dst[i] = hp[i];
    dst[i + 64] = hp[i + 64];
  }
}

// ---------------------------------------------------------------------------------------
// (a8) SiLU*mul: a = bf16(bf16(silu(g)) * u)             [V22] activation.py:137-141
// ---------------------------------------------------------------------------------------

__global__ void __launch_bounds__(256)
silu_mul_kernel(const __nv_bfloat16* __restrict__ gu, __nv_bfloat16* __restrict__ out, int inter, int64_t total_chunks) {
  pdl_wait();
  pdl_launch_dependents();
  const int cpr = inter >> 3;  // chunks per row
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total_chunks;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t t = idx / cpr;
    const int c = (int)(idx % cpr);
    const uint4 g = *reinterpret_cast<const uint4*>(gu + t * 2 * inter + c * 8);
    const uint4 u = *reinterpret_cast<const uint4*>(gu + t * 2 * inter + inter + c * 8);
    const uint32_t gw[4] = {g.x, g.y, g.z, g.w}, uw[4] = {u.x, u.y, u.z, u.w};
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
      o[j] = pack_bf16(silu_to_bf16(bf16lo(gw[j])) * bf16lo(uw[j]), silu_to_bf16(bf16hi(gw[j])) * bf16hi(uw[j]));
    *reinterpret_cast<uint4*>(out + t * inter + c * 8) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// ---------------------------------------------------------------------------------------
// (a10) greedy argmax per row, fp32 compare, ties -> lowest index.  [V22] sampler.py:90-91,235-236
// ---------------------------------------------------------------------------------------
constexpr int kArgmaxThreads = 1024;

__device__ __forceinline__ void take(float& bv, int& bi, float v, int i) {
  if (v > bv || (v == bv && i < bi)) {
    bv = v;
    bi = i;
  }
}

__global__ void __launch_bounds__(kArgmaxThreads)
argmax_kernel(const __nv_bfloat16* __restrict__ logits, int64_t ld, int vocab, int index_offset,
              int32_t* __restrict__ out_ids, float* __restrict__ out_max) {
  __shared__ float sv[kArgmaxThreads / 32];
  __shared__ int si[kArgmaxThreads / 32];
  pdl_wait();
  pdl_launch_dependents();
  const __nv_bfloat16* row = logits + (int64_t)blockIdx.x * ld;
  float bv = -INFINITY;
  int bi = 0x7fffffff;
  const int nchunk = vocab >> 3;
  for (int c = threadIdx.x; c < nchunk; c += kArgmaxThreads) {
    const uint4 q = *reinterpret_cast<const uint4*>(row + c * 8);
    const uint32_t qw[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      take(bv, bi, bf16lo(qw[j]), c * 8 + 2 * j);
      take(bv, bi, bf16hi(qw[j]), c * 8 + 2 * j + 1);
    }
  }
  for (int i = nchunk * 8 + threadIdx.x; i < vocab; i += kArgmaxThreads) take(bv, bi, __bfloat162float(row[i]), i);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    take(bv, bi, ov, oi);
  }
  if ((threadIdx.x & 31) == 0) {
    sv[threadIdx.x >> 5] = bv;
    si[threadIdx.x >> 5] = bi;
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    bv = sv[threadIdx.x];
    bi = si[threadIdx.x];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      take(bv, bi, ov, oi);
    }
    if (threadIdx.x == 0) {
      out_ids[blockIdx.x] = bi + index_offset;
      if (out_max) out_max[blockIdx.x] = bv;
    }
  }
}

// ---------------------------------------------------------------------------------------
// (a0) device-side advance of pure-decode rows: the sampled token becomes the next input, position
// and context grow by one, the slot follows the block table.  Lets a caller run several decode
// steps without a host round trip (inputs stay resident in HBM).
// ---------------------------------------------------------------------------------------
__global__ void advance_rows_kernel(int32_t* ids, int32_t* positions, int32_t* slots, int32_t* ctx_lens,
                                    const int32_t* __restrict__ block_table, int bt_stride,
                                    const int32_t* __restrict__ row_to_seq, const int32_t* __restrict__ sampled, int n,
                                    int block_size) {
  pdl_wait();
  pdl_launch_dependents();
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int p = positions[t] + 1;
  const int seq = row_to_seq ? row_to_seq[t] : t;
  ids[t] = sampled[t];
  positions[t] = p;
  ctx_lens[t] = p + 1;
  slots[t] = block_table[(int64_t)seq * bt_stride + p / block_size] * block_size + p % block_size;
}

}  // namespace

int advance_rows(int32_t* ids, int32_t* positions, int32_t* slots, int32_t* ctx_lens, const int32_t* block_table,
                 int bt_stride, const int32_t* row_to_seq, const int32_t* sampled, int n, int block_size,
                 cudaStream_t s) {
  if (n == 0) return B200_OK;
  B200_CUDA(launch(advance_rows_kernel, dim3((n + 127) / 128), dim3(128), 0, s, ids, positions, slots, ctx_lens,
                   block_table, bt_stride, row_to_seq, sampled, n, block_size));
  return B200_OK;
}

int embed(const int32_t* ids, const __nv_bfloat16* table, __nv_bfloat16* out, int T, int hidden, int vocab,
          cudaStream_t s, const int32_t* prev, int n_prev) {
  B200_REQUIRE(hidden % 8 == 0, B200_ERR_UNSUPPORTED, "embed: hidden %% 8 != 0");
  if (T == 0) return B200_OK;
  B200_CUDA(launch(embed_kernel, dim3(T), dim3(128), 0, s, ids, table, out, hidden, vocab, prev, n_prev));
  return B200_OK;
}

int rmsnorm(const __nv_bfloat16* x, __nv_bfloat16* res, const __nv_bfloat16* w, __nv_bfloat16* y, int T,
            int hidden, float eps, cudaStream_t s) {
  B200_REQUIRE(hidden % 128 == 0 && hidden <= 128 * kNormWarps * kNormMaxSlabs, B200_ERR_UNSUPPORTED,
               "rmsnorm: hidden=%d unsupported (multiple of 128, <= %d)", hidden, 128 * kNormWarps * kNormMaxSlabs);
  if (T == 0) return B200_OK;
  if (res)
    B200_CUDA(launch(rmsnorm_kernel<true>, dim3(T), dim3(kNormThreads), 0, s, x, res, w, y, hidden, eps));
  else
    B200_CUDA(launch(rmsnorm_kernel<false>, dim3(T), dim3(kNormThreads), 0, s, x, res, w, y, hidden, eps));
  return B200_OK;
}

int embed_stream(const int32_t* ids, const __nv_bfloat16* table, float* sres, float* ssq, int T, int hidden, int vocab,
                 cudaStream_t s, const int32_t* prev, int n_prev) {
  B200_REQUIRE(hidden % 128 == 0, B200_ERR_UNSUPPORTED, "embed_stream: hidden %% 128 != 0");
  if (T == 0) return B200_OK;
  B200_CUDA(launch(embed_stream_kernel, dim3(T), dim3(kNormThreads), 0, s, ids, table, sres, ssq, hidden, vocab, prev, n_prev));
  return B200_OK;
}

int rmsnorm_stream(const float* sres, const float* ssq, const int32_t* rows, const __nv_bfloat16* w, __nv_bfloat16* y,
                   int R, int hidden, float eps, cudaStream_t s) {
  B200_REQUIRE(hidden % 128 == 0, B200_ERR_UNSUPPORTED, "rmsnorm_stream: hidden %% 128 != 0");
  if (R == 0) return B200_OK;
  B200_CUDA(launch(rmsnorm_stream_kernel, dim3(R), dim3(kNormThreads), 0, s, sres, ssq, rows, w, y, hidden, eps));
  return B200_OK;
}

int rope_kv_append(__nv_bfloat16* qkv, const int32_t* positions, const int32_t* slots,
                   const __nv_bfloat16* cos_sin, __nv_bfloat16* kc, __nv_bfloat16* vc, int T, int hq,
                   int hkv, int block_size, cudaStream_t s) {
  if (T == 0) return B200_OK;
  B200_CUDA(launch(rope_kv_kernel, dim3(T, hq + 2 * hkv), dim3(64), 0, s, qkv, positions, slots, cos_sin, kc, vc,
                   hq, hkv, block_size));
  return B200_OK;
}

int silu_mul(const __nv_bfloat16* gu, __nv_bfloat16* out, int T, int inter, cudaStream_t s) {
  B200_REQUIRE(inter % 8 == 0, B200_ERR_UNSUPPORTED, "silu_mul: intermediate %% 8 != 0");
  if (T == 0) return B200_OK;
  const int64_t chunks = (int64_t)T * (inter / 8);
  int64_t grid64 = (chunks + 255) / 256;
  if (grid64 > (int64_t)kNumSMs * 8) grid64 = (int64_t)kNumSMs * 8;
  const int grid = (int)grid64;
  B200_CUDA(launch(silu_mul_kernel, dim3(grid), dim3(256), 0, s, gu, out, inter, chunks));
  return B200_OK;
}

int argmax_rows(const __nv_bfloat16* logits, int64_t ld, int rows, int vocab, int index_offset,
                int32_t* out_ids, float* out_max, cudaStream_t s) {
  B200_REQUIRE(ld % 8 == 0, B200_ERR_UNSUPPORTED, "argmax: ld %% 8 != 0");
  if (rows == 0) return B200_OK;
  B200_CUDA(launch(argmax_kernel, dim3(rows), dim3(kArgmaxThreads), 0, s, logits, ld, vocab, index_offset, out_ids,
                   out_max));
  return B200_OK;
}

}  // namespace b200
