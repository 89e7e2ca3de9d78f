// (a10, widened) Token sampling over bf16 logits rows: temperature, top-k, top-p, seeded.
//
// The hot path of the reference's clients is greedy ("temperature": 0, docs/accessing-deployed-models.md:176)
// and stays on argmax_kernel; this kernel serves requests that ask for anything else through the same
// OpenAI fields (docs/api-spec.yaml: temperature, top_p, seed; top_k is vLLM's extension).  Semantics follow
// vLLM's v1 sampler ([V22] vllm/v1/sample/sampler.py:90-91,235-256; ops/topk_topp_sampler.py): logits / T,
// top-k mask (values below the k-th largest are dropped, ties at it kept), then top-p on the renormalised
// rest (the smallest-probability tokens whose cumulative mass is <= 1 - p are dropped), then one draw from
// what is left.  A row with temperature 0 returns the argmax (lowest index on ties), bit-identical to
// argmax_kernel.
//
// One CTA per row.  The logits are bf16, so a token's sort key has 16 significant bits: the k-th largest
// value and the top-p cut are found by a two-level radix select (2 x 256-bin histograms in shared memory),
// with the probability mass accumulated in 2^-32 fixed point (integer atomics: the cut does not depend on
// the order of the additions, so a seed reproduces its tokens).  The draw is Gumbel-max: argmax over the
// kept tokens of logit / T - log(-log(u_i)), u_i from Philox-4x32-10 keyed by the request's seed with
// counter (token index, draw number): exact softmax sampling, no prefix sums, deterministic per seed.
// Bound: L2 (the row is read 3-7 times, 256 KiB per pass at V = 128 k).
#include <stdint.h>

#include "common.cuh"

namespace b200 {
namespace {

constexpr int kSampleThreads = 1024;

__device__ __forceinline__ uint32_t key16(__nv_bfloat16 v) {   // order-preserving: larger logit, larger key
  const uint32_t b = __bfloat16_as_ushort(v);
  return (b & 0x8000u) ? (~b & 0xffffu) : (b | 0x8000u);
}

__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
  const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
  const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
  c[0] = n0;
  c[1] = lo1;
  c[2] = n2;
  c[3] = lo0;
}
// 32 random bits -> uniform strictly inside (0, 1): the top 23 bits plus one half, exact in fp32 (24 significant
// bits), so that neither -log(u) nor -log(-log(u)) can be infinite.  (24 bits + 0.5 would round to 1.0 for the
// largest draw: that token would win the Gumbel-max whatever its logit.)
__host__ __device__ __forceinline__ float uniform_from_bits(uint32_t bits) {
  return ((float)(bits >> 9) + 0.5f) * (1.0f / 8388608.0f);
}
// uniform in (0, 1) for (seed, draw number, token index)
__device__ __forceinline__ float philox_uniform(uint64_t seed, uint64_t draw, uint32_t index) {
  uint32_t c[4] = {index, (uint32_t)draw, (uint32_t)(draw >> 32), 0x5a17u};
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    philox_round(c, k0, k1);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return uniform_from_bits(c[0]);
}

struct Best {
  float v;
  int i;
};
__device__ __forceinline__ Best better(Best a, Best b) {   // larger value, then lower index
  return (b.v > a.v || (b.v == a.v && b.i < a.i)) ? b : a;
}
__device__ __forceinline__ Best block_best(Best x, Best* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    Best y;
    y.v = __shfl_xor_sync(0xffffffffu, x.v, o);
    y.i = __shfl_xor_sync(0xffffffffu, x.i, o);
    x = better(x, y);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) sh[warp] = x;
  __syncthreads();
  if (warp == 0) {
    x = sh[lane];   // kSampleThreads / 32 == 32 warps
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      Best y;
      y.v = __shfl_xor_sync(0xffffffffu, x.v, o);
      y.i = __shfl_xor_sync(0xffffffffu, x.i, o);
      x = better(x, y);
    }
    if (lane == 0) sh[0] = x;
  }
  __syncthreads();
  const Best r = sh[0];
  __syncthreads();
  return r;
}

// Token i of row r lives at logits[(i / seg_len) * seg_stride + r * ld + i % seg_len]: one segment for a single
// GPU, one segment per rank for vocab-parallel logits gathered rank-major.
struct RowView {
  const __nv_bfloat16* base;
  int seg_len;
  int64_t seg_stride;
  __device__ __forceinline__ __nv_bfloat16 at(int i) const {
    return base[(int64_t)(i / seg_len) * seg_stride + (i % seg_len)];
  }
};

__global__ void __launch_bounds__(kSampleThreads)
sample_kernel(const __nv_bfloat16* __restrict__ logits, int64_t ld, int vocab, int seg_len, int64_t seg_stride,
              const float* __restrict__ temperature, const int32_t* __restrict__ top_k, const float* __restrict__ top_p,
              const float* __restrict__ min_p, const uint64_t* __restrict__ seeds, const uint64_t* __restrict__ draws,
              int32_t* __restrict__ out) {
  __shared__ Best sh_best[32];
  __shared__ unsigned int hist[256];
  __shared__ unsigned long long mass[256];
  __shared__ unsigned int sel[4];   // selected bin, count above it, ...
  __shared__ unsigned long long msel[2];
  pdl_wait();
  pdl_launch_dependents();
  const int row = blockIdx.x, tid = threadIdx.x;
  const RowView lv{logits + (int64_t)row * ld, seg_len, seg_stride};
  const float T = temperature ? temperature[row] : 0.f;

  // ---- greedy row, and the row maximum for everything else
  Best b{-INFINITY, 0x7fffffff};
  for (int i = tid; i < vocab; i += kSampleThreads) b = better(b, Best{__bfloat162float(lv.at(i)), i});
  b = block_best(b, sh_best);
  if (!(T > 0.f)) {
    if (tid == 0) out[row] = b.i;
    return;
  }
  const float inv_t = 1.0f / T, vmax = b.v;
  const int k = top_k ? top_k[row] : 0;
  const float p = top_p ? top_p[row] : 1.0f;
  uint32_t thr = 0;   // keep tokens with key16 >= thr
  // min_p ([V22] vllm/v1/sample/logits_processor: after the temperature, before top-k/top-p): keep tokens whose
  // probability is >= min_p x the largest one, i.e. (l - max) / T >= log(min_p).  The tokens it drops are the
  // lowest ones, so the k-th largest value of top-k is unaffected; top-p's mass is taken over what it leaves.
  const float mp = min_p ? min_p[row] : 0.f;
  const float min_logit = (mp > 0.f && mp <= 1.f) ? vmax + T * logf(mp) : -INFINITY;

  // ---- top-k: key of the k-th largest value (two-level radix select on the 16-bit key)
  if (k > 0 && k < vocab) {
    uint32_t prefix = 0, above = 0;
    for (int level = 0; level < 2; ++level) {
      for (int j = tid; j < 256; j += kSampleThreads) hist[j] = 0u;
      __syncthreads();
      for (int i = tid; i < vocab; i += kSampleThreads) {
        const uint32_t key = key16(lv.at(i));
        if (level == 0)
          atomicAdd(&hist[key >> 8], 1u);
        else if ((key >> 8) == prefix)
          atomicAdd(&hist[key & 255u], 1u);
      }
      __syncthreads();
      if (tid == 0) {
        uint32_t cum = above;
        int bin = 255;
        for (; bin > 0; --bin) {
          if (cum + hist[bin] >= (uint32_t)k) break;
          cum += hist[bin];
        }
        sel[0] = (uint32_t)bin;
        sel[1] = cum;
      }
      __syncthreads();
      if (level == 0) {
        prefix = sel[0];
        above = sel[1];
      } else {
        thr = (prefix << 8) | sel[0];
      }
      __syncthreads();
    }
  }

  // ---- top-p over what top-k left: drop from the bottom while the cumulative mass stays <= (1 - p) Z
  if (p < 1.0f) {
    const float keep = fmaxf(p, 0.f);
    uint32_t prefix = 0;
    unsigned long long below = 0, total = 0;
    for (int level = 0; level < 2; ++level) {
      for (int j = tid; j < 256; j += kSampleThreads) mass[j] = 0ull;
      __syncthreads();
      for (int i = tid; i < vocab; i += kSampleThreads) {
        const __nv_bfloat16 v = lv.at(i);
        const uint32_t key = key16(v);
        if (key < thr || __bfloat162float(v) < min_logit) continue;
        if (level == 1 && (key >> 8) != prefix) continue;
        const float e = __expf((__bfloat162float(v) - vmax) * inv_t);                    // in (0, 1]
        const unsigned long long fx = (unsigned long long)(e * 4294967296.0f);           // 2^-32 fixed point
        atomicAdd(&mass[level == 0 ? (key >> 8) : (key & 255u)], fx);
      }
      __syncthreads();
      if (tid == 0) {
        if (level == 0) {
          unsigned long long z = 0;
          for (int j = 0; j < 256; ++j) z += mass[j];
          total = z;
          msel[1] = z;
        } else {
          total = msel[1];
        }
        // largest mass that may be dropped: (1 - p) Z, in fixed point
        const unsigned long long drop = (unsigned long long)((double)total * (double)(1.0f - keep));
        unsigned long long cum = below;
        int bin = 0;
        for (; bin < 255; ++bin) {
          if (cum + mass[bin] > drop) break;
          cum += mass[bin];
        }
        sel[0] = (uint32_t)bin;
        msel[0] = cum;
      }
      __syncthreads();
      if (level == 0) {
        prefix = sel[0];
        below = msel[0];
      } else {
        const uint32_t cut = (prefix << 8) | sel[0];
        thr = max(thr, cut);
      }
      __syncthreads();
    }
  }

  // ---- one draw from the kept tokens: Gumbel-max
  const uint64_t seed = seeds ? seeds[row] : 0ull, draw = draws ? draws[row] : 0ull;
  Best s{-INFINITY, 0x7fffffff};
  for (int i = tid; i < vocab; i += kSampleThreads) {
    const __nv_bfloat16 v = lv.at(i);
    if (key16(v) < thr || __bfloat162float(v) < min_logit) continue;
    const float u = philox_uniform(seed, draw, (uint32_t)i);
    const float g = -logf(-logf(u));
    s = better(s, Best{(__bfloat162float(v) - vmax) * inv_t + g, i});
  }
  s = block_best(s, sh_best);
  if (tid == 0) out[row] = s.i < vocab ? s.i : b.i;
}

// ---- log-probabilities of chosen tokens + the top-N alternatives (OpenAI `logprobs` / `top_logprobs`) ----------
// log softmax over the RAW logits (vLLM's default logprobs_mode; [V22] vllm/v1/sample/sampler.py:90-118): for row r
//   out_lp[r] = l[tok[r]] - logsumexp(l),  top_ids/top_lp[r][0..n_top) = the n_top largest, value descending, ties by
// lowest index.  One CTA per row; the N-th largest value by the same two-level radix select as top-k.
constexpr int kTopCap = 96;   // candidates collected in shared memory (n_top <= 32 plus ties at the cut value)

__global__ void __launch_bounds__(kSampleThreads)
logprob_kernel(const __nv_bfloat16* __restrict__ logits, int64_t ld, int vocab, int seg_len, int64_t seg_stride,
               const int32_t* __restrict__ tok, int n_top, float* __restrict__ out_lp, int32_t* __restrict__ top_ids,
               float* __restrict__ top_lp) {
  __shared__ Best sh_best[32];
  __shared__ float sh_sum[32];
  __shared__ unsigned int hist[256];
  __shared__ unsigned int sel[2];
  __shared__ int cand_i[kTopCap];
  __shared__ float cand_v[kTopCap];
  __shared__ int n_cand, n_ties;
  pdl_wait();
  pdl_launch_dependents();
  const int row = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const RowView lv{logits + (int64_t)row * ld, seg_len, seg_stride};
  Best b{-INFINITY, 0x7fffffff};
  for (int i = tid; i < vocab; i += kSampleThreads) b = better(b, Best{__bfloat162float(lv.at(i)), i});
  b = block_best(b, sh_best);
  const float vmax = b.v;
  // sum of exp in a fixed order: per thread, then warp tree, then the 32 warp sums by one warp
  float se = 0.f;
  for (int i = tid; i < vocab; i += kSampleThreads) se += expf(__bfloat162float(lv.at(i)) - vmax);
  se = warp_sum(se);
  if (lane == 0) sh_sum[warp] = se;
  __syncthreads();
  if (warp == 0) {
    float t = warp_sum(sh_sum[lane]);
    if (lane == 0) sh_sum[0] = t;
  }
  __syncthreads();
  const float lse = vmax + logf(sh_sum[0]);
  if (tid == 0) {
    const int t = tok ? tok[row] : -1;
    out_lp[row] = (t >= 0 && t < vocab) ? __bfloat162float(lv.at(t)) - lse : -INFINITY;
  }
  if (n_top <= 0) return;
  // key of the n_top-th largest value
  uint32_t thr = 0;
  {
    uint32_t prefix = 0, above = 0;
    for (int level = 0; level < 2; ++level) {
      for (int j = tid; j < 256; j += kSampleThreads) hist[j] = 0u;
      __syncthreads();
      for (int i = tid; i < vocab; i += kSampleThreads) {
        const uint32_t key = key16(lv.at(i));
        if (level == 0)
          atomicAdd(&hist[key >> 8], 1u);
        else if ((key >> 8) == prefix)
          atomicAdd(&hist[key & 255u], 1u);
      }
      __syncthreads();
      if (tid == 0) {
        uint32_t cum = above;
        int bin = 255;
        for (; bin > 0; --bin) {
          if (cum + hist[bin] >= (uint32_t)n_top) break;
          cum += hist[bin];
        }
        sel[0] = (uint32_t)bin;
        sel[1] = cum;
      }
      __syncthreads();
      if (level == 0) {
        prefix = sel[0];
        above = sel[1];
      } else {
        thr = (prefix << 8) | sel[0];
      }
      __syncthreads();
    }
  }
  // candidates: everything above the cut (fewer than n_top), then the ties at the cut
  if (tid == 0) {
    n_cand = 0;
    n_ties = 0;
  }
  __syncthreads();
  for (int i = tid; i < vocab; i += kSampleThreads) {
    const __nv_bfloat16 v = lv.at(i);
    const uint32_t key = key16(v);
    if (key > thr) {
      const int at = atomicAdd(&n_cand, 1);
      if (at < kTopCap) {
        cand_i[at] = i;
        cand_v[at] = __bfloat162float(v);
      }
    }
  }
  __syncthreads();
  const int n_above = min(n_cand, kTopCap);
  for (int i = tid; i < vocab; i += kSampleThreads) {
    const __nv_bfloat16 v = lv.at(i);
    if (key16(v) == thr) {
      const int at = n_above + atomicAdd(&n_ties, 1);
      if (at < kTopCap) {
        cand_i[at] = i;
        cand_v[at] = __bfloat162float(v);
      }
    }
  }
  __syncthreads();
  if (tid == 0) {
    int n = min(n_above + n_ties, kTopCap);
    if (n_above + n_ties > kTopCap) {
      // more ties at the cut value than the buffer holds (degenerate logits): take the lowest indices by a scan
      n = n_above;
      for (int i = 0; i < vocab && n < kTopCap; ++i) {
        const __nv_bfloat16 v = lv.at(i);
        if (key16(v) == thr) {
          cand_i[n] = i;
          cand_v[n] = __bfloat162float(v);
          ++n;
        }
      }
    }
    // insertion sort: value descending, index ascending (n <= 96)
    for (int a = 1; a < n; ++a) {
      const int ci = cand_i[a];
      const float cv = cand_v[a];
      int p = a - 1;
      while (p >= 0 && (cand_v[p] < cv || (cand_v[p] == cv && cand_i[p] > ci))) {
        cand_i[p + 1] = cand_i[p];
        cand_v[p + 1] = cand_v[p];
        --p;
      }
      cand_i[p + 1] = ci;
      cand_v[p + 1] = cv;
    }
    for (int j = 0; j < n_top; ++j) {
      top_ids[(int64_t)row * n_top + j] = j < n ? cand_i[j] : -1;
      top_lp[(int64_t)row * n_top + j] = j < n ? cand_v[j] - lse : -INFINITY;
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// (a10, widened) logit_bias and repetition / frequency / presence penalties on the logit rows of a step, in the order
// of the image's sampler ([V22] vllm/v1/sample/sampler.py:366-416: logit_bias with the non-argmax-invariant
// processors, then apply_penalties = vllm/model_executor/layers/utils.py:51-89 with
// vllm/_custom_ops.py:360-373): for a token that occurs in the prompt or the output
//     x += bias;  x *= (x > 0 ? 1/rep : rep);  x -= freq * count_in_output;  x -= pres * (count_in_output > 0)
// in fp32.  Only tokens of the request's own history or bias table change, so the work is SPARSE: one entry per
// distinct (row, token), rows in CSR form.  The adjusted value is rounded back to the bf16 logits buffer (the one
// deviation from the image, which keeps the processed logits in fp32: R7) so that argmax, sampling and log-probabilities
// run on it unchanged.  Entries of one row must name distinct tokens (read-modify-write).
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
logit_adjust_kernel(__nv_bfloat16* __restrict__ logits, int64_t ld, int vocab_local, int index_offset,
                    const float* __restrict__ row_params, const int32_t* __restrict__ offsets,
                    const int32_t* __restrict__ ids, const int32_t* __restrict__ counts, const int32_t* __restrict__ seen,
                    const float* __restrict__ bias) {
  pdl_wait();
  pdl_launch_dependents();
  const int row = blockIdx.x;
  const float rep = row_params[3 * row], freq = row_params[3 * row + 1], pres = row_params[3 * row + 2];
  const float rep_inv = 1.0f / rep;
  for (int e = offsets[row] + threadIdx.x; e < offsets[row + 1]; e += blockDim.x) {
    const int id = ids[e] - index_offset;          // tensor parallel: this rank's vocabulary shard only
    if (id < 0 || id >= vocab_local) continue;
    __nv_bfloat16* px = logits + (int64_t)row * ld + id;
    float x = __bfloat162float(*px) + bias[e];
    if (seen[e]) x *= (x > 0.f ? rep_inv : rep);
    const int c = counts[e];
    x -= freq * (float)c;
    x -= pres * (c > 0 ? 1.0f : 0.0f);
    *px = __float2bfloat16_rn(x);
  }
}

}  // namespace

int logit_adjust_rows(__nv_bfloat16* logits, int64_t ld, int n_rows, int vocab_local, int index_offset, const float* row_params,
                      const int32_t* offsets, const int32_t* ids, const int32_t* counts, const int32_t* seen, const float* bias,
                      cudaStream_t s) {
  B200_REQUIRE(logits && row_params && offsets && ids && counts && seen && bias, B200_ERR_INVALID, "logit_adjust: null argument");
  if (n_rows == 0) return B200_OK;
  B200_CUDA(launch(logit_adjust_kernel, dim3(n_rows), dim3(256), (size_t)0, s, logits, ld, vocab_local, index_offset, row_params,
                   offsets, ids, counts, seen, bias));
  return B200_OK;
}

int logprob_rows(const __nv_bfloat16* logits, int64_t ld, int n_rows, int vocab, int seg_len, int64_t seg_stride,
                 const int32_t* tok, int n_top, float* out_lp, int32_t* top_ids, float* top_lp, cudaStream_t s) {
  B200_REQUIRE(logits && out_lp, B200_ERR_INVALID, "logprobs: null argument");
  B200_REQUIRE(n_top >= 0 && n_top <= 32 && (n_top == 0 || (top_ids && top_lp)), B200_ERR_INVALID,
               "logprobs: n_top=%d outside [0,32] or no output for it", n_top);
  B200_REQUIRE(vocab > 0 && seg_len > 0 && vocab % seg_len == 0, B200_ERR_INVALID, "logprobs: vocab=%d seg_len=%d", vocab,
               seg_len);
  if (n_rows == 0) return B200_OK;
  B200_CUDA(launch(logprob_kernel, dim3(n_rows), dim3(kSampleThreads), (size_t)0, s, logits, ld, vocab, seg_len, seg_stride,
                   tok, n_top, out_lp, top_ids, top_lp));
  return B200_OK;
}

int sample_rows(const __nv_bfloat16* logits, int64_t ld, int n_rows, int vocab, int seg_len, int64_t seg_stride,
                const float* temperature, const int32_t* top_k, const float* top_p, const float* min_p,
                const uint64_t* seeds, const uint64_t* draws, int32_t* out, cudaStream_t s) {
  B200_REQUIRE(logits && out, B200_ERR_INVALID, "sample: null argument");
  B200_REQUIRE(vocab > 0 && seg_len > 0 && vocab % seg_len == 0, B200_ERR_INVALID, "sample: vocab=%d seg_len=%d", vocab,
               seg_len);
  if (n_rows == 0) return B200_OK;
  B200_CUDA(launch(sample_kernel, dim3(n_rows), dim3(kSampleThreads), (size_t)0, s, logits, ld, vocab, seg_len, seg_stride,
                   temperature, top_k, top_p, min_p, seeds, draws, out));
  return B200_OK;
}

}  // namespace b200

// debug helper (not in the public header): the bits -> uniform mapping of the sampler, callable on the host so that a
// CPU test can pin its end points (0 < u < 1 for every 32-bit draw)
extern "C" __attribute__((visibility("default"))) float b200_debug_uniform_from_bits(uint32_t bits) {
  return b200::uniform_from_bits(bits);
}
