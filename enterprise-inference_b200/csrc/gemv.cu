// HBM-bound small-batch linear: y[T][n_out] = x[T][k] . w[n_out][k]^T (+bias), T <= 4 per pass.
//
// Roofline: HBM.  Algorithmic bytes per launch = n_out*k*2 (each weight read exactly once);
// x (T*k*2) and y are L2/L1 resident noise.  One warp owns R output rows; a lane streams
// 16-byte weight chunks with L1::no_allocate so that 8 x-rows worth of activations stay in L1.
// Weight loads of the first unrolled batch are issued BEFORE griddepcontrol.wait: weights do
// not depend on the previous kernel, so the fetch overlaps the predecessor's tail (PDL).
//
// Replaces: sgl-kernel weight_packed_linear / oneDNN bf16 matmul of the vLLM-CPU image
// ([V22] vllm/model_executor/layers/utils.py:277-299); rounding R2 of oracle/decode_oracle.py.
#include "common.cuh"

namespace b200 {

namespace {

constexpr int kGemvThreads = 256;
constexpr int kGemvWarps = kGemvThreads / 32;

template <int T>
__device__ __forceinline__ void fma8(float (&acc)[T], const uint4& w, const uint4 (&x)[T]) {
  const float w0 = bf16lo(w.x), w1 = bf16hi(w.x), w2 = bf16lo(w.y), w3 = bf16hi(w.y);
  const float w4 = bf16lo(w.z), w5 = bf16hi(w.z), w6 = bf16lo(w.w), w7 = bf16hi(w.w);
#pragma unroll
  for (int t = 0; t < T; ++t) {
    float a = acc[t];
    a = fmaf(w0, bf16lo(x[t].x), a);
    a = fmaf(w1, bf16hi(x[t].x), a);
    a = fmaf(w2, bf16lo(x[t].y), a);
    a = fmaf(w3, bf16hi(x[t].y), a);
    a = fmaf(w4, bf16lo(x[t].z), a);
    a = fmaf(w5, bf16hi(x[t].z), a);
    a = fmaf(w6, bf16lo(x[t].w), a);
    a = fmaf(w7, bf16hi(x[t].w), a);
    acc[t] = a;
  }
}

// T tokens, R rows per warp, U-way unrolled K loop (R*U 16-byte weight loads in flight per lane).
template <int T, int R, int U>
__global__ void __launch_bounds__(kGemvThreads)
gemv_kernel(const __nv_bfloat16* __restrict__ x, int64_t ldx, const __nv_bfloat16* __restrict__ w,
            const __nv_bfloat16* __restrict__ bias, __nv_bfloat16* __restrict__ y, int64_t ldy,
            int n_tokens, int n_out, int k) {
  const int lane = threadIdx.x & 31;
  const int warp = blockIdx.x * kGemvWarps + (threadIdx.x >> 5);
  const int row0 = warp * R;
  if (row0 >= n_out) {
    pdl_wait();
    return;
  }
  const int nchunk = k >> 3;  // 16-byte chunks per row
  const uint4* wrow[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int row = min(row0 + r, n_out - 1);
    wrow[r] = reinterpret_cast<const uint4*>(w + (int64_t)row * k);
  }
  const uint4* xrow[T];
#pragma unroll
  for (int t = 0; t < T; ++t)
    xrow[t] = reinterpret_cast<const uint4*>(x + (int64_t)min(t, n_tokens - 1) * ldx);

  float acc[R][T];
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int t = 0; t < T; ++t) acc[r][t] = 0.f;

  // first batch of weight loads, before waiting on the producer of x
  uint4 wv[U][R];
  int c = lane;
#pragma unroll
  for (int u = 0; u < U; ++u)
#pragma unroll
    for (int r = 0; r < R; ++r)
      if (c + u * 32 < nchunk) wv[u][r] = ld_stream(wrow[r] + c + u * 32);
  pdl_wait();
  pdl_launch_dependents();

  for (; c < nchunk; c += 32 * U) {
    uint4 xv[U][T];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int t = 0; t < T; ++t)
        if (c + u * 32 < nchunk) xv[u][t] = ld_ro(xrow[t] + c + u * 32);
    // consume this batch and fetch the next one
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (c + u * 32 < nchunk) {
#pragma unroll
        for (int r = 0; r < R; ++r) fma8<T>(acc[r], wv[u][r], xv[u]);
      }
      const int cn = c + 32 * U + u * 32;
      if (cn < nchunk) {
#pragma unroll
        for (int r = 0; r < R; ++r) wv[u][r] = ld_stream(wrow[r] + cn);
      }
    }
  }

#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int t = 0; t < T; ++t) acc[r][t] = warp_sum(acc[r][t]);

  if (lane == 0) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int row = row0 + r;
      if (row < n_out) {
        const float b = bias ? __bfloat162float(bias[row]) : 0.f;
#pragma unroll
        for (int t = 0; t < T; ++t)
          if (t < n_tokens) y[(int64_t)t * ldy + row] = __float2bfloat16_rn(acc[r][t] + b);
      }
    }
  }
}

template <int T, int R, int U>
int launch_gemv(const __nv_bfloat16* x, int64_t ldx, const __nv_bfloat16* w, const __nv_bfloat16* bias,
                __nv_bfloat16* y, int64_t ldy, int n_tokens, int n_out, int k, cudaStream_t s) {
  const int rows_per_block = kGemvWarps * R;
  const int grid = (n_out + rows_per_block - 1) / rows_per_block;
  B200_CUDA(launch(gemv_kernel<T, R, U>, dim3(grid), dim3(kGemvThreads), 0, s, x, ldx, w, bias, y, ldy,
                   n_tokens, n_out, k));
  return B200_OK;
}

}  // namespace

int linear_gemv(const __nv_bfloat16* x, int64_t ldx, const __nv_bfloat16* w, const __nv_bfloat16* bias,
                __nv_bfloat16* y, int64_t ldy, int T, int n_out, int k, cudaStream_t s) {
  B200_REQUIRE(k % 8 == 0 && ldx % 8 == 0, B200_ERR_UNSUPPORTED, "gemv: k and ldx must be multiples of 8 (k=%d)", k);
  B200_REQUIRE((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w)) % 16 == 0, B200_ERR_INVALID,
               "gemv: x and w must be 16-byte aligned");
  for (int t0 = 0; t0 < T; t0 += 4) {
    const int nt = min(4, T - t0);
    const __nv_bfloat16* xs = x + (int64_t)t0 * ldx;
    __nv_bfloat16* ys = y + (int64_t)t0 * ldy;
    int rc;
    if (nt == 1)
      rc = launch_gemv<1, 2, 4>(xs, ldx, w, bias, ys, ldy, nt, n_out, k, s);
    else if (nt == 2)
      rc = launch_gemv<2, 2, 4>(xs, ldx, w, bias, ys, ldy, nt, n_out, k, s);
    else
      rc = launch_gemv<4, 2, 2>(xs, ldx, w, bias, ys, ldy, nt, n_out, k, s);
    if (rc) return rc;
  }
  return B200_OK;
}

}  // namespace b200
