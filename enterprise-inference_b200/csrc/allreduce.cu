// (a11) One-shot tensor-parallel all-reduce over NVLink peer memory.
//
// NCCL's all-reduce costs 20-75 us per call at decode message sizes (16 KiB at T=1, 1 MiB at T=64 for
// h=8192), and a 70B step issues 160 of them: at TP8 that is 7-12 ms against a 2.7 ms weight stream
// (profiles/r01_bench_70b_tp_nccl.jsonl).  Here every rank's row-parallel linear writes its partial
// [T][h] bf16 straight into a buffer that all peers have mapped (CUDA IPC), and ONE kernel per rank
//   1. raises its flag in every peer's flag block (st.release.sys),
//   2. waits until all peers' flags for this buffer reached the same sequence number (ld.acquire.sys),
//   3. reads the partials of all ranks over NVLink, sums them in RANK ORDER in fp32 (so every rank
//      produces bit-identical results) and writes the bf16 sum to its local activation buffer.
// Two buffers alternate (O-proj uses 0, down-proj uses 1): a rank can only overwrite buffer b after its
// next all-reduce on the other buffer passed its barrier, i.e. after every peer finished reading b.
// The sequence number lives in device memory and is advanced by the last CTA, so the kernel can be
// replayed from a CUDA graph.  Bound: NVLink latency at T=1 (a few us), (world-1)*T*h*2 bytes of NVLink
// reads per rank at larger T.
//
// Replaces: the shm all-reduce of the vLLM-CPU image
// ([V22] vllm/distributed/device_communicators/cpu_communicator.py:229-276).
#include "common.cuh"

namespace b200 {

constexpr int kFlagKindStride = 2 * 8 * 32;   // uints between the "partial ready" and the "slice delivered" flags

namespace {

__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint4 ld_peer(const void* p) {
  uint4 r;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

__global__ void __launch_bounds__(256) allreduce_oneshot_kernel(ArParams p) {
  __shared__ unsigned int seq_s;
  // the sequence number of this buffer was last advanced by the collective two half-layers ago (complete: only the
  // directly preceding kernel can still be running under programmatic dependent launch): read it before the wait
  const int tid = threadIdx.x;
  if (tid == 0) seq_s = *reinterpret_cast<volatile unsigned int*>(p.state + p.buf_idx) + 1u;
  __syncthreads();
  pdl_wait();  // the partial of this rank was written by the previous kernel of the stream
  pdl_launch_dependents();
  const unsigned int seq = seq_s;
  const int slot = (p.buf_idx * 8 + p.rank) * 32;
  if (blockIdx.x == 0 && tid < p.world) {
    __threadfence_system();
    st_release_sys(p.flags[tid] + slot, seq);                       // "my partial for this buffer is ready"
  }
  if (tid < p.world) {
    const unsigned int* f = p.flags[p.rank] + (p.buf_idx * 8 + tid) * 32;
    while ((int)(ld_acquire_sys(f) - seq) < 0) __nanosleep(20);     // wrap-safe compare
  }
  __syncthreads();
  for (long long i = (long long)blockIdx.x * blockDim.x + tid; i < p.n_chunks; i += (long long)gridDim.x * blockDim.x) {
    // all peer loads are issued before the first use: ONE NVLink round trip per chunk, not one per rank
    uint4 pv[8];
#pragma unroll
    for (int r = 0; r < 8; ++r)
      if (r < p.world) pv[r] = ld_peer(reinterpret_cast<const uint4*>(p.bufs[r]) + i);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      if (r < p.world) {   // rank order: every rank computes the same bits
        const uint4 v = pv[r];
        acc[0] += bf16lo(v.x); acc[1] += bf16hi(v.x); acc[2] += bf16lo(v.y); acc[3] += bf16hi(v.y);
        acc[4] += bf16lo(v.z); acc[5] += bf16hi(v.z); acc[6] += bf16lo(v.w); acc[7] += bf16hi(v.w);
      }
    }
    reinterpret_cast<uint4*>(p.out)[i] = make_uint4(pack_bf16(acc[0], acc[1]), pack_bf16(acc[2], acc[3]),
                                                    pack_bf16(acc[4], acc[5]), pack_bf16(acc[6], acc[7]));
  }
  // the last CTA to finish advances the sequence number of this buffer (graph-replay safe)
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    const unsigned int t = atomicAdd(p.state + 2 + p.buf_idx, 1u);
    if (t == gridDim.x - 1) {
      p.state[2 + p.buf_idx] = 0u;
      *reinterpret_cast<volatile unsigned int*>(p.state + p.buf_idx) = seq;
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// Fused collective + compute: all-reduce of the row-parallel partials, residual add and RMSNorm in ONE
// kernel (one CTA per token row).  Same flag barrier as above; then every thread pulls its chunks of the
// row from all ranks over NVLink, sums them in rank order, rounds to bf16 (the all-reduced activation is
// bf16: identical numerics to all-reduce followed by the add+RMSNorm kernel), adds the residual in fp32
// and normalises (R1 of the oracle).  Replaces two launches per half-layer (160 per 70B step).
// ---------------------------------------------------------------------------------------------------
constexpr int kArNormThreads = 1024;  // one 16-byte chunk per thread at h = 8192: ONE NVLink round trip per row
constexpr int kArNormMaxChunks = 1;   // hidden <= 8192

__global__ void __launch_bounds__(kArNormThreads)
allreduce_add_rmsnorm_kernel(ArParams p, __nv_bfloat16* res, const __nv_bfloat16* __restrict__ w, __nv_bfloat16* y,
                             int hidden, float eps) {
  __shared__ unsigned int seq_s;
  __shared__ float red[kArNormThreads / 32];
  const int tid = threadIdx.x;
  if (tid == 0) seq_s = *reinterpret_cast<volatile unsigned int*>(p.state + p.buf_idx) + 1u;   // see allreduce_oneshot_kernel
  __syncthreads();
  pdl_wait();
  pdl_launch_dependents();
  const unsigned int seq = seq_s;
  if (blockIdx.x == 0 && tid < p.world) {
    __threadfence_system();
    st_release_sys(p.flags[tid] + (p.buf_idx * 8 + p.rank) * 32, seq);
  }
  if (tid < p.world) {
    const unsigned int* f = p.flags[p.rank] + (p.buf_idx * 8 + tid) * 32;
    while ((int)(ld_acquire_sys(f) - seq) < 0) __nanosleep(20);
  }
  __syncthreads();
  const int t = blockIdx.x;
  const int nchunk = hidden >> 3;
  uint4* rr = reinterpret_cast<uint4*>(res + (int64_t)t * hidden);
  float v[kArNormMaxChunks][8];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < kArNormMaxChunks; ++i) {
    const int c = tid + i * kArNormThreads;
    if (c < nchunk) {
      uint4 pv[8];
#pragma unroll
      for (int r = 0; r < 8; ++r)
        if (r < p.world) pv[r] = ld_peer(reinterpret_cast<const uint4*>(p.bufs[r] + (int64_t)t * hidden) + c);
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        if (r < p.world) {
          const uint4 q = pv[r];
          acc[0] += bf16lo(q.x); acc[1] += bf16hi(q.x); acc[2] += bf16lo(q.y); acc[3] += bf16hi(q.y);
          acc[4] += bf16lo(q.z); acc[5] += bf16hi(q.z); acc[6] += bf16lo(q.w); acc[7] += bf16hi(q.w);
        }
      }
      const uint4 b = rr[c];
      const uint32_t bw[4] = {b.x, b.y, b.z, b.w};
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        v[i][2 * j] = round_bf16(acc[2 * j]) + bf16lo(bw[j]);         // x = bf16(all-reduce); s = x + res in fp32
        v[i][2 * j + 1] = round_bf16(acc[2 * j + 1]) + bf16hi(bw[j]);
        o[j] = pack_bf16(v[i][2 * j], v[i][2 * j + 1]);
      }
      rr[c] = make_uint4(o[0], o[1], o[2], o[3]);                     // res' = bf16(s)
#pragma unroll
      for (int j = 0; j < 8; ++j) ss = fmaf(v[i][j], v[i][j], ss);
    }
  }
  ss = warp_sum(ss);
  if ((tid & 31) == 0) red[tid >> 5] = ss;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int i = 0; i < kArNormThreads / 32; ++i) tot += red[i];
  const float rinv = 1.0f / sqrtf(tot / (float)hidden + eps);
  const uint4* wr = reinterpret_cast<const uint4*>(w);
  uint4* yr = reinterpret_cast<uint4*>(y + (int64_t)t * hidden);
#pragma unroll
  for (int i = 0; i < kArNormMaxChunks; ++i) {
    const int c = tid + i * kArNormThreads;
    if (c < nchunk) {
      const uint4 wq = __ldg(wr + c);
      const uint32_t ww[4] = {wq.x, wq.y, wq.z, wq.w};
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float n0 = round_bf16(v[i][2 * j] * rinv), n1 = round_bf16(v[i][2 * j + 1] * rinv);
        o[j] = pack_bf16(n0 * bf16lo(ww[j]), n1 * bf16hi(ww[j]));
      }
      yr[c] = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    const unsigned int tk = atomicAdd(p.state + 2 + p.buf_idx, 1u);
    if (tk == gridDim.x - 1) {
      p.state[2 + p.buf_idx] = 0u;
      *reinterpret_cast<volatile unsigned int*>(p.state + p.buf_idx) = seq;
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// Two-shot form for larger messages (T > 16): the rows are dealt to the ranks (row t belongs to rank t mod world).
// After the same flag barrier the OWNER of a row pulls its partials from all ranks (1/world of the message instead of
// all of it), sums them in rank order, adds the residual (kept only on the owner: nobody else needs it), normalises,
// and writes the normed row into EVERY rank's gather buffer over NVLink; a second flag per rank ("my rows are
// delivered") closes the kernel, which ends only when all ranks' rows have arrived.  NVLink traffic per rank:
// 2 (world-1)/world of the message instead of (world-1) messages; collective + residual add + RMSNorm in one launch.
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kArNormThreads)
allreduce_twoshot_rmsnorm_kernel(ArParams p, int T, __nv_bfloat16* res, const __nv_bfloat16* __restrict__ w, int hidden,
                                 float eps) {
  __shared__ unsigned int seq_s;
  __shared__ float red[kArNormThreads / 32];
  const int tid = threadIdx.x;
  if (tid == 0) seq_s = *reinterpret_cast<volatile unsigned int*>(p.state + p.buf_idx) + 1u;   // see allreduce_oneshot_kernel
  __syncthreads();
  pdl_wait();
  pdl_launch_dependents();
  const unsigned int seq = seq_s;
  const int slot = (p.buf_idx * 8 + p.rank) * 32;
  if (blockIdx.x == 0 && tid < p.world) {
    __threadfence_system();
    st_release_sys(p.flags[tid] + slot, seq);                       // "my partial for this buffer is ready"
  }
  if (tid < p.world) {
    const unsigned int* f = p.flags[p.rank] + (p.buf_idx * 8 + tid) * 32;
    while ((int)(ld_acquire_sys(f) - seq) < 0) __nanosleep(20);
  }
  __syncthreads();
  const int t = p.rank + (int)blockIdx.x * p.world;                 // my row of this CTA (grid = ceil(rows owned), >= 1)
  const int nchunk = hidden >> 3;
  if (t < T) {
    const int c = tid;
    float v[8];
    float ss = 0.f;
    const bool act = c < nchunk;
    if (act) {
      uint4 pv[8];
#pragma unroll
      for (int r = 0; r < 8; ++r)
        if (r < p.world) pv[r] = ld_peer(reinterpret_cast<const uint4*>(p.bufs[r] + (int64_t)t * hidden) + c);
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        if (r < p.world) {
          const uint4 q = pv[r];
          acc[0] += bf16lo(q.x); acc[1] += bf16hi(q.x); acc[2] += bf16lo(q.y); acc[3] += bf16hi(q.y);
          acc[4] += bf16lo(q.z); acc[5] += bf16hi(q.z); acc[6] += bf16lo(q.w); acc[7] += bf16hi(q.w);
        }
      }
      uint4* rr = reinterpret_cast<uint4*>(res + (int64_t)t * hidden);
      const uint4 b = rr[c];
      const uint32_t bw[4] = {b.x, b.y, b.z, b.w};
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        v[2 * j] = round_bf16(acc[2 * j]) + bf16lo(bw[j]);
        v[2 * j + 1] = round_bf16(acc[2 * j + 1]) + bf16hi(bw[j]);
        o[j] = pack_bf16(v[2 * j], v[2 * j + 1]);
      }
      rr[c] = make_uint4(o[0], o[1], o[2], o[3]);
#pragma unroll
      for (int j = 0; j < 8; ++j) ss = fmaf(v[j], v[j], ss);
    }
    ss = warp_sum(ss);
    if ((tid & 31) == 0) red[tid >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
#pragma unroll
    for (int i = 0; i < kArNormThreads / 32; ++i) tot += red[i];
    const float rinv = 1.0f / sqrtf(tot / (float)hidden + eps);
    if (act) {
      const uint4 wq = __ldg(reinterpret_cast<const uint4*>(w) + c);
      const uint32_t ww[4] = {wq.x, wq.y, wq.z, wq.w};
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float n0 = round_bf16(v[2 * j] * rinv), n1 = round_bf16(v[2 * j + 1] * rinv);
        o[j] = pack_bf16(n0 * bf16lo(ww[j]), n1 * bf16hi(ww[j]));
      }
      const uint4 y = make_uint4(o[0], o[1], o[2], o[3]);
#pragma unroll
      for (int r = 0; r < 8; ++r)
        if (r < p.world) reinterpret_cast<uint4*>(p.gather[r] + (int64_t)t * hidden)[c] = y;   // all-gather by remote stores
    }
  }
  // the last CTA of this rank tells every rank that its rows are delivered, waits for everybody else's, and advances
  // the sequence number (graph-replay safe)
  __threadfence_system();
  __syncthreads();
  __shared__ unsigned int last_s;
  if (tid == 0) {
    const unsigned int tk = atomicAdd(p.state + 2 + p.buf_idx, 1u);
    last_s = (tk == gridDim.x - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (last_s) {
    if (tid < p.world) {
      __threadfence_system();
      st_release_sys(p.flags[tid] + kFlagKindStride + slot, seq);   // "my rows are in your gather buffer"
      const unsigned int* f = p.flags[p.rank] + kFlagKindStride + (p.buf_idx * 8 + tid) * 32;
      while ((int)(ld_acquire_sys(f) - seq) < 0) __nanosleep(20);
    }
    __syncthreads();
    if (tid == 0) {
      p.state[2 + p.buf_idx] = 0u;
      *reinterpret_cast<volatile unsigned int*>(p.state + p.buf_idx) = seq;
    }
  }
}

}  // namespace

int allreduce_twoshot_rmsnorm(const ArParams& p, int T, __nv_bfloat16* res, const __nv_bfloat16* w, int hidden, float eps,
                              cudaStream_t s) {
  B200_REQUIRE(hidden % 8 == 0 && hidden <= 8 * kArNormThreads, B200_ERR_UNSUPPORTED,
               "allreduce_twoshot_rmsnorm: hidden=%d unsupported", hidden);
  if (T == 0) return B200_OK;
  const int rows = T > p.rank ? (T - p.rank + p.world - 1) / p.world : 0;
  B200_CUDA(launch(allreduce_twoshot_rmsnorm_kernel, dim3(rows > 0 ? rows : 1), dim3(kArNormThreads), 0, s, p, T, res, w,
                   hidden, eps));
  return B200_OK;
}

int allreduce_add_rmsnorm(const ArParams& p, int T, __nv_bfloat16* res, const __nv_bfloat16* w, __nv_bfloat16* y, int hidden,
                          float eps, cudaStream_t s) {
  B200_REQUIRE(hidden % 8 == 0 && hidden <= 8 * kArNormThreads * kArNormMaxChunks, B200_ERR_UNSUPPORTED,
               "allreduce_add_rmsnorm: hidden=%d unsupported", hidden);
  if (T == 0) return B200_OK;
  B200_CUDA(launch(allreduce_add_rmsnorm_kernel, dim3(T), dim3(kArNormThreads), 0, s, p, res, w, y, hidden, eps));
  return B200_OK;
}

int allreduce_oneshot(const ArParams& p, cudaStream_t s) {
  if (p.n_chunks == 0) return B200_OK;
  long long g = (p.n_chunks + 255) / 256;
  if (g > 296) g = 296;
  B200_CUDA(launch(allreduce_oneshot_kernel, dim3((unsigned)g), dim3(256), 0, s, p));
  return B200_OK;
}

}  // namespace b200
