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

struct ArParams {
  const __nv_bfloat16* bufs[8];   // every rank's partial for this buffer index (peer-mapped)
  unsigned int* flags[8];         // every rank's flag block: [2 buffers][8 ranks] x 32 uints (128 B apart)
  unsigned int* state;            // local: counter[2] | ticket[2]
  __nv_bfloat16* out;
  long long n_chunks;             // 16-byte chunks
  int rank, world, buf_idx;
};

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
  pdl_wait();  // the partial of this rank was written by the previous kernel of the stream
  pdl_launch_dependents();
  const int tid = threadIdx.x;
  if (tid == 0) seq_s = *reinterpret_cast<volatile unsigned int*>(p.state + p.buf_idx) + 1u;
  __syncthreads();
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
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int r = 0; r < p.world; ++r) {
      const uint4 v = ld_peer(reinterpret_cast<const uint4*>(p.bufs[r]) + i);
      acc[0] += bf16lo(v.x); acc[1] += bf16hi(v.x); acc[2] += bf16lo(v.y); acc[3] += bf16hi(v.y);
      acc[4] += bf16lo(v.z); acc[5] += bf16hi(v.z); acc[6] += bf16lo(v.w); acc[7] += bf16hi(v.w);
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

}  // namespace

int allreduce_oneshot(const ArParams& p, cudaStream_t s) {
  if (p.n_chunks == 0) return B200_OK;
  long long g = (p.n_chunks + 255) / 256;
  if (g > 64) g = 64;
  B200_CUDA(launch(allreduce_oneshot_kernel, dim3((unsigned)g), dim3(256), 0, s, p));
  return B200_OK;
}

}  // namespace b200
