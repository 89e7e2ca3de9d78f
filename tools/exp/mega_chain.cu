// Persistent decode kernel for steps of at most 16 rows: a whole chain of linears (and, later, the whole layer
// stack) in ONE launch.
//
// Why: with one kernel per linear a decode step of 1..16 rows is a chain of ~160 dependent launches, and every
// boundary idles HBM for ~4 us (the predecessor's epilogue, grid completion, the successor's first dependent loads)
// against 5..36 us of streaming per kernel (profiles/r02_launches_bs1_summary.txt; the A/B runs in DESIGN.md §8 show
// that ring depth and co-residency knobs of the per-kernel design do not move it).  Here ONE CTA per SM stays resident
// for the whole chain: a producer thread streams weight tiles of phase after phase into a ring that fills the SM's
// shared memory (11 x 16 KiB per SM = 26 MB over the GPU, about 4 us of HBM time), so the weight stream of the NEXT
// linear runs while this one reduces, finalises and hands over through a grid barrier.
//
// Work split ("stream-K"): a phase (one linear, n_out x k) is m_tiles x n_kb units of one 128-row x 64-column weight
// tile (16 KiB); CTA c takes the contiguous unit range [c U/G, (c+1) U/G), i.e. every CTA streams the same number
// of bytes whatever the shape.  A run of units inside one slab is a SEGMENT: its MMAs accumulate in TMEM
// ([128 rows][16 tokens] fp32, double buffered).  A slab that is cut across CTAs is reduced through global memory:
// every part is written as fp32 [part][token][row], an atomic ticket finds the last arriver, which sums the parts in
// part order (ascending k: deterministic) and runs the epilogue.  Rounding: fp32 accumulation, ONE rounding (R2).
//
// Roles (320 threads): warp 0 = weight producer (TMA, one thread), warp 1 = TMEM owner + MMA issuer (one thread),
// warps 2-5 = operand builders (the [16 tokens][64] bf16 B tile of every unit, written where a SWIZZLE_128B TMA box
// would have put it), warps 6-9 = epilogue (thread = slab row, the 16 token columns in registers).
#include <cuda.h>
#include <stdlib.h>

#include <vector>

#include "common.cuh"

namespace b200 {

int get_tensor_map_2d(const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows, CUtensorMap* out);

// One linear of the chain.  x source: a bf16 buffer [T][k].  Epilogue: 0 = y[t][row] = bf16(acc + bias).
struct alignas(64) MegaPhase {
  CUtensorMap map_w;              // [n_out][k] bf16, box [128][64], SWIZZLE_128B
  const __nv_bfloat16* x;         // [T][ldx]
  const __nv_bfloat16* bias;      // [n_out] or null
  __nv_bfloat16* y;               // [T][ldy]
  long long ldx, ldy;
  int n_out, k, m_tiles, n_kb;
  int epi, pad0, pad1, pad2;
};

namespace {

constexpr int kMegaThreads = 320;
constexpr int kMN = 16;                       // token columns of the MMA
constexpr int kWBytes = 128 * 64 * 2;         // weight tile
constexpr int kXBytes = kMN * 64 * 2;         // operand tile
constexpr int kStage = kWBytes + kXBytes;     // 18 KiB
constexpr int kStages = 11;
constexpr int kMaxPhases = 192;
constexpr uint64_t kEvictFirst = 0x12F0000000000000ull;

struct PhaseHdr {   // the part of a phase every role needs, copied to shared memory once
  const __nv_bfloat16* x;
  const __nv_bfloat16* bias;
  __nv_bfloat16* y;
  long long ldx, ldy;
  int n_out, m_tiles, n_kb, epi;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// Bounded wait: a protocol error must end as a trap (a CUDA error on the host), never as a hung GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done = 0;
  for (long long spin = 0; !done; ++spin) {
    asm volatile(
        "{\n.reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n}\n"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (spin > (1ll << 26)) asm volatile("trap;");
  }
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_mma(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(acc)
      : "memory");
}
// K-major SWIZZLE_128B operand descriptor (see gemm_tc.cu::make_desc)
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3fff);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
constexpr uint32_t kIdesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kMN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

struct MegaParams {
  const MegaPhase* phases;
  int n_phases, T;
  float* partials;            // [G][2][16][128] fp32: a CTA has at most two cut segments per phase (its first and its last)
  unsigned int* tickets;      // [max m_tiles], zero on entry and on exit
  unsigned int* gbar;         // grid barrier counter, zero on entry
};

// unit range of CTA c in a phase of U units over G CTAs, and the CTA that owns unit u
__device__ __forceinline__ int unit_begin(int c, int U, int G) { return (int)((long long)c * U / G); }
__device__ __forceinline__ int cta_of_unit(int u, int U, int G) { return (int)((((long long)u + 1) * G + U - 1) / U) - 1; }

__global__ void __launch_bounds__(kMegaThreads, 1) mega_chain_kernel(MegaParams p) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* ring = smem;
  PhaseHdr* hdr = reinterpret_cast<PhaseHdr*>(ring + kStages * kStage);
  uint64_t* bars = reinterpret_cast<uint64_t*>(hdr + kMaxPhases);
  uint64_t* full_w = bars;                 // [S] weight tile landed (TMA)
  uint64_t* full_x = full_w + kStages;     // [S] operand tile written (builder warp)
  uint64_t* empty = full_x + kStages;      // [S] the MMAs that read the stage have completed
  uint64_t* acc_full = empty + kStages;    // [2] segment accumulated
  uint64_t* acc_free = acc_full + 2;       // [2] epilogue has drained the accumulator (4 warps)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_free + 2);
  volatile int* phase_ready = reinterpret_cast<volatile int*>(tmem_slot + 1);   // phases whose inputs are complete
  int* ticket_s = const_cast<int*>(phase_ready) + 1;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int G = gridDim.x, c = blockIdx.x, T = p.T;

  for (int i = threadIdx.x; i < p.n_phases; i += kMegaThreads) {
    const MegaPhase& ph = p.phases[i];
    hdr[i] = PhaseHdr{ph.x, ph.bias, ph.y, ph.ldx, ph.ldy, ph.n_out, ph.m_tiles, ph.n_kb, ph.epi};
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_w[s], 1);
      mbar_init(&full_x[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&acc_full[b], 1);
      mbar_init(&acc_free[b], 4);
    }
    *phase_ready = 1;   // phase 0 reads inputs that were complete before the launch
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(32));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== weight producer: runs ahead across phases (weights do not depend on anything) =====
    if (lane == 0) {
      int n = 0;
      for (int ph = 0; ph < p.n_phases; ++ph) {
        const int n_kb = hdr[ph].n_kb, U = hdr[ph].m_tiles * n_kb;
        const CUtensorMap* map = &p.phases[ph].map_w;
        const int u0 = unit_begin(c, U, G), u1 = unit_begin(c + 1, U, G);
        if (u0 < u1) asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
        for (int u = u0; u < u1; ++u, ++n) {
          const int st = n % kStages;
          if (n >= kStages) mbar_wait(&empty[st], ((n / kStages) - 1) & 1);
          mbar_expect_tx(&full_w[st], kWBytes);
          tma_load_2d(ring + st * kStage, map, (u % n_kb) * 64, (u / n_kb) * 128, &full_w[st], kEvictFirst);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      int n = 0, seg = 0;
      for (int ph = 0; ph < p.n_phases; ++ph) {
        const int n_kb = hdr[ph].n_kb, U = hdr[ph].m_tiles * n_kb;
        const int u0 = unit_begin(c, U, G), u1 = unit_begin(c + 1, U, G);
        for (int u = u0; u < u1; ++u, ++n) {
          const int st = n % kStages, kb = u % n_kb;
          const bool first = (u == u0) || kb == 0, last = (u + 1 == u1) || kb == n_kb - 1;
          const int buf = seg & 1;
          if (first && seg >= 2) mbar_wait(&acc_free[buf], ((seg >> 1) - 1) & 1);
          mbar_wait(&full_w[st], (n / kStages) & 1);
          mbar_wait(&full_x[st], (n / kStages) & 1);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(ring + st * kStage), b_addr = a_addr + kWBytes;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            tc_mma(tmem_base + buf * kMN, make_desc(a_addr + k * 32), make_desc(b_addr + k * 32), kIdesc, (first && k == 0) ? 0u : 1u);
          tc_commit(&empty[st]);
          if (last) {
            tc_commit(&acc_full[buf]);
            ++seg;
          }
        }
      }
    }
  } else if (warp < 6) {
    // ===== operand builders: warp bw writes the B tile of the units n = bw (mod 4) =====
    const int bw = warp - 2;
    int n = 0;
    for (int ph = 0; ph < p.n_phases; ++ph) {
      const int n_kb = hdr[ph].n_kb, U = hdr[ph].m_tiles * n_kb;
      const int u0 = unit_begin(c, U, G), u1 = unit_begin(c + 1, U, G);
      if (u0 < u1) {
        long long spin = 0;
        while (*phase_ready <= ph)
          if (++spin > (1ll << 28)) asm volatile("trap;");
        __threadfence_block();
      }
      const __nv_bfloat16* x = hdr[ph].x;
      const long long ldx = hdr[ph].ldx;
      for (int u = u0; u < u1; ++u, ++n) {
        if ((n & 3) != bw) continue;
        const int st = n % kStages, kb = u % n_kb;
        uint4 v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int task = lane + 32 * j, t = task >> 3, ch = task & 7;
          v[j] = t < T ? __ldcg(reinterpret_cast<const uint4*>(x + t * ldx + kb * 64 + ch * 8)) : make_uint4(0u, 0u, 0u, 0u);
        }
        if (n >= kStages) mbar_wait(&empty[st], ((n / kStages) - 1) & 1);
        const uint32_t bbase = smem_u32(ring + st * kStage + kWBytes);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int task = lane + 32 * j, t = task >> 3, ch = task & 7;
          const uint32_t dst = bbase + (uint32_t)t * 128u + (uint32_t)((ch ^ (t & 7)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(v[j].x), "r"(v[j].y), "r"(v[j].z), "r"(v[j].w) : "memory");
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(&full_x[st]);
      }
    }
  } else {
    // ===== epilogue: thread = slab row (TMEM lane), 16 token columns in registers =====
    const int quad = warp & 3, row = quad * 32 + lane, et = threadIdx.x - 6 * 32;
    int seg = 0;
    for (int ph = 0; ph < p.n_phases; ++ph) {
      const PhaseHdr h = hdr[ph];
      const int n_kb = h.n_kb, U = h.m_tiles * n_kb;
      const int u0 = unit_begin(c, U, G), u1 = unit_begin(c + 1, U, G);
      int u = u0;
      while (u < u1) {
        const int m = u / n_kb, ue = min(u1, (m + 1) * n_kb);   // this segment: units [u, ue) of slab m
        const int buf = seg & 1;
        mbar_wait(&acc_full[buf], (seg >> 1) & 1);
        tc_fence_after();
        uint32_t r[16];
        tmem_ld16(tmem_base + ((uint32_t)(quad * 32) << 16) + buf * kMN, r);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&acc_free[buf]);
        ++seg;
        float v[16];
#pragma unroll
        for (int t = 0; t < 16; ++t) v[t] = __uint_as_float(r[t]);
        const int first_cta = cta_of_unit(m * n_kb, U, G), last_cta = cta_of_unit((m + 1) * n_kb - 1, U, G);
        const int parts = last_cta - first_cta + 1;
        bool fin = true;
        if (parts > 1) {
          // slot 0: the segment that starts this CTA's range, slot 1: a later one (only the last can be cut)
          float* mine = p.partials + (((long long)c * 2 + (u == u0 ? 0 : 1)) * 16) * 128 + row;
#pragma unroll
          for (int t = 0; t < 16; ++t)
            if (t < T) __stcg(mine + t * 128, v[t]);
          __threadfence();
          asm volatile("bar.sync 1, 128;" ::: "memory");
          if (et == 0) *ticket_s = (int)atomicAdd(&p.tickets[m], 1u);
          asm volatile("bar.sync 1, 128;" ::: "memory");
          fin = (*ticket_s == parts - 1);
          if (fin) {
            __threadfence();
#pragma unroll
            for (int t = 0; t < 16; ++t) v[t] = 0.f;
            for (int q = 0; q < parts; ++q) {   // part order = ascending k: deterministic
              const int cc = first_cta + q;
              const int slot = (unit_begin(cc, U, G) / n_kb == m) ? 0 : 1;
              const float* src = p.partials + (((long long)cc * 2 + slot) * 16) * 128 + row;
#pragma unroll
              for (int t = 0; t < 16; ++t)
                if (t < T) v[t] += __ldcg(src + t * 128);
            }
            if (et == 0) p.tickets[m] = 0u;
          }
          asm volatile("bar.sync 1, 128;" ::: "memory");   // ticket_s is reused by the next segment
        }
        if (fin) {
          const int orow = m * 128 + row;
          if (orow < h.n_out) {
            const float b = h.bias ? __bfloat162float(h.bias[orow]) : 0.f;
#pragma unroll
            for (int t = 0; t < 16; ++t)
              if (t < T) h.y[t * h.ldy + orow] = __float2bfloat16_rn(v[t] + b);
          }
        }
        u = ue;
      }
      // grid barrier: every slab of this phase is final before anybody builds operands from its output
      __threadfence();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (et == 0) {
        atomicAdd(p.gbar, 1u);
        const unsigned int want = (unsigned int)G * (unsigned int)(ph + 1);
        long long spin = 0;
        unsigned int seen;
        do {
          asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(p.gbar) : "memory");
          if (++spin > (1ll << 26)) asm volatile("trap;");
        } while (seen < want);
        __threadfence_block();
        *phase_ready = ph + 2;
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(32));
  }
}

constexpr int kMegaSmem = kStages * kStage + kMaxPhases * (int)sizeof(PhaseHdr) + 1024 /* barriers */ + 1024 /* alignment */;

}  // namespace

// Device-side state of the chain runner (debug / microbenchmark entry: allocated on first use, per device).
struct MegaState {
  MegaPhase* d_phases = nullptr;
  float* partials = nullptr;
  unsigned int* tickets = nullptr;
  unsigned int* gbar = nullptr;
  int n_sm = 0;
};
static MegaState g_mega[64];

int mega_linear_chain(int n, const void* const* w, const int32_t* n_out, const int32_t* k, const void* x0, void* const* y, int T,
                      cudaStream_t s) {
  B200_REQUIRE(n > 0 && n <= kMaxPhases && T >= 1 && T <= kMN, B200_ERR_INVALID, "mega chain: %d phases / %d rows unsupported", n, T);
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < 64, B200_ERR_UNSUPPORTED, "mega chain: device index %d", dev);
  MegaState& st = g_mega[dev];
  if (!st.d_phases) {
    cudaDeviceProp prop;
    B200_CUDA(cudaGetDeviceProperties(&prop, dev));
    st.n_sm = prop.multiProcessorCount;
    B200_CUDA(cudaMalloc((void**)&st.d_phases, sizeof(MegaPhase) * kMaxPhases));
    B200_CUDA(cudaMalloc((void**)&st.partials, (size_t)st.n_sm * 2 * 16 * 128 * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&st.tickets, 4096 * sizeof(unsigned int)));
    B200_CUDA(cudaMalloc((void**)&st.gbar, 256));
    B200_CUDA(cudaMemset(st.tickets, 0, 4096 * sizeof(unsigned int)));
    B200_CUDA(cudaFuncSetAttribute(mega_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMegaSmem));
  }
  std::vector<MegaPhase> ph(n);
  for (int i = 0; i < n; ++i) {
    B200_REQUIRE(n_out[i] % 128 == 0 && k[i] % 64 == 0 && n_out[i] / 128 <= 4096, B200_ERR_UNSUPPORTED,
                 "mega chain: phase %d shape [%d x %d] (n_out %% 128, k %% 64)", i, n_out[i], k[i]);
    if (i > 0) B200_REQUIRE(k[i] <= n_out[i - 1], B200_ERR_INVALID, "mega chain: phase %d reads %d columns of a %d-wide input", i, k[i], n_out[i - 1]);
    memset(&ph[i], 0, sizeof(MegaPhase));
    int rc = get_tensor_map_2d(w[i], (uint64_t)n_out[i], (uint64_t)k[i], (uint64_t)k[i], 128, &ph[i].map_w);
    if (rc) return rc;
    ph[i].x = (const __nv_bfloat16*)(i == 0 ? x0 : y[i - 1]);
    ph[i].ldx = i == 0 ? k[0] : n_out[i - 1];
    ph[i].y = (__nv_bfloat16*)y[i];
    ph[i].ldy = n_out[i];
    ph[i].n_out = n_out[i];
    ph[i].k = k[i];
    ph[i].m_tiles = n_out[i] / 128;
    ph[i].n_kb = k[i] / 64;
  }
  B200_CUDA(cudaMemcpyAsync(st.d_phases, ph.data(), sizeof(MegaPhase) * n, cudaMemcpyHostToDevice, s));
  B200_CUDA(cudaStreamSynchronize(s));   // ph is a host temporary (debug entry: not on any hot path)
  B200_CUDA(cudaMemsetAsync(st.gbar, 0, 4, s));
  MegaParams p = {st.d_phases, n, T, st.partials, st.tickets, st.gbar};
  const bool pdl_prev = launch_opts().pdl;
  launch_opts().pdl = false;   // one CTA per SM must be resident together: never overlapped with a neighbour
  cudaError_t e = launch(mega_chain_kernel, dim3(st.n_sm), dim3(kMegaThreads), (size_t)kMegaSmem, s, p);
  launch_opts().pdl = pdl_prev;
  B200_CUDA(e);
  return B200_OK;
}

}  // namespace b200

// debug helper (not in the public header yet): a chain of plain linears y_i = x_i . w_i^T, x_{i+1} = y_i[:, :k_{i+1}]
extern "C" __attribute__((visibility("default"))) int b200_debug_mega_chain(int32_t n, const void* const* w, const int32_t* n_out,
                                                                           const int32_t* k, const void* x0, void* const* y,
                                                                           int32_t T, void* stream) {
  return b200::mega_linear_chain(n, w, n_out, k, x0, y, T, (cudaStream_t)stream);
}
