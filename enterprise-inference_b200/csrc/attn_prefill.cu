// (a6, prefill rows) Causal paged attention for the rows of prefill chunks on tcgen05.
//
// attn_tc.cu puts the 128 cached tokens of a tile on the UMMA M dimension and at most 16 query heads on N:
// right for decode (every cached byte is used once per step), but a 2048-row chunk then streams the
// sequence's K/V once per four rows and pays ~100 softmax instructions per (tile, column).  Here the QUERY
// side fills M instead, as in FlashAttention: one work item = up to 128/G consecutive rows of one sequence
// x the G query heads of one kv head = 128 MMA rows, and per 128-token tile
//   S[128 q][128 tok] = Q[128 q x 128 dim] . K_tile^T            (A = Q, B = K tile, both K-major)
//   O_t[128 q][128 dim] = P[128 q x 128 tok] . V_tile            (A = P K-major, B = V tile MN-major)
// so a sequence's K/V is read once per 128/G rows and the softmax is one thread per query row: the row
// max and sum need no shuffles, masking is one compare against the row's context length.
// P is handed to the tensor core as hi + lo bf16 (the oracle's R4 keeps P in fp32), O_t is folded into 128
// fp32 registers per thread with the running rescale (TMEM is never read-modify-written), S and O_t are
// double buffered in TMEM (2 x 128 + 2 x 128 = all 512 columns).
//
// Rows come in any mix (chunked prefill + decode rows in one step): a one-warp plan kernel cuts the step's
// rows into items (maximal runs of one sequence, at most 128/G rows each); a decode row becomes an item
// with one live row.  Bound: tensor pipe / MUFU (exp), not HBM: K/V bytes per item = ctx * 512 B.
//
// Replaces: the prefill half of the vLLM-CPU attention backend ([V22] vllm/v1/attention/backends/
// cpu_attn.py:377-396); chunked prefill is selected by the chart (xeon-values.yaml:80-83).
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"

namespace b200 {

int get_tensor_map_2d(const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows, CUtensorMap* out);

namespace {

constexpr int kTileTok = 128;
constexpr int kHalfBytes = 128 * 64 * 2;          // one [128 rows][64 cols] bf16 SWIZZLE_128B sub-tile: 16 KiB
constexpr int kTileBytes = 2 * kHalfBytes;        // a [128][128] bf16 operand tile: 32 KiB
constexpr int kSlots = 4;                         // K/V ring: K, V, K, V (two tiles in flight)
constexpr int kFaThreads = 352;                   // TMA warp, MMA warp, 8 softmax warps, Q-staging warp
constexpr uint32_t kFaTmemCols = 512;             // S0 | S1 | O0 | O1, 128 columns each
constexpr int kFaNeed = (1 /*Q*/ + 2 /*P hi, lo*/ + kSlots) * kTileBytes + 128 /* barriers */ + 2048 /* row max exchange */;
constexpr int kFaSmem = 232448;   // all of the SM's 227 KiB: leaves 896 B for aligning the base to 1024 (checked in the kernel)

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
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tc_mma(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(acc)
      : "memory");
}
// SWIZZLE_128B shared-memory matrix descriptor (same bit layout as in attn_tc.cu / gemm_tc.cu).
//  K-major  operand: rows are 128 B, 8-row groups SBO = 1024 B apart, LBO unused (16).
//  MN-major operand: 64 MN elements (128 B) x 8 K rows per swizzle atom; K groups SBO = 1024 B apart,
//                    the next 64-element MN atom LBO bytes away.
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3fff);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// bf16 x bf16 -> fp32, M = 128, N = 128; b_mn_major: B operand MN-major (the V tile, dims contiguous)
__host__ __device__ constexpr uint32_t fa_idesc(int b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)b_mn_major << 16) | ((uint32_t)(128 >> 3) << 17) |
         ((uint32_t)(128 >> 4) << 24);
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// byte offset of the 16-byte chunk holding columns [8*ch, 8*ch+8) of row m in a [128][128] bf16 operand
// stored as two K-major SWIZZLE_128B halves of 64 columns (Q and P tiles; same layout TMA gives K tiles)
__device__ __forceinline__ uint32_t tile_chunk_off(int m, int ch) {
  return (uint32_t)((ch >> 3) * kHalfBytes + (m >> 3) * 1024 + (m & 7) * 128 + (((ch & 7) ^ (m & 7)) << 4));
}

__device__ __forceinline__ float fast_exp2(float x) {   // x <= 0 or -inf here
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// debug (trace builds, -DB200_TRACE_BUILD): %globaltimer stamps of CTA 0's first softmax thread
__device__ unsigned long long g_fa_trace[256];
__device__ __forceinline__ void fa_trace(int slot) {
#ifdef B200_TRACE_BUILD
  if (blockIdx.x == 0 && slot < 256) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    g_fa_trace[slot] = t;
  }
#else
  (void)slot;
#endif
}

struct PlanItem {
  int row0, nrows, seq, ctx_max;
};

// ---- plan: rows -> items (one CTA; warp 0 scans, then one thread per item) -------------------------
// plan[0] = {n_items, 0, 0, 0}; plan[1 + i] = item i: rows [row0, row0 + nrows) of sequence seq, all of
// one sequence, nrows <= rb, ctx_max = their longest context.
__global__ void __launch_bounds__(256) attn_prefill_plan_kernel(const int32_t* __restrict__ row_to_seq,
                                                                 const int32_t* __restrict__ ctx_lens, int n_rows, int rb,
                                                                 PlanItem* __restrict__ plan) {
  __shared__ int s_items;
  pdl_wait();
  pdl_launch_dependents();
  const int lane = threadIdx.x & 31;
  if (threadIdx.x < 32) {
    int count = 0, seg_carry = 0;
#pragma unroll 4
    for (int base = 0; base < n_rows; base += 32) {   // the loads do not depend on the carried state: 4 rounds in flight
      const int r = base + lane;
      const bool valid = r < n_rows;
      const int seq = valid ? (row_to_seq ? row_to_seq[r] : r) : -1;
      const int prev = (valid && r > 0) ? (row_to_seq ? row_to_seq[r - 1] : r - 1) : -2;
      // start row of the segment (run of one sequence) that row r belongs to: inclusive max-scan
      int seg = (valid && (r == 0 || seq != prev)) ? r : -1;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, seg, o);
        if (lane >= o) seg = max(seg, v);
      }
      if (seg < 0) seg = seg_carry;
      const bool start = valid && ((r - seg) % rb == 0);
      const unsigned m = __ballot_sync(0xffffffffu, start);
      if (start) {
        const int idx = count + __popc(m & ((1u << lane) - 1u));
        plan[1 + idx].row0 = r;
        plan[1 + idx].seq = seq;
      }
      count += __popc(m);
      seg_carry = __shfl_sync(0xffffffffu, seg, 31);
    }
    if (lane == 0) {
      s_items = count;
      plan[0].row0 = count;
    }
  }
  __syncthreads();
  const int n_items = s_items;
  for (int i = threadIdx.x; i < n_items; i += blockDim.x) {
    const int row0 = plan[1 + i].row0;
    const int next = (i + 1 < n_items) ? plan[2 + i].row0 : n_rows;
    int cmax = 0;
    for (int r = row0; r < next; ++r) cmax = max(cmax, ctx_lens[r]);
    plan[1 + i].nrows = next - row0;
    plan[1 + i].ctx_max = cmax;
  }
}

// ---- the attention kernel ------------------------------------------------------------------------
template <int G>
__global__ void __launch_bounds__(kFaThreads, 1)
paged_prefill_fa_kernel(const __grid_constant__ CUtensorMap map_k, const __grid_constant__ CUtensorMap map_v,
                        const __nv_bfloat16* __restrict__ q, int64_t ldq, const int32_t* __restrict__ block_table,
                        int bt_stride, const int32_t* __restrict__ ctx_lens, __nv_bfloat16* __restrict__ out, int64_t ldo,
                        int hkv, int block_size, float scale, const PlanItem* __restrict__ plan) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  if ((int)(smem - smem_raw) + kFaNeed > kFaSmem) __trap();   // dynamic smem base less aligned than assumed
  unsigned char* qs = smem;                       // Q   [128 m][128 dim]
  unsigned char* p_hi = qs + kTileBytes;          // P   [128 m][128 tok], high bf16 term
  unsigned char* p_lo = p_hi + kTileBytes;        //                        low bf16 term
  unsigned char* kv = p_lo + kTileBytes;          // ring of kSlots tiles [128 tok][128 dim]
  uint64_t* kv_full = reinterpret_cast<uint64_t*>(kv + kSlots * kTileBytes);
  uint64_t* kv_free = kv_full + kSlots;
  uint64_t* s_full = kv_free + kSlots;            // [2]
  uint64_t* o_full = s_full + 2;                  // [2]
  uint64_t* p_ready = o_full + 2;                 // P(n) is in shared memory (8 arrivals: one per softmax warp)
  uint64_t* q_ready = p_ready + 1;                // Q of the current item is staged (32 arrivals)
  uint64_t* q_free = q_ready + 1;                 // the S MMAs of the current item no longer read Q
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(q_free + 1);
  int* pv_fresh = reinterpret_cast<int*>(tmem_slot + 1);   // 1: P.V of this tile starts a new O_t, 0: accumulates
  float* xmax = reinterpret_cast<float*>(kv_full + 16);   // [2 (tile parity)][2 halves][128 rows]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kSlots; ++s) {
      mbar_init(&kv_full[s], 1);
      mbar_init(&kv_free[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&s_full[s], 1);
      mbar_init(&o_full[s], 1);
    }
    mbar_init(p_ready, 8);
    mbar_init(q_ready, 32);
    mbar_init(q_free, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(kFaTmemCols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();   // q, the newest K/V rows and the plan come from the previous kernels
  pdl_launch_dependents();

  // Persistent CTA: work = (plan item, kv head), long items (the chunk's last rows) first.  The K/V ring,
  // the S/O double buffers and all barrier phases run on one tile counter across items.
  const int n_plan = plan[0].row0;
  const int n_work = n_plan * hkv;
  auto get_work = [&](int it, PlanItem& w, int& kvh) {
    kvh = it % hkv;
    w = plan[1 + (n_plan - 1 - it / hkv)];
  };

  if (warp == 0) {
    // ===== TMA producer: the whole warp walks the loop, the TMA instructions are predicated to one elected lane and
    // loop bounds / block ids are broadcast from lane 0, so that the operands stay in uniform registers =====
    {
      const bool issuer = elect_one();
      if (issuer) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_k) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_v) : "memory");
      }
      int n = 0;
      for (int it = blockIdx.x; it < n_work; it += gridDim.x) {
        PlanItem w;
        int kvh;
        get_work(it, w, kvh);
        const int nt = __shfl_sync(0xffffffffu, (w.ctx_max + kTileTok - 1) / kTileTok, 0);
        const int32_t* bt = block_table + (int64_t)w.seq * bt_stride;
        for (int i = 0; i < nt; ++i, ++n) {
          const int qk = 2 * n, qv = 2 * n + 1;
          const int sk = qk % kSlots, sv = qv % kSlots;
          const int tok0 = i * kTileTok;
          const int blk = __shfl_sync(0xffffffffu, bt[tok0 / block_size], 0);
          const int row0 = (blk * hkv + kvh) * block_size + (tok0 % block_size);
          if (qk >= kSlots) mbar_wait(&kv_free[sk], ((qk / kSlots) - 1) & 1);
          if (issuer) {
            mbar_expect_tx(&kv_full[sk], kTileBytes);
            tma_load_2d(kv + sk * kTileBytes, &map_k, 0, row0, &kv_full[sk]);
            tma_load_2d(kv + sk * kTileBytes + kHalfBytes, &map_k, 64, row0, &kv_full[sk]);
          }
          if (qv >= kSlots) mbar_wait(&kv_free[sv], ((qv / kSlots) - 1) & 1);
          if (issuer) {
            mbar_expect_tx(&kv_full[sv], kTileBytes);
            tma_load_2d(kv + sv * kTileBytes, &map_v, 0, row0, &kv_full[sv]);
            tma_load_2d(kv + sv * kTileBytes + kHalfBytes, &map_v, 64, row0, &kv_full[sv]);
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: S(n+1) is issued before waiting for P(n).  Whole warp in the loop, tcgen05 instructions
    // predicated to one elected lane, descriptors = uniform base + offset =====
    {
      constexpr uint32_t idesc_qk = fa_idesc(0);
      constexpr uint32_t idesc_pv = fa_idesc(1);
      const bool issuer = elect_one();
      const uint64_t desc_q = make_desc(smem_u32(qs), 16, 1024);
      const uint64_t desc_ph = make_desc(smem_u32(p_hi), 16, 1024);
      const uint64_t desc_pl = make_desc(smem_u32(p_lo), 16, 1024);
      const uint64_t desc_k0 = make_desc(smem_u32(kv), 16, 1024);
      const uint64_t desc_v0 = make_desc(smem_u32(kv), kHalfBytes, 1024);
      auto issue_s = [&](int n) {
        const int qk = 2 * n, sk = qk % kSlots;
        mbar_wait(&kv_full[sk], (qk / kSlots) & 1);
        tc_fence_after();
        const uint64_t dk = desc_k0 + (uint64_t)((sk * kTileBytes) >> 4);
        if (issuer) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {   // 16 head dims per step; +32 bytes inside the swizzle span, next half after 4
            const uint32_t o = (j >> 2) * kHalfBytes + (j & 3) * 32;
            tc_mma(tmem_base + (n & 1) * 128, desc_q + (o >> 4), dk + (o >> 4), idesc_qk, j ? 1u : 0u);
          }
          tc_commit(&kv_free[sk]);
          tc_commit(&s_full[n & 1]);
        }
        __syncwarp();
      };
      int n = 0, active = 0;
      for (int it = blockIdx.x; it < n_work; it += gridDim.x) {
        PlanItem w;
        int kvh;
        get_work(it, w, kvh);
        const int nt = __shfl_sync(0xffffffffu, (w.ctx_max + kTileTok - 1) / kTileTok, 0);
        if (nt == 0) continue;
        mbar_wait(q_ready, active & 1);
        tc_fence_after();
        issue_s(n);
        if (nt == 1 && issuer) tc_commit(q_free);
        for (int i = 0; i < nt; ++i, ++n) {
          if (i + 1 < nt) {
            issue_s(n + 1);
            if (i + 2 == nt && issuer) tc_commit(q_free);   // last S of this item issued
          }
          const int qv = 2 * n + 1, sv = qv % kSlots;
          mbar_wait(&kv_full[sv], (qv / kSlots) & 1);
          mbar_wait(p_ready, n & 1);
          tc_fence_after();
          // written before the P hand-over; every lane reads the same word
          const bool fresh = __shfl_sync(0xffffffffu, *reinterpret_cast<volatile int*>(pv_fresh), 0) != 0;
          const uint64_t dv = desc_v0 + (uint64_t)((sv * kTileBytes) >> 4);
          if (issuer) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {   // 16 tokens per step
              const uint32_t ao = (j >> 2) * kHalfBytes + (j & 3) * 32;
              const uint64_t db = dv + (uint64_t)((j * 2048) >> 4);   // MN-major: two 8-token groups
              tc_mma(tmem_base + 256, desc_ph + (ao >> 4), db, idesc_pv, (j || !fresh) ? 1u : 0u);
              tc_mma(tmem_base + 256, desc_pl + (ao >> 4), db, idesc_pv, 1u);
            }
            tc_commit(&kv_free[sv]);
            tc_commit(&o_full[n & 1]);
          }
          __syncwarp();
        }
        ++active;
      }
    }
  } else if (warp == 10) {
    // ===== Q stager: rows (row0 + m / G, head kvh * G + m % G) -> the swizzled K-major A operand =====
    int active = 0;
    for (int it = blockIdx.x; it < n_work; it += gridDim.x) {
      PlanItem w;
      int kvh;
      get_work(it, w, kvh);
      if (w.ctx_max <= 0) continue;
      if (active >= 1) mbar_wait(q_free, (active - 1) & 1);
#pragma unroll 8
      for (int i = lane; i < 128 * 16; i += 32) {
        const int m = i >> 4, ch = i & 15;
        const int r = m / G, g = m % G;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (r < w.nrows)
          v = *reinterpret_cast<const uint4*>(q + (int64_t)(w.row0 + r) * ldq + (kvh * G + g) * kHeadDim + ch * 8);
        *reinterpret_cast<uint4*>(qs + tile_chunk_off(m, ch)) = v;
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      mbar_arrive(q_ready);
      ++active;
    }
  } else {
    // ===== softmax / output: 256 threads; a PAIR of threads (warps w and w + 4, same TMEM lane) owns query
    // row m: thread `half` takes tokens [64 half, 64 half + 64) of S/P and head dims [64 half, +64) of O.
    // TMEM reads are the scarce resource (64 B/clk: a full 128 x 128 fp32 tile costs 1024 cycles), so S is
    // read ONCE per tile (kept in registers between the max and the exp) and O_t is NOT read per tile: the
    // P.V products accumulate in TMEM for as long as every row's reference max m_used is good enough
    // (tile max <= m_used + 8 in the log2 domain, i.e. p <= 256: exact, only the scaling is stale); a tile
    // where some row jumps further folds O_t into the registers, rescales and starts a new O_t. =====
    const int half = (warp - 2) >> 2;
    const int quad = warp & 3;
    const int m = quad * 32 + lane;
    const int r = m / G, g = m % G;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(half * 64);
    const float sc2 = scale * 1.4426950408889634f;   // scores in the log2 domain: p = exp2(s * sc2 - m_used)
    const bool lead = m == 0 && half == 0;
    if (lead) fa_trace(0);
    int n = 0;
    for (int it = blockIdx.x; it < n_work; it += gridDim.x) {
      PlanItem w;
      int kvh;
      get_work(it, w, kvh);
      const int nt = (w.ctx_max + kTileTok - 1) / kTileTok;
      if (nt == 0) continue;
      const bool alive = r < w.nrows;
      const int ctx_m = alive ? ctx_lens[w.row0 + r] : 0;
      float m_used = -INFINITY, l_run = 0.f;
      float acc[64];
#pragma unroll
      for (int d = 0; d < 64; ++d) acc[d] = 0.f;
      for (int i = 0; i < nt; ++i, ++n) {
        const int tok0 = i * kTileTok + half * 64;                 // first token of this thread's 64 columns
        const int n_live = min(max(ctx_m - tok0, 0), 64);          // causal: columns [0, n_live) count
        const int n_valid = min(kTileTok, w.ctx_max - i * kTileTok);   // tokens of the tile that exist at all
        // tcgen05.ld is warp-collective: which 32-column chunks are read is decided per warp
        const int n_live_w = __reduce_max_sync(0xffffffffu, n_live);
        const bool full = __reduce_min_sync(0xffffffffu, n_live) == 64;   // no masking needed in this warp
        if (lead && n < 40) fa_trace(8 + 6 * n);        // tile begins (waiting for S)
        mbar_wait(&s_full[n & 1], (n >> 1) & 1);
        if (lead && n < 40) fa_trace(9 + 6 * n);        // S landed
        tc_fence_after();
        const uint32_t s_addr = lane_addr + (n & 1) * 128;
        // S row (this thread's 64 columns), masked columns = -inf
        float sv[64];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          if (c * 32 < n_live_w) {
            uint32_t t32[32];
            tmem_ld32(s_addr + c * 32, t32);
            if (full) {
#pragma unroll
              for (int j = 0; j < 32; ++j) sv[c * 32 + j] = __uint_as_float(t32[j]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) sv[c * 32 + j] = (c * 32 + j < n_live) ? __uint_as_float(t32[j]) : -INFINITY;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) sv[c * 32 + j] = -INFINITY;
          }
        }
        float mx = -INFINITY;
#pragma unroll
        for (int j = 0; j < 64; ++j) mx = fmaxf(mx, sv[j]);
        // the pair's row max through shared memory; the same barrier ORs "some row must rescale"
        float* xm = xmax + (n & 1) * 256;
        xm[half * 128 + m] = mx;
        const int jump = (i == 0) || (mx * sc2 > m_used + 8.f);
        int any_jump;
        asm volatile(
            "{\n"
            ".reg .pred pi, po;\n"
            "setp.ne.b32 pi, %1, 0;\n"
            "bar.red.or.pred po, 1, 256, pi;\n"
            "selp.b32 %0, 1, 0, po;\n"
            "}\n"
            : "=r"(any_jump)
            : "r"(jump)
            : "memory");
        mx = fmaxf(mx, xm[(half ^ 1) * 128 + m]);
        if (lead && n < 40) fa_trace(10 + 6 * n);       // max + pair exchange done
        // the previous tile's P.V MMAs must have finished reading P before it is overwritten below
        if (i > 0) {
          mbar_wait(&o_full[(n - 1) & 1], ((n - 1) >> 1) & 1);
          tc_fence_after();
        }
        if (lead && n < 40) fa_trace(11 + 6 * n);       // previous P.V complete
        if (any_jump) {   // CTA-uniform
          // fold what has accumulated in TMEM since the last fold (same reference max as acc), then move
          // every row to its true running max
          if (i > 0) {
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              uint32_t o[32];
              tmem_ld32(lane_addr + 256 + c * 32, o);
#pragma unroll
              for (int j = 0; j < 32; ++j) acc[c * 32 + j] += __uint_as_float(o[j]);
            }
          }
          const float m_new = fmaxf(m_used, mx * sc2);
          const float alpha = (m_new > -INFINITY && m_used > -INFINITY) ? fast_exp2(m_used - m_new) : 1.f;
          if (i > 0) {
#pragma unroll
            for (int d = 0; d < 64; ++d) acc[d] *= alpha;
          }
          l_run *= alpha;
          m_used = m_new;
          if (lead) *reinterpret_cast<volatile int*>(pv_fresh) = 1;
        } else if (lead) {
          *reinterpret_cast<volatile int*>(pv_fresh) = 0;
        }
        if (lead && n < 40) fa_trace(12 + 6 * n);       // fold done
        // P = exp2(s * sc2 - m_used) as hi + lo bf16 (hi = p truncated to bf16, lo = the exact rest rounded):
        // this thread's 64 columns of row m of the A operand of P.V
        const float mref = m_used > -INFINITY ? m_used : 0.f;   // a row that has seen nothing: every s is -inf
        float lsum = 0.f;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t hi[16], lo[16];
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            const float p0 = fast_exp2(sv[c * 32 + j] * sc2 - mref);       // exp2(-inf) = 0 for masked columns
            const float p1 = fast_exp2(sv[c * 32 + j + 1] * sc2 - mref);
            lsum += p0 + p1;
            const uint32_t h0 = __float_as_uint(p0) & 0xffff0000u, h1 = __float_as_uint(p1) & 0xffff0000u;
            hi[j >> 1] = __byte_perm(h0, h1, 0x7632);
            lo[j >> 1] = pack_bf16(p0 - __uint_as_float(h0), p1 - __uint_as_float(h1));
          }
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint32_t off = tile_chunk_off(m, half * 8 + c * 4 + k);
            *reinterpret_cast<uint4*>(p_hi + off) = make_uint4(hi[4 * k], hi[4 * k + 1], hi[4 * k + 2], hi[4 * k + 3]);
            *reinterpret_cast<uint4*>(p_lo + off) = make_uint4(lo[4 * k], lo[4 * k + 1], lo[4 * k + 2], lo[4 * k + 3]);
          }
        }
        l_run += lsum;
        if (n_valid < kTileTok && half == 0) {
          // partial last tile: V rows past the context may hold anything (0 * NaN = NaN on the tensor core):
          // wait for the tile, then zero this thread's token row if it is past the end
          const int qv = 2 * n + 1, sv_slot = qv % kSlots;
          mbar_wait(&kv_full[sv_slot], (qv / kSlots) & 1);
          if (m >= n_valid) {
            unsigned char* vb = kv + sv_slot * kTileBytes;
#pragma unroll
            for (int h2 = 0; h2 < 2; ++h2)
#pragma unroll
              for (int cc = 0; cc < 8; ++cc)
                *reinterpret_cast<uint4*>(vb + h2 * kHalfBytes + m * 128 + cc * 16) = make_uint4(0u, 0u, 0u, 0u);
          }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic smem writes -> visible to the MMA
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(p_ready);
        if (lead && n < 40) fa_trace(13 + 6 * n);       // P handed over
      }
      // the pair's row sum, what is still in TMEM, then normalise and store 64 dims = 128 contiguous bytes
      {
        float* xl = xmax + (n & 1) * 256;   // the buffer the next tile would use: free until the barrier below
        xl[half * 128 + m] = l_run;
        asm volatile("bar.sync 1, 256;" ::: "memory");
        const float l_tot = l_run + xl[(half ^ 1) * 128 + m];
        mbar_wait(&o_full[(n - 1) & 1], ((n - 1) >> 1) & 1);
        tc_fence_after();
        const float inv = l_tot > 0.f ? 1.f / l_tot : 0.f;
        __nv_bfloat16* dst = out + (int64_t)(w.row0 + r) * ldo + (kvh * G + g) * kHeadDim + half * 64;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t o[32];
          tmem_ld32(lane_addr + 256 + c * 32, o);
          if (alive) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              uint32_t pk[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const int d = c * 32 + k * 8 + e * 2;
                const float v0 = (acc[d] + __uint_as_float(o[k * 8 + e * 2])) * inv;
                const float v1 = (acc[d + 1] + __uint_as_float(o[k * 8 + e * 2 + 1])) * inv;
                pk[e] = pack_bf16(v0, v1);
              }
              *reinterpret_cast<uint4*>(dst + c * 32 + k * 8) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            }
          }
        }
        tc_fence_before();
        asm volatile("bar.sync 1, 256;" ::: "memory");   // xl is rewritten by the next item; O_t has been read by all
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kFaTmemCols));
  }
}

template <int G>
int launch_fa(const CUtensorMap& mk, const CUtensorMap& mv, const __nv_bfloat16* q, int64_t ldq, const int32_t* block_table,
              int bt_stride, const int32_t* ctx_lens, __nv_bfloat16* out, int64_t ldo, int T, int hkv, int block_size,
              float scale, const PlanItem* plan, cudaStream_t s) {
  static bool attr_set[64] = {};
  B200_CUDA(ensure_dynamic_smem(paged_prefill_fa_kernel<G>, kFaSmem, attr_set));
  const int64_t max_work = (int64_t)T * hkv;   // upper bound on items x kv heads; the kernel reads the real count
  const int grid = (int)(max_work < kNumSMs ? max_work : kNumSMs);
  B200_CUDA(launch(paged_prefill_fa_kernel<G>, dim3(grid), dim3(kFaThreads), (size_t)kFaSmem, s, mk, mv, q, ldq, block_table,
                   bt_stride, ctx_lens, out, ldo, hkv, block_size, scale, plan));
  return B200_OK;
}

}  // namespace

int attn_prefill_read_trace(unsigned long long* host256) {
  B200_CUDA(cudaMemcpyFromSymbol(host256, g_fa_trace, 256 * sizeof(unsigned long long)));
  return B200_OK;
}

int64_t paged_attn_prefill_plan_bytes(int T) { return (int64_t)(T + 1) * (int64_t)sizeof(PlanItem); }

bool paged_attn_prefill_fa_supported(int hq, int hkv, int block_size) {
  if (hkv <= 0 || hq % hkv != 0 || block_size != kTileTok) return false;
  const int G = hq / hkv;
  return G == 1 || G == 2 || G == 4 || G == 5 || G == 8;
}

// Same contract as paged_attn (attn.cu); `ws` holds the plan (paged_attn_prefill_plan_bytes(T)).
// The plan depends on the step's rows only, not on the layer: a caller that runs many layers over the same
// rows builds it once (this function) and passes plan_ready = true to paged_attn_prefill_fa.
int paged_attn_prefill_plan(const int32_t* row_to_seq, const int32_t* ctx_lens, int T, int hq, int hkv, void* ws,
                            int64_t ws_bytes, cudaStream_t s) {
  B200_REQUIRE(hkv > 0 && hq % hkv == 0, B200_ERR_INVALID, "attn_prefill: hq=%d not a multiple of hkv=%d", hq, hkv);
  if (T == 0) return B200_OK;
  B200_REQUIRE(ws && ws_bytes >= paged_attn_prefill_plan_bytes(T), B200_ERR_INVALID,
               "attn_prefill: workspace too small (%lld < %lld)", (long long)ws_bytes,
               (long long)paged_attn_prefill_plan_bytes(T));
  B200_CUDA(launch(attn_prefill_plan_kernel, dim3(1), dim3(256), (size_t)0, s, row_to_seq, ctx_lens, T, 128 / (hq / hkv),
                   static_cast<PlanItem*>(ws)));
  return B200_OK;
}

int paged_attn_prefill_fa(const __nv_bfloat16* q, int64_t ldq, const __nv_bfloat16* kc, const __nv_bfloat16* vc,
                          const int32_t* block_table, int bt_stride, const int32_t* row_to_seq, const int32_t* ctx_lens,
                          __nv_bfloat16* out, int64_t ldo, int T, int hq, int hkv, int block_size, float scale, void* ws,
                          int64_t ws_bytes, bool plan_ready, cudaStream_t s) {
  B200_REQUIRE(paged_attn_prefill_fa_supported(hq, hkv, block_size), B200_ERR_UNSUPPORTED, "attn_prefill: unsupported shape");
  B200_REQUIRE(ldq % 8 == 0 && ldo % 8 == 0, B200_ERR_INVALID, "attn_prefill: ldq/ldo must be multiples of 8");
  B200_REQUIRE(((reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(out)) & 15) == 0, B200_ERR_INVALID,
               "attn_prefill: q and out must be 16-byte aligned");
  B200_REQUIRE(scale > 0.f, B200_ERR_INVALID, "attn_prefill: scale must be positive");
  if (T == 0) return B200_OK;
  B200_REQUIRE(ws && ws_bytes >= paged_attn_prefill_plan_bytes(T), B200_ERR_INVALID,
               "attn_prefill: workspace too small (%lld < %lld)", (long long)ws_bytes,
               (long long)paged_attn_prefill_plan_bytes(T));
  const int G = hq / hkv;
  PlanItem* plan = static_cast<PlanItem*>(ws);
  if (!plan_ready) {
    const int rc0 = paged_attn_prefill_plan(row_to_seq, ctx_lens, T, hq, hkv, ws, ws_bytes, s);
    if (rc0) return rc0;
  }
  CUtensorMap mk, mv;
  const uint64_t rows = 1ull << 31;   // upper limit for TMA's bounds logic; rows touched come from the block table
  int rc = get_tensor_map_2d(kc, rows, kHeadDim, kHeadDim, kTileTok, &mk);
  if (rc) return rc;
  rc = get_tensor_map_2d(vc, rows, kHeadDim, kHeadDim, kTileTok, &mv);
  if (rc) return rc;
  switch (G) {
    case 1: return launch_fa<1>(mk, mv, q, ldq, block_table, bt_stride, ctx_lens, out, ldo, T, hkv, block_size, scale, plan, s);
    case 2: return launch_fa<2>(mk, mv, q, ldq, block_table, bt_stride, ctx_lens, out, ldo, T, hkv, block_size, scale, plan, s);
    case 4: return launch_fa<4>(mk, mv, q, ldq, block_table, bt_stride, ctx_lens, out, ldo, T, hkv, block_size, scale, plan, s);
    case 5: return launch_fa<5>(mk, mv, q, ldq, block_table, bt_stride, ctx_lens, out, ldo, T, hkv, block_size, scale, plan, s);
    case 8: return launch_fa<8>(mk, mv, q, ldq, block_table, bt_stride, ctx_lens, out, ldo, T, hkv, block_size, scale, plan, s);
  }
  return B200_ERR_UNSUPPORTED;
}

}  // namespace b200
