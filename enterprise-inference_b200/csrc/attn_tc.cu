// (a6) Paged attention for single-query rows on the 5th-gen tensor cores (tcgen05 + TMEM).
//
// The CUDA-core kernel in attn.cu is compute bound (fp32 FMA + shuffles per cached token); this
// one keeps the math on the tensor pipe so that the kernel is bound by the KV stream again.
// Bound: HBM.  Algorithmic bytes per launch = sum over rows of ctx_len * n_kv_heads * 128 * 2 * 2.
//
// Shapes are swapped so that the 128 cached tokens of a tile fill the UMMA M dimension:
//   S^T[128 tok][16]  = K_tile[128 tok x 128 dim] . Q^T          (A = K, K-major;  B = Q padded to 16 heads)
//   O^T[128 dim][16]  = V_tile^T[128 dim x 128 tok] . P^T        (A = V, MN-major; B = P, 16 heads x 128 tok)
// K and V tiles arrive by TMA (cp.async.bulk.tensor, SWIZZLE_128B, two 64-dim halves each) straight
// from the paged cache — a (block, kv head) tile is one box — into a ring of 32 KiB slots (K, V, K, V, ...:
// three tiles in flight with six slots).  The softmax threads own one token (TMEM lane) each for S and
// one head-dim each for O: scores are read with tcgen05.ld, scaled, masked and exponentiated in fp32
// (online softmax, tile max per column by redux.sync + one exchange through smem), P is written back to
// smem as TWO bf16 terms (hi + lo, so that P keeps ~16 mantissa bits: R4 of the oracle keeps P in fp32)
// and the running output is rescaled in registers, so TMEM never needs a read-modify-write.  The live MMA
// columns are dealt to CG groups of four softmax warps (2 for decode, 4 for packed prefill rows) so that
// every scheduler has 2-4 warps to interleave.  GQA: the G query heads of the kv head share every K/V
// byte.  Split-KV + LSE merge as in attn.cu (same workspace layout, same merge kernel).  Prefill-dominated
// steps go to attn_prefill.cu (query rows on the MMA M) unless B200_ATTN_FA=0.
//
// Replaces: paged_attention_v1/v2 CPU kernels of the vLLM-CPU image
// ([V22] vllm/v1/attention/backends/cpu_attn.py:377-396).
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"

namespace b200 {

int get_tensor_map_2d(const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows, CUtensorMap* out);
void attn_split_plan_tc(int T, int hkv, int max_ctx, int* n_splits, int* tiles_per_split);
int launch_attn_merge(const float* part_o, const float* part_ml, __nv_bfloat16* out, int64_t ldo, int T, int hq,
                      int n_splits, cudaStream_t s);

namespace {

constexpr int kTileTok = 128;
constexpr int kHalfBytes = kTileTok * 64 * 2;     // one [128 tok][64 dim] swizzled sub-tile: 16 KiB
constexpr int kSlotBytes = 2 * kHalfBytes;        // one K tile or one V tile (lo|hi dims): 32 KiB
constexpr int kSmallTile = 16 * 64 * 2;           // one [16 heads][64] swizzled sub-tile: 2 KiB
// TMA warp, MMA warp, CG groups of 4 softmax warps, Q-staging warp
__host__ __device__ constexpr int attn_threads(int cg) { return (3 + 4 * cg) * 32; }
// column groups for NC live MMA columns: 2 warps per scheduler from 2 columns on, 4 from 8 on
__host__ __device__ constexpr int attn_col_groups(int nc) { return nc >= 8 ? 4 : (nc >= 2 ? 2 : 1); }
constexpr int kNPad = 16;                         // heads padded to the UMMA N
constexpr uint32_t kTmemCols = 64;                // S0 | S1 | O0 | O1, 16 columns each

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
// SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout).
//  K-major  operand: rows are 128 B, 8-row groups SBO = 1024 B apart, LBO unused (1).
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
// bf16 x bf16 -> fp32, M = 128, N = 16; a_major selects K-major (0) or MN-major (1) for A
__host__ __device__ constexpr uint32_t make_idesc(int a_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)(kNPad >> 3) << 17) |
         ((uint32_t)(128 >> 4) << 24);
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
template <int NL>
__device__ __forceinline__ void tmem_ld(uint32_t taddr, uint32_t (&r)[NL]) {
  if constexpr (NL == 1) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r[0]) : "r"(taddr));
  } else if constexpr (NL == 2) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(taddr));
  } else if constexpr (NL == 4) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(taddr));
  } else if constexpr (NL == 8) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr));
  } else {
    static_assert(NL == 16, "tmem_ld width");
    tmem_ld16(taddr, r);
    return;
  }
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// byte offset of element (row r < 16, column c < 128) inside a [16][128] bf16 operand stored as two
// K-major SWIZZLE_128B sub-tiles of 64 columns (Q and P operands)
__device__ __forceinline__ uint32_t small_off(int r, int c) {
  const int sub = c >> 6, cc = c & 63;
  const int chunk = (cc >> 3) ^ (r & 7);
  return sub * kSmallTile + (r >> 3) * 1024 + (r & 7) * 128 + chunk * 16 + (cc & 7) * 2;
}

// debug (B200_ATTN_TRACE=1): %globaltimer stamps of CTA 0's first softmax thread
__device__ unsigned long long g_attn_trace[64];
__device__ int g_attn_trace_on;
__device__ __forceinline__ void attn_trace(int slot) {
#ifndef B200_TRACE_BUILD
  (void)slot;   // compiled out of the product build: even a predicated global load sits on the per-item path
  return;
#endif
  if (g_attn_trace_on && blockIdx.x == 0 && slot < 64) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    g_attn_trace[slot] = t;
  }
}

struct Smem {
  // dynamic, 1024-aligned: kv[kKvStages][kStageBytes] | q[2*kSmallTile] | p_hi[2][2*kSmallTile] | p_lo[2][...]
  // followed by the barriers and small exchange areas (see carve-up in the kernel)
};

// NSLOT: 32 KiB ring slots, alternately a K tile and a V tile (tile n: K in slot 2n % NSLOT, V in slot
// (2n+1) % NSLOT).  3 slots (96 KiB) let a CTA share its SM with a CTA of the neighbouring linear kernel,
// so that under programmatic dependent launch their prologues and first tiles overlap this kernel's tail
// and vice versa; 6 slots (192 KiB in flight) are used for long contexts where the stream is what matters.
template <int G, int R, int NSLOT, int CG>
__global__ void __launch_bounds__(attn_threads(CG), 1)
paged_attn_tc_kernel(const __grid_constant__ CUtensorMap map_k, const __grid_constant__ CUtensorMap map_v,
                     const __nv_bfloat16* __restrict__ q, int64_t ldq, const int32_t* __restrict__ block_table,
                     int bt_stride, const int32_t* __restrict__ row_to_seq, const int32_t* __restrict__ ctx_lens,
                     __nv_bfloat16* __restrict__ out, int64_t ldo, int hkv, int block_size, int tiles_per_split,
                     int n_splits, int n_rows, float scale, float* __restrict__ part_o, float* __restrict__ part_ml,
                     int early) {
  extern __shared__ unsigned char smem_raw[];
  if (threadIdx.x == 64) attn_trace(0);
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* kv = smem;
  unsigned char* qs = kv + NSLOT * kSlotBytes;
  unsigned char* p_hi = qs + 2 * 2 * kSmallTile;          // qs: [2 buffers][2 sub-tiles]; p: [2 buffers][2 sub-tiles]
  unsigned char* p_lo = p_hi + 2 * 2 * kSmallTile;
  uint64_t* bars = reinterpret_cast<uint64_t*>(p_lo + 2 * 2 * kSmallTile);
  uint64_t* kv_full = bars;                  // [NSLOT] TMA landed
  uint64_t* kv_free = kv_full + 6;           // [NSLOT] the MMAs that read the slot have completed
  uint64_t* s_full = kv_free + 6;            // [2]
  uint64_t* p_ready = s_full + 2;            // [2]
  uint64_t* o_full = p_ready + 2;            // [2] one per O buffer
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 6);   // o_full[2..3] q_ready[2], o_full[4..5] q_free[2]
  float* xmax_all = reinterpret_cast<float*>(tmem_slot + 2);   // [CG groups][2 (tile parity)][4 warps][16 columns]
  float* xsum = xmax_all + 4 * 128;                             // [CG groups][4 warps][16 columns]

  constexpr int NC = G * R;   // live columns of the N = 16 MMA
  static_assert(NC <= kNPad, "row group does not fit the MMA N");
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int hq = hkv * G;
  uint64_t* q_ready = o_full + 2;   // [2], 32 arrivals: Q of work item j is staged in qs[j & 1]
  uint64_t* q_free = o_full + 4;    // [2], tcgen05.commit: the S MMAs of item j no longer read qs[j & 1]

  // Persistent CTA: work items (row group, kv head, split) are taken round-robin; the K/V ring, the S/P
  // double buffers and all barrier phases run on ONE tile counter across items, so the TMA producer
  // streams the next item's tiles while the softmax threads finish the current one.
  //
  // A row group is R consecutive query rows: their R*G (<= 16) query heads fill the columns of the
  // N = 16 MMA.  When the R rows belong to ONE sequence (a prefill chunk) the group is one pass over that
  // sequence's K/V with a per-column causal mask (1/R of the cache traffic of row-at-a-time); otherwise
  // (decode rows of different sequences, a chunk boundary) it degrades to R single-row passes ("runs").
  const int n_groups = (n_rows + R - 1) / R;
  const int n_items = n_splits * hkv * n_groups;
  struct Item {
    int r0, kvh, split, tile_begin, nrun, total_tiles, blk0;
    int ctx[R];            // context length per row (0: row absent)
    int run_tiles[R];      // tiles of run u inside this split
    int run_ctx[R];        // longest context of run u
    const int32_t* run_bt[R];
  };
  auto get_item = [&](int it) {
    Item w;   // split-major: all first chunks, then all second chunks, ... (long items first)
    const int n_base = hkv * n_groups;
    w.split = it / n_base;
    w.kvh = (it % n_base) % hkv;
    w.r0 = ((it % n_base) / hkv) * R;
    w.tile_begin = w.split * tiles_per_split;
    int seq[R];
    bool same = true;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int row = w.r0 + r;
      const bool valid = row < n_rows;
      w.ctx[r] = valid ? ctx_lens[row] : 0;
      seq[r] = valid ? (row_to_seq ? row_to_seq[row] : row) : -1;
      if (valid && seq[r] != seq[0]) same = false;
    }
    w.nrun = same ? 1 : R;
    w.total_tiles = 0;
#pragma unroll
    for (int u = 0; u < R; ++u) {
      int cmax = 0, sq = -1;
      if (same) {
        if (u == 0) {
          sq = seq[0];
#pragma unroll
          for (int r = 0; r < R; ++r) cmax = max(cmax, w.ctx[r]);
        }
      } else {
        cmax = w.ctx[u];
        sq = seq[u];
      }
      const int total = (cmax + kTileTok - 1) / kTileTok;
      w.run_ctx[u] = cmax;
      w.run_tiles[u] = sq < 0 ? 0 : max(min(w.tile_begin + tiles_per_split, total) - w.tile_begin, 0);
      w.run_bt[u] = block_table + (int64_t)max(sq, 0) * bt_stride;
      w.total_tiles += w.run_tiles[u];
    }
    // first tile's block id, fetched with the rest of the metadata (one dependent load less at the item switch)
    w.blk0 = w.run_tiles[0] > 0 ? w.run_bt[0][(w.tile_begin * kTileTok) / block_size] : 0;
    return w;
  };
  // The first item's metadata is requested BEFORE the set-up below and before griddepcontrol.wait: its
  // dependent global loads (~1 us) overlap the barrier/TMEM set-up.  Safe under programmatic dependent
  // launch: context lengths and block tables are written at the start of a step, at least two kernels
  // earlier, and a kernel only launches once its predecessor has passed its own griddepcontrol.wait.
  // (Decode kernels only: with R > 1 the compiler keeps a hoisted Item on the stack, and the stores of
  // the look-ahead fetch then wait for their loads -- 1.6 us per item; those roles fetch their own.)
  Item first = {};
  if (R == 1 && (int)blockIdx.x < n_items) first = get_item(blockIdx.x);
  auto first_item = [&]() -> Item {
    if constexpr (R == 1) {
      return first;
    } else {
      Item z = {};
      if ((int)blockIdx.x < n_items) z = get_item(blockIdx.x);
      return z;
    }
  };

  int pre_tiles = 0;   // tiles of the first item already requested (meaningful in thread 0 = the TMA thread)
  if (threadIdx.x == 0) {
    for (int s = 0; s < NSLOT; ++s) {
      mbar_init(&kv_full[s], 1);
      mbar_init(&kv_free[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&s_full[s], 1);
      mbar_init(&p_ready[s], 4 * CG);   // one arrival per softmax warp
      mbar_init(&o_full[s], 1);
      mbar_init(&q_ready[s], 32);
      mbar_init(&q_free[s], 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    // Decode-only steps (`early`): every row is its own sequence and the previous kernel wrote only
    // position ctx-1 of it, so the tiles that end before that position are from earlier steps and can be
    // requested right now -- before the set-up below and before griddepcontrol.wait.
    if (R == 1 && early && first.total_tiles > 0) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      const int lim = min(NSLOT / 2, first.run_tiles[0]);
      int blk[NSLOT / 2];
#pragma unroll
      for (int i = 0; i < NSLOT / 2; ++i)
        blk[i] = (i > 0 && i < lim) ? first.run_bt[0][((first.tile_begin + i) * kTileTok) / block_size] : first.blk0;
#pragma unroll
      for (int i = 0; i < NSLOT / 2; ++i) {
        const int tok0 = (first.tile_begin + i) * kTileTok;
        if (i >= lim || i != pre_tiles || tok0 + kTileTok > first.run_ctx[0] - 1) continue;
        const int row0 = (blk[i] * hkv + first.kvh) * block_size + (tok0 % block_size);
        mbar_expect_tx(&kv_full[2 * i], kSlotBytes);
        tma_load_2d(kv + (2 * i) * kSlotBytes, &map_k, 0, row0, &kv_full[2 * i]);
        tma_load_2d(kv + (2 * i) * kSlotBytes + kHalfBytes, &map_k, 64, row0, &kv_full[2 * i]);
        mbar_expect_tx(&kv_full[2 * i + 1], kSlotBytes);
        tma_load_2d(kv + (2 * i + 1) * kSlotBytes, &map_v, 0, row0, &kv_full[2 * i + 1]);
        tma_load_2d(kv + (2 * i + 1) * kSlotBytes + kHalfBytes, &map_v, 64, row0, &kv_full[2 * i + 1]);
        ++pre_tiles;
      }
    }
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(kTmemCols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  // zero the small operand tiles (Q, P hi/lo) once: rows >= G must not feed NaNs into the unused columns
  for (int i = threadIdx.x; i < (4 + 8) * kSmallTile / 16; i += attn_threads(CG))
    reinterpret_cast<uint4*>(qs)[i] = make_uint4(0u, 0u, 0u, 0u);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();               // q and the newest K/V rows come from the previous kernel; the set-up above overlapped its tail
  pdl_launch_dependents();
  if (threadIdx.x == 64) attn_trace(1);

  // Every role walks the same item sequence; an item's metadata (context lengths, block-table rows:
  // dependent global loads, ~1 us each) is fetched one item AHEAD (`nxt`) so that an item switch never
  // stalls the TMA stream.

  if (warp == 0) {
    // ===== TMA producer (whole warp in the loop, the TMA instructions predicated to one lane; loop bounds and block
    // ids are broadcast from lane 0 so that the compiler knows them warp-uniform) =====
    {
      const bool issuer = elect_one();
      pre_tiles = __shfl_sync(0xffffffffu, pre_tiles, 0);
      if (issuer) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_k) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_v) : "memory");
      }
      int n = 0;
      Item nxt = first_item();
      if (R > 1 && (int)blockIdx.x < n_items) nxt = get_item(blockIdx.x);
      for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
        const Item w = nxt;
        if (it + (int)gridDim.x < n_items) nxt = get_item(it + gridDim.x);
        const int nrun = __shfl_sync(0xffffffffu, w.nrun, 0);
#pragma unroll
        for (int u = 0; u < R; ++u) {
          if (u >= nrun) break;
          const int nt = __shfl_sync(0xffffffffu, w.run_tiles[u], 0);
          for (int i = 0; i < nt; ++i, ++n) {
            const int qk = 2 * n, qv = 2 * n + 1;          // ring sequence numbers of this tile's K and V
            const int sk = qk % NSLOT, sv = qv % NSLOT;
            const int tok0 = (w.tile_begin + i) * kTileTok;
            if (n < pre_tiles) continue;   // requested before the set-up (decode-only steps)
            const int blk = __shfl_sync(0xffffffffu, (u == 0 && i == 0) ? w.blk0 : w.run_bt[u][tok0 / block_size], 0);
            const int row0 = (blk * hkv + w.kvh) * block_size + (tok0 % block_size);
            if (qk >= NSLOT) mbar_wait(&kv_free[sk], ((qk / NSLOT) - 1) & 1);
            if (issuer) {
              mbar_expect_tx(&kv_full[sk], kSlotBytes);
              tma_load_2d(kv + sk * kSlotBytes, &map_k, 0, row0, &kv_full[sk]);
              tma_load_2d(kv + sk * kSlotBytes + kHalfBytes, &map_k, 64, row0, &kv_full[sk]);
            }
            if (qv >= NSLOT) mbar_wait(&kv_free[sv], ((qv / NSLOT) - 1) & 1);
            if (issuer) {
              mbar_expect_tx(&kv_full[sv], kSlotBytes);
              tma_load_2d(kv + sv * kSlotBytes, &map_v, 0, row0, &kv_full[sv]);
              tma_load_2d(kv + sv * kSlotBytes + kHalfBytes, &map_v, 64, row0, &kv_full[sv]);
            }
            __syncwarp();
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer, software pipelined: S(n+1) is issued before waiting for P(n).
    // The tiles of an item (all its runs) form one flat sequence: Q is the same for all of them.
    // The WHOLE warp walks the loop and waits on the barriers; only the tcgen05.mma / commit instructions are
    // predicated to one lane.  Everything they consume (descriptors, TMEM addresses) is then warp-uniform by
    // construction and lives in uniform registers: ~4 instructions per MMA instead of the ~20 (register -> uniform
    // register broadcast loops) that a single-lane branch costs, on the one thread every tile waits for. =====
    {
      constexpr uint32_t idesc_qk = make_idesc(0);
      constexpr uint32_t idesc_pv = make_idesc(1);
      const bool issuer = elect_one();
      // descriptor = base | (address >> 4): the base bits are constants, the address part is added per operand
      const uint64_t desc_k0 = make_desc(smem_u32(kv), 16, 1024);
      const uint64_t desc_v0 = make_desc(smem_u32(kv), kHalfBytes, 1024);
      const uint64_t desc_q0 = make_desc(smem_u32(qs), 16, 1024);
      const uint64_t desc_ph0 = make_desc(smem_u32(p_hi), 16, 1024);
      const uint64_t desc_pl0 = make_desc(smem_u32(p_lo), 16, 1024);
      uint64_t desc_q = desc_q0;
      auto issue_s = [&](int n) {
        const int qk = 2 * n, sk = qk % NSLOT;
        mbar_wait(&kv_full[sk], (qk / NSLOT) & 1);
        tc_fence_after();
        const uint64_t dk = desc_k0 + (uint64_t)((sk * kSlotBytes) >> 4);
        if (issuer) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t ao = (j >> 2) * kHalfBytes + (j & 3) * 32;
            const uint32_t bo = (j >> 2) * kSmallTile + (j & 3) * 32;
            tc_mma(tmem_base + (n & 1) * 16, dk + (ao >> 4), desc_q + (bo >> 4), idesc_qk, j ? 1u : 0u);
          }
          tc_commit(&kv_free[sk]);   // the K slot is free as soon as these MMAs have read it
          tc_commit(&s_full[n & 1]);
        }
        __syncwarp();
      };
      int n = 0, active = 0;
      Item nxt = first_item();
      if (R > 1 && (int)blockIdx.x < n_items) nxt = get_item(blockIdx.x);
      for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
        const int nt = __shfl_sync(0xffffffffu, nxt.total_tiles, 0);   // warp-uniform, and known to be
        if (it + (int)gridDim.x < n_items) nxt = get_item(it + gridDim.x);
        if (nt == 0) continue;
        mbar_wait(&q_ready[active & 1], (active >> 1) & 1);   // this item's Q is in shared memory
        tc_fence_after();
        desc_q = desc_q0 + (uint64_t)(((active & 1) * 2 * kSmallTile) >> 4);
        issue_s(n);
        if (nt == 1 && issuer) tc_commit(&q_free[active & 1]);
        for (int i = 0; i < nt; ++i, ++n) {
          if (i + 1 < nt) {
            issue_s(n + 1);
            if (i + 2 == nt && issuer) tc_commit(&q_free[active & 1]);   // last S of this item issued
          }
          const int qv = 2 * n + 1, sv = qv % NSLOT;
          mbar_wait(&kv_full[sv], (qv / NSLOT) & 1);
          mbar_wait(&p_ready[n & 1], (n >> 1) & 1);
          tc_fence_after();
          const uint64_t dv = desc_v0 + (uint64_t)((sv * kSlotBytes) >> 4);
          const uint64_t dph = desc_ph0 + (uint64_t)(((n & 1) * 2 * kSmallTile) >> 4);
          const uint64_t dpl = desc_pl0 + (uint64_t)(((n & 1) * 2 * kSmallTile) >> 4);
          if (issuer) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {  // 16 tokens per step: two 8-token K groups of the MN-major V tile
              const uint32_t bo = (j >> 2) * kSmallTile + (j & 3) * 32;
              tc_mma(tmem_base + 32 + (n & 1) * 16, dv + (uint64_t)((j * 2048) >> 4), dph + (bo >> 4), idesc_pv, j ? 1u : 0u);
              tc_mma(tmem_base + 32 + (n & 1) * 16, dv + (uint64_t)((j * 2048) >> 4), dpl + (bo >> 4), idesc_pv, 1u);
            }
            tc_commit(&kv_free[sv]);  // the V slot is no longer read
            tc_commit(&o_full[n & 1]);
          }
          __syncwarp();
        }
        ++active;
      }
    }
  } else if (warp == 2 + 4 * CG) {
    // ===== Q stager: the group's R*G query heads -> qs[j & 1] as the swizzled K-major B operand of
    // S^T = K . Q^T (column c = row c / G, head c % G), one item ahead of the MMA warp =====
    int active = 0;
    Item nxt = first_item();
    for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
      const Item w = nxt;
      if (it + (int)gridDim.x < n_items) nxt = get_item(it + gridDim.x);
      if (w.total_tiles == 0) continue;
      if (active >= 2) mbar_wait(&q_free[active & 1], ((active >> 1) - 1) & 1);
      unsigned char* qb = qs + (active & 1) * 2 * kSmallTile;
      for (int i = lane; i < NC * 16; i += 32) {
        const int c = i >> 4, ch = i & 15;   // 16-byte chunk ch of column c
        const int row = w.r0 + c / G;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (row < n_rows) v = *reinterpret_cast<const uint4*>(q + (int64_t)row * ldq + (w.kvh * G + c % G) * kHeadDim + ch * 8);
        *reinterpret_cast<uint4*>(qb + small_off(c, ch * 8)) = v;
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      mbar_arrive(&q_ready[active & 1]);
      ++active;
    }
  } else {
    // ===== softmax / output: CG column groups of 128 threads; thread = token (S lanes) and = head dim
    // (O lanes).  One warp per scheduler cannot hide its own latencies (the per-tile chain is ~100
    // dependent instructions per column), so the NC live columns are dealt to CG groups of four warps
    // (one per TMEM lane quadrant): the groups run the same tile concurrently on different columns and
    // share nothing but the P hand-over barrier. =====
    const int grp = (warp - 2) >> 2;
    const int quad = warp & 3;
    const int tau = quad * 32 + lane;   // TMEM lane
    constexpr int NCG = (NC + CG - 1) / CG;                                              // columns per group
    constexpr int NL = NCG <= 1 ? 1 : (NCG <= 2 ? 2 : (NCG <= 4 ? 4 : (NCG <= 8 ? 8 : 16)));   // tcgen05.ld width
    const int c0 = grp * NCG;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0;
    float* xmax_grp = xmax_all + grp * 128;   // [2 (tile parity)][4 quads][16]
    float* xsum_grp = xsum + grp * 64;        // [4 quads][16]
    const uint32_t bar_id = 1 + grp;
    const bool lead = tau == 0 && grp == 0;   // trace stamps
    uint32_t poff[NCG];                       // this thread's element of column c0+j in a P operand tile
    int col_row[NCG], col_head[NCG];
#pragma unroll
    for (int j = 0; j < NCG; ++j) {
      const int c = min(c0 + j, kNPad - 1);
      poff[j] = small_off(c, tau);
      col_row[j] = c / G;
      col_head[j] = c % G;
    }
    int n = 0, item_no = 0;
    Item nxt = first_item();
    for (int it = blockIdx.x; it < n_items; it += gridDim.x, ++item_no) {
      const Item w = nxt;
      if (it + (int)gridDim.x < n_items) nxt = get_item(it + gridDim.x);
      if (lead && item_no < 6) attn_trace(2 + 3 * item_no);
      float m_run[NCG], l_run[NCG], acc[NCG], alpha_prev[NCG];
      int col_ctx[NCG];   // context of the column's row; 0 for a dead column (past NC, or row absent)
#pragma unroll
      for (int j = 0; j < NCG; ++j) {
        m_run[j] = -INFINITY;
        l_run[j] = 0.f;
        acc[j] = 0.f;
        alpha_prev[j] = 1.f;
        col_ctx[j] = 0;
        // mask-select (not an indexed read: a dynamic index would push the whole Item onto the stack)
#pragma unroll
        for (int r = 0; r < R; ++r) col_ctx[j] |= w.ctx[r] & -(int)(col_row[j] == r && c0 + j < NC);
      }
      int done = 0;   // tiles of this item already handed to the MMA warp
#pragma unroll
      for (int u = 0; u < R; ++u) {
        if (u >= w.nrun) break;
        for (int i = 0; i < w.run_tiles[u]; ++i, ++n, ++done) {
          const int qv = 2 * n + 1, sv = qv % NSLOT;
          const int tok0 = (w.tile_begin + i) * kTileTok;
          const int n_valid = min(kTileTok, w.run_ctx[u] - tok0);
          mbar_wait(&s_full[n & 1], (n >> 1) & 1);
          if (done == 0 && lead && item_no < 6) attn_trace(3 + 3 * item_no);
          if (lead && n < 10) attn_trace(20 + 4 * n);       // S(n) landed
          tc_fence_after();
          uint32_t r[NL];
          tmem_ld<NL>(lane_addr + (n & 1) * 16, r);
          float sc[NCG];
          float* xmax = xmax_grp + (n & 1) * 64;   // double buffered: one group-wide barrier per tile is enough
#pragma unroll
          for (int j = 0; j < NCG; ++j) {
            // column (row col_row, head col_head): live in this run, and causal: token < that row's context
            const bool live = (w.nrun == 1 || col_row[j] == u) && (tok0 + tau < col_ctx[j]);
            sc[j] = live ? __uint_as_float(r[j]) * scale : -INFINITY;
          }
          // tile max per column over the 128 tokens
#pragma unroll
          for (int j = 0; j < NCG; ++j) {
            const float wm = warp_max_redux(sc[j]);
            if (lane == 0) xmax[quad * 16 + j] = wm;
          }
          asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
          if (lead && n < 10) attn_trace(21 + 4 * n);       // tile max exchanged
          float alpha[NCG];
          unsigned char* ph = p_hi + (n & 1) * 2 * kSmallTile;
          unsigned char* pl = p_lo + (n & 1) * 2 * kSmallTile;
#pragma unroll
          for (int j = 0; j < NCG; ++j) {
            const float tm = fmaxf(fmaxf(xmax[j], xmax[16 + j]), fmaxf(xmax[32 + j], xmax[48 + j]));
            const float m_new = fmaxf(m_run[j], tm);
            // a column with nothing seen so far (other run, or causally empty tile) keeps its state
            const bool seen = m_new > -INFINITY;
            alpha[j] = seen ? __expf(m_run[j] - m_new) : 1.f;   // exp(-inf) = 0 on the column's first tile
            m_run[j] = m_new;
            const float pv = seen ? __expf(sc[j] - m_new) : 0.f;   // 0 for masked tokens
            l_run[j] = l_run[j] * alpha[j] + pv;
            const float hi = round_bf16(pv);
            const float lo = pv - hi;
            if (c0 + j < kNPad) {
              *reinterpret_cast<__nv_bfloat16*>(ph + poff[j]) = __float2bfloat16_rn(hi);
              *reinterpret_cast<__nv_bfloat16*>(pl + poff[j]) = __float2bfloat16_rn(lo);
            }
          }
          if (n_valid < kTileTok && grp == 0) {
            // partial last tile: V rows past the context may hold anything (0 * NaN = NaN on the tensor
            // core) — wait for the tile, then zero this thread's row if it is past the end
            mbar_wait(&kv_full[sv], (qv / NSLOT) & 1);
            if (tau >= n_valid) {
              unsigned char* vb = kv + sv * kSlotBytes;
#pragma unroll
              for (int h2 = 0; h2 < 2; ++h2)
#pragma unroll
                for (int cc = 0; cc < 8; ++cc)
                  *reinterpret_cast<uint4*>(vb + h2 * kHalfBytes + tau * 128 + cc * 16) = make_uint4(0u, 0u, 0u, 0u);
            }
          }
          if (lead && n < 10) attn_trace(22 + 4 * n);       // P(n) stored
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic smem writes -> visible to the MMA
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&p_ready[n & 1]);
          if (lead && n < 10) attn_trace(23 + 4 * n);       // P(n) handed over
          // Fold the PREVIOUS tile's O_tile^T[dim = tau][column] into the running output (O is double
          // buffered in TMEM): the P.V MMAs of this tile run while the next tile's softmax is computed.
          if (done > 0) {
            mbar_wait(&o_full[(n - 1) & 1], ((n - 1) >> 1) & 1);
            tc_fence_after();
            tmem_ld<NL>(lane_addr + 32 + ((n - 1) & 1) * 16, r);
#pragma unroll
            for (int j = 0; j < NCG; ++j) acc[j] = acc[j] * alpha_prev[j] + __uint_as_float(r[j]);
          }
#pragma unroll
          for (int j = 0; j < NCG; ++j) alpha_prev[j] = alpha[j];
        }
      }
      if (done > 0) {   // the last tile's O
        uint32_t r[NL];
        mbar_wait(&o_full[(n - 1) & 1], ((n - 1) >> 1) & 1);
        tc_fence_after();
        tmem_ld<NL>(lane_addr + 32 + ((n - 1) & 1) * 16, r);
#pragma unroll
        for (int j = 0; j < NCG; ++j) acc[j] = acc[j] * alpha_prev[j] + __uint_as_float(r[j]);
      }
      // total l per column: sum the per-token partial sums over the 128 threads
#pragma unroll
      for (int j = 0; j < NCG; ++j) {
        const float ws = warp_sum(l_run[j]);
        if (lane == 0) xsum_grp[quad * 16 + j] = ws;
      }
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
      const int d = tau;
#pragma unroll
      for (int j = 0; j < NCG; ++j) {
        const int row = w.r0 + col_row[j];
        if (c0 + j < NC && row < n_rows) {
          const float l = xsum_grp[j] + xsum_grp[16 + j] + xsum_grp[32 + j] + xsum_grp[48 + j];
          const int head = w.kvh * G + col_head[j];
          if (n_splits == 1) {
            out[(int64_t)row * ldo + head * kHeadDim + d] = __float2bfloat16_rn(l > 0.f ? acc[j] / l : 0.f);
          } else {
            const int64_t pi = ((int64_t)row * hq + head) * n_splits + w.split;
            part_o[pi * kHeadDim + d] = acc[j];
            if (d == 0) {
              part_ml[pi * 2] = m_run[j];
              part_ml[pi * 2 + 1] = l;
            }
          }
        }
      }
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");  // xsum is rewritten by the next item
      if (lead && item_no < 6) attn_trace(4 + 3 * item_no);
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols));
  }
}

constexpr int attn_tc_smem(int nslot) { return nslot * kSlotBytes + (4 + 8) * kSmallTile + 4096 /* barriers, exchange */ + 1024 /* alignment */; }

template <int G, int R, int NSLOT, int CG>
int launch_attn_tc3(const CUtensorMap& mk, const CUtensorMap& mv, const __nv_bfloat16* q, int64_t ldq,
                    const int32_t* block_table, int bt_stride, const int32_t* row_to_seq, const int32_t* ctx_lens,
                    __nv_bfloat16* out, int64_t ldo, int T, int hkv, int block_size, int n_splits, int tps, float scale,
                    float* part_o, float* part_ml, int early, cudaStream_t s) {
  static bool attr_set[64] = {};
  B200_CUDA(ensure_dynamic_smem(paged_attn_tc_kernel<G, R, NSLOT, CG>, attn_tc_smem(NSLOT), attr_set));
  static const int trace_on = getenv("B200_ATTN_TRACE") ? atoi(getenv("B200_ATTN_TRACE")) : 0;
  if (trace_on) cudaMemcpyToSymbol(g_attn_trace_on, &trace_on, sizeof(int));
  const int64_t items = (int64_t)n_splits * hkv * ((T + R - 1) / R);
  // persistent: at most one CTA per SM, items round-robin.  Items cost about the same, so the kernel ends
  // when the CTAs with ceil(items/grid) items end.  Keep that round count, but spread the items evenly:
  // 512 items run as 128 CTAs x 4 (each CTA gets a larger share of HBM while all of them stream) rather
  // than 148 CTAs of which 68 take a fourth round while 80 idle.
  int grid = (int)(items < kNumSMs ? items : kNumSMs);
  if (items > kNumSMs) {
    const int64_t rounds = (items + kNumSMs - 1) / kNumSMs;
    grid = (int)((items + rounds - 1) / rounds);
  }
  B200_CUDA(launch(paged_attn_tc_kernel<G, R, NSLOT, CG>, dim3(grid), dim3(attn_threads(CG)), (size_t)attn_tc_smem(NSLOT), s,
                   mk, mv, q, ldq, block_table, bt_stride, row_to_seq, ctx_lens, out, ldo, hkv, block_size, tps, n_splits,
                   T, scale, part_o, part_ml, early));
  return B200_OK;
}

// B200_ATTN_CG=1: one softmax group (A/B against the column groups)
template <int G, int R, int NSLOT>
int launch_attn_tc2(const CUtensorMap& mk, const CUtensorMap& mv, const __nv_bfloat16* q, int64_t ldq,
                    const int32_t* block_table, int bt_stride, const int32_t* row_to_seq, const int32_t* ctx_lens,
                    __nv_bfloat16* out, int64_t ldo, int T, int hkv, int block_size, int n_splits, int tps, float scale,
                    float* part_o, float* part_ml, int early, cudaStream_t s) {
  static const int cg_env = getenv("B200_ATTN_CG") ? atoi(getenv("B200_ATTN_CG")) : 0;
  constexpr int CG = attn_col_groups(G * R);
  if (cg_env == 1 || CG == 1)
    return launch_attn_tc3<G, R, NSLOT, 1>(mk, mv, q, ldq, block_table, bt_stride, row_to_seq, ctx_lens, out, ldo, T, hkv,
                                           block_size, n_splits, tps, scale, part_o, part_ml, early, s);
  return launch_attn_tc3<G, R, NSLOT, CG>(mk, mv, q, ldq, block_table, bt_stride, row_to_seq, ctx_lens, out, ldo, T, hkv,
                                          block_size, n_splits, tps, scale, part_o, part_ml, early, s);
}

template <int G, int R>
int launch_attn_tc(int nslot, const CUtensorMap& mk, const CUtensorMap& mv, const __nv_bfloat16* q, int64_t ldq,
                   const int32_t* block_table, int bt_stride, const int32_t* row_to_seq, const int32_t* ctx_lens,
                   __nv_bfloat16* out, int64_t ldo, int T, int hkv, int block_size, int n_splits, int tps, float scale,
                   float* part_o, float* part_ml, int early, cudaStream_t s) {
  if (nslot == 3)
    return launch_attn_tc2<G, R, 3>(mk, mv, q, ldq, block_table, bt_stride, row_to_seq, ctx_lens, out, ldo, T, hkv,
                                    block_size, n_splits, tps, scale, part_o, part_ml, early, s);
  return launch_attn_tc2<G, R, 6>(mk, mv, q, ldq, block_table, bt_stride, row_to_seq, ctx_lens, out, ldo, T, hkv,
                                  block_size, n_splits, tps, scale, part_o, part_ml, early, s);
}

}  // namespace

int attn_tc_read_trace(unsigned long long* host64) {
  B200_CUDA(cudaMemcpyFromSymbol(host64, g_attn_trace, 64 * sizeof(unsigned long long)));
  return B200_OK;
}

bool paged_attn_tc_supported(int hq, int hkv, int block_size) {
  if (hkv <= 0 || hq % hkv) return false;
  const int G = hq / hkv;
  return block_size % kTileTok == 0 && (G == 1 || G == 2 || G == 4 || G == 5 || G == 8);
}

// rows per group when packing query rows of one sequence into the 16 MMA columns
int attn_tc_pack_rows(int G) { return G <= 4 ? 4 : (G == 5 ? 3 : 2); }

int paged_attn_tc(const __nv_bfloat16* q, int64_t ldq, const __nv_bfloat16* kc, const __nv_bfloat16* vc,
                  const int32_t* block_table, int bt_stride, const int32_t* row_to_seq, const int32_t* ctx_lens,
                  __nv_bfloat16* out, int64_t ldo, int T, int hq, int hkv, int block_size, int max_ctx, float scale,
                  void* ws, int64_t ws_bytes, int pack, cudaStream_t s) {
  B200_REQUIRE(paged_attn_tc_supported(hq, hkv, block_size), B200_ERR_UNSUPPORTED, "attn_tc: unsupported shape");
  B200_REQUIRE(ldq % 8 == 0, B200_ERR_INVALID, "attn_tc: ldq must be a multiple of 8");
  if (T == 0) return B200_OK;
  const int G = hq / hkv;
  const int R = pack > 0 ? attn_tc_pack_rows(G) : 1;
  static const int early_env = getenv("B200_ATTN_EARLY") ? atoi(getenv("B200_ATTN_EARLY")) : 1;
  const int early = (pack < 0 && early_env) ? 1 : 0;   // pack < 0: decode-only step (see the kernel)
  // 6 slots = 192 KiB in flight.  A 3-slot ring (B200_ATTN_SLOTS=3) would let this kernel share an SM with the
  // neighbouring linear's CTA under PDL, but measured: no gain in the step at 2 tiles per row and 19 % slower
  // in isolation at 5 tiles per row (64 x 640), so it is not the default.
  static const int nslot_env = getenv("B200_ATTN_SLOTS") ? atoi(getenv("B200_ATTN_SLOTS")) : 0;
  const int nslot = nslot_env == 3 ? 3 : 6;
  (void)max_ctx;
  int n_splits, tps;
  attn_split_plan_tc((T + R - 1) / R, hkv, max_ctx, &n_splits, &tps);
  float* part_o = nullptr;
  float* part_ml = nullptr;
  if (n_splits > 1) {
    const int64_t need = (int64_t)T * hq * n_splits * (kHeadDim + 2) * sizeof(float);
    B200_REQUIRE(ws && ws_bytes >= need, B200_ERR_INVALID, "attn_tc: workspace too small (%lld < %lld)",
                 (long long)ws_bytes, (long long)need);
    part_o = static_cast<float*>(ws);
    part_ml = part_o + (int64_t)T * hq * n_splits * kHeadDim;
  }
  // The cache is addressed as rows of 128 dims; the row bound is only an upper limit for TMA's
  // out-of-bounds logic (rows actually touched come from the block table).
  CUtensorMap mk, mv;
  const uint64_t rows = 1ull << 31;
  int rc = get_tensor_map_2d(kc, rows, kHeadDim, kHeadDim, kTileTok, &mk);
  if (rc) return rc;
  rc = get_tensor_map_2d(vc, rows, kHeadDim, kHeadDim, kTileTok, &mv);
  if (rc) return rc;
#define B200_ATTN_TC_CASE(g, r)                                                                                        \
  rc = launch_attn_tc<g, r>(nslot, mk, mv, q, ldq, block_table, bt_stride, row_to_seq, ctx_lens, out, ldo, T, hkv, block_size, \
                            n_splits, tps, scale, part_o, part_ml, early, s)
  if (pack <= 0) {
    switch (G) {
      case 1: B200_ATTN_TC_CASE(1, 1); break;
      case 2: B200_ATTN_TC_CASE(2, 1); break;
      case 4: B200_ATTN_TC_CASE(4, 1); break;
      case 5: B200_ATTN_TC_CASE(5, 1); break;
      case 8: B200_ATTN_TC_CASE(8, 1); break;
      default: rc = B200_ERR_UNSUPPORTED;
    }
  } else {
    switch (G) {
      case 1: B200_ATTN_TC_CASE(1, 4); break;
      case 2: B200_ATTN_TC_CASE(2, 4); break;
      case 4: B200_ATTN_TC_CASE(4, 4); break;
      case 5: B200_ATTN_TC_CASE(5, 3); break;
      case 8: B200_ATTN_TC_CASE(8, 2); break;
      default: rc = B200_ERR_UNSUPPORTED;
    }
  }
#undef B200_ATTN_TC_CASE
  if (rc) return rc;
  if (n_splits > 1) return launch_attn_merge(part_o, part_ml, out, ldo, T, hq, n_splits, s);
  return B200_OK;
}

}  // namespace b200
