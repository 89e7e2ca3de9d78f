// tcgen05 / TMEM weight-streaming linear: y[T][n_out] = x[T][k] . w[n_out][k]^T (+bias), T >= 5.
//
// Decode GEMMs are skinny (T = 8..256 token rows against n_out = 4096..128256 weight rows), so the
// operands are SWAPPED: a 128-row slab of W is the UMMA "A" operand (M = 128, fills the tensor
// core's M), the token rows are "B" (N = 16..256), and the fp32 accumulator D[128 features][N
// tokens] lives in TMEM (lane = feature, column = token).  Bound: HBM — algorithmic bytes per
// launch = n_out*k*2, each weight element crosses HBM once; the tensor pipe needs
// max(128,M)*N/256 cycles per K=16 step = 32 cycles for N = 64 against 4 KiB of W, i.e. it can
// absorb 128 B/clk/SM where HBM delivers ~23 B/clk/SM, so tensor utilisation is expected to be
// low (~15-20 %) and HBM GB/s is the figure of merit.
//
// Pipeline per CTA (192 threads): warp 0 = TMA producer (cp.async.bulk.tensor, SWIZZLE_128B,
// W with an evict-first L2 policy, x evict-last), warp 1 = TMEM allocator + single-thread
// tcgen05.mma issuer (tcgen05.commit releases smem stages and signals the epilogue), warps 2-5 =
// epilogue (tcgen05.ld 32x32b -> bias -> bf16 -> global).  A CTA owns (128-row slab, token tile,
// K split).  Split-K is deterministic: partials go to an fp32 workspace, the LAST CTA of a slab
// (atomic ticket) sums them in split order, so results do not depend on scheduling.
// W tiles of the first stages are requested BEFORE griddepcontrol.wait (they do not depend on the
// previous kernel) so that under programmatic dependent launch the weight stream never drains.
//
// Replaces: sgl-kernel weight_packed_linear (AMX) / oneDNN bf16 matmul of the vLLM-CPU image
// ([V22] vllm/model_executor/layers/utils.py:277-299); rounding R2 of oracle/decode_oracle.py.
#include <cuda.h>
#include <stdlib.h>

#include <mutex>
#include <unordered_map>

#include "common.cuh"

namespace b200 {

namespace {

constexpr int kBlockM = 128;   // weight rows per CTA = UMMA M
constexpr int kBlockK = 64;    // bf16 elements per stage row = 128 bytes = one swizzle span
constexpr int kUmmaK = 16;
constexpr int kTcThreads = 192;
constexpr int kCounterBytes = 64 * 1024;  // slab tickets at the head of the workspace (16384 slabs x token tiles)
constexpr uint64_t kEvictFirst = 0x12F0000000000000ull;
constexpr uint64_t kEvictLast = 0x14F0000000000000ull;
constexpr uint64_t kEvictNormal = 0x1000000000000000ull;

#ifndef B200_STAGES16
#define B200_STAGES16 5
#endif
#ifndef B200_STAGES64
#define B200_STAGES64 4
#endif
// N = 256 (prefill-sized token tiles): 2 stages of 48 KiB so that TWO CTAs share an SM (2 x 256 TMEM columns):
// one CTA's epilogue then overlaps the other's main loop; 4 stages are in flight per SM as before.
#ifndef B200_STAGES256
#define B200_STAGES256 2
#endif
__host__ __device__ constexpr int stages_for(int N) { return N <= 16 ? B200_STAGES16 : (N <= 32 ? 4 : (N <= 64 ? B200_STAGES64 : (N <= 128 ? 3 : B200_STAGES256))); }
__host__ __device__ constexpr int stage_bytes(int N) { return kBlockM * kBlockK * 2 + N * kBlockK * 2; }
constexpr int kTailBytes = 1024;   // barriers, TMEM slot, ticket | per-token 1/rms [64] (fused-norm consumer)
// Decode-sized tiles (N <= 32) exchange their cluster split-K partials by PUSH: every CTA stores the tokens owner r
// finalises straight into r's receive buffer [token][source rank][128 rows] (fp32) with st.async, whose transaction
// bytes complete r's mbarrier: no cluster barrier and no remote-load round trip between the last MMA and the output.
__host__ __device__ constexpr int recv_bytes(int N) { return N <= 32 ? (N + 8) * kBlockM * 4 : 0; }

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
// mbarrier wait that also acquires what peers of the cluster released before arriving (remote arrivals)
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAITC_LOOP:\n"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n"
      "@p bra.uni WAITC_DONE;\n"
      "bra.uni WAITC_LOOP;\n"
      "WAITC_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ uint32_t map_to_rank(uint32_t local_smem_addr, int rank) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_smem_addr), "r"(rank));
  return remote;
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar,
                                            uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc], bf16 x bf16 -> fp32, issued by ONE thread
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
// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor layout):
// start>>4 [0,14) | LBO>>4 [16,30) (ignored for swizzled K-major, 1) | SBO>>4 [32,46) = 1024 B between
// 8-row groups | version=1 [46,48) | layout_type=2 (SWIZZLE_128B) [61,64)
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3fff);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// cute::UMMA::InstrDescriptor: c_format F32 (1)<<4 | a_format BF16 (1)<<7 | b_format BF16 (1)<<10 |
// a_major K (0)<<15 | b_major K (0)<<16 | N>>3 <<17 | M>>4 <<24
__host__ __device__ constexpr uint32_t make_idesc(int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(kBlockM >> 4) << 24);
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

struct TcParams {
  const __nv_bfloat16* bias;
  __nv_bfloat16* y;
  int64_t ldy;
  int T, n_out, n_kblocks, splits;
  float* partials;      // [splits][tok_tiles*N][n_out_pad] fp32
  unsigned int* tickets;  // [m_tiles * tok_tiles], zero on entry, zero on exit
  int n_out_pad;
  int cluster;  // 1: the `splits` CTAs of a slab form a thread-block cluster and reduce through DSMEM
  // MODE_ROPE_KV (slab == one 128-wide head of the fused qkv projection)
  const int32_t* positions;
  const int32_t* slots;
  const __nv_bfloat16* cos_sin;
  __nv_bfloat16* kc;
  __nv_bfloat16* vc;
  int hq, hkv, block_size;
  unsigned long long* trace;  // debug (B200_TC_TRACE=1): %globaltimer stamps of CTA 0, else null
  // Residual stream in fp32 (RMSNorm folded into the linears, T <= 64):
  //   producer (MODE kResidual, O / down projection): s = bf16(acc) + bf16(sres) in fp32 -> sres, and the sum of
  //     squares of the slab's 128 values of s -> ssq[tok][m_tile]              (n_out = hidden)
  //   consumer (NI, QKV / gate_up / lm_head): the B operand is bf16(bf16(s * rinv) * norm_w) built in shared memory
  //     from sres, rinv = 1/sqrt(sum_slabs ssq / hidden + eps)                  (k = hidden)
  float* sres;
  float* ssq;
  const __nv_bfloat16* norm_w;
  const int32_t* rows;   // consumer: token t of the launch reads stream row rows[t] (lm_head's logit rows), or null
  int nslab;             // hidden / 128
  float eps;
};

__device__ __forceinline__ void tc_trace(const TcParams& p, int slot) {
#ifndef B200_TRACE_BUILD
  (void)p;
  (void)slot;
  return;
#endif
  if (p.trace && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    p.trace[slot] = t;
  }
}

// Epilogue modes (template parameter):
//  kPlain    y[t][row] = bf16(acc + bias)
//  kSiluMul  the slab holds 64 gate rows then the 64 matching up rows (B200_W_GATE_UP_INTERLEAVED):
//            y[t][m_tile*64 + i] = bf16(bf16(silu(bf16(g_i))) * bf16(u_i))          (R5)
//  kRopeKv   the slab is one head of the fused qkv projection: NeoX RoPE in bf16 arithmetic (R3) on q
//            and k heads, q rows to y, k/v rows straight into the paged cache at slots[t]
//  kResidual the linear's output is added to the fp32 residual stream (see TcParams::sres): no y is written
enum { kPlain = 0, kSiluMul = 1, kRopeKv = 2, kResidual = 3 };


// One thread finalises 4 consecutive slab rows (r4..r4+3, lane = r4/4) of one token.
template <int MODE>
__device__ __forceinline__ void finalize4(const TcParams& p, int m_tile, int lane, int tok, float4 a, const float (&bb)[4]) {
  const int r4 = lane * 4;
  float v[4] = {round_bf16(a.x + bb[0]), round_bf16(a.y + bb[1]), round_bf16(a.z + bb[2]), round_bf16(a.w + bb[3])};
  if (MODE == kPlain) {
    const int row = m_tile * kBlockM + r4;
    __nv_bfloat16* dst = p.y + (int64_t)tok * p.ldy + row;
    if (row + 3 < p.n_out && (p.ldy % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.y) & 7) == 0)) {
      *reinterpret_cast<uint2*>(dst) = make_uint2(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]));
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if (row + e < p.n_out) dst[e] = __float2bfloat16_rn(v[e]);
    }
  } else if (MODE == kSiluMul) {
    // lanes 0..15 hold 4 gate rows, lanes 16..31 the 4 matching up rows: after the exchange every lane has
    // both, and each half-warp finalises two of the four features (all 32 lanes busy: the accurate
    // expf + division of R5 is what this epilogue costs at large T)
    float x[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) x[e] = __shfl_xor_sync(0xffffffffu, v[e], 16);
    const bool lo = lane < 16;
    const float g0 = lo ? v[0] : x[2], g1 = lo ? v[1] : x[3];
    const float u0 = lo ? x[0] : v[2], u1 = lo ? x[1] : v[3];
    const int f = m_tile * 64 + (lane & 15) * 4 + (lo ? 0 : 2);                  // output feature
    const float o0 = silu_to_bf16(g0) * u0, o1 = silu_to_bf16(g1) * u1;
    __nv_bfloat16* dst = p.y + (int64_t)tok * p.ldy + f;
    if ((p.ldy % 2 == 0) && ((reinterpret_cast<uintptr_t>(p.y) & 3) == 0)) {
      *reinterpret_cast<uint32_t*>(dst) = pack_bf16(o0, o1);
    } else {
      dst[0] = __float2bfloat16_rn(o0);
      dst[1] = __float2bfloat16_rn(o1);
    }
  } else if (MODE == kResidual) {
    // x = bf16(acc + bias); s = x + bf16(s_old) in fp32 (R1: the residual is kept in bf16, the sum is not rounded
    // before the variance and the normalisation); sum of squares in a fixed order: lane's 4 rows, then the warp tree
    const int row = m_tile * kBlockM + r4;
    float4* sp = reinterpret_cast<float4*>(p.sres + (int64_t)tok * p.n_out + row);
    const float4 so = *sp;
    const float4 sn = make_float4(v[0] + round_bf16(so.x), v[1] + round_bf16(so.y), v[2] + round_bf16(so.z),
                                  v[3] + round_bf16(so.w));
    *sp = sn;
    float ss = fmaf(sn.x, sn.x, 0.f);
    ss = fmaf(sn.y, sn.y, ss);
    ss = fmaf(sn.z, sn.z, ss);
    ss = fmaf(sn.w, sn.w, ss);
    ss = warp_sum(ss);
    if (lane == 0) p.ssq[(int64_t)tok * p.nslab + m_tile] = ss;
  } else {  // kRopeKv
    const int head = m_tile;
    float o[4];
    if (head < p.hq + p.hkv) {
      float other[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) other[e] = __shfl_xor_sync(0xffffffffu, v[e], 16);
      const int i0 = r4 & 63;
      const __nv_bfloat16* cs = p.cos_sin + (int64_t)p.positions[tok] * kHeadDim;
      const uint2 cq = *reinterpret_cast<const uint2*>(cs + i0);
      const uint2 sq = *reinterpret_cast<const uint2*>(cs + 64 + i0);
      const float c[4] = {bf16lo(cq.x), bf16hi(cq.x), bf16lo(cq.y), bf16hi(cq.y)};
      const float sn[4] = {bf16lo(sq.x), bf16hi(sq.x), bf16lo(sq.y), bf16hi(sq.y)};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        // lanes 0..15 hold x1 (o1 = x1*c - x2*s), lanes 16..31 hold x2 (o2 = x2*c + x1*s); every op rounds
        const float a0 = round_bf16(v[e] * c[e]), b0 = round_bf16(other[e] * sn[e]);
        o[e] = lane < 16 ? round_bf16(a0 - b0) : round_bf16(a0 + b0);
      }
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e) o[e] = v[e];
    }
    __nv_bfloat16* dst;
    if (head < p.hq) {
      dst = p.y + (int64_t)tok * p.ldy + head * kHeadDim + r4;
    } else {
      const int slot = p.slots[tok];
      if (slot < 0) return;
      const bool is_k = head < p.hq + p.hkv;
      const int kvh = is_k ? head - p.hq : head - p.hq - p.hkv;
      const int64_t blk = slot / p.block_size, off = slot % p.block_size;
      dst = (is_k ? p.kc : p.vc) + ((blk * p.hkv + kvh) * p.block_size + off) * kHeadDim + r4;
    }
    *reinterpret_cast<uint2*>(dst) = make_uint2(pack_bf16(o[0], o[1]), pack_bf16(o[2], o[3]));
  }
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// Fused-norm consumer: the four epilogue warps build the B operand (the normalised token rows) in the ring while the
// TMA warp streams W.  Warp wq owns the K blocks i = wq (mod 4): it reads the fp32 stream rows (L2), applies R1
//   n = bf16(s * rinv), y = bf16(n * w),   rinv = 1/sqrt(sum over slabs of ssq / hidden + eps) (slabs in ascending order)
// and stores the bf16 row chunks where a SWIZZLE_128B TMA box would have put them (16-byte chunk c of row t at
// t*128 + ((c ^ (t & 7)) << 4)); rows >= T keep whatever the ring held (their accumulator columns are never read).
template <int N, int S>
__device__ __forceinline__ void norm_b_tiles(const TcParams& p, unsigned char* smem, uint64_t* full, uint64_t* empty,
                                             float* rinv_s, int tok0, int kb0, int nkb, int wq, int lane) {
  constexpr int kStageBytes = stage_bytes(N);
  constexpr int kABytes = kBlockM * kBlockK * 2;
  constexpr int U = 4;
  const int te = wq * 32 + lane;
  const int tmax = min(N, p.T - tok0);
  const int hidden = p.nslab * kBlockM;
  const int ntask = tmax * 8;   // (token row, 16-byte chunk of the 64-wide K block)
  float4 a[U][2];
  uint4 wv[U];
  auto load = [&](int i, int task0) {
    const int kcol = (kb0 + i) * kBlockK;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int task = task0 + 32 * u;
      if (task < ntask) {
        const int t = task >> 3, c = task & 7;
        const int row = p.rows ? __ldg(p.rows + tok0 + t) : tok0 + t;
        const float4* src = reinterpret_cast<const float4*>(p.sres + (int64_t)row * hidden + kcol + c * 8);
        a[u][0] = __ldcg(src);
        a[u][1] = __ldcg(src + 1);
        wv[u] = __ldg(reinterpret_cast<const uint4*>(p.norm_w + kcol + c * 8));
      }
    }
  };
  // the first tasks' stream rows are requested BEFORE the variance is summed: the two L2 round trips (ssq, sres)
  // that follow griddepcontrol.wait overlap instead of following each other
  if (wq < nkb) load(wq, lane);
  if (te < tmax) {
    const int row = p.rows ? __ldg(p.rows + tok0 + te) : tok0 + te;
    const float* q = p.ssq + (int64_t)row * p.nslab;
    float tot = 0.f;
#pragma unroll 8
    for (int i = 0; i < p.nslab; ++i) tot += __ldcg(q + i);
    rinv_s[te] = 1.0f / sqrtf(tot / (float)hidden + p.eps);
  }
  asm volatile("bar.sync 1, 128;" ::: "memory");
  for (int i = wq; i < nkb; i += 4) {
    const int st = i % S;
    const uint32_t bbase = smem_u32(smem + st * kStageBytes + kABytes);
    bool need_wait = i >= S;
    for (int task0 = lane; task0 < ntask; task0 += 32 * U) {
      if (i != wq || task0 != lane) load(i, task0);
      if (need_wait) {   // the MMAs that read this stage's previous tile have completed
        mbar_wait(&empty[st], ((i / S) - 1) & 1);
        need_wait = false;
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int task = task0 + 32 * u;
        if (task < ntask) {
          const int t = task >> 3, c = task & 7;
          const float r = rinv_s[t];
          const float sv[8] = {a[u][0].x, a[u][0].y, a[u][0].z, a[u][0].w, a[u][1].x, a[u][1].y, a[u][1].z, a[u][1].w};
          const uint32_t ww[4] = {wv[u].x, wv[u].y, wv[u].z, wv[u].w};
          uint32_t o[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float n0 = round_bf16(sv[2 * j] * r), n1 = round_bf16(sv[2 * j + 1] * r);
            o[j] = pack_bf16(n0 * bf16lo(ww[j]), n1 * bf16hi(ww[j]));
          }
          const uint32_t dst = bbase + (uint32_t)t * 128u + (uint32_t)((c ^ (t & 7)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3]) : "memory");
        }
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> visible to tcgen05.mma
    __syncwarp();
    if (lane == 0) mbar_arrive(&full[st]);
  }
}

template <int N, int MODE, int S, bool NI>
__global__ void __launch_bounds__(kTcThreads)
linear_tc_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_x, TcParams p) {
  static_assert(!NI || N <= 64, "the fused-norm consumer serves decode-sized token tiles");
  constexpr int kStageBytes = stage_bytes(N);
  constexpr uint32_t kTmaBytes = NI ? kBlockM * kBlockK * 2 : kStageBytes;   // NI: only W arrives by TMA
  constexpr int kABytes = kBlockM * kBlockK * 2;
  constexpr int kTmemCols = N < 32 ? 32 : N;
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + S * kStageBytes);
  uint64_t* empty = full + S;
  uint64_t* tmem_full = empty + S;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
  int* ticket_slot = reinterpret_cast<int*>(tmem_slot + 1);
  float* rinv_s = reinterpret_cast<float*>(smem + S * kStageBytes + 256);   // [64]
  constexpr bool kPush = recv_bytes(N) > 0;
  uint64_t* recv_bar = reinterpret_cast<uint64_t*>(smem + S * kStageBytes + 512);
  float* recv = reinterpret_cast<float*>(smem + S * kStageBytes + kTailBytes);   // [token][source rank][128]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // Several token tiles (prefill): the token tile is the FAST index, so that the CTAs that share a weight slab
  // run in the same wave (the slab comes from HBM once, then from L2; x stays in L2 throughout).
  const int lin = blockIdx.x + gridDim.x * blockIdx.z;
  const int m_tile = gridDim.z > 1 ? lin / (int)gridDim.z : (int)blockIdx.x, split = blockIdx.y;
  const int tok_tile = gridDim.z > 1 ? lin % (int)gridDim.z : 0;
  const uint64_t w_policy = gridDim.z > 1 ? kEvictNormal : kEvictFirst;   // a slab read once is evicted first
  if (threadIdx.x == 0) tc_trace(p, 0);   // CTA start
  // K range of this split (balanced: the first `rem` splits get one more block)
  const int base = p.n_kblocks / p.splits, rem = p.n_kblocks % p.splits;
  const int kb0 = split * base + min(split, rem);
  const int nkb = base + (split < rem ? 1 : 0);

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(&full[s], NI ? 2 : 1);   // NI: the TMA producer and the warp that wrote the stage's B tile
      mbar_init(&empty[s], 1);
    }
    mbar_init(tmem_full, 1);
    if (kPush && p.cluster) {
      // the peers' st.async stores complete transaction bytes on this barrier: S2 partials of every token this CTA
      // finalises (tokens split, split + S2, ... of the tile), 128 fp32 rows each
      const int tmax = min(N, p.T - tok_tile * N);
      const int owned = split < tmax ? (tmax - split + p.splits - 1) / p.splits : 0;
      mbar_init(recv_bar, 1);
      if (owned) mbar_expect_tx(recv_bar, (uint32_t)(owned * p.splits * kBlockM * 4));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(kTmemCols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const bool push = kPush && p.cluster;
  // push exchange: a peer may store into this CTA's receive buffer and arrive on its barrier only once the barrier
  // exists -- every thread arrives here and waits (long since satisfied) before its first remote access / at the end
  if (push && warp != 0) asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");   // warp 0: after its first loads

  if (warp == 0) {
    // ===== TMA producer (whole warp in the loop, the TMA instructions predicated to one lane: see the MMA issuer) =====
    const bool issuer = elect_one();
    const int pre = min(S, nkb);
    if (issuer) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
      // weights first: they do not depend on the producer of x
      for (int i = 0; i < pre; ++i) {
        mbar_expect_tx(&full[i], kTmaBytes);
        tma_load_2d(smem + i * kStageBytes, &map_w, (kb0 + i) * kBlockK, m_tile * kBlockM, &full[i], w_policy);
      }
      tc_trace(p, 1);   // W of the first stages requested
    }
    __syncwarp();
    if (push) asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    pdl_wait();
    if (issuer) {
      tc_trace(p, 2);   // previous kernel complete
      pdl_launch_dependents();
      if constexpr (!NI) {
        for (int i = 0; i < pre; ++i)
          tma_load_2d(smem + i * kStageBytes + kABytes, &map_x, (kb0 + i) * kBlockK, tok_tile * N, &full[i], kEvictLast);
      }
    }
    __syncwarp();
    for (int i = pre; i < nkb; ++i) {
      const int st = i % S;
      mbar_wait(&empty[st], ((i / S) - 1) & 1);
      if (issuer) {
        mbar_expect_tx(&full[st], kTmaBytes);
        tma_load_2d(smem + st * kStageBytes, &map_w, (kb0 + i) * kBlockK, m_tile * kBlockM, &full[st], w_policy);
        if constexpr (!NI)
          tma_load_2d(smem + st * kStageBytes + kABytes, &map_x, (kb0 + i) * kBlockK, tok_tile * N, &full[st], kEvictLast);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the whole warp walks the loop and waits; only the tcgen05 instructions are predicated to
    // one lane, so that their operands (descriptors, TMEM address) are warp-uniform by construction and stay in
    // uniform registers (a single-lane branch pays a register -> uniform-register broadcast loop per MMA) =====
    pdl_wait();
    {
      constexpr uint32_t idesc = make_idesc(N);
      const bool issuer = elect_one();
      const uint64_t desc0 = make_desc(smem_u32(smem));
      for (int i = 0; i < nkb; ++i) {
        const int st = i % S;
        mbar_wait(&full[st], (i / S) & 1);
        if (i == 0 && issuer) tc_trace(p, 3);   // first stage landed
        tc_fence_after();
        const uint64_t da = desc0 + (uint64_t)((st * kStageBytes) >> 4);
        const uint64_t db = da + (uint64_t)(kABytes >> 4);
        if (issuer) {
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k) {
            // advancing K inside the 128-byte swizzle span = +32 bytes on the start address
            tc_mma(tmem_base, da + (uint64_t)((k * kUmmaK * 2) >> 4), db + (uint64_t)((k * kUmmaK * 2) >> 4), idesc,
                   (i | k) ? 1u : 0u);
          }
          tc_commit(&empty[st]);  // arrives when the MMAs above have finished reading this stage
        }
        __syncwarp();
      }
      if (issuer) {
        tc_trace(p, 4);           // last MMA issued
        tc_commit(tmem_full);     // accumulator complete
      }
    }
  } else {
    // ===== epilogue: 4 warps, TMEM lane quadrant = warp % 4 =====
    const int quad = warp & 3;
    const int srow = quad * 32 + lane;                      // slab row owned while draining TMEM
    const int tok0 = tok_tile * N;
    const int wq = warp - 2;
    pdl_wait();
    if constexpr (NI) norm_b_tiles<N, S>(p, smem, full, empty, rinv_s, tok0, kb0, nkb, wq, lane);
    if (nkb > 0) {
      mbar_wait(tmem_full, 0);
      tc_fence_after();
    }
    if (warp == 2 && lane == 0) tc_trace(p, 5);   // accumulator complete (epilogue starts)
    const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16);
    const bool direct = p.splits == 1 || p.cluster;   // the fp32 tile is staged in this CTA's smem
    // All MMAs have completed (tmem_full), so the pipeline stages are free: stage the fp32 tile
    // [token][128 rows] there for the direct case, or spill it to the split-K workspace.
    float* stg = reinterpret_cast<float*>(smem);
    constexpr bool kChunked = S * kStageBytes < N * kBlockM * 4;   // N = 256 with 2 stages: 128 tokens at a time
    if constexpr (kChunked) {
      // (host side: this instantiation is only launched with splits == 1 and no cluster)
      constexpr int kEpiTok = 128;
      static_assert(kEpiTok * kBlockM * 4 <= S * kStageBytes, "staging area");
      const int r4 = lane * 4;
      float bb[4] = {0.f, 0.f, 0.f, 0.f};
      if (p.bias) {
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (m_tile * kBlockM + r4 + e < p.n_out) bb[e] = __bfloat162float(p.bias[m_tile * kBlockM + r4 + e]);
      }
      for (int cb = 0; cb < N && tok0 + cb < p.T; cb += kEpiTok) {
#pragma unroll
        for (int c = 0; c < kEpiTok; c += 16) {
          uint32_t r[16];
          if (nkb > 0) {
            tmem_ld16(taddr + cb + c, r);
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) r[j] = 0u;
          }
#pragma unroll
          for (int j = 0; j < 16; ++j) stg[(c + j) * kBlockM + srow] = __uint_as_float(r[j]);
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        const int tmax = min(kEpiTok, p.T - tok0 - cb);
        if (MODE == kRopeKv) {
          for (int t = wq + 4 * lane; t < tmax; t += 128) {
            const char* row = reinterpret_cast<const char*>(p.cos_sin + (int64_t)p.positions[tok0 + cb + t] * kHeadDim);
            asm volatile("prefetch.global.L1 [%0];" ::"l"(row));
            asm volatile("prefetch.global.L1 [%0];" ::"l"(row + 128));
          }
        }
        for (int t = wq; t < tmax; t += 16) {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int tt = t + 4 * u;
            if (tt < tmax)   // warp-uniform: the shuffles inside finalize4 stay convergent
              finalize4<MODE>(p, m_tile, lane, tok0 + cb + tt, *reinterpret_cast<const float4*>(stg + tt * kBlockM + r4), bb);
          }
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");   // the staging area is reused by the next 128 tokens
      }
    } else {
#pragma unroll
    for (int c = 0; c < N; c += 16) {
      uint32_t r[16];
      if (nkb > 0) {
        tmem_ld16(taddr + c, r);
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) r[j] = 0u;
      }
      if (push) {
        if constexpr (kPush) {
          // token t of the tile goes to CTA t % splits, which finalises it: slot [t / splits][this rank][row]
          if (c == 0) asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");   // the peers' barriers exist
          const int tmax = min(N, p.T - tok0);
          const uint32_t recv_addr = smem_u32(recv), recv_bar_addr = smem_u32(recv_bar);
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int t = c + j;
            if (t < tmax) {
              const int owner = t % p.splits, tl = t / p.splits;
              const uint32_t dst = map_to_rank(recv_addr + (uint32_t)(((tl * p.splits + split) * kBlockM + srow) * 4), owner);
              const uint32_t bar = map_to_rank(recv_bar_addr, owner);
              asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.u32 [%0], %1, [%2];" ::"r"(dst),
                           "r"(r[j]), "r"(bar)
                           : "memory");
            }
          }
        }
      } else if (direct) {
#pragma unroll
        for (int j = 0; j < 16; ++j) stg[(c + j) * kBlockM + srow] = __uint_as_float(r[j]);
      } else {
        float* dst = p.partials + ((int64_t)split * gridDim.z * N + tok0 + c) * p.n_out_pad + m_tile * kBlockM + srow;
#pragma unroll
        for (int j = 0; j < 16; ++j) dst[(int64_t)j * p.n_out_pad] = __uint_as_float(r[j]);
      }
    }
    if (warp == 2 && lane == 0) tc_trace(p, 6);   // tile staged
    bool finalize = true;
    const int tile_id = m_tile * gridDim.z + tok_tile;
    if (push) {
      if constexpr (kPush) {
        finalize = false;
        const int tmax = min(N, p.T - tok0), S2 = p.splits, r4 = lane * 4;
        if (threadIdx.x == 64) tc_trace(p, 7);    // partials pushed (each store signals the owner's barrier itself)
        if (split < tmax) {
          // this CTA finalises tokens split, split + S2, ...: all S2 partials of them are in its own shared memory once
          // their transaction bytes have arrived; summed in rank order (deterministic, same order as the pull path)
          mbar_wait_cluster(recv_bar, 0);
          if (threadIdx.x == 64) tc_trace(p, 10);   // every partial of this CTA's tokens has landed
          float bb[4] = {0.f, 0.f, 0.f, 0.f};
          if (p.bias) {
#pragma unroll
            for (int e = 0; e < 4; ++e)
              if (m_tile * kBlockM + r4 + e < p.n_out) bb[e] = __bfloat162float(p.bias[m_tile * kBlockM + r4 + e]);
          }
          for (int tl = wq; split + S2 * tl < tmax; tl += 4) {
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int src = 0; src < S2; ++src) {
              const float4 v = *reinterpret_cast<const float4*>(recv + (tl * S2 + src) * kBlockM + r4);
              acc.x += v.x;
              acc.y += v.y;
              acc.z += v.z;
              acc.w += v.w;
            }
            finalize4<MODE>(p, m_tile, lane, tok0 + split + S2 * tl, acc, bb);   // warp-uniform loop
          }
        }
        if (threadIdx.x == 64) tc_trace(p, 11);   // this CTA's share finalized
      }
    } else if (p.cluster) {
      finalize = false;   // done after the cluster barrier below, by every CTA of the cluster for its share
    } else if (direct) {
      asm volatile("bar.sync 1, 128;" ::: "memory");
    } else {
      // deterministic split-K through global memory: the last CTA of this (slab, token tile) reduces
      __threadfence();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (warp == 2 && lane == 0) *ticket_slot = (int)atomicAdd(&p.tickets[tile_id], 1u);
      asm volatile("bar.sync 1, 128;" ::: "memory");
      finalize = (*ticket_slot == p.splits - 1);
      if (finalize) __threadfence();
    }
    if (finalize) {
      // 4 warps x 32 lanes: lane owns 4 consecutive slab rows, warp wq owns tokens wq, wq+4, ...
      const int r4 = lane * 4;
      const int tmax = min(N, p.T - tok0);
      float bb[4] = {0.f, 0.f, 0.f, 0.f};
      if (p.bias) {
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (m_tile * kBlockM + r4 + e < p.n_out) bb[e] = __bfloat162float(p.bias[m_tile * kBlockM + r4 + e]);
      }
      const int64_t split_stride = (int64_t)gridDim.z * N * p.n_out_pad;
      if (MODE == kRopeKv && tmax > 16) {
        // prefill-sized tiles: the cos/sin row of a token is a dependent load (position -> row); request
        // the rows of all of this warp's tokens now so that the loop below finds them in L1/L2
        for (int t = wq + 4 * lane; t < tmax; t += 128) {
          const char* row = reinterpret_cast<const char*>(p.cos_sin + (int64_t)p.positions[tok0 + t] * kHeadDim);
          asm volatile("prefetch.global.L1 [%0];" ::"l"(row));
          asm volatile("prefetch.global.L1 [%0];" ::"l"(row + 128));
        }
      }
      for (int t = wq; t < tmax; t += 16) {
        float4 acc[4];
        if (p.splits == 1) {
#pragma unroll
          for (int u = 0; u < 4; ++u)
            acc[u] = (t + 4 * u < tmax) ? *reinterpret_cast<const float4*>(stg + (t + 4 * u) * kBlockM + r4)
                                        : make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
#pragma unroll
          for (int u = 0; u < 4; ++u) acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
          const float* src = p.partials + (int64_t)(tok0 + t) * p.n_out_pad + m_tile * kBlockM + r4;
#pragma unroll 3
          for (int sp = 0; sp < p.splits; ++sp) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              if (t + 4 * u < tmax) {
                const float4 v = __ldcg(reinterpret_cast<const float4*>(src + sp * split_stride + (int64_t)(4 * u) * p.n_out_pad));
                acc[u].x += v.x;
                acc[u].y += v.y;
                acc[u].z += v.z;
                acc[u].w += v.w;
              }
            }
          }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int tt = t + 4 * u;
          // warp-uniform condition: the shuffles inside finalize4 stay convergent
          if (tt < tmax) finalize4<MODE>(p, m_tile, lane, tok0 + tt, acc[u], bb);
        }
      }
      if (p.splits > 1 && warp == 2 && lane == 0) p.tickets[tile_id] = 0u;  // leave the tickets zero for the next launch
    }
    }   // !kChunked
    tc_fence_before();
  }
  if (push) {
    // the other warps still owe the wait of the start-up cluster barrier (the epilogue warps waited before pushing)
    if (warp < 2) asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  } else if (threadIdx.x == 64) {
    tc_trace(p, 7);   // before the cluster phase
  }
  if (p.cluster && !push) {
    // ---- split-K over a thread-block cluster: every CTA staged its fp32 partial tile [token][128] in
    // its own shared memory; after a cluster barrier CTA `rank` finalises the tokens t = rank (mod S),
    // summing the S partials in rank order (deterministic) straight out of its peers' smem (DSMEM).
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (threadIdx.x == 64) tc_trace(p, 10);   // every CTA of the cluster has staged its partial tile
    if (warp >= 2) {
      const int S2 = p.splits, rank = split, wq = warp - 2, r4 = lane * 4;
      const int tok0 = tok_tile * N;
      const int tmax = min(N, p.T - tok0);
      float bb[4] = {0.f, 0.f, 0.f, 0.f};
      if (p.bias) {
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (m_tile * kBlockM + r4 + e < p.n_out) bb[e] = __bfloat162float(p.bias[m_tile * kBlockM + r4 + e]);
      }
      const uint32_t stg_addr = smem_u32(smem);
      // two tokens per pass: all 2 x S remote 16-byte loads are issued before the first use, so the
      // DSMEM latency (~1 us) is paid once per pass instead of once per peer
      for (int t = rank + S2 * wq; t < tmax; t += 8 * S2) {
        const int t1 = t + 4 * S2;
        float4 v[2][8];
#pragma unroll
        for (int pr = 0; pr < 8; ++pr) {
          if (pr < S2) {
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              const int tt = u ? t1 : t;
              const uint32_t local = stg_addr + (uint32_t)((tt < tmax ? tt : t) * kBlockM + r4) * 4u;
              uint32_t remote;
              asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(pr));
              asm volatile("ld.shared::cluster.v4.f32 {%0,%1,%2,%3}, [%4];"
                           : "=f"(v[u][pr].x), "=f"(v[u][pr].y), "=f"(v[u][pr].z), "=f"(v[u][pr].w)
                           : "r"(remote)
                           : "memory");
            }
          }
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int tt = u ? t1 : t;
          float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int pr = 0; pr < 8; ++pr) {
            if (pr < S2) {   // rank order: deterministic and identical to the global-memory path
              acc.x += v[u][pr].x;
              acc.y += v[u][pr].y;
              acc.z += v[u][pr].z;
              acc.w += v[u][pr].w;
            }
          }
          if (tt < tmax) finalize4<MODE>(p, m_tile, lane, tok0 + tt, acc, bb);   // warp-uniform condition
        }
      }
    }
    if (threadIdx.x == 64) tc_trace(p, 11);   // this CTA's share finalized
    // nobody may exit (and release its shared memory) while a peer still reads it
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  }
  if (threadIdx.x == 64) tc_trace(p, 8);   // reduce + finalize done
  __syncthreads();
  if (threadIdx.x == 0) tc_trace(p, 9);    // CTA end
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols));
  }
}

// ---- host side ---------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
    else
      cudaGetLastError();
  }
  return fn;
}

struct MapKey {
  const void* ptr;
  uint64_t rows, cols, ld;
  uint32_t box_rows;
  bool operator==(const MapKey& o) const {
    return ptr == o.ptr && rows == o.rows && cols == o.cols && ld == o.ld && box_rows == o.box_rows;
  }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    return std::hash<const void*>()(k.ptr) ^ (k.rows * 1315423911ull) ^ (k.cols << 7) ^ (k.ld << 17) ^ k.box_rows;
  }
};

// 2D bf16 tensor map [rows][cols] (row stride ld elements), box [box_rows][64], SWIZZLE_128B, zero OOB fill.
int get_map(const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows, CUtensorMap* out) {
  static std::mutex mu;
  static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
  MapKey key{ptr, rows, cols, ld, box_rows};
  std::lock_guard<std::mutex> lk(mu);
  auto it = cache.find(key);
  if (it != cache.end()) {
    *out = it->second;
    return B200_OK;
  }
  EncodeTiledFn fn = encode_fn();
  B200_REQUIRE(fn != nullptr, B200_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)kBlockK, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMap m;
  CUresult r = fn(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, B200_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for [%llu x %llu] ld %llu", (int)r,
               (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld);
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, m);
  *out = m;
  return B200_OK;
}

unsigned long long* g_trace = nullptr;
unsigned long long* trace_buffer() {
  static const bool on = getenv("B200_TC_TRACE") && atoi(getenv("B200_TC_TRACE"));
  if (on && !g_trace) {
    if (cudaMalloc((void**)&g_trace, 16 * sizeof(unsigned long long)) != cudaSuccess) g_trace = nullptr;
  }
  return on ? g_trace : nullptr;
}

inline int n_tile_for(int T) { return T <= 16 ? 16 : (T <= 32 ? 32 : (T <= 64 ? 64 : (T <= 128 ? 128 : 256))); }

struct Plan {
  int N, m_tiles, tok_tiles, n_kblocks, splits, n_out_pad;
};

Plan make_plan(int T, int n_out, int k) {
  Plan pl;
  pl.N = n_tile_for(T);
  pl.m_tiles = (n_out + kBlockM - 1) / kBlockM;
  pl.tok_tiles = (T + pl.N - 1) / pl.N;
  pl.n_kblocks = (k + kBlockK - 1) / kBlockK;
  pl.n_out_pad = pl.m_tiles * kBlockM;
  // Split K until about one CTA per SM exists (two kernels of the step can then be co-resident
  // under programmatic dependent launch: the next one prefetches its first weight stages while
  // this one drains).
  const int slots = kNumSMs;
  const int min_kb = 4;
  const int tiles = pl.m_tiles * pl.tok_tiles;
  int s = 1;
  if (tiles < slots) {
    // enough splits for AT LEAST one CTA per SM (rounded up): a CTA's stream is bounded by its own bytes in flight, so
    // SMs without a CTA are bandwidth left unused.  Measured against rounding to nearest: one TP8 rank's share of 70B
    // 6.95 -> 6.75 ms/step at batch 64 and 5.01 -> 4.76 at batch 1 (O / down: 64 slabs x 3 instead of x 2), 8B batch 64
    // 4.63 -> 4.60 (QKV: 48 slabs x 4 instead of x 3); rounding down had been 3 % slower.
    s = (slots + tiles - 1) / tiles;
    s = min(s, max(pl.n_kblocks / min_kb, 1));        // at least min_kb K blocks (64 KiB of W each) per split
    s = max(1, min(s, 32));
    // Never more splits than a cluster can reduce through DSMEM (8): the global-memory split-K (partials + tickets)
    // that larger counts fall back to is far slower — one TP8 rank's share of 70B, whose 10 QKV slabs were cut 15 ways:
    // 8.41 ms/step at batch 64 against 7.05 with 8 splits, 5.11 against 4.94 at batch 1.  (Clusters of 16 through the
    // non-portable size: 4.92 / 7.25 ms, no better.)  B200_TC_CLUSTER=0 still reaches the global path for A/B runs.
    static const int cluster_env = getenv("B200_TC_CLUSTER") ? atoi(getenv("B200_TC_CLUSTER")) : 1;
    if (cluster_env) s = min(s, 8);
  }
  pl.splits = s;
  return pl;
}

template <int N, int MODE, int S, bool NI>
int launch_tc3(const CUtensorMap& mw, const CUtensorMap& mx, const TcParams& p, const Plan& pl, cudaStream_t s) {
  static bool attr_set[64] = {};
  const int smem = S * stage_bytes(N) + 1024 + kTailBytes + recv_bytes(N);
  B200_CUDA(ensure_dynamic_smem(linear_tc_kernel<N, MODE, S, NI>, smem, attr_set));
  B200_CUDA(launch_cluster(linear_tc_kernel<N, MODE, S, NI>, dim3(pl.m_tiles, pl.splits, pl.tok_tiles), dim3(kTcThreads),
                           (size_t)smem, s, p.cluster ? dim3(1, pl.splits, 1) : dim3(1, 1, 1), mw, mx, p));
  return B200_OK;
}

template <int N, int MODE, bool NI>
int launch_tc2(const CUtensorMap& mw, const CUtensorMap& mx, const TcParams& p, const Plan& pl, cudaStream_t s) {
  if constexpr (N == 256) {
    // the 2-stage kernel (two CTAs per SM, chunked epilogue) has no room for a whole fp32 tile: split-K
    // launches of a single token tile keep 4 stages
    if (pl.splits > 1 || p.cluster) return launch_tc3<N, MODE, 4, NI>(mw, mx, p, pl, s);
  }
  return launch_tc3<N, MODE, stages_for(N), NI>(mw, mx, p, pl, s);
}

template <int N>
int launch_tc(int mode, bool ni, const CUtensorMap& mw, const CUtensorMap& mx, const TcParams& p, const Plan& pl,
              cudaStream_t s) {
  if constexpr (N <= 64) {   // the fp32 residual stream serves decode-sized steps only
    if (mode == kResidual) return launch_tc2<N, kResidual, false>(mw, mx, p, pl, s);
    if (ni) {
      switch (mode) {
        case kSiluMul: return launch_tc2<N, kSiluMul, true>(mw, mx, p, pl, s);
        case kRopeKv: return launch_tc2<N, kRopeKv, true>(mw, mx, p, pl, s);
        default: return launch_tc2<N, kPlain, true>(mw, mx, p, pl, s);
      }
    }
  }
  B200_REQUIRE(mode != kResidual && !ni, B200_ERR_UNSUPPORTED, "linear_tc: the fused-norm modes need T <= 64 (T tile %d)", N);
  switch (mode) {
    case kSiluMul: return launch_tc2<N, kSiluMul, false>(mw, mx, p, pl, s);
    case kRopeKv: return launch_tc2<N, kRopeKv, false>(mw, mx, p, pl, s);
    default: return launch_tc2<N, kPlain, false>(mw, mx, p, pl, s);
  }
}

}  // namespace

// debug: copy the 16 %globaltimer stamps of the last traced launch (B200_TC_TRACE=1) to the host
int linear_tc_read_trace(unsigned long long* host16) {
  if (!g_trace) return B200_ERR_STATE;
  B200_CUDA(cudaMemcpy(host16, g_trace, 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  return B200_OK;
}

// shared with attn_tc.cu: [rows][cols] bf16 (row stride ld), box [box_rows][64], SWIZZLE_128B
int get_tensor_map_2d(const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows, CUtensorMap* out) {
  return get_map(ptr, rows, cols, ld, box_rows, out);
}

bool linear_tc_supported(int T, int n_out, int k) {
  if (T < 1 || n_out < 1) return false;
  if (k % 8 != 0) return false;  // TMA global strides are multiples of 16 bytes
  const Plan pl = make_plan(T, n_out, k);
  return (int64_t)pl.m_tiles * pl.tok_tiles * 4 <= kCounterBytes && pl.tok_tiles <= 65535 && pl.splits <= 65535;
}

int64_t linear_tc_workspace_bytes(int T, int n_out, int k) {
  if (T < 1) T = 1;
  const Plan pl = make_plan(T, n_out, k);
  const int64_t part = pl.splits > 1 ? (int64_t)pl.splits * pl.tok_tiles * pl.N * pl.n_out_pad * 4 : 0;
  return kCounterBytes + part + 256;
}

int linear_tc(const __nv_bfloat16* x, int64_t ldx, const __nv_bfloat16* w, const __nv_bfloat16* bias,
              __nv_bfloat16* y, int64_t ldy, int T, int n_out, int k, void* ws, int64_t ws_bytes,
              const LinearExtra* ex, cudaStream_t s) {
  B200_REQUIRE(k % 8 == 0, B200_ERR_UNSUPPORTED, "linear_tc: k must be a multiple of 8");
  B200_REQUIRE(reinterpret_cast<uintptr_t>(w) % 16 == 0, B200_ERR_INVALID, "linear_tc: w must be 16-byte aligned");
  const Plan pl = make_plan(T, n_out, k);
  const int64_t need = linear_tc_workspace_bytes(T, n_out, k);
  static const int cluster_env = getenv("B200_TC_CLUSTER") ? atoi(getenv("B200_TC_CLUSTER")) : 1;
  const bool use_cluster = cluster_env != 0 && pl.splits > 1 && pl.splits <= 8;
  B200_REQUIRE(pl.splits == 1 || use_cluster || (ws && ws_bytes >= need), B200_ERR_INVALID,
               "linear_tc: workspace too small (%lld < %lld)", (long long)ws_bytes, (long long)need);
  const int mode = ex ? ex->mode : 0;
  TcParams p = {};
  p.cluster = use_cluster ? 1 : 0;
  p.trace = trace_buffer();
  if (mode == kSiluMul) {
    B200_REQUIRE(n_out % kBlockM == 0, B200_ERR_UNSUPPORTED, "linear_tc(silu_mul): n_out=%d must be a multiple of 128", n_out);
  } else if (mode == kRopeKv) {
    B200_REQUIRE(ex->positions && ex->slots && ex->cos_sin && ex->kc && ex->vc && ex->block_size > 0 &&
                     n_out == (ex->hq + 2 * ex->hkv) * kHeadDim,
                 B200_ERR_INVALID, "linear_tc(rope_kv): bad arguments (n_out=%d, hq=%d, hkv=%d)", n_out, ex->hq, ex->hkv);
    B200_REQUIRE(ldy % 4 == 0 && (reinterpret_cast<uintptr_t>(y) & 7) == 0, B200_ERR_INVALID,
                 "linear_tc(rope_kv): y must be 8-byte aligned with ldy %% 4 == 0");
    p.positions = ex->positions;
    p.slots = ex->slots;
    p.cos_sin = ex->cos_sin;
    p.kc = ex->kc;
    p.vc = ex->vc;
    p.hq = ex->hq;
    p.hkv = ex->hkv;
    p.block_size = ex->block_size;
  } else if (mode == kResidual) {
    B200_REQUIRE(ex->sres && ex->ssq && n_out % kBlockM == 0 && pl.N <= 64, B200_ERR_INVALID,
                 "linear_tc(residual): needs the fp32 stream, n_out %% 128 == 0 and T <= 64 (n_out=%d T=%d)", n_out, T);
    p.sres = ex->sres;
    p.ssq = ex->ssq;
    p.nslab = n_out / kBlockM;
  } else {
    B200_REQUIRE(mode == kPlain, B200_ERR_INVALID, "linear_tc: unknown epilogue mode %d", mode);
  }
  const bool ni = ex && ex->norm_in;
  if (ni) {
    B200_REQUIRE(mode != kResidual && ex->sres && ex->ssq && ex->norm_w && k % 128 == 0 && pl.N <= 64 && pl.tok_tiles == 1,
                 B200_ERR_INVALID, "linear_tc(norm_in): needs the fp32 stream, k %% 128 == 0 and T <= 64 (k=%d T=%d)", k, T);
    p.sres = ex->sres;
    p.ssq = ex->ssq;
    p.norm_w = ex->norm_w;
    p.rows = ex->rows;
    p.nslab = k / kBlockM;
    p.eps = ex->eps;
  }
  CUtensorMap mw, mx;
  int rc = get_map(w, (uint64_t)n_out, (uint64_t)k, (uint64_t)k, kBlockM, &mw);
  if (rc) return rc;
  if (ni) {
    mx = mw;   // unused: the B operand is built in shared memory
  } else {
    B200_REQUIRE(x && reinterpret_cast<uintptr_t>(x) % 16 == 0 && ldx % 8 == 0, B200_ERR_INVALID, "linear_tc: bad x");
    rc = get_map(x, (uint64_t)T, (uint64_t)k, (uint64_t)ldx, (uint32_t)pl.N, &mx);
    if (rc) return rc;
  }
  p.bias = bias;
  p.y = y;
  p.ldy = ldy;
  p.T = T;
  p.n_out = n_out;
  p.n_kblocks = pl.n_kblocks;
  p.splits = pl.splits;
  p.tickets = reinterpret_cast<unsigned int*>(ws);
  p.partials = ws ? reinterpret_cast<float*>(static_cast<char*>(ws) + kCounterBytes) : nullptr;
  p.n_out_pad = pl.n_out_pad;
  switch (pl.N) {
    case 16: return launch_tc<16>(mode, ni, mw, mx, p, pl, s);
    case 32: return launch_tc<32>(mode, ni, mw, mx, p, pl, s);
    case 64: return launch_tc<64>(mode, ni, mw, mx, p, pl, s);
    case 128: return launch_tc<128>(mode, ni, mw, mx, p, pl, s);
    default: return launch_tc<256>(mode, ni, mw, mx, p, pl, s);
  }
}

}  // namespace b200
