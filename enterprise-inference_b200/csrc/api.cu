// The C-ABI of include/b200_decode.h: error plumbing, the per-rank context (weights, caller-owned
// KV cache, private activations/scratch, CUDA graphs, NCCL communicator) and the decode step that
// strings the kernels together.  No torch types anywhere: plain pointers and sizes.
#include <dlfcn.h>
#include <nccl.h>  // types only; the library is resolved at run time from the process (torch's copy)
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <unordered_map>
#include <vector>

#include "common.cuh"

namespace b200 {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
  return B200_ERR_CUDA;
}

LaunchOpts& launch_opts() {
  static thread_local LaunchOpts o;
  return o;
}
int& launch_counter() {
  static thread_local int c = 0;
  return c;
}

// ---- NCCL, resolved from the already-loaded library -------------------------------------------
struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
};

static NcclApi& nccl() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return api;
  tried = true;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) return api;
  api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(h, "ncclGetUniqueId");
  api.CommInitRank = (decltype(api.CommInitRank))dlsym(h, "ncclCommInitRank");
  api.CommDestroy = (decltype(api.CommDestroy))dlsym(h, "ncclCommDestroy");
  api.AllReduce = (decltype(api.AllReduce))dlsym(h, "ncclAllReduce");
  api.AllGather = (decltype(api.AllGather))dlsym(h, "ncclAllGather");
  api.GetErrorString = (decltype(api.GetErrorString))dlsym(h, "ncclGetErrorString");
  api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllReduce && api.AllGather;
  return api;
}

#define B200_NCCL(expr)                                                                      \
  do {                                                                                       \
    ncclResult_t _r = (expr);                                                                \
    if (_r != ncclSuccess) {                                                                 \
      set_error("NCCL error %d (%s): %s", (int)_r,                                           \
                nccl().GetErrorString ? nccl().GetErrorString(_r) : "?", #expr);             \
      return B200_ERR_NCCL;                                                                  \
    }                                                                                        \
  } while (0)

// Pick the winner among the ranks' (max, index) pairs: highest value, lowest index on ties.
__global__ void tp_pick_kernel(const float* __restrict__ gathered, int world, int rows, int32_t* __restrict__ out) {
  pdl_wait();
  pdl_launch_dependents();
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  float bv = -INFINITY;
  int bi = 0x7fffffff;
  for (int w = 0; w < world; ++w) {
    const float v = gathered[((int64_t)w * rows + r) * 2];
    const int i = __float_as_int(gathered[((int64_t)w * rows + r) * 2 + 1]);
    if (v > bv || (v == bv && i < bi)) {
      bv = v;
      bi = i;
    }
  }
  out[r] = bi;
}
__global__ void tp_pack_kernel(const float* __restrict__ mx, const int32_t* __restrict__ idx, int rows,
                               float* __restrict__ packed) {
  pdl_wait();
  pdl_launch_dependents();
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  packed[r * 2] = mx[r];
  packed[r * 2 + 1] = __int_as_float(idx[r]);
}

}  // namespace b200

using namespace b200;

// ------------------------------------------------------------------------------------------------
struct LayerW {
  const __nv_bfloat16 *ln1 = nullptr, *qkv = nullptr, *qkv_b = nullptr, *o = nullptr, *ln2 = nullptr, *gu = nullptr,
                      *down = nullptr;
  bool gu_interleaved = false;  // gu holds 128-row slabs of 64 gate rows + the 64 matching up rows
};

struct GraphKey {
  uint64_t parts[16];   // every pointer and size the captured graph baked in
  bool operator==(const GraphKey& o) const { return memcmp(parts, o.parts, sizeof(parts)) == 0; }
};
struct GraphKeyHash {
  size_t operator()(const GraphKey& k) const {
    uint64_t h = 0x1234;
    for (uint64_t v : k.parts) h ^= v + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2);
    return (size_t)h;
  }
};

struct b200_ctx {
  b200_model_cfg cfg;
  int hq, hkv, inter, vocab_local, qkv_w;  // local (per TP rank) sizes
  int max_logit_rows;
  std::vector<LayerW> layers;
  const __nv_bfloat16 *embed = nullptr, *final_norm = nullptr, *lm_head = nullptr, *cos_sin = nullptr;
  int64_t max_pos = 0;
  __nv_bfloat16 *kc = nullptr, *vc = nullptr;
  int64_t n_blocks = 0;
  // private device buffers
  void* pool = nullptr;
  int64_t pool_bytes = 0;
  __nv_bfloat16 *x, *res, *normed, *qkv_buf, *attn_out, *gu_buf, *act, *lm_in, *logits;
  float *sres, *ssq;   // fp32 residual stream [T][h] and its per-slab sums of squares [T][h/128] (norm folded into the linears)
  void* attn_ws;
  int64_t attn_ws_bytes;
  void* lin_ws;
  int64_t lin_ws_bytes;
  int32_t* d_meta;
  int64_t meta_ints;
  int32_t* d_tokens;
  float* d_max;
  float* d_pairs;     // [rows][2] local (max, idx)
  float* d_gathered;  // [world][rows][2]
  void* d_sample;     // sampling parameters of the rows, packed (b200_ctx_sample_last)
  void* d_logprob;    // [rows] chosen ids | [rows] logprob | [rows][32] top ids | [rows][32] top logprobs
  void* d_adjust;     // packed penalties / logit_bias of the step's rows (b200_ctx_adjust_last)
  __nv_bfloat16* logits_full;  // TP only: [world][rows][vocab_local], gathered for the sampler
  // graphs
  bool graphs = true;
  std::unordered_map<GraphKey, cudaGraphExec_t, GraphKeyHash> graph_cache;
  int last_launches = 0;
  // taps
  int tap_layer = -1;
  void *tap_hidden = nullptr, *tap_res = nullptr;
  // TP
  ncclComm_t comm = nullptr;
  // all-reduce over peer memory (allreduce.cu): one IPC-shareable allocation per rank
  //   [ partial 0 | partial 1 | gather 0 | gather 1 | flag block (2 x 2 x 8 x 128 B) | state (counter[2], ticket[2]) ]
  char* ar_mem = nullptr;
  int64_t ar_buf_bytes = 0, ar_total_bytes = 0;
  char* ar_peer[8] = {};
  bool ar_ready = false;
  int gemv_max_t = 0;  // rows <= this use the CUDA-core GEMV; 0: the tcgen05 path serves every T (measured faster)
  bool pdl = true;
  bool fuse = true;                    // epilogue fusion (RoPE+KV, SiLU*mul) on the tcgen05 path
  // Steps of at most this many rows fold the RMSNorms into the linears (0: off).  Measured on 8B: batch 1 3.05 -> 3.00
  // ms/step; at 64 rows every CTA of a consumer re-normalises 1 MiB of fp32 rows out of L2 and the step slows from 4.6
  // to 6.3 ms, so the fold stops at the 16-row tile.
  int fuse_norm_max_t = 16;
};

static int64_t align_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }
constexpr int64_t kAdjustMaxInts = 4ll << 20;   // 16 MiB: ~1 M (row, token) entries of penalties / logit_bias per step
constexpr int64_t kArFlagBytes = 2 * 2 * 8 * 128;   // "partial ready" and "rows delivered" flags: [kind][buffer][rank] x 128 B

extern "C" const char* b200_last_error(void) { return g_err; }
extern "C" int b200_abi_version(void) { return 1; }

extern "C" int b200_device_ok(void) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  cudaDeviceProp p;
  if (cudaGetDeviceProperties(&p, dev) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return p.major == 10 ? 1 : 0;
}

extern "C" int b200_ctx_create(const b200_model_cfg* cfg, b200_ctx** out) {
  B200_REQUIRE(cfg && out, B200_ERR_INVALID, "ctx_create: null argument");
  B200_REQUIRE(cfg->head_dim == kHeadDim, B200_ERR_UNSUPPORTED, "ctx_create: head_dim must be 128 (got %d)", cfg->head_dim);
  B200_REQUIRE(cfg->tp_world >= 1 && cfg->tp_rank >= 0 && cfg->tp_rank < cfg->tp_world, B200_ERR_INVALID,
               "ctx_create: bad tp rank/world %d/%d", cfg->tp_rank, cfg->tp_world);
  B200_REQUIRE(cfg->n_q_heads % cfg->tp_world == 0 && cfg->n_kv_heads % cfg->tp_world == 0 &&
                   cfg->intermediate % cfg->tp_world == 0 && cfg->vocab % cfg->tp_world == 0,
               B200_ERR_UNSUPPORTED, "ctx_create: heads/intermediate/vocab must divide by tp_world=%d", cfg->tp_world);
  B200_REQUIRE(cfg->hidden % 8 == 0 && cfg->intermediate % (8 * cfg->tp_world) == 0 && cfg->max_tokens > 0 &&
                   cfg->n_layers > 0 && cfg->block_size > 0 && cfg->max_blocks_per_seq > 0,
               B200_ERR_INVALID, "ctx_create: bad sizes");
  b200_ctx* c = new b200_ctx();
  c->cfg = *cfg;
  const int W = cfg->tp_world;
  c->hq = cfg->n_q_heads / W;
  c->hkv = cfg->n_kv_heads / W;
  c->inter = cfg->intermediate / W;
  c->vocab_local = cfg->vocab / W;
  c->qkv_w = (c->hq + 2 * c->hkv) * kHeadDim;
  c->max_logit_rows = cfg->max_tokens < 512 ? cfg->max_tokens : 512;
  c->layers.resize(cfg->n_layers);
  if (const char* e = getenv("B200_GEMV_MAX_T")) c->gemv_max_t = atoi(e);
  if (const char* e = getenv("B200_GRAPHS")) c->graphs = atoi(e) != 0;
  if (const char* e = getenv("B200_PDL")) c->pdl = atoi(e) != 0;
  if (const char* e = getenv("B200_FUSE")) c->fuse = atoi(e) != 0;
  if (const char* e = getenv("B200_FUSE_NORM_MAX_T")) c->fuse_norm_max_t = atoi(e);

  const int64_t T = cfg->max_tokens, h = cfg->hidden, R = c->max_logit_rows;
  const int max_ctx = cfg->max_blocks_per_seq * cfg->block_size;
  c->attn_ws_bytes = 0;
  for (int t = 1; t <= T; t = t < 64 ? t + 1 : t * 2) {  // the bound is not monotone in T: scan
    const int64_t b = paged_attn_workspace_bytes(t, c->hq, max_ctx, cfg->block_size);
    if (b > c->attn_ws_bytes) c->attn_ws_bytes = b;
  }
  c->attn_ws_bytes = std::max<int64_t>(c->attn_ws_bytes, paged_attn_workspace_bytes((int)T, c->hq, max_ctx, cfg->block_size));
  int64_t lw = 0;
  const int outs[5] = {c->qkv_w, (int)h, 2 * c->inter, (int)h, c->vocab_local};
  const int ks[5] = {(int)h, c->hq * kHeadDim, (int)h, c->inter, (int)h};
  for (int t = 1; t <= T; t = (t < 16 ? 16 : (t % 2 == 0 ? t + 1 : (t - 1) * 2))) {  // 1,16,17,32,33,64,65,...
    for (int i = 0; i < 5; ++i) lw = std::max<int64_t>(lw, linear_tc_workspace_bytes(t, outs[i], ks[i]));
  }
  for (int i = 0; i < 5; ++i) lw = std::max<int64_t>(lw, linear_tc_workspace_bytes((int)T, outs[i], ks[i]));
  c->lin_ws_bytes = lw;
  c->meta_ints = 6 * T + (int64_t)T * cfg->max_blocks_per_seq;

  struct Item {
    void** p;
    int64_t bytes;
  };
  std::vector<Item> items = {
      {(void**)&c->x, T * h * 2},
      {(void**)&c->res, T * h * 2},
      {(void**)&c->normed, T * h * 2},
      {(void**)&c->sres, std::min<int64_t>(T, 64) * h * 4},
      {(void**)&c->ssq, std::min<int64_t>(T, 64) * (h / 128 + 1) * 4},
      {(void**)&c->qkv_buf, T * c->qkv_w * 2},
      {(void**)&c->attn_out, T * c->hq * kHeadDim * 2},
      {(void**)&c->gu_buf, T * 2 * c->inter * 2},
      {(void**)&c->act, T * c->inter * 2},
      {(void**)&c->lm_in, R * h * 2},
      {(void**)&c->logits, R * (int64_t)c->vocab_local * 2},
      {(void**)&c->attn_ws, c->attn_ws_bytes},
      {(void**)&c->lin_ws, c->lin_ws_bytes},
      {(void**)&c->d_meta, c->meta_ints * 4},
      {(void**)&c->d_tokens, R * 4},
      {(void**)&c->d_max, R * 4},
      {(void**)&c->d_pairs, R * 8},
      {(void**)&c->d_gathered, R * 8 * W},
      {(void**)&c->d_sample, R * 40},
      {(void**)&c->d_logprob, R * (8 + 32 * 8)},
      {(void**)&c->d_adjust, kAdjustMaxInts * 4},
      {(void**)&c->logits_full, W > 1 ? R * (int64_t)c->vocab_local * 2 * W : 16},
  };
  int64_t total = 0;
  for (auto& it : items) total += align_up(std::max<int64_t>(it.bytes, 16), 1024);
  cudaError_t e = cudaMalloc(&c->pool, total);
  if (e != cudaSuccess) {
    delete c;
    return cuda_fail(e, "cudaMalloc(ctx pool)", __FILE__, __LINE__);
  }
  e = cudaMemset(c->pool, 0, total);
  if (e != cudaSuccess) {
    cudaFree(c->pool);
    delete c;
    return cuda_fail(e, "cudaMemset(ctx pool)", __FILE__, __LINE__);
  }
  c->pool_bytes = total;
  int64_t off = 0;
  for (auto& it : items) {
    *it.p = static_cast<char*>(c->pool) + off;
    off += align_up(std::max<int64_t>(it.bytes, 16), 1024);
  }
  if (W > 1) {
    c->ar_buf_bytes = align_up(T * h * 2, 1024);
    c->ar_total_bytes = 4 * c->ar_buf_bytes + kArFlagBytes + 256;
    e = cudaMalloc((void**)&c->ar_mem, c->ar_total_bytes);
    if (e == cudaSuccess) e = cudaMemset(c->ar_mem, 0, c->ar_total_bytes);
    if (e != cudaSuccess) {
      cudaFree(c->pool);
      delete c;
      return cuda_fail(e, "cudaMalloc(all-reduce buffers)", __FILE__, __LINE__);
    }
  }
  *out = c;
  return B200_OK;
}

extern "C" int b200_tp_ipc_export(b200_ctx* c, void* host_handle64) {
  B200_REQUIRE(c && host_handle64, B200_ERR_INVALID, "tp_ipc_export: null");
  B200_REQUIRE(c->ar_mem, B200_ERR_STATE, "tp_ipc_export: tp_world == 1");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t hnd;
  B200_CUDA(cudaIpcGetMemHandle(&hnd, c->ar_mem));
  memcpy(host_handle64, &hnd, 64);
  return B200_OK;
}

extern "C" int b200_tp_ipc_import(b200_ctx* c, const void* host_handles, int32_t world) {
  B200_REQUIRE(c && host_handles, B200_ERR_INVALID, "tp_ipc_import: null");
  B200_REQUIRE(c->ar_mem && world == c->cfg.tp_world && world <= 8, B200_ERR_STATE, "tp_ipc_import: bad world %d", world);
  for (int r = 0; r < world; ++r) {
    if (r == c->cfg.tp_rank) {
      c->ar_peer[r] = c->ar_mem;
      continue;
    }
    cudaIpcMemHandle_t hnd;
    memcpy(&hnd, static_cast<const char*>(host_handles) + 64 * r, 64);
    void* ptr = nullptr;
    B200_CUDA(cudaIpcOpenMemHandle(&ptr, hnd, cudaIpcMemLazyEnablePeerAccess));
    c->ar_peer[r] = static_cast<char*>(ptr);
  }
  c->ar_ready = true;
  return B200_OK;
}

extern "C" int b200_ctx_destroy(b200_ctx* c) {
  if (!c) return B200_OK;
  for (int r = 0; r < 8; ++r)
    if (c->ar_peer[r] && c->ar_peer[r] != c->ar_mem) cudaIpcCloseMemHandle(c->ar_peer[r]);
  if (c->ar_mem) cudaFree(c->ar_mem);
  for (auto& kv : c->graph_cache) cudaGraphExecDestroy(kv.second);
  if (c->comm && nccl().ok) nccl().CommDestroy(c->comm);
  if (c->pool) cudaFree(c->pool);
  delete c;
  return B200_OK;
}

extern "C" int64_t b200_ctx_workspace_bytes(const b200_ctx* c) { return c ? c->pool_bytes : 0; }

extern "C" int b200_set_weight(b200_ctx* c, int32_t layer, int32_t kind, const void* w, int64_t rows, int64_t cols) {
  B200_REQUIRE(c && w, B200_ERR_INVALID, "set_weight: null argument");
  B200_REQUIRE(reinterpret_cast<uintptr_t>(w) % 16 == 0, B200_ERR_INVALID, "set_weight: pointer not 16-byte aligned");
  const int64_t h = c->cfg.hidden;
  const bool per_layer = !(kind == B200_W_EMBED || kind == B200_W_FINAL_NORM || kind == B200_W_LM_HEAD);
  if (per_layer) B200_REQUIRE(layer >= 0 && layer < c->cfg.n_layers, B200_ERR_INVALID, "set_weight: layer %d out of range", layer);
  int64_t er = 0, ec = 0;
  const __nv_bfloat16* p = static_cast<const __nv_bfloat16*>(w);
  switch (kind) {
    case B200_W_EMBED: er = c->cfg.vocab; ec = h; break;
    case B200_W_LN1: case B200_W_LN2: case B200_W_FINAL_NORM: er = h; ec = 1; break;
    case B200_W_QKV: er = c->qkv_w; ec = h; break;
    case B200_W_QKV_BIAS: er = c->qkv_w; ec = 1; break;
    case B200_W_O: er = h; ec = c->hq * kHeadDim; break;
    case B200_W_GATE_UP: case B200_W_GATE_UP_INTERLEAVED: er = 2 * c->inter; ec = h; break;
    case B200_W_DOWN: er = h; ec = c->inter; break;
    case B200_W_LM_HEAD: er = c->vocab_local; ec = h; break;
    default: B200_REQUIRE(false, B200_ERR_INVALID, "set_weight: unknown kind %d", kind);
  }
  B200_REQUIRE(rows == er && cols == ec, B200_ERR_INVALID, "set_weight: kind %d expects [%lld,%lld], got [%lld,%lld]", kind,
               (long long)er, (long long)ec, (long long)rows, (long long)cols);
  switch (kind) {
    case B200_W_EMBED: c->embed = p; break;
    case B200_W_FINAL_NORM: c->final_norm = p; break;
    case B200_W_LM_HEAD: c->lm_head = p; break;
    case B200_W_LN1: c->layers[layer].ln1 = p; break;
    case B200_W_LN2: c->layers[layer].ln2 = p; break;
    case B200_W_QKV: c->layers[layer].qkv = p; break;
    case B200_W_QKV_BIAS: c->layers[layer].qkv_b = p; break;
    case B200_W_O: c->layers[layer].o = p; break;
    case B200_W_GATE_UP: c->layers[layer].gu = p; c->layers[layer].gu_interleaved = false; break;
    case B200_W_GATE_UP_INTERLEAVED:
      B200_REQUIRE(c->inter % 64 == 0, B200_ERR_UNSUPPORTED, "interleaved gate_up needs intermediate_local %% 64 == 0");
      c->layers[layer].gu = p; c->layers[layer].gu_interleaved = true; break;
    case B200_W_DOWN: c->layers[layer].down = p; break;
  }
  return B200_OK;
}

extern "C" int b200_set_kv_cache(b200_ctx* c, void* k_cache, void* v_cache, int64_t n_blocks) {
  B200_REQUIRE(c && k_cache && v_cache && n_blocks > 0, B200_ERR_INVALID, "set_kv_cache: bad argument");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(k_cache) | reinterpret_cast<uintptr_t>(v_cache)) % 128 == 0, B200_ERR_INVALID,
               "set_kv_cache: caches must be 128-byte aligned");
  c->kc = static_cast<__nv_bfloat16*>(k_cache);
  c->vc = static_cast<__nv_bfloat16*>(v_cache);
  c->n_blocks = n_blocks;
  return B200_OK;
}

extern "C" int b200_set_rope_cache(b200_ctx* c, const void* cos_sin, int64_t max_pos) {
  B200_REQUIRE(c && cos_sin && max_pos > 0, B200_ERR_INVALID, "set_rope_cache: bad argument");
  c->cos_sin = static_cast<const __nv_bfloat16*>(cos_sin);
  c->max_pos = max_pos;
  return B200_OK;
}

extern "C" int b200_nccl_unique_id(void* host_id128) {
  B200_REQUIRE(host_id128, B200_ERR_INVALID, "nccl_unique_id: null");
  B200_REQUIRE(nccl().ok, B200_ERR_NCCL, "libnccl.so.2 not found in the process");
  ncclUniqueId id;
  B200_NCCL(nccl().GetUniqueId(&id));
  memcpy(host_id128, &id, sizeof(id));
  return B200_OK;
}

extern "C" int b200_nccl_init(b200_ctx* c, const void* host_id128) {
  B200_REQUIRE(c && host_id128, B200_ERR_INVALID, "nccl_init: null");
  B200_REQUIRE(nccl().ok, B200_ERR_NCCL, "libnccl.so.2 not found in the process");
  ncclUniqueId id;
  memcpy(&id, host_id128, sizeof(id));
  B200_NCCL(nccl().CommInitRank(&c->comm, c->cfg.tp_world, id, c->cfg.tp_rank));
  return B200_OK;
}

extern "C" int b200_allreduce_bf16(b200_ctx* c, void* buf, int64_t count, void* stream) {
  B200_REQUIRE(c && buf, B200_ERR_INVALID, "allreduce: null");
  if (c->cfg.tp_world == 1) return B200_OK;
  B200_REQUIRE(c->comm, B200_ERR_STATE, "allreduce: b200_nccl_init not called");
  ++launch_counter();
  B200_NCCL(nccl().AllReduce(buf, buf, (size_t)count, ncclBfloat16, ncclSum, c->comm, (cudaStream_t)stream));
  return B200_OK;
}

extern "C" int b200_set_launch_pdl(int32_t enabled) {
  launch_opts().pdl = enabled != 0;
  return B200_OK;
}

extern "C" int b200_ctx_set_graphs(b200_ctx* c, int32_t enabled) {
  B200_REQUIRE(c, B200_ERR_INVALID, "null ctx");
  c->graphs = enabled != 0;
  return B200_OK;
}
extern "C" int b200_ctx_last_step_launches(const b200_ctx* c) { return c ? c->last_launches : 0; }
extern "C" int b200_ctx_set_tap(b200_ctx* c, int32_t layer, void* out_hidden, void* out_residual) {
  B200_REQUIRE(c, B200_ERR_INVALID, "null ctx");
  c->tap_layer = layer;
  c->tap_hidden = out_hidden;
  c->tap_res = out_residual;
  return B200_OK;
}

// ---- linear dispatch ----------------------------------------------------------------------------
static bool tc_path(int T, int n_out, int k, int gemv_max_t) {
  return T > gemv_max_t && linear_tc_supported(T, n_out, k);
}

static int linear_any(const __nv_bfloat16* x, int64_t ldx, const __nv_bfloat16* w, const __nv_bfloat16* bias,
                      __nv_bfloat16* y, int64_t ldy, int T, int n_out, int k, int impl, int gemv_max_t, void* ws,
                      int64_t ws_bytes, const LinearExtra* ex, cudaStream_t s) {
  if (T == 0) return B200_OK;
  if (impl == 0) impl = tc_path(T, n_out, k, gemv_max_t) ? 2 : 1;
  if (impl == 1) {
    B200_REQUIRE(!ex || (ex->mode == 0 && !ex->norm_in), B200_ERR_UNSUPPORTED, "linear: fused epilogues need the tcgen05 path");
    return linear_gemv(x, ldx, w, bias, y, ldy, T, n_out, k, s);
  }
  B200_REQUIRE(linear_tc_supported(T, n_out, k), B200_ERR_UNSUPPORTED,
               "linear: tcgen05 path does not support T=%d n_out=%d k=%d", T, n_out, k);
  return linear_tc(x, ldx, w, bias, y, ldy, T, n_out, k, ws, ws_bytes, ex, s);
}

namespace b200 { int linear_tc_read_trace(unsigned long long* host16); int attn_tc_read_trace(unsigned long long* host64);
                 int attn_prefill_read_trace(unsigned long long* host256); }
// debug helper (not in the public header): tile timeline of CTA 0 of the last prefill attention (trace builds)
extern "C" __attribute__((visibility("default"))) int b200_debug_prefill_trace(unsigned long long* host256) {
  return b200::attn_prefill_read_trace(host256);
}
extern "C" __attribute__((visibility("default"))) int b200_debug_attn_trace(unsigned long long* host64) {
  return b200::attn_tc_read_trace(host64);
}
// debug helper (not in the public header): timestamps of CTA 0 of the last tcgen05 linear, B200_TC_TRACE=1
extern "C" __attribute__((visibility("default"))) int b200_debug_linear_trace(unsigned long long* host16) {
  return b200::linear_tc_read_trace(host16);
}

extern "C" int64_t b200_linear_workspace_bytes(int32_t T, int32_t n_out, int32_t k) {
  return linear_tc_workspace_bytes(T, n_out, k);
}

extern "C" int b200_linear_bf16(const void* x, int64_t ldx, const void* w, const void* bias, void* y, int64_t ldy,
                                int32_t T, int32_t n_out, int32_t k, int32_t impl, void* ws, int64_t ws_bytes,
                                void* stream) {
  B200_REQUIRE(x && w && y && T >= 0 && n_out > 0 && k > 0, B200_ERR_INVALID, "linear: bad argument");
  int gm = 0;
  if (const char* e = getenv("B200_GEMV_MAX_T")) gm = atoi(e);
  return linear_any((const __nv_bfloat16*)x, ldx, (const __nv_bfloat16*)w, (const __nv_bfloat16*)bias,
                    (__nv_bfloat16*)y, ldy, T, n_out, k, impl, gm, ws, ws_bytes, nullptr, (cudaStream_t)stream);
}

extern "C" int b200_linear_fused_bf16(const void* x, int64_t ldx, const void* w, const void* bias, void* y, int64_t ldy,
                                      int32_t T, int32_t n_out, int32_t k, int32_t mode, const int32_t* positions,
                                      const int32_t* slot_mapping, const void* cos_sin, void* k_cache_layer,
                                      void* v_cache_layer, int32_t n_q_heads, int32_t n_kv_heads, int32_t block_size,
                                      void* ws, int64_t ws_bytes, void* stream) {
  B200_REQUIRE(x && w && y && T >= 0 && n_out > 0 && k > 0, B200_ERR_INVALID, "linear_fused: bad argument");
  B200_REQUIRE(mode == 1 || mode == 2, B200_ERR_INVALID, "linear_fused: mode must be 1 (silu*mul) or 2 (rope+kv)");
  LinearExtra ex;
  ex.mode = mode;
  ex.positions = positions;
  ex.slots = slot_mapping;
  ex.cos_sin = (const __nv_bfloat16*)cos_sin;
  ex.kc = (__nv_bfloat16*)k_cache_layer;
  ex.vc = (__nv_bfloat16*)v_cache_layer;
  ex.hq = n_q_heads;
  ex.hkv = n_kv_heads;
  ex.block_size = block_size;
  return linear_any((const __nv_bfloat16*)x, ldx, (const __nv_bfloat16*)w, (const __nv_bfloat16*)bias,
                    (__nv_bfloat16*)y, ldy, T, n_out, k, 2, 0, ws, ws_bytes, &ex, (cudaStream_t)stream);
}

extern "C" int b200_linear_residual_bf16(const void* x, int64_t ldx, const void* w, const void* bias, float* sres, float* ssq,
                                         int32_t T, int32_t n_out, int32_t k, void* ws, int64_t ws_bytes, void* stream) {
  B200_REQUIRE(x && w && sres && ssq && T >= 0 && n_out > 0 && k > 0, B200_ERR_INVALID, "linear_residual: bad argument");
  LinearExtra ex;
  ex.mode = 3;
  ex.sres = sres;
  ex.ssq = ssq;
  return linear_any((const __nv_bfloat16*)x, ldx, (const __nv_bfloat16*)w, (const __nv_bfloat16*)bias, nullptr, n_out, T,
                    n_out, k, 2, 0, ws, ws_bytes, &ex, (cudaStream_t)stream);
}

extern "C" int b200_linear_normed_bf16(const float* sres, const float* ssq, const void* norm_w, float eps, const int32_t* rows,
                                       const void* w, const void* bias, void* y, int64_t ldy, int32_t T, int32_t n_out,
                                       int32_t k, int32_t mode, const int32_t* positions, const int32_t* slot_mapping,
                                       const void* cos_sin, void* k_cache_layer, void* v_cache_layer, int32_t n_q_heads,
                                       int32_t n_kv_heads, int32_t block_size, void* ws, int64_t ws_bytes, void* stream) {
  B200_REQUIRE(sres && ssq && norm_w && w && y && T >= 0 && n_out > 0 && k > 0, B200_ERR_INVALID, "linear_normed: bad argument");
  B200_REQUIRE(mode >= 0 && mode <= 2, B200_ERR_INVALID, "linear_normed: mode must be 0 (plain), 1 (silu*mul) or 2 (rope+kv)");
  LinearExtra ex;
  ex.mode = mode;
  ex.norm_in = true;
  ex.sres = const_cast<float*>(sres);
  ex.ssq = const_cast<float*>(ssq);
  ex.norm_w = (const __nv_bfloat16*)norm_w;
  ex.rows = rows;
  ex.eps = eps;
  ex.positions = positions;
  ex.slots = slot_mapping;
  ex.cos_sin = (const __nv_bfloat16*)cos_sin;
  ex.kc = (__nv_bfloat16*)k_cache_layer;
  ex.vc = (__nv_bfloat16*)v_cache_layer;
  ex.hq = n_q_heads;
  ex.hkv = n_kv_heads;
  ex.block_size = block_size;
  return linear_any(nullptr, k, (const __nv_bfloat16*)w, (const __nv_bfloat16*)bias, (__nv_bfloat16*)y, ldy, T, n_out, k, 2, 0,
                    ws, ws_bytes, &ex, (cudaStream_t)stream);
}

extern "C" int b200_embed_stream(const int32_t* ids, const void* table, float* sres, float* ssq, int32_t T, int32_t hidden,
                                 int32_t vocab, void* stream) {
  B200_REQUIRE(ids && table && sres && ssq, B200_ERR_INVALID, "embed_stream: null");
  return embed_stream(ids, (const __nv_bfloat16*)table, sres, ssq, T, hidden, vocab, (cudaStream_t)stream);
}

extern "C" int b200_rmsnorm_stream(const float* sres, const float* ssq, const int32_t* rows, const void* w, void* y,
                                   int32_t n_rows, int32_t hidden, float eps, void* stream) {
  B200_REQUIRE(sres && ssq && w && y, B200_ERR_INVALID, "rmsnorm_stream: null");
  return rmsnorm_stream(sres, ssq, rows, (const __nv_bfloat16*)w, (__nv_bfloat16*)y, n_rows, hidden, eps, (cudaStream_t)stream);
}

extern "C" int b200_rmsnorm(const void* x, void* res, const void* w, void* y, int32_t T, int32_t hidden, float eps,
                            void* stream) {
  B200_REQUIRE(x && w && y, B200_ERR_INVALID, "rmsnorm: null");
  return rmsnorm((const __nv_bfloat16*)x, (__nv_bfloat16*)res, (const __nv_bfloat16*)w, (__nv_bfloat16*)y, T, hidden,
                 eps, (cudaStream_t)stream);
}

extern "C" int b200_embed(const int32_t* ids, const void* table, void* out, int32_t T, int32_t hidden, int32_t vocab,
                          void* stream) {
  B200_REQUIRE(ids && table && out, B200_ERR_INVALID, "embed: null");
  return embed(ids, (const __nv_bfloat16*)table, (__nv_bfloat16*)out, T, hidden, vocab, (cudaStream_t)stream);
}

extern "C" int b200_rope_kv_append(void* qkv, const int32_t* positions, const int32_t* slot_mapping, const void* cos_sin,
                                   void* k_cache_layer, void* v_cache_layer, int32_t T, int32_t hq, int32_t hkv,
                                   int32_t block_size, void* stream) {
  B200_REQUIRE(qkv && positions && slot_mapping && cos_sin && k_cache_layer && v_cache_layer, B200_ERR_INVALID,
               "rope_kv_append: null");
  return rope_kv_append((__nv_bfloat16*)qkv, positions, slot_mapping, (const __nv_bfloat16*)cos_sin,
                        (__nv_bfloat16*)k_cache_layer, (__nv_bfloat16*)v_cache_layer, T, hq, hkv, block_size,
                        (cudaStream_t)stream);
}

extern "C" int64_t b200_paged_attn_workspace_bytes(int32_t T, int32_t hq, int32_t max_ctx, int32_t block_size) {
  return paged_attn_workspace_bytes(T, hq, max_ctx, block_size);
}

extern "C" int b200_paged_attn_decode(const void* q, int64_t ldq, const void* kc, const void* vc,
                                      const int32_t* block_table, int32_t bt_stride, const int32_t* row_to_seq,
                                      const int32_t* ctx_lens, void* out, int64_t ldo, int32_t T, int32_t hq,
                                      int32_t hkv, int32_t block_size, int32_t max_ctx, float scale, void* ws,
                                      int64_t ws_bytes, void* stream) {
  B200_REQUIRE(q && kc && vc && block_table && ctx_lens && out, B200_ERR_INVALID, "paged_attn: null");
  return paged_attn((const __nv_bfloat16*)q, ldq, (const __nv_bfloat16*)kc, (const __nv_bfloat16*)vc, block_table,
                    bt_stride, row_to_seq, ctx_lens, (__nv_bfloat16*)out, ldo, T, hq, hkv, block_size, max_ctx, scale,
                    ws, ws_bytes, (cudaStream_t)stream);
}

extern "C" int b200_paged_attn_prefill(const void* q, int64_t ldq, const void* kc, const void* vc,
                                       const int32_t* block_table, int32_t bt_stride, const int32_t* row_to_seq,
                                       const int32_t* ctx_lens, void* out, int64_t ldo, int32_t T, int32_t hq,
                                       int32_t hkv, int32_t block_size, int32_t max_ctx, float scale, void* ws,
                                       int64_t ws_bytes, void* stream) {
  B200_REQUIRE(q && kc && vc && block_table && ctx_lens && out, B200_ERR_INVALID, "paged_attn_prefill: null");
  return paged_attn_ex((const __nv_bfloat16*)q, ldq, (const __nv_bfloat16*)kc, (const __nv_bfloat16*)vc, block_table,
                       bt_stride, row_to_seq, ctx_lens, (__nv_bfloat16*)out, ldo, T, hq, hkv, block_size, max_ctx, scale,
                       ws, ws_bytes, 1, (cudaStream_t)stream);
}

extern "C" int b200_silu_mul(const void* gu, void* out, int32_t T, int32_t inter, void* stream) {
  B200_REQUIRE(gu && out, B200_ERR_INVALID, "silu_mul: null");
  return silu_mul((const __nv_bfloat16*)gu, (__nv_bfloat16*)out, T, inter, (cudaStream_t)stream);
}

extern "C" int b200_argmax(const void* logits, int64_t ld, int32_t rows, int32_t vocab, int32_t index_offset,
                           int32_t* out_ids, float* out_max, void* stream) {
  B200_REQUIRE(logits && out_ids, B200_ERR_INVALID, "argmax: null");
  return argmax_rows((const __nv_bfloat16*)logits, ld, rows, vocab, index_offset, out_ids, out_max,
                     (cudaStream_t)stream);
}

extern "C" int b200_advance_decode_rows(int32_t* ids, int32_t* positions, int32_t* slot_mapping, int32_t* ctx_lens,
                                        const int32_t* block_table, int32_t bt_stride, const int32_t* row_to_seq,
                                        const int32_t* sampled, int32_t n_rows, int32_t block_size, void* stream) {
  B200_REQUIRE(ids && positions && slot_mapping && ctx_lens && block_table && sampled && block_size > 0, B200_ERR_INVALID,
               "advance_decode_rows: bad argument");
  return advance_rows(ids, positions, slot_mapping, ctx_lens, block_table, bt_stride, row_to_seq, sampled, n_rows,
                      block_size, (cudaStream_t)stream);
}

// Row-parallel linears write into partial_dst(); tp_reduce() turns the partials of all ranks into c->x.
// The one-shot peer-memory kernel wins where the all-reduce is latency bound (small messages); for large
// messages NCCL's bandwidth-optimal algorithms take over.  Measured at TP8 (70B): 16 KiB messages 6.9 ms per step
// one-shot vs 9.7 ms NCCL; 1 MiB messages 11.9 vs 15.0 ms — so the switch-over sits at 2 MiB (B200_PEER_AR_MAX_KB).
static bool use_peer_ar(const b200_ctx* c, int T) {
  static const int64_t max_kb = getenv("B200_PEER_AR_MAX_KB") ? atoi(getenv("B200_PEER_AR_MAX_KB")) : 2048;
  return c->ar_ready && (int64_t)T * c->cfg.hidden * 2 <= max_kb * 1024;
}
static __nv_bfloat16* partial_dst(b200_ctx* c, int buf_idx, int T) {
  return use_peer_ar(c, T) ? reinterpret_cast<__nv_bfloat16*>(c->ar_mem + (int64_t)buf_idx * c->ar_buf_bytes) : c->x;
}
static ArParams ar_params(b200_ctx* c, int buf_idx, int T) {
  ArParams p = {};
  for (int r = 0; r < c->cfg.tp_world; ++r) {
    p.bufs[r] = reinterpret_cast<const __nv_bfloat16*>(c->ar_peer[r] + (int64_t)buf_idx * c->ar_buf_bytes);
    p.flags[r] = reinterpret_cast<unsigned int*>(c->ar_peer[r] + 4 * c->ar_buf_bytes);
    p.gather[r] = reinterpret_cast<__nv_bfloat16*>(c->ar_peer[r] + (int64_t)(2 + buf_idx) * c->ar_buf_bytes);
  }
  p.state = reinterpret_cast<unsigned int*>(c->ar_mem + 4 * c->ar_buf_bytes + kArFlagBytes);
  p.out = c->x;
  p.n_chunks = (int64_t)T * c->cfg.hidden / 8;
  p.rank = c->cfg.tp_rank;
  p.world = c->cfg.tp_world;
  p.buf_idx = buf_idx;
  return p;
}
static int tp_reduce(b200_ctx* c, int buf_idx, int T, cudaStream_t s) {
  const int64_t count = (int64_t)T * c->cfg.hidden;
  if (!use_peer_ar(c, T)) return b200_allreduce_bf16(c, c->x, count, s);   // NCCL
  return allreduce_oneshot(ar_params(c, buf_idx, T), s);
}
// all-reduce of buffer buf_idx + residual add + RMSNorm; *normed receives where the normed rows are.  Peer memory:
// ONE kernel does collective + add + norm — up to 16 rows every rank pulls everything (one flag round, latency bound),
// above that the rows are dealt to the ranks (two-shot: 1/world of the NVLink traffic, the normed rows land in this
// rank's gather buffer).  NCCL or a single rank: the all-reduce followed by the add+RMSNorm kernel.
static int tp_reduce_norm(b200_ctx* c, int buf_idx, int T, const __nv_bfloat16* w, const __nv_bfloat16** normed, cudaStream_t s) {
  static const int fuse_env = getenv("B200_FUSE_AR_NORM") ? atoi(getenv("B200_FUSE_AR_NORM")) : 1;   // 70B TP8 batch 1: 6.41 ms/step fused vs 6.88 ms as two kernels
  static const int twoshot_env = getenv("B200_AR_TWOSHOT") ? atoi(getenv("B200_AR_TWOSHOT")) : 1;
  const b200_model_cfg& m = c->cfg;
  *normed = c->normed;
  const bool peer = m.tp_world > 1 && fuse_env && c->tap_layer < 0 && use_peer_ar(c, T) && m.hidden <= 8192;
  // Measured (70B, [64 x 8192] rows): TP8 two-shot 21.8 us vs one-shot 24.4 us per call; TP2 two-shot 18.8 us, where the
  // one-shot pull moves a single MiB: the second flag round only pays once the pull would move several MiB per rank.
  if (peer && twoshot_env && T > 16 && (int64_t)(m.tp_world - 1) * T * m.hidden * 2 >= (4 << 20)) {
    const ArParams p = ar_params(c, buf_idx, T);
    *normed = p.gather[m.tp_rank];
    return allreduce_twoshot_rmsnorm(p, T, c->res, w, m.hidden, m.rms_eps, s);
  }
  if (peer)   // one CTA per row pulls that row from every rank: one flag round
    return allreduce_add_rmsnorm(ar_params(c, buf_idx, T), T, c->res, w, c->normed, m.hidden, m.rms_eps, s);
  if (m.tp_world > 1) {
    int rc = tp_reduce(c, buf_idx, T, s);
    if (rc) return rc;
  }
  return rmsnorm(c->x, c->res, w, c->normed, T, m.hidden, m.rms_eps, s);
}

// Per-kernel entry to the step's tensor-parallel collective (tests, timing): `partial` ([T][hidden] bf16, this rank's
// row-parallel partial sums) is copied into the exchange buffer the O (buf_idx 0) / down (1) projection writes, then the
// same code path as inside b200_decode_step runs: with norm_w the all-reduce + residual add + RMSNorm (res updated in
// place, y = normed rows), without it the plain all-reduce into y.
extern "C" int b200_tp_allreduce(b200_ctx* c, int32_t buf_idx, const void* partial, int32_t T, void* res, const void* norm_w,
                                 void* y, void* stream) {
  B200_REQUIRE(c && T > 0 && T <= c->cfg.max_tokens && (buf_idx == 0 || buf_idx == 1), B200_ERR_INVALID,
               "tp_allreduce: bad argument");
  B200_REQUIRE(c->cfg.tp_world > 1 && c->comm, B200_ERR_STATE, "tp_allreduce: needs tp_world > 1 and b200_nccl_init");
  B200_REQUIRE(norm_w || y, B200_ERR_INVALID, "tp_allreduce: the plain all-reduce needs an output");
  cudaStream_t s = (cudaStream_t)stream;
  const size_t bytes = (size_t)T * c->cfg.hidden * 2;
  if (partial) B200_CUDA(cudaMemcpyAsync(partial_dst(c, buf_idx, T), partial, bytes, cudaMemcpyDeviceToDevice, s));
  int rc;
  if (norm_w) {
    // res == NULL / y == NULL: stay on the ctx's own residual and normed rows (timing: the collective kernel alone)
    if (res) B200_CUDA(cudaMemcpyAsync(c->res, res, bytes, cudaMemcpyDeviceToDevice, s));
    const __nv_bfloat16* normed = nullptr;
    rc = tp_reduce_norm(c, buf_idx, T, (const __nv_bfloat16*)norm_w, &normed, s);
    if (rc) return rc;
    if (res) B200_CUDA(cudaMemcpyAsync(res, c->res, bytes, cudaMemcpyDeviceToDevice, s));   // two-shot: only this rank's rows (t % world == rank) are current
    if (y) B200_CUDA(cudaMemcpyAsync(y, normed, bytes, cudaMemcpyDeviceToDevice, s));
  } else {
    rc = tp_reduce(c, buf_idx, T, s);
    if (rc) return rc;
    B200_CUDA(cudaMemcpyAsync(y, c->x, bytes, cudaMemcpyDeviceToDevice, s));
  }
  return B200_OK;
}

// ---- the step --------------------------------------------------------------------------------------
struct StepArgs {
  const int32_t *ids, *positions, *slots, *block_table, *row_to_seq, *ctx_lens, *logit_rows;
  int bt_stride, T, max_ctx, R;
  int32_t* out_tokens;
  void* out_logits;
  int n_seqs = 0;   // sequences in the step when the caller knows it (host-buffer entry), else 0
  // host-buffer entry: the tokens the previous step left in the ctx; an id -(j+1) means "row j of those"
  const int32_t* prev_tokens = nullptr;
  int n_prev = 0;
};

static int enqueue_step(b200_ctx* c, const StepArgs& a, cudaStream_t s) {
  const b200_model_cfg& m = c->cfg;
  const int T = a.T, h = m.hidden;
  const bool tp = m.tp_world > 1;
  const float scale = 1.0f / sqrtf((float)kHeadDim);
  const int64_t kv_layer_elems = c->n_blocks * c->hkv * (int64_t)m.block_size * kHeadDim;
  int rc;
#define STEP(expr)        \
  do {                    \
    rc = (expr);          \
    if (rc) return rc;    \
  } while (0)

  const int qn = c->qkv_w, on = h, ok = c->hq * kHeadDim, gn = 2 * c->inter, dn = h, dk = c->inter;
  // TP: the all-reduce after the down projection is folded into the norm that consumes it (next layer's ln1 or the
  // final norm); with a debug tap it runs right after the projection so that the tap sees reduced activations
  const bool defer_ar = tp && c->tap_layer < 0;
  // Decode-sized steps on one GPU keep the residual stream in fp32 and fold every RMSNorm into its neighbours: the O
  // and down projections add their output into the stream and leave per-slab sums of squares, the QKV, gate_up and
  // lm_head linears normalise their input rows on the fly (same rounding points as the norm kernel, R1, and the same
  // summation order: bit-identical to the unfused step).  Two launches fewer per layer.
  const bool fnorm = !tp && c->fuse && c->tap_layer < 0 && T <= std::min(c->fuse_norm_max_t, 64) && h % 128 == 0 &&
                     tc_path(T, qn, h, c->gemv_max_t) && tc_path(T, on, ok, c->gemv_max_t) &&
                     tc_path(T, gn, h, c->gemv_max_t) && tc_path(T, dn, dk, c->gemv_max_t) &&
                     tc_path(a.R, c->vocab_local, h, c->gemv_max_t);
  auto norm_in = [&](LinearExtra& ex, const __nv_bfloat16* w) {
    ex.norm_in = true;
    ex.sres = c->sres;
    ex.ssq = c->ssq;
    ex.norm_w = w;
    ex.eps = m.rms_eps;
  };
  // Prefill-heavy steps (many more rows than sequences) pack consecutive rows of a sequence into one pass
  // over its K/V; decode steps keep one row per work item.
  // Decode-only steps (every row its own sequence: one new token each) are flagged with -1: the attention may
  // then request the cache tiles older than that token before the QKV kernel has completed.
  int attn_pack = a.row_to_seq && (a.n_seqs > 0 ? (T >= 4 * a.n_seqs && T >= 32) : T >= 512);
  if (!attn_pack && (a.row_to_seq == nullptr || a.n_seqs == T)) attn_pack = -1;
  if (attn_pack > 0) {
    // the prefill attention's row plan depends on the step, not on the layer: build it once
    const int built = paged_attn_prefill_plan_step(a.row_to_seq, a.ctx_lens, T, c->hq, c->hkv, m.block_size, c->attn_ws,
                                                   c->attn_ws_bytes, s);
    if (built < 0) return built;
    if (built == 1) attn_pack = 2;
  }
  const __nv_bfloat16* normed = c->normed;   // where the current normed rows are (the two-shot all-reduce delivers them elsewhere)
  if (fnorm)
    STEP(embed_stream(a.ids, c->embed, c->sres, c->ssq, T, h, m.vocab, s, a.prev_tokens, a.n_prev));
  else
    STEP(embed(a.ids, c->embed, c->res, T, h, m.vocab, s, a.prev_tokens, a.n_prev));
  for (int l = 0; l < m.n_layers; ++l) {
    const LayerW& L = c->layers[l];
    if (!fnorm) {
      if (l == 0)
        STEP(rmsnorm(c->res, nullptr, L.ln1, c->normed, T, h, m.rms_eps, s));
      else
        STEP(defer_ar ? tp_reduce_norm(c, 1, T, L.ln1, &normed, s) : rmsnorm(c->x, c->res, L.ln1, c->normed, T, h, m.rms_eps, s));
    }
    __nv_bfloat16* kcl = c->kc + (int64_t)l * kv_layer_elems;
    __nv_bfloat16* vcl = c->vc + (int64_t)l * kv_layer_elems;
    {  // QKV (+ RoPE + KV append in the epilogue when on the tcgen05 path)
      LinearExtra ex;
      const bool tcp = tc_path(T, qn, h, c->gemv_max_t);
      if (tcp && c->fuse) {
        ex.mode = 2;
        ex.positions = a.positions;
        ex.slots = a.slots;
        ex.cos_sin = c->cos_sin;
        ex.kc = kcl;
        ex.vc = vcl;
        ex.hq = c->hq;
        ex.hkv = c->hkv;
        ex.block_size = m.block_size;
      }
      if (fnorm) norm_in(ex, L.ln1);
      STEP(linear_any(normed, h, L.qkv, L.qkv_b, c->qkv_buf, qn, T, qn, h, 0, c->gemv_max_t, c->lin_ws,
                      c->lin_ws_bytes, &ex, s));
      if (ex.mode != 2)
        STEP(rope_kv_append(c->qkv_buf, a.positions, a.slots, c->cos_sin, kcl, vcl, T, c->hq, c->hkv, m.block_size, s));
    }
    STEP(paged_attn_ex(c->qkv_buf, c->qkv_w, kcl, vcl, a.block_table, a.bt_stride, a.row_to_seq, a.ctx_lens, c->attn_out,
                       c->hq * kHeadDim, T, c->hq, c->hkv, m.block_size, a.max_ctx, scale, c->attn_ws, c->attn_ws_bytes,
                       attn_pack, s));
    {  // O
      LinearExtra ex;
      if (fnorm) {
        ex.mode = 3;
        ex.sres = c->sres;
        ex.ssq = c->ssq;
      }
      STEP(linear_any(c->attn_out, ok, L.o, nullptr, tp ? partial_dst(c, 0, T) : c->x, h, T, on, ok, 0, c->gemv_max_t,
                      c->lin_ws, c->lin_ws_bytes, &ex, s));
    }
    if (!fnorm) STEP(tp_reduce_norm(c, 0, T, L.ln2, &normed, s));   // (all-reduce +) residual add + RMSNorm
    {  // gate_up (+ SiLU*mul in the epilogue over interleaved slabs)
      LinearExtra ex;
      const bool tcp = tc_path(T, gn, h, c->gemv_max_t);
      if (fnorm) norm_in(ex, L.ln2);
      if (L.gu_interleaved) {
        B200_REQUIRE(tcp, B200_ERR_UNSUPPORTED, "interleaved gate_up weights need the tcgen05 path");
        ex.mode = 1;
        STEP(linear_any(normed, h, L.gu, nullptr, c->act, c->inter, T, gn, h, 0, c->gemv_max_t, c->lin_ws,
                        c->lin_ws_bytes, &ex, s));
      } else {
        STEP(linear_any(normed, h, L.gu, nullptr, c->gu_buf, gn, T, gn, h, 0, c->gemv_max_t, c->lin_ws,
                        c->lin_ws_bytes, &ex, s));
        STEP(silu_mul(c->gu_buf, c->act, T, c->inter, s));
      }
    }
    {  // down
      LinearExtra ex;
      if (fnorm) {
        ex.mode = 3;
        ex.sres = c->sres;
        ex.ssq = c->ssq;
      }
      STEP(linear_any(c->act, dk, L.down, nullptr, tp ? partial_dst(c, 1, T) : c->x, h, T, dn, dk, 0, c->gemv_max_t,
                      c->lin_ws, c->lin_ws_bytes, &ex, s));
    }
    if (tp && !defer_ar) STEP(tp_reduce(c, 1, T, s));   // else: folded into the next norm (ln1 of layer l+1 / final norm)
    if (l == c->tap_layer && c->tap_hidden) {
      B200_CUDA(cudaMemcpyAsync(c->tap_hidden, c->x, (size_t)T * h * 2, cudaMemcpyDeviceToDevice, s));
      if (c->tap_res) B200_CUDA(cudaMemcpyAsync(c->tap_res, c->res, (size_t)T * h * 2, cudaMemcpyDeviceToDevice, s));
    }
  }
  if (fnorm) {
    // lm_head normalises its own rows (through the logit-row indirection) while the rows are few; with many logit
    // rows a thousand slabs would each redo the same rows, so the norm runs once as a kernel of its own
    LinearExtra ex;
    const __nv_bfloat16* lm_in = c->lm_in;
    if (a.R <= 16) {
      norm_in(ex, c->final_norm);
      ex.rows = a.logit_rows;
    } else {
      STEP(rmsnorm_stream(c->sres, c->ssq, a.logit_rows, c->final_norm, c->lm_in, a.R, h, m.rms_eps, s));
    }
    STEP(linear_any(lm_in, h, c->lm_head, nullptr, c->logits, c->vocab_local, a.R, c->vocab_local, h, 0, c->gemv_max_t,
                    c->lin_ws, c->lin_ws_bytes, &ex, s));
  } else {
    STEP(defer_ar ? tp_reduce_norm(c, 1, T, c->final_norm, &normed, s) : rmsnorm(c->x, c->res, c->final_norm, c->normed, T, h, m.rms_eps, s));
    const __nv_bfloat16* lm_in = normed;
    if (a.logit_rows) {
      STEP(embed(a.logit_rows, normed, c->lm_in, a.R, h, T, s));  // row gather
      lm_in = c->lm_in;
    }
    STEP(linear_any(lm_in, h, c->lm_head, nullptr, c->logits, c->vocab_local, a.R, c->vocab_local, h, 0, c->gemv_max_t,
                    c->lin_ws, c->lin_ws_bytes, nullptr, s));
  }
  if (!tp) {
    STEP(argmax_rows(c->logits, c->vocab_local, a.R, c->vocab_local, 0, a.out_tokens, nullptr, s));
    if (a.out_logits)
      B200_CUDA(cudaMemcpyAsync(a.out_logits, c->logits, (size_t)a.R * c->vocab_local * 2, cudaMemcpyDeviceToDevice, s));
  } else {
    // vocab-parallel greedy: per-rank (max, global index), all-gather of R pairs, pick
    STEP(argmax_rows(c->logits, c->vocab_local, a.R, c->vocab_local, m.tp_rank * c->vocab_local, c->d_tokens, c->d_max, s));
    B200_CUDA(launch(tp_pack_kernel, dim3((a.R + 127) / 128), dim3(128), 0, s, (const float*)c->d_max,
                     (const int32_t*)c->d_tokens, a.R, c->d_pairs));
    ++launch_counter();
    B200_NCCL(nccl().AllGather(c->d_pairs, c->d_gathered, (size_t)a.R * 2, ncclFloat32, c->comm, s));
    B200_CUDA(launch(tp_pick_kernel, dim3((a.R + 127) / 128), dim3(128), 0, s, (const float*)c->d_gathered, m.tp_world,
                     a.R, a.out_tokens));
    if (a.out_logits) {
      // [world][R][vocab_local] rank-major; the host side re-interleaves (test/debug path only)
      ++launch_counter();
      B200_NCCL(nccl().AllGather(c->logits, a.out_logits, (size_t)a.R * c->vocab_local, ncclBfloat16, c->comm, s));
    }
  }
#undef STEP
  return B200_OK;
}

static int pow2_bucket(int v) {
  int p = 128;
  while (p < v) p <<= 1;
  return p;
}

static int run_step(b200_ctx* c, const StepArgs& a, cudaStream_t s) {
  const bool pdl_prev = launch_opts().pdl;   // the caller's per-thread setting is restored on every exit
  B200_REQUIRE(c->embed && c->final_norm && c->lm_head && c->cos_sin && c->kc, B200_ERR_STATE,
               "decode_step: weights / rope cache / kv cache not all set");
  for (int l = 0; l < c->cfg.n_layers; ++l) {
    const LayerW& L = c->layers[l];
    B200_REQUIRE(L.ln1 && L.qkv && L.o && L.ln2 && L.gu && L.down && (!c->cfg.qkv_bias || L.qkv_b), B200_ERR_STATE,
                 "decode_step: layer %d weights incomplete", l);
  }
  B200_REQUIRE(a.T > 0 && a.T <= c->cfg.max_tokens, B200_ERR_INVALID, "decode_step: n_tokens=%d outside (0,%d]", a.T,
               c->cfg.max_tokens);
  B200_REQUIRE(a.R > 0 && a.R <= c->max_logit_rows && a.R <= a.T, B200_ERR_INVALID,
               "decode_step: n_logit_rows=%d outside (0,min(%d,n_tokens)]", a.R, c->max_logit_rows);
  B200_REQUIRE(a.max_ctx > 0 && a.max_ctx <= c->cfg.max_blocks_per_seq * c->cfg.block_size, B200_ERR_INVALID,
               "decode_step: max_ctx_len=%d outside (0,%d]", a.max_ctx, c->cfg.max_blocks_per_seq * c->cfg.block_size);
  if (c->cfg.tp_world > 1) B200_REQUIRE(c->comm, B200_ERR_STATE, "decode_step: b200_nccl_init not called");

  const bool use_graph = c->graphs && s != nullptr && c->tap_layer < 0;
  launch_opts().pdl = getenv("B200_PDL") ? atoi(getenv("B200_PDL")) != 0 : c->pdl;
  int rc = B200_OK;
  if (!use_graph) {
    launch_counter() = 0;
    rc = enqueue_step(c, a, s);
    c->last_launches = launch_counter();
    launch_opts().pdl = pdl_prev;
    return rc;
  }
  const GraphKey key{{(uint64_t)a.ids,        (uint64_t)a.positions,  (uint64_t)a.slots,
                      (uint64_t)a.block_table, (uint64_t)a.row_to_seq, (uint64_t)a.ctx_lens,
                      (uint64_t)a.logit_rows,  (uint64_t)a.out_tokens, (uint64_t)a.out_logits,
                      (uint64_t)a.bt_stride,   (uint64_t)a.T,          (uint64_t)pow2_bucket(a.max_ctx),
                      (uint64_t)a.R,           (uint64_t)launch_opts().pdl, (uint64_t)a.n_seqs, (uint64_t)a.prev_tokens}};
  auto it = c->graph_cache.find(key);
  if (it == c->graph_cache.end()) {
    StepArgs b = a;
    b.max_ctx = pow2_bucket(a.max_ctx);
    cudaGraph_t g = nullptr;
    B200_CUDA(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
    launch_counter() = 0;
    rc = enqueue_step(c, b, s);
    const int n = launch_counter();
    cudaError_t e = cudaStreamEndCapture(s, &g);
    launch_opts().pdl = pdl_prev;
    if (rc) {
      if (g) cudaGraphDestroy(g);
      return rc;
    }
    if (e != cudaSuccess) return cuda_fail(e, "cudaStreamEndCapture", __FILE__, __LINE__);
    cudaGraphExec_t ge = nullptr;
    e = cudaGraphInstantiate(&ge, g, 0);
    cudaGraphDestroy(g);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGraphInstantiate", __FILE__, __LINE__);
    if (c->graph_cache.size() > 256) {  // bound the cache: drop everything, shapes re-capture on demand
      for (auto& kv : c->graph_cache) cudaGraphExecDestroy(kv.second);
      c->graph_cache.clear();
    }
    it = c->graph_cache.emplace(key, ge).first;
    c->last_launches = n;
  }
  launch_opts().pdl = pdl_prev;
  B200_CUDA(cudaGraphLaunch(it->second, s));
  return B200_OK;
}

// (a10, widened) sampling ---------------------------------------------------------------------------------
extern "C" int b200_sample(const void* logits, int64_t ld, int32_t n_rows, int32_t vocab, const float* temperature,
                           const int32_t* top_k, const float* top_p, const float* min_p, const uint64_t* seeds,
                           const uint64_t* draws, int32_t* out_ids, void* stream) {
  B200_REQUIRE(n_rows >= 0 && vocab > 0 && ld >= vocab, B200_ERR_INVALID, "sample: bad sizes (rows=%d vocab=%d ld=%lld)",
               n_rows, vocab, (long long)ld);
  return sample_rows((const __nv_bfloat16*)logits, ld, n_rows, vocab, vocab, 0, temperature, top_k, top_p, min_p, seeds,
                     draws, out_ids, (cudaStream_t)stream);
}

// The rows are the logit rows of the LAST step of this ctx (its logits are still in the ctx's buffer).
// host_params: n_rows x {float temperature; int32 top_k; float top_p; uint32 pad; uint64 seed; uint64 draw} as
// six arrays back to back: temperature[n] | top_k[n] | top_p[n] | min_p[n] | seed[n] | draw[n]  (32 n bytes).
extern "C" int b200_ctx_sample_last(b200_ctx* c, int32_t n_rows, const void* host_params, int32_t* host_out_ids,
                                    void* stream) {
  B200_REQUIRE(c && host_params && host_out_ids, B200_ERR_INVALID, "sample_last: null argument");
  B200_REQUIRE(n_rows > 0 && n_rows <= c->max_logit_rows, B200_ERR_INVALID, "sample_last: n_rows=%d outside (0,%d]", n_rows,
               c->max_logit_rows);
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t n = n_rows;
  B200_CUDA(cudaMemcpyAsync(c->d_sample, host_params, (size_t)n * 32, cudaMemcpyHostToDevice, s));
  const char* base = static_cast<const char*>(c->d_sample);
  const float* temperature = reinterpret_cast<const float*>(base);
  const int32_t* top_k = reinterpret_cast<const int32_t*>(base + 4 * n);
  const float* top_p = reinterpret_cast<const float*>(base + 8 * n);
  const float* min_p = reinterpret_cast<const float*>(base + 12 * n);   // 0 = off
  const uint64_t* seeds = reinterpret_cast<const uint64_t*>(base + 16 * n);
  const uint64_t* draws = reinterpret_cast<const uint64_t*>(base + 24 * n);
  const int W = c->cfg.tp_world;
  int rc;
  if (W > 1) {
    // vocab-parallel logits: gather the shards rank-major; every rank then draws the same token (the kernel
    // is deterministic given the seed)
    B200_NCCL(nccl().AllGather(c->logits, c->logits_full, (size_t)n * c->vocab_local, ncclBfloat16, c->comm, s));
    rc = sample_rows(c->logits_full, c->vocab_local, n_rows, c->vocab_local * W, c->vocab_local, n * c->vocab_local,
                     temperature, top_k, top_p, min_p, seeds, draws, c->d_tokens, s);
  } else {
    rc = sample_rows(c->logits, c->vocab_local, n_rows, c->vocab_local, c->vocab_local, 0, temperature, top_k, top_p, min_p,
                     seeds, draws, c->d_tokens, s);
  }
  if (rc) return rc;
  B200_CUDA(cudaMemcpyAsync(host_out_ids, c->d_tokens, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
  return B200_OK;
}

extern "C" int b200_logit_adjust(void* logits, int64_t ld, int32_t n_rows, int32_t vocab, const float* row_params,
                                 const int32_t* offsets, const int32_t* ids, const int32_t* counts, const int32_t* seen,
                                 const float* bias, void* stream) {
  B200_REQUIRE(n_rows >= 0 && vocab > 0 && ld >= vocab, B200_ERR_INVALID, "logit_adjust: bad sizes (rows=%d vocab=%d ld=%lld)", n_rows,
               vocab, (long long)ld);
  return logit_adjust_rows((__nv_bfloat16*)logits, ld, n_rows, vocab, 0, row_params, offsets, ids, counts, seen, bias,
                           (cudaStream_t)stream);
}

// Penalties / logit_bias over the logit rows of the LAST step (still in the ctx's buffer), before b200_ctx_sample_last.
// host_packed (int32 words): float row_params[3 n] | int32 offsets[n + 1] | int32 ids[E] | int32 counts[E] | int32 seen[E]
// | float bias[E], E = offsets[n] = n_entries.  With tensor parallelism every rank adjusts its own vocabulary shard.
extern "C" int b200_ctx_adjust_last(b200_ctx* c, int32_t n_rows, const void* host_packed, int32_t n_entries, void* stream) {
  B200_REQUIRE(c && host_packed, B200_ERR_INVALID, "adjust_last: null argument");
  B200_REQUIRE(n_rows > 0 && n_rows <= c->max_logit_rows && n_entries >= 0, B200_ERR_INVALID, "adjust_last: n_rows=%d outside (0,%d]",
               n_rows, c->max_logit_rows);
  const int64_t n = n_rows, E = n_entries, words = 3 * n + n + 1 + 4 * E;
  B200_REQUIRE(words <= kAdjustMaxInts, B200_ERR_UNSUPPORTED, "adjust_last: %lld entries exceed the %lld-word buffer", (long long)E,
               (long long)kAdjustMaxInts);
  cudaStream_t s = (cudaStream_t)stream;
  B200_CUDA(cudaMemcpyAsync(c->d_adjust, host_packed, (size_t)words * 4, cudaMemcpyHostToDevice, s));
  const int32_t* w = static_cast<const int32_t*>(c->d_adjust);
  const float* row_params = reinterpret_cast<const float*>(w);
  const int32_t* offsets = w + 3 * n;
  const int32_t* ids = offsets + n + 1;
  return logit_adjust_rows(c->logits, c->vocab_local, n_rows, c->vocab_local, c->cfg.tp_rank * c->vocab_local, row_params, offsets, ids,
                           ids + E, ids + 2 * E, reinterpret_cast<const float*>(ids + 3 * E), s);
}

extern "C" int b200_logprobs(const void* logits, int64_t ld, int32_t n_rows, int32_t vocab, const int32_t* token_ids,
                             int32_t n_top, float* out_logprob, int32_t* out_top_ids, float* out_top_logprobs, void* stream) {
  B200_REQUIRE(n_rows >= 0 && vocab > 0 && ld >= vocab, B200_ERR_INVALID, "logprobs: bad sizes (rows=%d vocab=%d ld=%lld)",
               n_rows, vocab, (long long)ld);
  return logprob_rows((const __nv_bfloat16*)logits, ld, n_rows, vocab, vocab, 0, token_ids, n_top, out_logprob, out_top_ids,
                      out_top_logprobs, (cudaStream_t)stream);
}

// Log-probabilities over the logit rows of the LAST step (see b200_ctx_sample_last for the TP gather).
// host_out: float logprob[n] | int32 top_ids[n][n_top] | float top_logprobs[n][n_top], contiguous.
extern "C" int b200_ctx_logprobs_last(b200_ctx* c, int32_t n_rows, const int32_t* host_token_ids, int32_t n_top,
                                      void* host_out, void* stream) {
  B200_REQUIRE(c && host_token_ids && host_out, B200_ERR_INVALID, "logprobs_last: null argument");
  B200_REQUIRE(n_rows > 0 && n_rows <= c->max_logit_rows && n_top >= 0 && n_top <= 32, B200_ERR_INVALID,
               "logprobs_last: n_rows=%d outside (0,%d] or n_top=%d outside [0,32]", n_rows, c->max_logit_rows, n_top);
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t n = n_rows;
  char* base = static_cast<char*>(c->d_logprob);
  int32_t* d_tok = reinterpret_cast<int32_t*>(base);
  float* d_lp = reinterpret_cast<float*>(base + 4 * n);
  int32_t* d_ids = reinterpret_cast<int32_t*>(base + 8 * n);
  float* d_tlp = reinterpret_cast<float*>(base + 8 * n + 4 * n * n_top);
  B200_CUDA(cudaMemcpyAsync(d_tok, host_token_ids, (size_t)n * 4, cudaMemcpyHostToDevice, s));
  const int W = c->cfg.tp_world;
  int rc;
  if (W > 1) {
    B200_NCCL(nccl().AllGather(c->logits, c->logits_full, (size_t)n * c->vocab_local, ncclBfloat16, c->comm, s));
    rc = logprob_rows(c->logits_full, c->vocab_local, n_rows, c->vocab_local * W, c->vocab_local, n * c->vocab_local, d_tok,
                      n_top, d_lp, d_ids, d_tlp, s);
  } else {
    rc = logprob_rows(c->logits, c->vocab_local, n_rows, c->vocab_local, c->vocab_local, 0, d_tok, n_top, d_lp, d_ids, d_tlp, s);
  }
  if (rc) return rc;
  B200_CUDA(cudaMemcpyAsync(host_out, d_lp, (size_t)n * 4 * (1 + 2 * n_top), cudaMemcpyDeviceToHost, s));
  return B200_OK;
}

extern "C" int b200_decode_step(b200_ctx* c, const int32_t* ids, const int32_t* positions, const int32_t* slot_mapping,
                                const int32_t* block_table, int32_t bt_stride, const int32_t* row_to_seq,
                                const int32_t* ctx_lens, int32_t n_tokens, int32_t max_ctx_len,
                                const int32_t* logit_rows, int32_t n_logit_rows, int32_t* out_token_ids,
                                void* out_logits, void* stream) {
  B200_REQUIRE(c && ids && positions && slot_mapping && block_table && ctx_lens && out_token_ids, B200_ERR_INVALID,
               "decode_step: null argument");
  StepArgs a{ids, positions, slot_mapping, block_table, row_to_seq, ctx_lens, logit_rows, bt_stride, n_tokens,
             max_ctx_len, logit_rows ? n_logit_rows : n_tokens, out_token_ids, out_logits};
  return run_step(c, a, (cudaStream_t)stream);
}

extern "C" int b200_decode_step_host(b200_ctx* c, const int32_t* host_packed, int32_t n_tokens, int32_t n_seqs,
                                     int32_t bt_stride, int32_t max_ctx_len, int32_t n_logit_rows,
                                     int32_t* host_out_token_ids, void* stream) {
  B200_REQUIRE(c && host_packed && host_out_token_ids, B200_ERR_INVALID, "decode_step_host: null argument");
  B200_REQUIRE(n_tokens > 0 && n_tokens <= c->cfg.max_tokens && n_seqs > 0 && n_seqs <= n_tokens && bt_stride > 0 &&
                   bt_stride <= c->cfg.max_blocks_per_seq && n_logit_rows > 0 && n_logit_rows <= n_tokens,
               B200_ERR_INVALID, "decode_step_host: bad sizes (T=%d seqs=%d bt_stride=%d rows=%d)", n_tokens, n_seqs,
               bt_stride, n_logit_rows);
  cudaStream_t s = (cudaStream_t)stream;
  // ONE host->device copy of the dense packed buffer; the device pointers are functions of
  // (T, rows, seqs, bt_stride) only, so steps of the same shape replay the same captured graph.
  const int64_t T = n_tokens;
  int32_t* d = c->d_meta;
  const int64_t n_ints = 5 * T + n_logit_rows + (int64_t)n_seqs * bt_stride;
  B200_REQUIRE(n_ints <= c->meta_ints, B200_ERR_INVALID, "decode_step_host: metadata larger than the ctx buffer");
  B200_CUDA(cudaMemcpyAsync(d, host_packed, (size_t)n_ints * 4, cudaMemcpyHostToDevice, s));
  StepArgs a{d, d + T, d + 2 * T, d + 5 * T + n_logit_rows, d + 4 * T, d + 3 * T, d + 5 * T, bt_stride, n_tokens,
             max_ctx_len, n_logit_rows, c->d_tokens, nullptr};
  a.n_seqs = n_seqs;
  a.prev_tokens = c->d_tokens;
  a.n_prev = c->max_logit_rows;
  int rc = run_step(c, a, s);
  if (rc) return rc;
  B200_CUDA(cudaMemcpyAsync(host_out_token_ids, c->d_tokens, (size_t)n_logit_rows * 4, cudaMemcpyDeviceToHost, s));
  return B200_OK;
}
