"""Device state of one TP rank and one scheduler step through the C-ABI.

Mirrors the worker/model-runner of the image the chart deploys (one worker process per TP rank,
/root/reference/core/helm-charts/vllm/xeon-values.yaml:78-79; [V22] vllm/v1/worker/*model_runner*):
it owns the weights, the paged KV cache and the step inputs, and runs `execute_model` — here
`b200_decode_step_host` — once per scheduler iteration.  torch allocates; CUDA computes.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import _abi
from .config import ModelConfig, rope_cos_sin


@dataclass
class StepBatch:
    """Step inputs.  One entry per token row; block_tables one row per sequence."""
    ids: List[int] = field(default_factory=list)
    positions: List[int] = field(default_factory=list)
    slots: List[int] = field(default_factory=list)
    ctx_lens: List[int] = field(default_factory=list)
    row_to_seq: List[int] = field(default_factory=list)
    logit_rows: List[int] = field(default_factory=list)
    block_tables: List[Sequence[int]] = field(default_factory=list)

    @property
    def n_tokens(self):
        return len(self.ids)

    @property
    def max_ctx(self):
        c = self.ctx_lens
        return int(c.max()) if isinstance(c, np.ndarray) else max(c)


class ModelRunner:
    def __init__(self, cfg: ModelConfig, device: Optional[torch.device] = None, *, max_tokens: int = 2048,
                 max_model_len: int = 8192, block_size: int = 128, n_blocks: Optional[int] = None,
                 kv_cache_gib: Optional[float] = None, tp_rank: int = 0, tp_world: int = 1,
                 graphs: bool = True):
        if not torch.cuda.is_available():
            raise RuntimeError("ModelRunner needs a CUDA device (B200, sm_100a); there is no CPU path")
        self.lib = _abi.lib()
        self.cfg = cfg
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        torch.cuda.set_device(self.device)
        if not self.lib.b200_device_ok():
            raise RuntimeError("the current CUDA device is not compute capability 10.x (B200)")
        self.block_size = block_size
        self.max_tokens = max_tokens
        self.max_model_len = max_model_len
        self.max_blocks_per_seq = (max_model_len + block_size - 1) // block_size
        self.tp_rank, self.tp_world = tp_rank, tp_world
        self.hq = cfg.n_q_heads // tp_world
        self.hkv = cfg.n_kv_heads // tp_world
        self.inter = cfg.intermediate // tp_world
        self.vocab_local = cfg.vocab // tp_world
        self.stream = torch.cuda.Stream(device=self.device)

        mc = _abi.ModelCfg(hidden=cfg.hidden, n_layers=cfg.n_layers, n_q_heads=cfg.n_q_heads,
                           n_kv_heads=cfg.n_kv_heads, head_dim=cfg.head_dim, intermediate=cfg.intermediate,
                           vocab=cfg.vocab, qkv_bias=int(cfg.qkv_bias), rms_eps=cfg.rms_eps,
                           block_size=block_size, max_tokens=max_tokens,
                           max_blocks_per_seq=self.max_blocks_per_seq, tp_rank=tp_rank, tp_world=tp_world)
        h = C.c_void_p()
        _abi.check(self.lib.b200_ctx_create(C.byref(mc), C.byref(h)))
        self.ctx = h
        _abi.check(self.lib.b200_ctx_set_graphs(self.ctx, int(graphs)))

        # KV cache [L][n_blocks][hkv][block][128] bf16, zero-filled (a masked tail never reads NaNs)
        per_block = cfg.n_layers * self.hkv * block_size * cfg.head_dim * 2 * 2
        if n_blocks is None:
            if kv_cache_gib is None:
                free, _ = torch.cuda.mem_get_info(self.device)
                kv_cache_gib = max(free / 2 ** 30 * 0.5, 0.25)
            n_blocks = max(int(kv_cache_gib * 2 ** 30) // per_block, 2)
        self.n_blocks = n_blocks
        shape = (cfg.n_layers, n_blocks, self.hkv, block_size, cfg.head_dim)
        self.k_cache = torch.zeros(shape, dtype=torch.bfloat16, device=self.device)
        self.v_cache = torch.zeros(shape, dtype=torch.bfloat16, device=self.device)
        _abi.check(self.lib.b200_set_kv_cache(self.ctx, self.k_cache.data_ptr(), self.v_cache.data_ptr(), n_blocks))
        self.cos_sin = rope_cos_sin(cfg, max(max_model_len, 16)).to(self.device)
        _abi.check(self.lib.b200_set_rope_cache(self.ctx, self.cos_sin.data_ptr(), self.cos_sin.shape[0]))

        self._weights: List[torch.Tensor] = []   # keeps the borrowed device tensors alive
        meta_ints = 6 * max_tokens + max_tokens * self.max_blocks_per_seq
        # two pinned (metadata, tokens) pairs: step_async alternates them, so the host can pack step n+1 while the
        # copies of step n are still in flight; the synchronous calls use pair 0
        self._h_meta2 = [torch.empty(meta_ints, dtype=torch.int32).pin_memory() for _ in range(2)]
        self._h_out2 = [torch.empty(max(max_tokens, 1), dtype=torch.int32).pin_memory() for _ in range(2)]
        self._done2 = [torch.cuda.Event() for _ in range(2)]
        self._flip = 0
        self._h_meta, self._h_out = self._h_meta2[0], self._h_out2[0]
        self._h_meta_np = self._h_meta.numpy()
        self._h_out_np = self._h_out.numpy()
        self.h2d_bytes_last = 0
        self.d2h_bytes_last = 0
        self.max_logit_rows = min(max_tokens, 512)
        self._h_sample = torch.zeros(self.max_logit_rows * 32, dtype=torch.uint8).pin_memory()
        self._h_sample_np = self._h_sample.numpy()
        self._h_lp_tok = torch.zeros(self.max_logit_rows, dtype=torch.int32).pin_memory()
        self._h_lp_out = torch.zeros(self.max_logit_rows * (1 + 2 * 32), dtype=torch.float32).pin_memory()

    # ---- weights ---------------------------------------------------------------------------
    def _set(self, layer: int, kind: int, t: torch.Tensor):
        t = t.to(device=self.device, dtype=torch.bfloat16).contiguous()
        self._weights.append(t)
        rows, cols = (t.shape[0], t.shape[1]) if t.dim() == 2 else (t.shape[0], 1)
        _abi.check(self.lib.b200_set_weight(self.ctx, layer, kind, t.data_ptr(), rows, cols))

    def load_weights(self, w: Dict[str, torch.Tensor]):
        """HF-named tensors, fused and sharded for this TP rank by weights.shard_weights."""
        from .weights import shard_weights
        for layer, kind, t in shard_weights(w, self.cfg, self.tp_rank, self.tp_world):
            self._set(layer, kind, t)

    def init_random_weights(self, seed: int = 1234, std: float = 0.02):
        """Random-init N(0, std^2) weights generated on the device (benchmarks; no checkpoint
        is reachable without network).  Norm weights are 1."""
        c = self.cfg
        g = torch.Generator(device=self.device)
        g.manual_seed(seed + 7919 * self.tp_rank)

        def rnd(*shape):
            t = torch.empty(shape, dtype=torch.bfloat16, device=self.device)
            t.normal_(0.0, std, generator=g)
            return t

        ones = torch.ones(c.hidden, dtype=torch.bfloat16, device=self.device)
        ge = torch.Generator(device=self.device)
        ge.manual_seed(seed)  # the embedding is replicated: same bytes on every rank
        emb = torch.empty((c.vocab, c.hidden), dtype=torch.bfloat16, device=self.device)
        emb.normal_(0.0, std, generator=ge)
        self._set(0, _abi.W_EMBED, emb)
        qkv_rows = (self.hq + 2 * self.hkv) * c.head_dim
        for l in range(c.n_layers):
            self._set(l, _abi.W_LN1, ones)
            self._set(l, _abi.W_QKV, rnd(qkv_rows, c.hidden))
            if c.qkv_bias:
                self._set(l, _abi.W_QKV_BIAS, rnd(qkv_rows))
            self._set(l, _abi.W_O, rnd(c.hidden, self.hq * c.head_dim))
            self._set(l, _abi.W_LN2, ones)
            self._set(l, _abi.W_GATE_UP_INTERLEAVED if self.inter % 64 == 0 else _abi.W_GATE_UP,
                      rnd(2 * self.inter, c.hidden))
            self._set(l, _abi.W_DOWN, rnd(c.hidden, self.inter))
        self._set(0, _abi.W_FINAL_NORM, ones)
        self._set(0, _abi.W_LM_HEAD, rnd(self.vocab_local, c.hidden))

    # ---- TP --------------------------------------------------------------------------------
    def init_tp(self, group=None, peer_allreduce: bool = None):
        """Create the ctx's NCCL communicator; the 128-byte id travels over torch.distributed."""
        if self.tp_world == 1:
            return
        import os
        import torch.distributed as dist
        if peer_allreduce is None:
            peer_allreduce = os.environ.get("B200_PEER_ALLREDUCE", "1") != "0"
        buf = (C.c_char * 128)()
        if self.tp_rank == 0:
            _abi.check(self.lib.b200_nccl_unique_id(buf))
        obj = [bytes(buf)]
        dist.broadcast_object_list(obj, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        idb = (C.c_char * 128).from_buffer_copy(obj[0])
        _abi.check(self.lib.b200_nccl_init(self.ctx, idb))
        if peer_allreduce:
            # exchange the CUDA-IPC handles of the all-reduce buffers: from here on the per-layer
            # all-reduces are one peer-memory kernel per rank instead of NCCL
            h = (C.c_char * 64)()
            _abi.check(self.lib.b200_tp_ipc_export(self.ctx, h))
            handles = [None] * self.tp_world
            dist.all_gather_object(handles, bytes(h), group=group)
            allh = (C.c_char * (64 * self.tp_world)).from_buffer_copy(b"".join(handles))
            _abi.check(self.lib.b200_tp_ipc_import(self.ctx, allh, self.tp_world))
            dist.barrier(group=group)

    # ---- one step ---------------------------------------------------------------------------
    def pack(self, b: StepBatch, into: Optional[np.ndarray] = None):
        """Write the dense packed int32 buffer b200_decode_step_host expects into pinned memory."""
        T, R, S = b.n_tokens, len(b.logit_rows), len(b.block_tables)
        tables = b.block_tables
        bt_stride = tables.shape[1] if isinstance(tables, np.ndarray) else max(max(len(r) for r in tables), 1)
        a = self._h_meta_np if into is None else into
        a[0:T] = b.ids
        a[T:2 * T] = b.positions
        a[2 * T:3 * T] = b.slots
        a[3 * T:4 * T] = b.ctx_lens
        a[4 * T:5 * T] = b.row_to_seq
        a[5 * T:5 * T + R] = b.logit_rows
        o = 5 * T + R
        bt = a[o:o + S * bt_stride].reshape(S, bt_stride)
        if isinstance(tables, np.ndarray):
            bt[:] = tables
        else:
            bt.fill(0)
            for i, row in enumerate(tables):
                bt[i, :len(row)] = row
        return T, R, S, bt_stride, o + S * bt_stride

    def step(self, b: StepBatch, sync: bool = True) -> np.ndarray:
        """Host-buffer step (the serving path and bench.py's e2e): pinned H2D, step, D2H."""
        T, R, S, bt_stride, n_ints = self.pack(b)
        return self.step_packed(T, R, S, bt_stride, b.max_ctx, n_ints, sync)

    def step_packed(self, T: int, R: int, S: int, bt_stride: int, max_ctx: int, n_ints: int, sync: bool = True) -> np.ndarray:
        """The step over whatever is in the pinned packed buffer (filled by pack(), or by a broadcast
        from rank 0 on a tensor-parallel worker)."""
        _abi.check(self.lib.b200_decode_step_host(self.ctx, self._h_meta.data_ptr(), T, S, bt_stride, max_ctx, R,
                                                  self._h_out.data_ptr(), self.stream.cuda_stream))
        self.h2d_bytes_last, self.d2h_bytes_last = n_ints * 4, R * 4
        if sync:
            self.stream.synchronize()
        return self._h_out_np[:R]

    def step_async(self, b: StepBatch) -> int:
        """Enqueue a host-buffer step and return at once; step_result(handle) waits for its tokens.  At most two steps
        may be outstanding.  Ids of -(j+1) in `b` mean "the token of logit row j of the step enqueued just before"."""
        k = self._flip
        self._flip ^= 1
        T, R, S, bt_stride, n_ints = self.pack(b, self._h_meta2[k].numpy())
        _abi.check(self.lib.b200_decode_step_host(self.ctx, self._h_meta2[k].data_ptr(), T, S, bt_stride, b.max_ctx, R,
                                                  self._h_out2[k].data_ptr(), self.stream.cuda_stream))
        self._done2[k].record(self.stream)
        self.h2d_bytes_last, self.d2h_bytes_last = n_ints * 4, R * 4
        return k | (R << 1)

    def step_result(self, handle: int) -> np.ndarray:
        k, R = handle & 1, handle >> 1
        self._done2[k].synchronize()
        return self._h_out2[k].numpy()[:R]

    def step_device(self, ids, positions, slots, block_table, ctx_lens, max_ctx, *, row_to_seq=None,
                    logit_rows=None, out_tokens=None, out_logits=None, stream=None):
        """Device-pointer step (inputs already resident in HBM): bench.py's `value`, and tests."""
        T = ids.numel()
        R = logit_rows.numel() if logit_rows is not None else T
        if out_tokens is None:
            out_tokens = torch.empty(R, dtype=torch.int32, device=self.device)
        s = (stream or self.stream).cuda_stream
        _abi.check(self.lib.b200_decode_step(
            self.ctx, ids.data_ptr(), positions.data_ptr(), slots.data_ptr(), block_table.data_ptr(),
            block_table.stride(0), _abi.ptr(row_to_seq), ctx_lens.data_ptr(), T, int(max_ctx),
            _abi.ptr(logit_rows), R, out_tokens.data_ptr(), _abi.ptr(out_logits), s))
        return out_tokens

    def pack_sampling(self, temperature, top_k, top_p, seeds, draws, min_p=None) -> int:
        """Per-row sampling parameters of the last step's logit rows into the pinned buffer
        (layout of b200_ctx_sample_last); returns the row count."""
        n = len(temperature)
        if not 0 < n <= self.max_logit_rows:
            raise ValueError(f"sample_last: {n} rows outside (0, {self.max_logit_rows}]")
        a = self._h_sample_np
        a[0:4 * n].view(np.float32)[:] = temperature
        a[4 * n:8 * n].view(np.int32)[:] = top_k
        a[8 * n:12 * n].view(np.float32)[:] = top_p
        a[12 * n:16 * n].view(np.float32)[:] = 0.0 if min_p is None else min_p
        a[16 * n:24 * n].view(np.uint64)[:] = np.asarray(seeds, dtype=np.uint64)
        a[24 * n:32 * n].view(np.uint64)[:] = np.asarray(draws, dtype=np.uint64)
        return n

    def sample_packed(self, n: int, sync: bool = True) -> np.ndarray:
        _abi.check(self.lib.b200_ctx_sample_last(self.ctx, n, self._h_sample.data_ptr(), self._h_out.data_ptr(),
                                                 self.stream.cuda_stream))
        if sync:
            self.stream.synchronize()
        return self._h_out_np[:n]

    def sample_last(self, temperature, top_k, top_p, seeds, draws, sync: bool = True, min_p=None) -> np.ndarray:
        """Tokens for the logit rows of the LAST step under per-row (temperature, top_k, top_p, seed, draw);
        temperature 0 rows return the step's greedy token.  The greedy hot path never calls this."""
        return self.sample_packed(self.pack_sampling(temperature, top_k, top_p, seeds, draws, min_p), sync)

    ADJUST_MAX_WORDS = 4 << 20    # b200_ctx_adjust_last's buffer (16 MiB)

    def pack_adjust(self, rows) -> tuple:
        """rows: one entry per logit row of the last step, None (untouched) or a dict(prompt, output, rep, freq, pres,
        bias={token: value}).  Writes the packed buffer of b200_ctx_adjust_last into pinned memory (allocated on first
        use); returns (n_rows, n_entries).  One entry per distinct token of a row's prompt, output and bias table."""
        n = len(rows)
        if not 0 < n <= self.max_logit_rows:
            raise ValueError(f"adjust_last: {n} rows outside (0, {self.max_logit_rows}]")
        if getattr(self, "_h_adjust", None) is None:
            self._h_adjust = torch.zeros(self.ADJUST_MAX_WORDS, dtype=torch.int32)
            if torch.cuda.is_available():
                self._h_adjust = self._h_adjust.pin_memory()
            self._h_adjust_np = self._h_adjust.numpy()
        ids, counts, seen, bias, offsets = [], [], [], [], [0]
        params = np.empty(3 * n, dtype=np.float32)
        for i, r in enumerate(rows):
            if r is None:
                params[3 * i:3 * i + 3] = (1.0, 0.0, 0.0)
                offsets.append(offsets[-1])
                continue
            params[3 * i:3 * i + 3] = (r["rep"], r["freq"], r["pres"])
            out_t, out_c = np.unique(np.asarray(r["output"], dtype=np.int64), return_counts=True)
            hist = np.union1d(np.asarray(r["prompt"], dtype=np.int64), out_t)
            b = r.get("bias") or {}
            toks = np.union1d(hist, np.fromiter(b.keys(), dtype=np.int64, count=len(b))) if b else hist
            c = np.zeros(len(toks), dtype=np.int32)
            c[np.searchsorted(toks, out_t)] = out_c
            bb = np.zeros(len(toks), dtype=np.float32)
            if b:
                bb[np.searchsorted(toks, np.fromiter(b.keys(), dtype=np.int64, count=len(b)))] = np.fromiter(b.values(), dtype=np.float32, count=len(b))
            ids.append(toks.astype(np.int32))
            counts.append(c)
            seen.append(np.isin(toks, hist).astype(np.int32))
            bias.append(bb)
            offsets.append(offsets[-1] + len(toks))
        E = offsets[-1]
        words = 3 * n + n + 1 + 4 * E
        if words > self.ADJUST_MAX_WORDS:
            raise ValueError(f"adjust_last: {E} (row, token) entries exceed the step's buffer")
        a = self._h_adjust_np
        a[:3 * n].view(np.float32)[:] = params
        a[3 * n:4 * n + 1] = offsets
        o = 4 * n + 1
        if E:
            a[o:o + E] = np.concatenate(ids)
            a[o + E:o + 2 * E] = np.concatenate(counts)
            a[o + 2 * E:o + 3 * E] = np.concatenate(seen)
            a[o + 3 * E:o + 4 * E].view(np.float32)[:] = np.concatenate(bias)
        return n, E

    def adjust_packed(self, n: int, n_entries: int):
        _abi.check(self.lib.b200_ctx_adjust_last(self.ctx, n, self._h_adjust.data_ptr(), n_entries, self.stream.cuda_stream))

    def adjust_last(self, rows):
        """logit_bias / penalties on the logit rows of the LAST step, in place, before sample_last (see pack_adjust)."""
        self.adjust_packed(*self.pack_adjust(rows))

    def logprobs_last(self, token_ids, n_top: int = 0):
        """Log-probabilities (log softmax of the raw logits) of `token_ids` under the LAST step's logit rows, plus
        the n_top most likely tokens per row: (logprob[n], top_ids[n, n_top], top_logprobs[n, n_top])."""
        n = len(token_ids)
        if not 0 < n <= self.max_logit_rows or not 0 <= n_top <= 32:
            raise ValueError(f"logprobs_last: {n} rows / n_top {n_top} out of range")
        self._h_lp_tok.numpy()[:n] = token_ids
        _abi.check(self.lib.b200_ctx_logprobs_last(self.ctx, n, self._h_lp_tok.data_ptr(), n_top, self._h_lp_out.data_ptr(),
                                                   self.stream.cuda_stream))
        self.stream.synchronize()
        out = self._h_lp_out.numpy()
        lp = out[:n].copy()
        ids = out[n:n + n * n_top].view(np.int32).reshape(n, n_top).copy()
        tlp = out[n + n * n_top:n + 2 * n * n_top].reshape(n, n_top).copy()
        return lp, ids, tlp

    def last_step_launches(self) -> int:
        return int(self.lib.b200_ctx_last_step_launches(self.ctx))

    def set_graphs(self, on: bool):
        _abi.check(self.lib.b200_ctx_set_graphs(self.ctx, int(on)))

    def close(self):
        if getattr(self, "ctx", None):
            torch.cuda.synchronize(self.device)
            self.lib.b200_ctx_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
