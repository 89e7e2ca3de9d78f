#!/usr/bin/env python
"""bench.py — the decode hot path on B200, next to the reference's CPU path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl b200|reference]

A "step" is one scheduler iteration of the decode path over one batch of sequences (every
sequence emits one token).  Default workload = BASELINE.json configs[1]: Llama-3.1-8B-Instruct
bf16, continuous batch of 64 requests, 512-token prompts / 256 output tokens, random-init weights,
synthetic prompt KV (prefill flops are not part of the decode metric).  8B fits one
GPU and does not shard (SURVEY.md §8e "replicas only"), so the headline of --gpus N is N independent
replicas (weak scaling, no data-path collective).

Printed line (rank 0): metric/value/unit per the driver contract plus
  roofline      dominant kernel's achieved HBM GB/s (CUDA events, live) vs MEASURED_PEAKS.json
  cpu_baseline  the oracle (CPU port of the reference image's algorithm) timed on this host
  e2e           same metric through the engine API with pinned HOST buffers (every step: H2D of its metadata, the
                step, D2H of its tokens; the engine enqueues step n+1 before it reads step n, engine.py)
  also          (default workload, N = 1) the other single-GPU configs of BASELINE.json, a few dozen steps each:
                8B batch 1 (configs[0], the north_star target) and 8B batch 8 at 32 k context (configs[2])
  tp            (default workload) the tensor-parallel path at TP = N: a parity check of the sharded step against
                the committed oracle fixture (peer-memory and NCCL all-reduce), then Llama-3.1-70B decode at
                batch 64 and batch 1 (configs[4]) with the all-reduce timed on its own; at N = 1 that is 70B on one GPU
"""
from __future__ import annotations

import argparse
import gc
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "enterprise-inference_b200"

OUT_LEN = {"llama8b-bs64": 256, "llama8b-bs1": 128, "llama8b-ctx32k": 64, "llama70b-tp": 256, "llama70b-tp-bs1": 128,
           "qwen32b-tp": 256, "llama70b-shard8": 256, "llama70b-shard8-bs1": 128, "tiny": 16}   # generated tokens per request: a finished request is replaced by a new one

WORKLOADS = {
    # name: (model, batch, ctx_start, description)
    "llama8b-bs64": ("llama-3.1-8b", 64, 512,
                     "Llama-3.1-8B-Instruct bf16 decode, continuous batch of 64, 512-in/256-out (BASELINE configs[1])"),
    "llama8b-bs1": ("llama-3.1-8b", 1, 128, "Llama-3.1-8B-Instruct bf16 decode, batch 1, 128-in/128-out (configs[0])"),
    "llama8b-ctx32k": ("llama-3.1-8b", 8, 32768 - 64, "Llama-3.1-8B-Instruct bf16 decode, batch 8 at 32k context (configs[2])"),
    "llama70b-tp": ("llama-3.1-70b", 64, 512, "Llama-3.1-70B-Instruct bf16 decode, TP=N, batch 64 (configs[4])"),
    "llama70b-tp-bs1": ("llama-3.1-70b", 1, 128, "Llama-3.1-70B-Instruct bf16 decode, TP=N, batch 1, 128-in/128-out (configs[4])"),
    "qwen32b-tp": ("qwen2-32b", 32, 512, "DeepSeek-R1-Distill-Qwen-32B bf16 decode, TP=N, batch 32 (configs[3])"),
    "llama70b-shard8": ("llama-3.1-70b-tp8-shard", 64, 512, "one rank's share of Llama-3.1-70B at TP8 on one GPU, no collectives, batch 64 (profiling)"),
    "llama70b-shard8-bs1": ("llama-3.1-70b-tp8-shard", 1, 128, "one rank's share of Llama-3.1-70B at TP8 on one GPU, no collectives, batch 1 (profiling)"),
    "tiny": ("tiny", 4, 16, "tiny config (self-test)"),
}
KV_BLOCK = 128


_RESULT_FD = None


def claim_stdout():
    """Rank 0 prints ONE JSON line on stdout: route everything else that libraries write to fd 1
    (NCCL's version banner, torchrun chatter) to stderr and keep the real stdout for the result."""
    global _RESULT_FD
    if _RESULT_FD is None:
        sys.stdout.flush()
        _RESULT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def timed_mean_ctx(workload: str, steps: int, warmup: int) -> float:
    """Mean context length over the timed steps.  ONE function for both arms: step n (warm-up included) runs at
    context ctx_start + (n mod out_len) + 1, requests are replaced after out_len tokens."""
    ctx0, out_len = WORKLOADS[workload][2], OUT_LEN[workload]
    return sum(ctx0 + (n % out_len) + 1 for n in range(warmup, warmup + steps)) / max(steps, 1)


def config_for(workload: str, steps: int, warmup: int, world: int, tp: bool) -> dict:
    """The `config` object of the JSON line, identical in both arms for the same command line."""
    model, B, ctx0, desc = WORKLOADS[workload]
    cfg = importlib.import_module(PKG + ".config").get_config(model)
    mean_ctx = timed_mean_ctx(workload, steps, warmup)
    step_bytes = (cfg.weight_bytes() + B * mean_ctx * cfg.kv_bytes_per_token()) / (world if tp else 1)
    return {"workload": desc, "batch_per_gpu": B, "ctx_start": ctx0, "mean_ctx": mean_ctx,
            "parallelism": (f"tp{world}" if tp else f"replicas x{world}"), "kv_block": KV_BLOCK,
            "l2": "inputs larger than L2: %.1f GB of weights+KV streamed per step" % (step_bytes / 1e9)}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append((time.monotonic(), ln.strip()))

    def stop(self, t0: float, t1: float):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [l for (t, l) in self.lines if t0 - 0.05 <= t <= t1 + 0.15] or [l for (_, l) in self.lines]
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in rows:
            f = [x.strip() for x in l.split(",")]
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
                for n, v in zip(names, f[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the oracle on the host cores
# ------------------------------------------------------------------------------------------------
def cpu_decode_baseline(model: str, batch: int, ctx: int, steps: int, warmup: int, budget_s: float):
    """Times the CPU oracle (oracle/decode_oracle.py, the port of the algorithm the reference's
    vLLM-CPU image runs; the image itself is unavailable: no network, no container runtime) for
    `steps` decode steps of the workload.  Returns (tok_s, ms_per_step, steps_run, info)."""
    import torch
    from oracle import decode_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = O.CONFIGS[model]
    g = torch.Generator().manual_seed(1234)

    def rnd(*shape):
        return torch.empty(shape, dtype=torch.bfloat16).normal_(0.0, 0.02, generator=g)

    # One layer's random weights aliased across all layers: bounds host RAM and init time; a layer
    # (436 MB for 8B) is far larger than the host LLC, so every layer still streams from DRAM.
    layer = {"input_layernorm.weight": torch.ones(cfg.hidden, dtype=torch.bfloat16),
             "post_attention_layernorm.weight": torch.ones(cfg.hidden, dtype=torch.bfloat16),
             "self_attn.q_proj.weight": rnd(cfg.q_size, cfg.hidden), "self_attn.k_proj.weight": rnd(cfg.kv_size, cfg.hidden),
             "self_attn.v_proj.weight": rnd(cfg.kv_size, cfg.hidden), "self_attn.o_proj.weight": rnd(cfg.hidden, cfg.q_size),
             "mlp.gate_proj.weight": rnd(cfg.intermediate, cfg.hidden), "mlp.up_proj.weight": rnd(cfg.intermediate, cfg.hidden),
             "mlp.down_proj.weight": rnd(cfg.hidden, cfg.intermediate)}
    if cfg.qkv_bias:
        for x, n in (("q", cfg.q_size), ("k", cfg.kv_size), ("v", cfg.kv_size)):
            layer[f"self_attn.{x}_proj.bias"] = rnd(n)
    w = {"model.embed_tokens.weight": rnd(cfg.vocab, cfg.hidden), "model.norm.weight": torch.ones(cfg.hidden, dtype=torch.bfloat16),
         "lm_head.weight": rnd(cfg.vocab, cfg.hidden)}
    for l in range(cfg.n_layers):
        for k, v in layer.items():
            w[f"model.layers.{l}.{k}"] = v
    cap = ctx + steps + warmup + 2
    om = O.OracleModel(cfg, w, capacity=cap)
    # packed qkv / gate_up copies are per layer in OracleModel: alias them too
    for L in om.layers[1:]:
        L.update(om.layers[0])
    # synthetic prompt KV: ONE [batch][cap][kv heads][128] tensor serves as K and as V of every layer (bounds host RAM;
    # 84 MB at batch 64 x ctx 640, so part of it may stay in the LLC: that flatters the CPU, say so in `sample`).
    # The sequences are rows of it, so the attention of a step is one batched SDPA call, not a Python loop.
    kv = torch.zeros(batch, cap, cfg.n_kv_heads, cfg.head_dim, dtype=torch.bfloat16)
    kv[:, :ctx].normal_(0.0, 0.5, generator=g)
    states = []
    for b in range(batch):
        st = O.SeqState()
        st.K = [kv[b]] * cfg.n_layers
        st.V = [kv[b]] * cfg.n_layers
        st.n = ctx
        states.append(st)
    toks = torch.randint(0, cfg.vocab, (batch,), generator=g).tolist()
    times, t_begin = [], time.monotonic()
    with torch.no_grad():
        for i in range(warmup + steps):
            t0 = time.monotonic()
            logits = om.step_batch(states, toks, fast_linear=True, kv_batch=(kv, kv))
            toks = O.greedy(logits).tolist()
            dt = time.monotonic() - t0
            if i >= warmup:
                times.append(dt)
            done = i + 1
            if i + 1 >= min(warmup, 1) + 1 and (time.monotonic() - t_begin) + dt > budget_s:
                break
    if not times:
        times = [dt]
    ms = 1e3 * sum(times) / len(times)
    info = {"cores": cores, "kind": "port",
            "sample": (f"{len(times)} decode step(s) of {batch} seqs at ctx {ctx} (+{max(done - len(times), 0)} warm-up), oracle "
                       f"step_batch with bf16 oneDNN linears and one batched bf16 SDPA call per layer, torch {torch.__version__}, "
                       f"{cores} threads; one layer's random weights aliased across layers (layer >> LLC); one tensor is K "
                       "and V of every layer (may partly sit in the LLC: flatters the CPU); reference image vllm-cpu v0.10.2 "
                       "unavailable on this box")}
    return batch / ms * 1e3, ms, len(times), info


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    model, batch, ctx0, desc = WORKLOADS[args.workload]
    tp = "-tp" in args.workload
    W = max(args.warmup, 3)          # the B200 arm's effective warm-up: both arms see the same contexts
    config = config_for(args.workload, args.steps, W, world, tp)
    # the CPU steps run at the B200 arm's mean timed context (the port grows by one token per step: centre it)
    ctx = max(int(round(config["mean_ctx"] - args.warmup - (args.steps - 1) / 2 - 1)), 1)
    tok_s, ms, steps_run, info = cpu_decode_baseline(model, batch, ctx, args.steps, args.warmup, budget_s=200.0)
    line = {"impl": "reference", "metric": "output tok/s (decode)", "value": round(tok_s, 3), "unit": "tok/s",
            "n_gpus": args.gpus, "steps": steps_run, "warmup": args.warmup, "ms_per_step": round(ms, 2),
            "p50_tpot_ms": round(ms, 2), "higher_is_better": True, "scaling": "strong" if tp else "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic (random-init weights N(0,0.02), synthetic prompt KV, greedy, ignore_eos)",
            "config": config,
            "cpu_baseline": {"value": round(tok_s, 3), "unit": "tok/s", **info},
            "e2e": {"value": round(tok_s, 3), "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# ------------------------------------------------------------------------------------------------
# the B200 arm
# ------------------------------------------------------------------------------------------------
class Env:
    """Process-wide state of the B200 arm: rank/world, torch.distributed, the library."""

    def __init__(self):
        import torch
        self.torch = torch
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dist = None
        if self.world > 1:
            # rank 0 prints ONE JSON line on stdout: keep NCCL's version banner off it
            if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
                os.environ["NCCL_DEBUG"] = "WARN"
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist
        self.Cfg = importlib.import_module(PKG + ".config")
        self.R = importlib.import_module(PKG + ".runner")
        self.E = importlib.import_module(PKG + ".engine")
        self.abi = importlib.import_module(PKG + "._abi")
        self.lib = self.abi.lib()
        self.hbm_peak, self.peak_src = peaks()

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, ms: float) -> float:
        if self.dist is None:
            return ms
        t = self.torch.tensor([ms], device="cuda", dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def free_gib(self) -> float:
        gc.collect()
        self.torch.cuda.empty_cache()
        return self.torch.cuda.mem_get_info()[0] / 2 ** 30


def measure(env: Env, workload: str, K: int, W: int, *, roofline: bool, e2e: bool, clocks: bool, ar_timing: bool = False):
    """One workload on this process group: the device-resident timed loop (value), optionally the dominant kernel's
    roofline, the end-to-end loop through the engine API, clock sampling, and the all-reduce on its own."""
    torch, abi, lib = env.torch, env.abi, env.lib
    rank, world = env.rank, env.world
    model, B, ctx0, desc = WORKLOADS[workload]
    tp = "-tp" in workload
    cfg = env.Cfg.get_config(model)
    out_len = OUT_LEN[workload]
    max_len = ctx0 + out_len + 2          # bounded for any --steps: requests are recycled after out_len tokens
    bs = KV_BLOCK
    blocks_per_seq = (max_len + bs - 1) // bs
    n_blocks = B * blocks_per_seq + 2
    r = env.R.ModelRunner(cfg, max_tokens=max(B, 8), max_model_len=blocks_per_seq * bs, n_blocks=n_blocks,
                          tp_rank=rank if tp else 0, tp_world=world if tp else 1)
    r.init_random_weights(seed=1234)
    if tp:
        r.init_tp()
    gk = torch.Generator(device="cuda").manual_seed(99 + rank)
    r.k_cache.normal_(0.0, 0.5, generator=gk)   # synthetic prompt KV
    r.v_cache.normal_(0.0, 0.5, generator=gk)
    res = {"workload": workload, "desc": desc, "batch": B}

    # ---- (1) device-resident: inputs already in HBM, the sampled token feeds the next step ----
    bt = torch.arange(B * blocks_per_seq, dtype=torch.int32, device="cuda").view(B, blocks_per_seq).contiguous()
    ids = torch.randint(0, cfg.vocab, (B,), dtype=torch.int32, device="cuda")
    pos = torch.full((B,), ctx0 - 1, dtype=torch.int32, device="cuda")
    slots = (bt[:, (ctx0 - 1) // bs] * bs + (ctx0 - 1) % bs).to(torch.int32).contiguous()
    ctxl = (pos + 1).contiguous()
    out = torch.empty(B, dtype=torch.int32, device="cuda")
    stream = r.stream
    step_no = [0]
    init = [t.clone() for t in (pos, slots, ctxl)]
    ctx_sum = [0.0, 0]

    def dev_step():
        k = step_no[0] % out_len
        if k == 0 and step_no[0] > 0:
            # every request has produced out_len tokens: the batch is replaced by new requests whose prompt
            # KV is (the same synthetic) cache content — positions/slots/context restart, the model does not
            for dst, src in zip((pos, slots, ctxl), init):
                dst.copy_(src)
        max_ctx = ctx0 + k + 1
        ctx_sum[0] += max_ctx
        ctx_sum[1] += 1
        r.step_device(ids, pos, slots, bt, ctxl, max_ctx, out_tokens=out, stream=stream)
        abi.check(lib.b200_advance_decode_rows(ids.data_ptr(), pos.data_ptr(), slots.data_ptr(), ctxl.data_ptr(),
                                               bt.data_ptr(), blocks_per_seq, None, out.data_ptr(), B, bs,
                                               stream.cuda_stream))
        step_no[0] += 1

    with torch.cuda.stream(stream):
        for _ in range(W):
            dev_step()
    env.barrier()
    ctx_sum[0], ctx_sum[1] = 0.0, 0
    sampler = ClockSampler(env.local) if clocks else None
    if sampler:
        sampler.start()
        time.sleep(0.25)
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
    t_wall0 = time.monotonic()
    with torch.cuda.stream(stream):
        evs[0].record(stream)
        for i in range(K):
            dev_step()
            evs[i + 1].record(stream)
    env.barrier()
    t_wall1 = time.monotonic()
    if sampler:
        res["clocks"] = sampler.stop(t_wall0, t_wall1)
    total_ms = env.max_over_ranks(evs[0].elapsed_time(evs[K]))
    per_step = [evs[i].elapsed_time(evs[i + 1]) for i in range(K)]
    launches_per_step = r.last_step_launches() + 1
    n_eff = 1 if tp else world
    mean_ctx = ctx_sum[0] / max(ctx_sum[1], 1)
    assert abs(mean_ctx - timed_mean_ctx(workload, K, W)) < 1e-6, "the two arms must see the same contexts"
    step_bytes = (cfg.weight_bytes() + B * mean_ctx * cfg.kv_bytes_per_token()) / (world if tp else 1)
    step_gbs = step_bytes / (total_ms / K) / 1e6
    res.update(value=n_eff * B * K / total_ms * 1e3, ms_per_step=total_ms / K, p50_tpot_ms=statistics.median(per_step),
               steps=K, launches_per_step=launches_per_step, mean_ctx=mean_ctx, n_eff=n_eff,
               whole_step={"bytes": int(step_bytes), "GBps": round(step_gbs, 1), "frac": round(step_gbs / env.hbm_peak, 4)})

    # ---- (2) roofline of the dominant kernel: the gate_up projection (31% of the weight bytes), ----
    # launched back to back over the layers' distinct weights (7.5 GB >> L2), CUDA events on its stream
    if roofline:
        gu = [t for t in r._weights if t.dim() == 2 and t.shape[0] == 2 * r.inter and t.shape[1] == cfg.hidden]
        x = torch.randn(B, cfg.hidden, device="cuda", dtype=torch.bfloat16)
        y = torch.empty(B, 2 * r.inter, device="cuda", dtype=torch.bfloat16)
        wsb = int(lib.b200_linear_workspace_bytes(B, 2 * r.inter, cfg.hidden))
        wsp = torch.zeros(max(wsb, 16), dtype=torch.uint8, device="cuda")
        reps = max(1, 64 // len(gu))

        def gu_pass():
            for _ in range(reps):
                for wt in gu:
                    abi.check(lib.b200_linear_bf16(x.data_ptr(), cfg.hidden, wt.data_ptr(), None, y.data_ptr(), 2 * r.inter, B,
                                                   2 * r.inter, cfg.hidden, 0, wsp.data_ptr(), wsb, stream.cuda_stream))
        abi.check(lib.b200_set_launch_pdl(1))   # same launch attribute as inside the step
        with torch.cuda.stream(stream):
            gu_pass()
            env.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            gu_pass()
            e1.record(stream)
        env.barrier()
        abi.check(lib.b200_set_launch_pdl(0))
        k_ms = e0.elapsed_time(e1) / (reps * len(gu))
        k_bytes = 2 * r.inter * cfg.hidden * 2
        achieved = k_bytes / k_ms / 1e6
        pasted = model == "llama-3.1-8b" and B == 64
        res["roofline"] = {"bound": "hbm", "kernel": "linear gate_up [2I x h] weight stream, T=%d" % B,
                           "achieved": round(achieved, 1), "peak": env.hbm_peak, "unit": "GB/s",
                           "frac": round(achieved / env.hbm_peak, 4), "traffic": 238.9e6 if pasted else None,
                           "traffic_source": ("ncu --set full capture of this kernel at this shape, pasted from "
                                              "profiles/r01_ncu_linear_gate_up.md (dram read 235.45 MB + write 3.4 MB); "
                                              "not measured live") if pasted else None,
                           "peak_source": env.peak_src, "bytes_per_launch": k_bytes, "us_per_launch": round(k_ms * 1e3, 2),
                           "timed": "back to back over the layers' distinct weights, outside the step",
                           "whole_step": res["whole_step"]}
        del x, y, wsp

    # ---- (2b) the tensor-parallel collective on its own: all-reduce + add + RMSNorm of one half-layer ----
    if ar_timing and tp and world > 1:
        h = cfg.hidden
        nw = torch.ones(h, device="cuda", dtype=torch.bfloat16)
        n_it = 200

        def ar_pass():
            for i in range(n_it):   # no staging copies: the kernel runs on whatever the exchange buffers hold
                abi.check(lib.b200_tp_allreduce(r.ctx, i & 1, None, B, None, nw.data_ptr(), None, stream.cuda_stream))
        with torch.cuda.stream(stream):
            ar_pass()
            env.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            ar_pass()
            e1.record(stream)
        env.barrier()
        us = env.max_over_ranks(e0.elapsed_time(e1)) / n_it * 1e3
        res["allreduce_us_per_call"] = round(us, 2)
        res["allreduce_note"] = ("all-reduce + residual add + RMSNorm of one [%d x %d] bf16 half-layer as the step runs it "
                                 "(one kernel), timed back to back outside the step; %d calls per step" % (B, h, 2 * cfg.n_layers))
        del nw

    # ---- (3) end to end through the engine API: host scheduling, pinned H2D, step, D2H ----
    if e2e:
        E = env.E
        eng = E.LLMEngine(r, max_num_batched_tokens=max(B, 8), max_num_seqs=max(B, 8), max_model_len=blocks_per_seq * bs)
        gp = torch.Generator().manual_seed(1000 + rank)
        n_req = [0]

        def admit():
            while len(eng.scheduler.running) < B:
                prompt = torch.randint(0, cfg.vocab, (ctx0,), generator=gp).tolist()
                req = E.Request(f"r{n_req[0]}", prompt, max_tokens=out_len, ignore_eos=True)
                n_req[0] += 1
                eng.add_prefilled(req, ctx0 - 1)   # prompt KV is the synthetic cache content

        def e2e_step():
            admit()                                # continuous batching: finished requests are replaced
            eng.step()

        for _ in range(W):
            e2e_step()
        env.barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record(stream)
        for _ in range(K):
            e2e_step()
        a1.record(stream)
        env.barrier()
        e2e_ms = env.max_over_ranks(a0.elapsed_time(a1))
        res["e2e"] = {"value": round(n_eff * B * K / e2e_ms * 1e3, 1), "unit": "tok/s", "h2d_bytes_per_step": int(r.h2d_bytes_last),
                      "d2h_bytes_per_step": int(r.d2h_bytes_last), "ms_per_step": round(e2e_ms / K, 4)}
    r.close()
    del r, bt, ids, pos, slots, ctxl, out, init
    env.free_gib()
    return res


def brief(m: dict) -> dict:
    """A measured workload as an entry of `also` / `tp`."""
    o = {"workload": m["desc"], "batch": m["batch"], "steps": m["steps"], "ms_per_step": round(m["ms_per_step"], 4),
         "p50_tpot_ms": round(m["p50_tpot_ms"], 4), "tok_s": round(m["value"], 1), "mean_ctx": m["mean_ctx"],
         "launches_per_step": m["launches_per_step"], "whole_step": m["whole_step"]}
    for k in ("e2e", "allreduce_us_per_call", "allreduce_note"):
        if k in m:
            o[k] = m[k]
    return o


def tp_parity(env: Env) -> dict:
    """The sharded step at TP = world against the committed oracle fixture (tests/golden/tiny-g8x8.npz: teacher-forced
    logits of oracle/decode_oracle.py, generated by oracle/make_golden.py): a 40-row prefill step (one-shot all-reduce)
    and single-row decode steps (fused all-reduce + add + RMSNorm), with the peer-memory kernels and with NCCL.
    Bars as tests/test_tp_gpu.py: softmax max-abs error <= 2e-3, vocab-parallel pick == argmax of the gathered logits,
    greedy token == the oracle's wherever the oracle's top-2 margin is >= 0.08.  Only the seeded weight / prompt
    generators of oracle/ are imported here; the expected logits come from the fixture."""
    import numpy as np
    torch = env.torch
    from oracle import decode_oracle as O
    name = "tiny-g8x8"
    g = np.load(os.path.join(ROOT, "tests", "golden", f"{name}.npz"))
    ocfg = O.CONFIGS[name]
    w = O.make_weights(ocfg, seed=int(g["seed"]), norm_jitter=float(g["norm_jitter"]))
    if O.weights_digest(w) != str(g["digest"]):
        return {"parity": "not run: the host RNG stream differs from the fixture's"}
    ids = g["prompt"].tolist() + g["tokens"].tolist()[:-1]
    want = torch.from_numpy(g["tf_logits"]).float()
    n_pre, n_dec = len(g["prompt"]), 8
    out = {"model": name, "tp": env.world, "rows": n_pre + n_dec}
    worst = 0.0
    try:
        for peer in (True, False):
            r = env.R.ModelRunner(env.Cfg.ModelConfig(**ocfg.to_dict()), max_tokens=64, max_model_len=512, n_blocks=8,
                                  tp_rank=env.rank, tp_world=env.world)
            r.load_weights(w)
            r.init_tp(peer_allreduce=peer)
            bt = torch.tensor([[3]], dtype=torch.int32, device="cuda")
            rows = []
            for (s, e) in [(0, n_pre)] + [(n_pre + i, n_pre + i + 1) for i in range(n_dec)]:
                n = e - s
                pos = torch.arange(s, e, dtype=torch.int32, device="cuda")
                logits = torch.empty(env.world, n, ocfg.vocab // env.world, dtype=torch.bfloat16, device="cuda")
                toks = r.step_device(torch.tensor(ids[s:e], dtype=torch.int32).cuda(), pos, 3 * KV_BLOCK + pos, bt, pos + 1, e,
                                     row_to_seq=torch.zeros(n, dtype=torch.int32).cuda(), out_logits=logits)
                torch.cuda.synchronize()
                full = logits.permute(1, 0, 2).reshape(n, ocfg.vocab).float().cpu()
                if toks.cpu().tolist() != full.argmax(-1).tolist():
                    raise AssertionError(f"vocab-parallel pick != argmax at rows {s}..{e}")
                rows.append(full)
            got = torch.cat(rows)
            ref = want[: got.shape[0]]
            err = (torch.softmax(got, -1) - torch.softmax(ref, -1)).abs().max().item()
            dis = got.argmax(-1) != ref.argmax(-1)
            if err > 2e-3:
                raise AssertionError(f"softmax max-abs error {err:.2e} ({'peer' if peer else 'nccl'} all-reduce)")
            if dis.any() and not (O.top2_margin(ref)[dis] < 0.08).all():
                raise AssertionError("greedy token differs from the oracle's outside a near-tie")
            worst = max(worst, err)
            r.close()
            del r
            env.free_gib()
        out.update(parity="ok", softmax_max_abs_err=worst, allreduce_paths=["peer-memory", "nccl"])
    except Exception as ex:   # a failed check is a result, not a crash of the bench
        out.update(parity=f"FAILED: {ex}")
    # every rank must agree that it passed
    ok = torch.tensor([1 if out.get("parity") == "ok" else 0], device="cuda")
    if env.dist is not None:
        env.dist.all_reduce(ok, op=env.dist.ReduceOp.MIN)
    if int(ok.item()) == 0 and out.get("parity") == "ok":
        out["parity"] = "FAILED on another rank"
    return out


def run_b200(args):
    env = Env()
    K, W = args.steps, args.warmup
    tp = "-tp" in args.workload
    m = measure(env, args.workload, K, W, roofline=True, e2e=True, clocks=True, ar_timing=tp)
    default = args.workload == "llama8b-bs64" and not args.headline_only
    also, tpr = [], None
    if default and env.world == 1:
        for wl, k in (("llama8b-bs1", 64), ("llama8b-ctx32k", 24)):
            try:
                also.append(brief(measure(env, wl, k, 4, roofline=False, e2e=True, clocks=False)))
            except Exception as ex:
                also.append({"workload": WORKLOADS[wl][3], "error": str(ex)[:300]})
                env.free_gib()
    if default:
        tpr = {"tp": env.world}
        if env.world > 1:
            tpr["parity"] = tp_parity(env)
        # 70B bf16 is 141 GB of weights: per GPU 141 / TP GB + KV; at TP = 1 it must fit next to 16 GB of KV cache
        need_gib = (env.Cfg.get_config("llama-3.1-70b").weight_bytes() / env.world + 20e9) / 2 ** 30
        free_gib = env.free_gib()
        if free_gib < need_gib:
            tpr["llama70b"] = f"does not fit: needs ~{need_gib:.0f} GiB per GPU at TP={env.world}, {free_gib:.0f} GiB free"
        else:
            for wl, k in (("llama70b-tp", 32), ("llama70b-tp-bs1", 48)) + ((("qwen32b-tp", 32),) if env.world in (2, 4) else ()):
                try:
                    tpr[wl] = brief(measure(env, wl, k, 4, roofline=False, e2e=False, clocks=False, ar_timing=True))
                except Exception as ex:
                    tpr[wl] = {"error": str(ex)[:300]}
                    env.free_gib()

    cpu = None
    if env.rank == 0 and env.world == 1 and not args.no_cpu_baseline:
        model, B = WORKLOADS[args.workload][0], WORKLOADS[args.workload][1]
        try:
            tok_s, ms, n, info = cpu_decode_baseline(model, B, int(round(m["mean_ctx"])), steps=2, warmup=1, budget_s=30.0)
            cpu = {"value": round(tok_s, 3), "unit": "tok/s", "ms_per_step": round(ms, 1), **info}
        except Exception as ex:  # host RAM etc.
            cpu = {"value": None, "unit": "tok/s", "cores": os.cpu_count(), "kind": "port", "sample": f"not run: {ex}"}

    if env.rank == 0:
        line = {"metric": "output tok/s (decode)", "value": round(m["value"], 1), "unit": "tok/s", "n_gpus": env.world,
                "steps": K, "warmup": W, "ms_per_step": round(m["ms_per_step"], 4),
                "p50_tpot_ms": round(m["p50_tpot_ms"], 4), "higher_is_better": True,
                "scaling": "strong" if tp else "weak", "vs_baseline": None, "dtype": "bf16",
                "data": "synthetic (random-init weights N(0,0.02), synthetic prompt KV, greedy, ignore_eos)",
                "config": config_for(args.workload, K, W, env.world, tp),
                "clocks": m["clocks"], "gpu_launches": m["launches_per_step"] * K, "launches_per_step": m["launches_per_step"],
                "roofline": m["roofline"], "e2e": m["e2e"]}
        if "allreduce_us_per_call" in m:
            line["allreduce_us_per_call"] = m["allreduce_us_per_call"]
        if also:
            line["also"] = also
        if tpr is not None:
            line["tp"] = tpr
        if cpu is not None:
            line["cpu_baseline"] = cpu
        emit(line)
    if env.dist is not None:
        env.dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=256)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="llama8b-bs64", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--headline-only", action="store_true", help="skip the `also` and `tp` sections of the default run")
    args = ap.parse_args()
    claim_stdout()
    if args.impl == "reference":
        run_reference(args)
    else:
        args.warmup = max(args.warmup, 3)
        run_b200(args)


if __name__ == "__main__":
    main()
