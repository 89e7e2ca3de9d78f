"""Tensor-parallel execution for the server: one process per GPU (the chart's
`--distributed_executor_backend mp`, /root/reference/core/helm-charts/vllm/xeon-values.yaml:78-79).

Rank 0 owns the scheduler and the HTTP front end.  Every scheduler iteration it packs the step inputs
into its pinned buffer and broadcasts them (header + int32 payload, gloo) to the worker ranks, which run
the same `b200_decode_step_host` on their weight shards; the per-layer all-reduces and the vocab-parallel
greedy pick happen on the device (NCCL / peer memory), so every rank ends the step with the same tokens.
"""
from __future__ import annotations

import os
from typing import Optional

import numpy as np
import torch
import torch.distributed as dist

HEADER = 8   # T, R, S, bt_stride, max_ctx, n_ints, opcode, reserved
OP_STEP, OP_STOP, OP_SAMPLE, OP_LOGPROBS, OP_ADJUST = 1, 2, 3, 4, 5


class TPLeader:
    """Wraps rank 0's ModelRunner: `step(batch)` drives all ranks."""

    def __init__(self, runner, group=None):
        self.runner, self.group = runner, group
        self.n_blocks, self.block_size = runner.n_blocks, runner.block_size
        self.max_tokens, self.max_model_len, self.cfg = runner.max_tokens, runner.max_model_len, runner.cfg
        self._hdr = torch.zeros(HEADER, dtype=torch.int32)

    def step(self, batch):
        r = self.runner
        T, R, S, bt_stride, n_ints = r.pack(batch)
        self._hdr[:7] = torch.tensor([T, R, S, bt_stride, batch.max_ctx, n_ints, OP_STEP], dtype=torch.int32)
        dist.broadcast(self._hdr, src=0, group=self.group)
        dist.broadcast(r._h_meta[:n_ints], src=0, group=self.group)
        return r.step_packed(T, R, S, bt_stride, batch.max_ctx, n_ints)

    def sample_last(self, temperature, top_k, top_p, seeds, draws, min_p=None):
        """Sampling over the last step's logits on every rank (the vocab shards are all-gathered on the device)."""
        r = self.runner
        n = r.pack_sampling(temperature, top_k, top_p, seeds, draws, min_p)   # min_p rides in the same buffer
        self._hdr[:7] = torch.tensor([n, 0, 0, 0, 0, 0, OP_SAMPLE], dtype=torch.int32)
        dist.broadcast(self._hdr, src=0, group=self.group)
        dist.broadcast(r._h_sample[:32 * n], src=0, group=self.group)
        return r.sample_packed(n)

    def adjust_last(self, rows):
        """logit_bias / penalties on the last step's logits: every rank adjusts its own vocabulary shard."""
        r = self.runner
        n, E = r.pack_adjust(rows)
        words = 4 * n + 1 + 4 * E
        self._hdr[:7] = torch.tensor([n, E, 0, 0, 0, words, OP_ADJUST], dtype=torch.int32)
        dist.broadcast(self._hdr, src=0, group=self.group)
        dist.broadcast(r._h_adjust[:words], src=0, group=self.group)
        r.adjust_packed(n, E)

    def logprobs_last(self, token_ids, n_top: int = 0):
        """Log-probabilities over the last step's logits: a collective like sample_last (rank 0 returns the result)."""
        r = self.runner
        n = len(token_ids)
        self._hdr[:7] = torch.tensor([n, n_top, 0, 0, 0, 0, OP_LOGPROBS], dtype=torch.int32)
        dist.broadcast(self._hdr, src=0, group=self.group)
        toks = torch.tensor([int(t) for t in token_ids], dtype=torch.int32)
        dist.broadcast(toks, src=0, group=self.group)
        return r.logprobs_last(toks.tolist(), n_top)

    def stop(self):
        self._hdr[6] = OP_STOP
        dist.broadcast(self._hdr, src=0, group=self.group)

    @property
    def h2d_bytes_last(self):
        return self.runner.h2d_bytes_last


def worker_loop(runner, group=None) -> None:
    """Ranks > 0: receive each step from rank 0 and execute it until told to stop."""
    hdr = torch.zeros(HEADER, dtype=torch.int32)
    while True:
        dist.broadcast(hdr, src=0, group=group)
        T, R, S, bt_stride, max_ctx, n_ints, op = (int(v) for v in hdr[:7])
        if op == OP_STOP:
            return
        if op == OP_LOGPROBS:
            toks = torch.zeros(T, dtype=torch.int32)
            dist.broadcast(toks, src=0, group=group)
            runner.logprobs_last(toks.tolist(), R)
            continue
        if op == OP_SAMPLE:
            dist.broadcast(runner._h_sample[:32 * T], src=0, group=group)
            runner.sample_packed(T)
            continue
        if op == OP_ADJUST:
            if getattr(runner, "_h_adjust", None) is None:
                runner._h_adjust = torch.zeros(runner.ADJUST_MAX_WORDS, dtype=torch.int32).pin_memory()
                runner._h_adjust_np = runner._h_adjust.numpy()
            dist.broadcast(runner._h_adjust[:n_ints], src=0, group=group)
            runner.adjust_packed(T, R)
            continue
        dist.broadcast(runner._h_meta[:n_ints], src=0, group=group)
        runner.step_packed(T, R, S, bt_stride, max_ctx, n_ints)


def _worker_main(rank: int, world: int, port: int, build_runner, ready) -> None:
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    runner = build_runner(rank, world)
    runner.init_tp()
    ready.put(rank)
    worker_loop(runner)
    runner.close()
    dist.destroy_process_group()


def launch_workers(world: int, build_runner, port: Optional[int] = None):
    """Spawn ranks 1..world-1, initialise this process as rank 0; returns (leader, processes)."""
    import torch.multiprocessing as mp
    port = port or (29700 + os.getpid() % 2000)
    ctx = mp.get_context("spawn")
    ready = ctx.Queue()
    procs = [ctx.Process(target=_worker_main, args=(r, world, port, build_runner, ready), daemon=True) for r in range(1, world)]
    for p in procs:
        p.start()
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK="0", WORLD_SIZE=str(world), LOCAL_RANK="0")
    torch.cuda.set_device(0)
    dist.init_process_group("gloo", rank=0, world_size=world)
    runner = build_runner(0, world)
    runner.init_tp()
    for _ in procs:
        ready.get(timeout=600)
    return TPLeader(runner), procs


def shutdown(leader: TPLeader, procs) -> None:
    """Stop the workers, tear down this rank, then reap the worker processes (in this order: the
    process-group teardown is collective, so rank 0 must not wait for the workers before joining it)."""
    leader.stop()
    leader.runner.close()
    dist.destroy_process_group()
    for p in procs:
        p.join(timeout=60)
