"""Host-side engine: KV block manager + continuous-batching scheduler with chunked prefill.

Mirrors the scheduler of the image the chart deploys, as parameterised by the chart's flags
(/root/reference/core/helm-charts/vllm/xeon-values.yaml:70-89): --block-size 128,
--max-num-batched-tokens 2048 (token budget of one step), --max-num-seqs 256,
--enable_chunked_prefill, --max-model-len 8192.  Greedy only on the device
(clients send "temperature": 0, /root/reference/docs/accessing-deployed-models.md:176).
Pure host logic: unit-tested on CPU with a fake runner.
"""
from __future__ import annotations

import os
import time
from collections import deque
from dataclasses import dataclass, field
from typing import Deque, Dict, List, Optional

import numpy as np

from .runner import StepBatch


class OutOfBlocks(RuntimeError):
    pass


class BlockManager:
    """Free-list allocator of fixed-size KV blocks; block tables are per sequence."""

    def __init__(self, n_blocks: int, block_size: int):
        self.n_blocks, self.block_size = n_blocks, block_size
        self.free: List[int] = list(range(n_blocks - 1, -1, -1))

    def blocks_needed(self, n_tokens: int) -> int:
        return (n_tokens + self.block_size - 1) // self.block_size

    def can_grow(self, table: List[int], n_tokens: int) -> bool:
        return self.blocks_needed(n_tokens) - len(table) <= len(self.free)

    def grow(self, table: List[int], n_tokens: int) -> None:
        need = self.blocks_needed(n_tokens) - len(table)
        if need > len(self.free):
            raise OutOfBlocks(f"need {need} blocks, {len(self.free)} free")
        for _ in range(need):
            table.append(self.free.pop())

    def release(self, table: List[int]) -> None:
        self.free.extend(reversed(table))
        table.clear()

    @property
    def usage(self) -> float:
        return 1.0 - len(self.free) / max(self.n_blocks, 1)


@dataclass(eq=False)   # identity semantics: requests live in lists/sets of the scheduler
class Request:
    request_id: str
    prompt_ids: List[int]
    max_tokens: int = 16
    ignore_eos: bool = False
    eos_token_ids: tuple = ()
    stop_token_ids: tuple = ()
    min_tokens: int = 0
    # sampling (docs/api-spec.yaml: temperature, top_p, seed; top_k is vLLM's extension).  temperature 0 =
    # greedy = the device's argmax; anything else goes through b200_ctx_sample_last after the step.
    temperature: float = 0.0
    top_k: int = 0
    top_p: float = 1.0
    min_p: float = 0.0                  # 0: off
    seed: Optional[int] = None          # None: drawn once when the request is added
    logprobs: Optional[int] = None      # None: off; N >= 0: logprob of each emitted token + the N most likely
    # logit_bias / penalties (docs/api-spec.yaml:275-311): applied to the step's logits before the token is chosen
    repetition_penalty: float = 1.0
    frequency_penalty: float = 0.0
    presence_penalty: float = 0.0
    logit_bias: Optional[Dict[int, float]] = None
    arrival: float = field(default_factory=time.monotonic)
    # state
    output_ids: List[int] = field(default_factory=list)
    n_computed: int = 0                 # tokens whose KV is in the cache
    block_table: List[int] = field(default_factory=list)
    row: int = -1                       # row of the scheduler's block-table mirror while running
    finished: bool = False
    finish_reason: Optional[str] = None
    first_token_time: Optional[float] = None
    finish_time: Optional[float] = None
    token_times: List[float] = field(default_factory=list)
    logprob_data: List[tuple] = field(default_factory=list)   # per emitted token: (logprob, [(id, logprob), ...])
    error: Optional[str] = None         # set when the engine failed this request (finish_reason "error")
    # a step that will produce this request's next token has been enqueued but not read back: the token counts as
    # present for scheduling, and its id is the (negative) code the device resolves (-(j+1): logit row j of that step)
    pending: int = 0

    @property
    def adjusts(self) -> bool:
        return bool(self.logit_bias) or self.repetition_penalty != 1.0 or self.frequency_penalty != 0.0 or self.presence_penalty != 0.0

    @property
    def plain(self) -> bool:
        """The step's own argmax is the token: nothing has to run on the host between the step and the next one."""
        return self.temperature <= 0.0 and self.logprobs is None and not self.adjusts

    @property
    def all_ids(self) -> List[int]:
        return self.prompt_ids + self.output_ids + ([self.pending] if self.pending else [])

    @property
    def n_total(self) -> int:
        return len(self.prompt_ids) + len(self.output_ids) + (1 if self.pending else 0)

    def token_at(self, p: int) -> int:
        n = len(self.prompt_ids)
        if p < n:
            return self.prompt_ids[p]
        return self.output_ids[p - n] if p - n < len(self.output_ids) else self.pending


@dataclass
class Scheduled:
    req: Request
    n_new: int          # token rows this step
    samples: bool       # the last row produces a token


class Scheduler:
    def __init__(self, blocks: BlockManager, *, max_num_batched_tokens: int = 2048, max_num_seqs: int = 256,
                 max_model_len: int = 8192, max_logit_rows: int = 512):
        self.blocks = blocks
        self.max_tokens = max_num_batched_tokens
        self.max_seqs = min(max_num_seqs, max_logit_rows)
        self.max_model_len = max_model_len
        self.waiting: Deque[Request] = deque()
        self.running: List[Request] = []
        # numpy mirror of the running sequences' block tables (one row each): the decode fast path
        # gathers from it instead of walking Python lists
        self.max_blocks = blocks.blocks_needed(max_model_len) + 1
        self.bt_np = np.zeros((self.max_seqs + 1, self.max_blocks), dtype=np.int32)
        self.free_rows: List[int] = list(range(self.max_seqs, -1, -1))
        self.failed: List[Request] = []      # requests the scheduler gave up on; LLMEngine.step reports them as done

    def _grow(self, req: Request, n_tokens: int) -> None:
        old = len(req.block_table)
        self.blocks.grow(req.block_table, n_tokens)
        if req.row < 0:
            req.row = self.free_rows.pop()
            old = 0
        new = len(req.block_table)
        if new > old:
            self.bt_np[req.row, old:new] = req.block_table[old:new]

    def _release(self, req: Request) -> None:
        self.blocks.release(req.block_table)
        if req.row >= 0:
            self.free_rows.append(req.row)
            req.row = -1

    def add(self, req: Request) -> None:
        if len(req.prompt_ids) == 0:
            raise ValueError("empty prompt")
        if len(req.prompt_ids) + req.max_tokens > self.max_model_len:
            raise ValueError(f"prompt ({len(req.prompt_ids)}) + max_tokens ({req.max_tokens}) exceeds "
                             f"max-model-len {self.max_model_len}")
        if self.blocks.blocks_needed(len(req.prompt_ids) + req.max_tokens) > self.blocks.n_blocks:
            raise ValueError(f"prompt ({len(req.prompt_ids)}) + max_tokens ({req.max_tokens}) does not fit the KV cache "
                             f"({self.blocks.n_blocks} blocks of {self.blocks.block_size} tokens)")
        self.waiting.append(req)

    def _fail(self, req: Request, why: str) -> None:
        """Give up on ONE request (its blocks and row are returned); the engine and every other request go on."""
        if req in self.running:
            self.running.remove(req)
        if req in self.waiting:
            self.waiting.remove(req)
        self._release(req)
        req.pending = 0
        req.finished, req.finish_reason, req.error, req.finish_time = True, "error", why, time.monotonic()
        self.failed.append(req)

    def abort(self, req: Request) -> bool:
        """Client went away / stop string hit: drop the request wherever it is (vLLM aborts on disconnect)."""
        if req.finished:
            return False
        if req in self.running:
            self.running.remove(req)
        elif req in self.waiting:
            self.waiting.remove(req)
        else:
            return False
        self._release(req)
        req.pending = 0
        req.finished, req.finish_reason, req.finish_time = True, "abort", time.monotonic()
        return True

    def has_work(self) -> bool:
        return bool(self.waiting or self.running)

    def schedule(self) -> List[Scheduled]:
        """Decode rows first (one per running, fully-prefilled sequence), then prefill chunks of the
        running and newly admitted sequences until the token budget is spent."""
        budget = self.max_tokens
        out: List[Scheduled] = []
        bs = self.blocks.block_size
        # running: decode or continue prefill
        for req in list(self.running):
            if budget == 0:
                break
            if req.row < 0:               # preempted earlier in this pass
                continue
            if req.pending and len(req.output_ids) + 1 >= req.max_tokens:
                continue                  # its last token is in flight: nothing more to compute
            remaining = req.n_total - req.n_computed
            n_new = remaining if remaining < budget else budget
            # the common decode row stays inside its last block: no allocator call at all
            if req.n_computed + n_new > len(req.block_table) * bs:
                if not self.blocks.can_grow(req.block_table, req.n_computed + n_new):
                    # preempt the youngest running request (recompute later), vLLM's default policy
                    victim = self.running[-1]
                    if victim is req and len(self.running) == 1:
                        self._fail(req, "the sequence no longer fits the KV cache")   # add() bounds this; belt and braces
                        continue
                    self._preempt(victim)
                    if victim is req:
                        continue
                    if not self.blocks.can_grow(req.block_table, req.n_computed + n_new):
                        continue
                self._grow(req, req.n_computed + n_new)
            out.append(Scheduled(req, n_new, n_new == remaining))
            budget -= n_new
        # admit
        while self.waiting and budget > 0 and len(self.running) < self.max_seqs:
            req = self.waiting[0]
            n_new = min(req.n_total - req.n_computed, budget)
            if not self.blocks.can_grow(req.block_table, req.n_computed + n_new):
                if not self.running:
                    # nothing is running, so no block will ever come back: waiting would spin forever
                    self._fail(req, f"a chunk of {n_new} tokens does not fit the KV cache ({len(self.blocks.free)} blocks free)")
                    continue
                break
            self.waiting.popleft()
            self._grow(req, req.n_computed + n_new)
            self.running.append(req)
            out.append(Scheduled(req, n_new, n_new == req.n_total - req.n_computed))
            budget -= n_new
        return out

    def _preempt(self, req: Request) -> None:
        self.running.remove(req)
        self._release(req)
        req.n_computed = 0
        self.waiting.appendleft(req)

    def build_batch(self, sched: List[Scheduled]) -> StepBatch:
        bs = self.blocks.block_size
        n = len(sched)
        if all(s.n_new == 1 and s.samples for s in sched):
            # decode fast path: one row per sequence, vectorised
            rows = np.fromiter((s.req.row for s in sched), dtype=np.int64, count=n)
            pos = np.fromiter((s.req.n_computed for s in sched), dtype=np.int32, count=n)
            ids = np.fromiter((s.req.token_at(s.req.n_computed) for s in sched), dtype=np.int32, count=n)
            ar = np.arange(n, dtype=np.int32)
            stride = int(pos.max()) // bs + 1
            return StepBatch(ids=ids, positions=pos, slots=self.bt_np[rows, pos // bs] * bs + pos % bs,
                             ctx_lens=pos + 1, row_to_seq=ar, logit_rows=ar, block_tables=self.bt_np[rows, :stride])
        b = StepBatch()
        for si, s in enumerate(sched):
            r = s.req
            ids = r.all_ids
            for p in range(r.n_computed, r.n_computed + s.n_new):
                b.ids.append(ids[p])
                b.positions.append(p)
                b.slots.append(r.block_table[p // bs] * bs + p % bs)
                b.ctx_lens.append(p + 1)
                b.row_to_seq.append(si)
            if s.samples:
                b.logit_rows.append(len(b.ids) - 1)
            b.block_tables.append(r.block_table)
        return b

    def mark_launched(self, sched: List[Scheduled]) -> None:
        """The step has been enqueued: its rows count as computed, and every sampling row owes one token (code -(k+1):
        logit row k of that step) that resolve() fills in."""
        k = 0
        for s in sched:
            r = s.req
            if s.samples:
                k += 1
            if r.finished:      # finished (stop token, abort) while an earlier step was in flight: a wasted row
                continue
            r.n_computed += s.n_new
            if s.samples:
                r.pending = -k

    def resolve(self, sched: List[Scheduled], sampled: List[int], now: Optional[float] = None) -> List[Request]:
        """Apply a launched step's tokens; returns the requests that finished."""
        now = time.monotonic() if now is None else now
        done, k = [], 0
        for s in sched:
            if not s.samples:
                continue
            tok = int(sampled[k])
            k += 1
            r = s.req
            if r.finished:
                continue
            r.pending = 0
            r.output_ids.append(tok)
            r.token_times.append(now)
            if r.first_token_time is None:
                r.first_token_time = now
            n_out = len(r.output_ids)
            reason = None
            if n_out >= r.min_tokens:
                if tok in r.stop_token_ids:
                    reason = "stop"
                elif not r.ignore_eos and tok in r.eos_token_ids:
                    reason = "stop"
            if reason is None and n_out >= r.max_tokens:
                reason = "length"
            if reason:
                r.finished, r.finish_reason, r.finish_time = True, reason, now
                if r in self.running:
                    self.running.remove(r)
                elif r in self.waiting:      # preempted while its token was in flight
                    self.waiting.remove(r)
                self._release(r)
                done.append(r)
        return done

    def commit(self, sched: List[Scheduled], sampled: List[int], now: Optional[float] = None) -> List[Request]:
        """mark_launched + resolve in one go (the synchronous step)."""
        self.mark_launched(sched)
        return self.resolve(sched, sampled, now)


class LLMEngine:
    """add_request / step loop around a ModelRunner (or anything with .step(StepBatch))."""

    def __init__(self, runner, *, max_num_batched_tokens: int = 2048, max_num_seqs: int = 256,
                 max_model_len: int = 8192):
        self.runner = runner
        self.blocks = BlockManager(runner.n_blocks, runner.block_size)
        self.scheduler = Scheduler(self.blocks, max_num_batched_tokens=min(max_num_batched_tokens, runner.max_tokens),
                                   max_num_seqs=max_num_seqs, max_model_len=min(max_model_len, runner.max_model_len))
        self.requests: Dict[str, Request] = {}
        self.steps = 0
        # Overlap: while every live request is `plain` (greedy, no logprobs, no logit adjustment) the engine enqueues
        # step n+1 before it reads step n's tokens -- the rows' ids are codes the device resolves from the tokens step n
        # left there -- so scheduling, packing and the token read-back of one step hide behind the next one
        # (vLLM's async scheduling).  A runner without step_async (the TP leader, test doubles) keeps the
        # synchronous step; B200_ENGINE_ASYNC=0 forces it.
        self.overlap = hasattr(runner, "step_async") and os.environ.get("B200_ENGINE_ASYNC", "1") != "0"
        self._special: set = set()       # ids of the live requests that are not plain
        self._inflight = None            # (scheduled rows, runner handle) of the enqueued, unread step

    def add_request(self, req: Request) -> None:
        if req.temperature > 0.0 and req.seed is None:
            req.seed = int.from_bytes(os.urandom(8), "little")
        if req.seed is not None:
            req.seed = int(req.seed) & (2 ** 64 - 1)      # any Python int (vLLM takes int64 seeds, negatives too): 64 bits of key
        req.top_k = min(max(int(req.top_k), 0), self.runner.cfg.vocab) if hasattr(self.runner, "cfg") else max(int(req.top_k), 0)
        self.scheduler.add(req)
        self._remember(req)

    def _remember(self, req: Request) -> None:
        self.requests[req.request_id] = req
        if not req.plain:
            self._special.add(req.request_id)

    def _forget(self, request_id: str) -> None:
        self.requests.pop(request_id, None)
        self._special.discard(request_id)

    def abort_request(self, request_id: str) -> Optional[Request]:
        req = self.requests.get(request_id)
        if req is None or not self.scheduler.abort(req):
            return None
        self._forget(request_id)
        return req

    def add_prefilled(self, req: Request, n_computed: int) -> None:
        """Benchmark/test utility: admit a request whose first n_computed tokens are taken as already
        in the KV cache (synthetic prompt KV) — blocks are allocated, nothing is computed."""
        assert 0 <= n_computed < req.n_total
        self.scheduler._grow(req, max(n_computed, 1))
        req.n_computed = n_computed
        self.scheduler.running.append(req)
        self._remember(req)

    def has_work(self) -> bool:
        return self.scheduler.has_work() or self._inflight is not None

    def _take_failed(self) -> List[Request]:
        failed, self.scheduler.failed = self.scheduler.failed, []
        for r in failed:
            self._forget(r.request_id)
        return failed

    def _resolve_inflight(self) -> List[Request]:
        sched, handle = self._inflight
        self._inflight = None
        toks = self.runner.step_result(handle)          # waits for that step only; a later one keeps running
        done = self.scheduler.resolve(sched, toks)
        for r in done:
            self._forget(r.request_id)
        return done

    def step(self) -> List[Request]:
        """One engine iteration; returns the requests that finished in it.  In overlap mode an iteration enqueues the
        next step and then reads the previous one, so a token is reported one call after its step was enqueued (at the
        same wall-clock time: the read waits for exactly that step)."""
        if self.overlap and not self._special:
            sched = self.scheduler.schedule()
            handle = None
            if sched:
                batch = self.scheduler.build_batch(sched)
                if not len(batch.logit_rows):
                    batch.logit_rows.append(batch.n_tokens - 1)
                handle = self.runner.step_async(batch)
                self.steps += 1
            done = self._resolve_inflight() if self._inflight is not None else []
            if sched:
                self.scheduler.mark_launched(sched)
                self._inflight = (sched, handle)
            return done + self._take_failed()
        if self._inflight is not None:
            # a request that needs the host between step and token has arrived: finish the step in flight first
            return self._resolve_inflight() + self._take_failed()
        sched = self.scheduler.schedule()
        if not sched:
            return self._take_failed()
        batch = self.scheduler.build_batch(sched)
        if len(batch.logit_rows):
            sampled = self.runner.step(batch)
        else:  # a prefill chunk that samples nothing still needs one logit row for the ABI
            batch.logit_rows.append(batch.n_tokens - 1)
            self.runner.step(batch)
            sampled = []
        self.steps += 1
        sampled = list(sampled)
        adjusted = bool(sampled) and any(s.samples and s.req.adjusts for s in sched)
        if adjusted:
            # logit_bias / penalties: a sparse kernel over the logits the step left on the device, before any token is
            # chosen from them (the step's own argmax of these rows is discarded: the sampler below re-takes it)
            rows = [s.req for s in sched if s.samples]
            self.runner.adjust_last([dict(prompt=r.prompt_ids, output=r.output_ids, rep=r.repetition_penalty,
                                          freq=r.frequency_penalty, pres=r.presence_penalty, bias=r.logit_bias)
                                     if r.adjusts else None for r in rows])
        if sampled and (adjusted or any(s.samples and s.req.temperature > 0.0 for s in sched)):
            # some rows do not want the step's argmax: one more kernel over the logits the step left on the device
            # (greedy rows get the argmax of those logits).  Draw number = tokens generated so far: a seed replays.
            rows = [s.req for s in sched if s.samples]
            sampled = list(self.runner.sample_last([r.temperature for r in rows], [r.top_k for r in rows],
                                                   [r.top_p for r in rows], [r.seed or 0 for r in rows],
                                                   [len(r.output_ids) for r in rows],
                                                   min_p=[r.min_p for r in rows]))
        if sampled and any(s.samples and s.req.logprobs is not None for s in sched):
            # log softmax of the raw logits for the tokens just chosen (+ the N most likely), same logits buffer
            rows = [s.req for s in sched if s.samples]
            n_top = min(max((r.logprobs or 0) for r in rows), 32)
            lp, ids, tlp = self.runner.logprobs_last([int(t) for t in sampled], n_top)
            for i, r in enumerate(rows):
                if r.logprobs is not None:
                    k = min(r.logprobs, n_top)
                    r.logprob_data.append((float(lp[i]), [(int(ids[i, j]), float(tlp[i, j])) for j in range(k)]))
        done = self.scheduler.commit(sched, sampled)
        for r in done:
            self._forget(r.request_id)
        return done + self._take_failed()

    def generate(self, prompts: List[List[int]], max_tokens: int, ignore_eos: bool = True) -> List[List[int]]:
        reqs = [Request(f"gen-{i}", list(p), max_tokens=max_tokens, ignore_eos=ignore_eos) for i, p in enumerate(prompts)]
        for r in reqs:
            self.add_request(r)
        while self.has_work():
            self.step()
        return [r.output_ids for r in reqs]
