"""CPU: block manager + continuous-batching scheduler (host logic) with a fake runner."""
import numpy as np
import pytest

from conftest import pkg

E = pkg("engine")


class FakeRunner:
    """Deterministic 'model': next token = (sum of the sequence so far) % 97, computed from the
    step inputs exactly as a real runner would see them (checks positions/slots/ctx/block tables)."""
    def __init__(self, n_blocks=16, block_size=4, max_tokens=16, max_model_len=64):
        self.n_blocks, self.block_size, self.max_tokens, self.max_model_len = n_blocks, block_size, max_tokens, max_model_len
        self.kv = {}          # slot -> token id
        self.batches = []

    def step(self, b):
        self.batches.append(b)
        assert b.n_tokens <= self.max_tokens
        for i in range(b.n_tokens):
            assert b.ctx_lens[i] == b.positions[i] + 1
            bt = b.block_tables[b.row_to_seq[i]]
            assert b.slots[i] == bt[b.positions[i] // self.block_size] * self.block_size + b.positions[i] % self.block_size
            self.kv[b.slots[i]] = b.ids[i]
        out = []
        for r in b.logit_rows:
            bt = b.block_tables[b.row_to_seq[r]]
            ctx = [self.kv[bt[p // self.block_size] * self.block_size + p % self.block_size] for p in range(b.ctx_lens[r])]
            out.append(sum(ctx) % 97)
        return np.array(out, dtype=np.int32)


class AsyncFakeRunner(FakeRunner):
    """The same 'model' behind the runner's overlap interface: step_async enqueues (ids of -(j+1) are resolved from the
    tokens of the step enqueued just before, as the device does), step_result hands the tokens over.  It checks the
    engine's side of the contract: at most two steps outstanding, results read oldest first."""
    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.last = np.zeros(0, dtype=np.int32)
        self.outstanding = []
        self.results = {}
        self.n_async = 0
        self.n_coded = 0

    def step_async(self, b):
        ids = np.array(b.ids, dtype=np.int64).copy()
        for i, t in enumerate(ids):
            if t < 0:
                assert -t - 1 < len(self.last), "code refers to a logit row the previous step did not have"
                ids[i] = self.last[-t - 1]
                self.n_coded += 1
        b.ids = ids
        out = FakeRunner.step(self, b)
        self.last = out
        self.n_async += 1
        h = self.n_async
        self.outstanding.append(h)
        assert len(self.outstanding) <= 2, "more than two steps in flight"
        self.results[h] = out
        return h

    def step_result(self, h):
        assert self.outstanding and self.outstanding[0] == h, "results must be read oldest first"
        self.outstanding.pop(0)
        return self.results.pop(h)

    def step(self, b):
        assert not self.outstanding, "a synchronous step while another one is in flight"
        out = super().step(b)
        self.last = out
        return out


def expect(prompt, n):
    ids = list(prompt)
    for _ in range(n):
        ids.append(sum(ids) % 97)
    return ids[len(prompt):]


def test_block_manager():
    bm = E.BlockManager(4, 128)
    t = []
    bm.grow(t, 1); assert len(t) == 1
    bm.grow(t, 128); assert len(t) == 1
    bm.grow(t, 129); assert len(t) == 2
    assert not bm.can_grow(t, 128 * 5)
    with pytest.raises(E.OutOfBlocks):
        bm.grow(t, 128 * 5)
    bm.release(t)
    assert len(bm.free) == 4 and bm.usage == 0.0


def test_continuous_batching_chunked_prefill_matches_sequential():
    fr = FakeRunner(n_blocks=64, block_size=4, max_tokens=16)
    eng = E.LLMEngine(fr, max_num_batched_tokens=16, max_num_seqs=8, max_model_len=64)
    prompts = [[1, 2, 3], list(range(5, 40)), [7] * 9, [11, 13]]   # the 35-token prompt needs 3 chunks
    outs = eng.generate(prompts, max_tokens=6)
    for p, o in zip(prompts, outs):
        assert o == expect(p, 6)
    assert len(eng.blocks.free) == 64                       # everything released
    assert any(len(b.block_tables) > 1 for b in fr.batches)  # really batched
    assert all(b.n_tokens <= 16 for b in fr.batches)


def test_preemption_recomputes_and_still_matches():
    fr = FakeRunner(n_blocks=6, block_size=4, max_tokens=32, max_model_len=32)
    eng = E.LLMEngine(fr, max_num_batched_tokens=32, max_num_seqs=4, max_model_len=32)
    prompts = [[3, 1, 4, 1, 5, 9, 2], [2, 7, 1, 8, 2, 8], [1, 6, 1, 8]]
    outs = eng.generate(prompts, max_tokens=8)                # 6 blocks * 4 = 24 slots < total demand
    for p, o in zip(prompts, outs):
        assert o == expect(p, 8)


def test_stop_conditions_and_limits():
    fr = FakeRunner()
    eng = E.LLMEngine(fr, max_num_batched_tokens=16, max_num_seqs=4, max_model_len=64)
    first = expect([1, 2, 3], 1)[0]
    r = E.Request("a", [1, 2, 3], max_tokens=10, eos_token_ids=(first,))
    eng.add_request(r)
    while eng.has_work():
        eng.step()
    assert r.output_ids == [first] and r.finish_reason == "stop"
    r2 = E.Request("b", [1, 2, 3], max_tokens=4, eos_token_ids=(first,), ignore_eos=True)
    eng.add_request(r2)
    while eng.has_work():
        eng.step()
    assert len(r2.output_ids) == 4 and r2.finish_reason == "length"
    with pytest.raises(ValueError):
        eng.add_request(E.Request("c", [], max_tokens=1))
    with pytest.raises(ValueError):
        eng.add_request(E.Request("d", [1] * 60, max_tokens=10))


def test_abort_returns_blocks_and_row_wherever_the_request_is():
    fr = FakeRunner(n_blocks=8, block_size=4, max_tokens=8, max_model_len=32)
    eng = E.LLMEngine(fr, max_num_batched_tokens=8, max_num_seqs=2, max_model_len=32)
    a, b, c = (E.Request(n, [1, 2, 3, 4, 5], max_tokens=12) for n in "abc")
    for r in (a, b, c):
        eng.add_request(r)
    eng.step()                                  # a (and part of b) admitted; c waits (max_num_seqs = 2)
    assert a in eng.scheduler.running and c in eng.scheduler.waiting
    rows_before = len(eng.scheduler.free_rows)
    assert eng.abort_request("a") is a and a.finished and a.finish_reason == "abort" and a.block_table == []
    assert len(eng.scheduler.free_rows) == rows_before + 1
    assert eng.abort_request("c") is c and c not in eng.scheduler.waiting
    assert eng.abort_request("a") is None and eng.abort_request("nope") is None
    while eng.has_work():
        eng.step()
    assert b.output_ids == expect([1, 2, 3, 4, 5], 12) and len(eng.blocks.free) == 8


def test_a_request_that_can_never_be_admitted_fails_alone():
    """with nothing running and the head of the queue too big for the free blocks the scheduler used to
    return [] forever (has_work() stayed true: a busy spin).  Now add() rejects what can never fit, and a request that
    stops fitting is failed by itself while the others go on."""
    fr = FakeRunner(n_blocks=5, block_size=4, max_tokens=16, max_model_len=64)
    eng = E.LLMEngine(fr, max_num_batched_tokens=16, max_num_seqs=4, max_model_len=64)
    with pytest.raises(ValueError, match="KV cache"):
        eng.add_request(E.Request("big", [1] * 20, max_tokens=8))          # 28 tokens > 20 slots
    # a request that passed add() but whose chunk cannot be placed although nothing runs (blocks held elsewhere)
    held = []
    eng.blocks.grow(held, 12)                    # someone else holds 3 of the 5 blocks
    stuck = E.Request("stuck", [1] * 9, max_tokens=2)                      # its first chunk needs 3 blocks, 2 are free
    ok = E.Request("ok", [1, 2, 3], max_tokens=5)                          # 8 tokens = 2 blocks
    eng.add_request(stuck)
    eng.add_request(ok)
    done = []
    for _ in range(40):
        if not eng.has_work():
            break
        done += eng.step()
    assert not eng.has_work(), "the engine must not spin on a request that cannot be admitted"
    assert stuck.finished and stuck.finish_reason == "error" and "KV cache" in stuck.error and stuck in done
    assert ok.output_ids == expect([1, 2, 3], 5)


def test_seed_and_top_k_are_brought_into_range():
    fr = FakeRunner()
    fr.cfg = type("C", (), {"vocab": 97})()
    eng = E.LLMEngine(fr, max_num_batched_tokens=16, max_num_seqs=4, max_model_len=64)
    r = E.Request("s", [1, 2], max_tokens=1, temperature=1.0, seed=-5, top_k=2 ** 40)
    eng.add_request(r)
    assert r.seed == 2 ** 64 - 5 and r.top_k == 97
    np.asarray([r.seed], dtype=np.uint64), np.asarray([r.top_k], dtype=np.int32)   # what pack_sampling does


from hypothesis import given, settings, strategies as st, HealthCheck


@settings(max_examples=60, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(
    prompts=st.lists(st.lists(st.integers(0, 96), min_size=1, max_size=20), min_size=1, max_size=7),
    max_new=st.lists(st.integers(1, 9), min_size=7, max_size=7),
    arrivals=st.lists(st.integers(0, 6), min_size=7, max_size=7),
    n_blocks=st.integers(8, 24), budget=st.sampled_from([4, 8, 16]), max_seqs=st.integers(1, 5),
)
def test_scheduler_property_any_arrival_pattern_matches_sequential(prompts, max_new, arrivals, n_blocks, budget, max_seqs):
    """Whatever the arrival pattern, token budget, sequence cap and cache size (preemptions included): every
    request gets exactly the tokens a sequential run would give it, a step never exceeds the budget or the
    sequence cap, no cache block is shared by two running requests, and every block comes back."""
    bs = 4
    need = max((len(p) + max_new[i] + bs - 1) // bs for i, p in enumerate(prompts))
    if need > n_blocks:          # a single sequence must fit (the engine raises OutOfBlocks otherwise, tested elsewhere)
        n_blocks = need
    fr = FakeRunner(n_blocks=n_blocks, block_size=bs, max_tokens=budget, max_model_len=64)
    eng = E.LLMEngine(fr, max_num_batched_tokens=budget, max_num_seqs=max_seqs, max_model_len=64)
    reqs = [E.Request(f"r{i}", list(p), max_tokens=max_new[i], ignore_eos=True) for i, p in enumerate(prompts)]
    pending = sorted(range(len(reqs)), key=lambda i: arrivals[i])
    step = 0
    while pending or eng.has_work():
        while pending and arrivals[pending[0]] <= step:
            eng.add_request(reqs[pending.pop(0)])
        if eng.has_work():
            n_before = len(fr.batches)
            eng.step()
            for b in fr.batches[n_before:]:
                assert b.n_tokens <= budget and len(b.block_tables) <= max_seqs
            used = [blk for r in eng.scheduler.running for blk in r.block_table]
            assert len(used) == len(set(used)), "a cache block is owned by two running requests"
        step += 1
        assert step < 2000
    for r, p, n in zip(reqs, prompts, max_new):
        assert r.finished and r.output_ids == expect(p, n)
    assert len(eng.blocks.free) == n_blocks


def test_overlap_mode_matches_sequential_and_really_overlaps():
    """With a runner that has step_async the engine enqueues step n+1 before reading step n: same tokens as the
    sequential run, decode rows carry device-side token codes, every block comes back."""
    fr = AsyncFakeRunner(n_blocks=64, block_size=4, max_tokens=16)
    eng = E.LLMEngine(fr, max_num_batched_tokens=16, max_num_seqs=8, max_model_len=64)
    assert eng.overlap
    prompts = [[1, 2, 3], list(range(5, 40)), [7] * 9, [11, 13]]
    outs = eng.generate(prompts, max_tokens=6)
    for p, o in zip(prompts, outs):
        assert o == expect(p, 6)
    assert len(eng.blocks.free) == 64 and not eng.requests and eng._inflight is None
    assert fr.n_async == len(fr.batches) and fr.n_coded >= 4 * 4     # every decode row after the first came as a code


def test_overlap_mode_preemption_and_unpredicted_stop():
    fr = AsyncFakeRunner(n_blocks=6, block_size=4, max_tokens=32, max_model_len=32)
    eng = E.LLMEngine(fr, max_num_batched_tokens=32, max_num_seqs=4, max_model_len=32)
    prompts = [[3, 1, 4, 1, 5, 9, 2], [2, 7, 1, 8, 2, 8], [1, 6, 1, 8]]
    outs = eng.generate(prompts, max_tokens=8)
    for p, o in zip(prompts, outs):
        assert o == expect(p, 8)
    assert len(eng.blocks.free) == 6
    # a stop token the host cannot foresee: the request is already part of the next enqueued step when its stop token is
    # read; that row is wasted, its token discarded, and the neighbour is unaffected
    fr = AsyncFakeRunner()
    eng = E.LLMEngine(fr, max_num_batched_tokens=16, max_num_seqs=4, max_model_len=64)
    want = expect([1, 2, 3], 10)
    a = E.Request("a", [1, 2, 3], max_tokens=10, stop_token_ids=(want[3],))
    b = E.Request("b", [5, 6], max_tokens=10, ignore_eos=True)
    eng.add_request(a)
    eng.add_request(b)
    while eng.has_work():
        eng.step()
    assert a.output_ids == want[:4] and a.finish_reason == "stop" and a.pending == 0
    assert b.output_ids == expect([5, 6], 10) and b.finish_reason == "length"
    assert len(eng.blocks.free) == fr.n_blocks and not eng.requests


def test_overlap_mode_abort_in_flight_and_special_request_drains():
    fr = AsyncFakeRunner(n_blocks=32)
    fr.cfg = type("C", (), {"vocab": 97})()
    fr.sample_last = lambda t, k, p, s, d, min_p=None: fr.last      # 'sampling' = the step's own tokens
    eng = E.LLMEngine(fr, max_num_batched_tokens=16, max_num_seqs=4, max_model_len=64)
    a = E.Request("a", [1, 2, 3], max_tokens=12, ignore_eos=True)
    b = E.Request("b", [4, 5], max_tokens=12, ignore_eos=True)
    eng.add_request(a)
    eng.add_request(b)
    for _ in range(3):
        eng.step()
    assert eng._inflight is not None and a.pending < 0
    gone = eng.abort_request("a")                       # its row of the step in flight is discarded
    assert gone is a and a.pending == 0 and all(t >= 0 for t in a.output_ids)
    n_a = len(a.output_ids)
    eng.step()
    assert len(a.output_ids) == n_a
    # a sampling request arrives: the step in flight is read first, then the engine steps synchronously
    c = E.Request("c", [9, 9], max_tokens=3, temperature=0.5, seed=1, ignore_eos=True)
    eng.add_request(c)
    n_async = fr.n_async
    while not c.finished:
        eng.step()
    assert fr.n_async == n_async and eng._inflight is None     # no overlapped step while c was alive
    while eng.has_work():
        eng.step()
    assert fr.n_async > n_async                                 # and overlap again afterwards
    assert b.output_ids == expect([4, 5], 12) and c.output_ids == expect([9, 9], 3)
    assert len(eng.blocks.free) == 32 and not eng.requests and not eng._special


@settings(max_examples=60, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(
    prompts=st.lists(st.lists(st.integers(0, 96), min_size=1, max_size=20), min_size=1, max_size=7),
    max_new=st.lists(st.integers(1, 9), min_size=7, max_size=7),
    arrivals=st.lists(st.integers(0, 6), min_size=7, max_size=7),
    stops=st.lists(st.integers(0, 96), min_size=0, max_size=12),
    n_blocks=st.integers(8, 24), budget=st.sampled_from([4, 8, 16]), max_seqs=st.integers(1, 5),
)
def test_overlap_property_any_arrival_pattern_matches_sequential(prompts, max_new, arrivals, stops, n_blocks, budget, max_seqs):
    """The property above for the overlapped engine, with stop tokens the host cannot foresee thrown in."""
    bs = 4
    need = max((len(p) + max_new[i] + bs - 1) // bs for i, p in enumerate(prompts))
    n_blocks = max(n_blocks, need)
    fr = AsyncFakeRunner(n_blocks=n_blocks, block_size=bs, max_tokens=budget, max_model_len=64)
    eng = E.LLMEngine(fr, max_num_batched_tokens=budget, max_num_seqs=max_seqs, max_model_len=64)
    reqs = [E.Request(f"r{i}", list(p), max_tokens=max_new[i], stop_token_ids=tuple(stops)) for i, p in enumerate(prompts)]
    pending = sorted(range(len(reqs)), key=lambda i: arrivals[i])
    step = 0
    while pending or eng.has_work():
        while pending and arrivals[pending[0]] <= step:
            eng.add_request(reqs[pending.pop(0)])
        if eng.has_work():
            n_before = len(fr.batches)
            eng.step()
            for b in fr.batches[n_before:]:
                assert b.n_tokens <= budget and len(b.block_tables) <= max_seqs
            used = [blk for r in eng.scheduler.running for blk in r.block_table]
            assert len(used) == len(set(used)), "a cache block is owned by two running requests"
        step += 1
        assert step < 2000
    for r, p, n in zip(reqs, prompts, max_new):
        want = expect(p, n)
        cut = next((i + 1 for i, t in enumerate(want) if t in stops), n)
        assert r.finished and r.pending == 0 and r.output_ids == want[:cut]
    assert len(eng.blocks.free) == n_blocks and eng._inflight is None


def test_pack_adjust_builds_one_entry_per_distinct_token():
    """runner.pack_adjust (host side of b200_ctx_adjust_last): CSR rows over the distinct tokens of prompt, output and bias
    table, output counts, the seen flag (prompt or output), untouched rows as empty rows with neutral parameters."""
    R = pkg("runner")
    stub = type("Stub", (), {"max_logit_rows": 8, "ADJUST_MAX_WORDS": R.ModelRunner.ADJUST_MAX_WORDS, "_h_adjust": None})()
    rows = [dict(prompt=[5, 9, 5], output=[9, 9, 2], rep=1.3, freq=0.5, pres=-1.0, bias={7: 2.5, 9: -1.0}), None,
            dict(prompt=[1], output=[], rep=1.0, freq=0.0, pres=0.0, bias={4: 1.0})]
    n, E = R.ModelRunner.pack_adjust(stub, rows)
    a = stub._h_adjust_np
    assert (n, E) == (3, 6)
    params = a[:9].view(np.float32).tolist()
    assert params == pytest.approx([1.3, 0.5, -1.0, 1.0, 0.0, 0.0, 1.0, 0.0, 0.0])
    assert a[9:13].tolist() == [0, 4, 4, 6]
    o = 13
    assert a[o:o + E].tolist() == [2, 5, 7, 9, 1, 4]                       # ids, sorted per row
    assert a[o + E:o + 2 * E].tolist() == [1, 0, 0, 2, 0, 0]               # occurrences in the output
    assert a[o + 2 * E:o + 3 * E].tolist() == [1, 1, 0, 1, 1, 0]           # in prompt or output (7 and 4 are bias-only)
    assert a[o + 3 * E:o + 4 * E].view(np.float32).tolist() == pytest.approx([0.0, 0.0, 2.5, -1.0, 0.0, 1.0])
    with pytest.raises(ValueError):
        R.ModelRunner.pack_adjust(stub, [None] * 9)
