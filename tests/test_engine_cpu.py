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
