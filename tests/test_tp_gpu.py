"""GPU (needs >= 2 devices; skipped otherwise): tensor-parallel decode through the C-ABI with the
ctx's own NCCL communicator — TP=2 logits and greedy tokens against the single-rank oracle."""
import os

import pytest
import torch

from conftest import pkg
from oracle import decode_oracle as O

pytestmark = pytest.mark.gpu


def _worker(rank, world, name, port, out_q, peer):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)      # only carries the NCCL unique id
    C, R = pkg("config"), pkg("runner")
    ocfg = O.CONFIGS[name]
    w = O.make_weights(ocfg, seed=1234, norm_jitter=0.1)
    r = R.ModelRunner(C.ModelConfig(**ocfg.to_dict()), max_tokens=64, max_model_len=512, n_blocks=8,
                      tp_rank=rank, tp_world=world)
    r.load_weights(w)
    r.init_tp(peer_allreduce=peer)
    ids = O.make_prompt(ocfg, 5, 40)
    T = len(ids)
    bt = torch.tensor([[3]], dtype=torch.int32, device="cuda")
    pos = torch.arange(T, dtype=torch.int32, device="cuda")
    logits = torch.empty(world, T, ocfg.vocab // world, dtype=torch.bfloat16, device="cuda")
    for graphs in (False, True):
        r.set_graphs(graphs)
        toks = r.step_device(torch.tensor(ids, dtype=torch.int32).cuda(), pos, 3 * 128 + pos, bt, pos + 1, T,
                             row_to_seq=torch.zeros(T, dtype=torch.int32).cuda(), out_logits=logits)
        torch.cuda.synchronize()
    full = logits.permute(1, 0, 2).reshape(T, ocfg.vocab).float().cpu()
    # the collective of a half-layer on its own (b200_tp_allreduce): plain all-reduce and the fused
    # all-reduce + add + RMSNorm, at the row counts that select the fused kernel (<= 16) and the one-shot kernel
    abi, h = pkg("_abi"), ocfg.hidden
    for i, Tr in enumerate((1, 16, 40, 64)):
        g = torch.Generator().manual_seed(100 + Tr)
        parts = [torch.randn(Tr, h, generator=g).to(torch.bfloat16) for _ in range(world)]
        res0 = torch.randn(Tr, h, generator=g).to(torch.bfloat16)
        nw = (1 + 0.1 * torch.randn(h, generator=g)).to(torch.bfloat16)
        total = parts[0].float()
        for p in parts[1:]:
            total = total + p.float()                      # rank order, fp32, one rounding
        want_x = total.to(torch.bfloat16)
        mine, y = parts[rank].cuda(), torch.empty(Tr, h, dtype=torch.bfloat16, device="cuda")
        abi.check(r.lib.b200_tp_allreduce(r.ctx, 0, mine.data_ptr(), Tr, None, None, y.data_ptr(), r.stream.cuda_stream))
        r.stream.synchronize()
        assert torch.equal(y.cpu(), want_x), f"all-reduce of {Tr} rows differs from the rank-ordered fp32 sum"
        resd, nwd = res0.cuda(), nw.cuda()
        abi.check(r.lib.b200_tp_allreduce(r.ctx, 1, mine.data_ptr(), Tr, resd.data_ptr(), nwd.data_ptr(), y.data_ptr(),
                                          r.stream.cuda_stream))
        r.stream.synchronize()
        want_y, want_res = O.add_rms_norm(want_x, res0, nw, ocfg.rms_eps)
        # above 16 rows the peer-memory path deals the rows to the ranks: a rank keeps the residual of ITS rows only
        mine_rows = slice(rank, None, world) if (peer and Tr > 16) else slice(None)
        assert torch.equal(resd.cpu()[mine_rows], want_res[mine_rows]), f"residual after the fused all-reduce + norm ({Tr} rows)"
        d = (y.cpu().float() - want_y.float()).abs()
        ulp = want_y.float().abs().clamp_min(2.0 ** -126).log2().floor().exp2() * 2.0 ** -7
        assert (d <= torch.maximum(2 * ulp, torch.tensor(2.0 ** -9))).all() and (d > 0).float().mean() <= 0.01
    if rank == 0:
        out_q.put((full, toks.cpu()))
    dist.barrier()
    r.close()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("peer", [True, False], ids=["peer-memory-allreduce", "nccl-allreduce"])
def test_tp2_decode_step_matches_oracle(peer):
    import torch.multiprocessing as mp
    name = "tiny-g8"
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 2000 + (7 if peer else 0)
    procs = [ctx.Process(target=_worker, args=(r, 2, name, port, q, peer)) for r in range(2)]
    for p in procs:
        p.start()
    full, toks = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    cfg = O.CONFIGS[name]
    w = O.make_weights(cfg, seed=1234, norm_jitter=0.1)
    om = O.OracleModel(cfg, w, capacity=64)
    want = om.forward(om.new_seq(), O.make_prompt(cfg, 5, 40)).float()
    assert (torch.softmax(full, -1) - torch.softmax(want, -1)).abs().max() <= 2e-3
    assert toks.tolist() == full.argmax(-1).tolist()      # vocab-parallel pick == argmax of the gathered logits
    dis = toks != want.argmax(-1)
    assert (O.top2_margin(want)[dis] < 0.08).all()


def _build_tiny_g8(rank, world):
    C, R = pkg("config"), pkg("runner")
    ocfg = O.CONFIGS["tiny-g8"]
    r = R.ModelRunner(C.ModelConfig(**ocfg.to_dict()), max_tokens=64, max_model_len=512, n_blocks=16, tp_rank=rank, tp_world=world)
    r.load_weights(O.make_weights(ocfg, seed=1234, norm_jitter=0.1))
    return r


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_tp2_serving_executor_generates_oracle_tokens():
    """The server's multi-process executor: rank 0 schedules and broadcasts each step, a spawned worker
    rank executes it on its shard; greedy tokens through the engine match the oracle."""
    TP, E = pkg("tp_exec"), pkg("engine")
    leader, procs = TP.launch_workers(2, _build_tiny_g8, port=29800 + os.getpid() % 1000)
    try:
        eng = E.LLMEngine(leader, max_num_batched_tokens=64, max_num_seqs=4, max_model_len=512)
        cfg = O.CONFIGS["tiny-g8"]
        prompts = [O.make_prompt(cfg, 7, 20), O.make_prompt(cfg, 8, 70)]
        outs = eng.generate(prompts, max_tokens=12)
        om = O.OracleModel(cfg, O.make_weights(cfg, seed=1234, norm_jitter=0.1), capacity=128)
        for p, got in zip(prompts, outs):
            want, logits = om.generate(p, 12, keep_logits=True)
            for i, (a, b) in enumerate(zip(got, want)):
                if a != b:
                    assert float(O.top2_margin(logits[i])) < 0.08
                    break
        # sampling and logprobs are collectives over the vocab shards: seeded draws replay, top_k = 1 and the
        # top-1 logprob agree with the greedy token
        def run(specs):
            eng2 = E.LLMEngine(leader, max_num_batched_tokens=64, max_num_seqs=4, max_model_len=512)
            reqs = [E.Request(f"s{i}", list(prompts[0]), max_tokens=8, ignore_eos=True, **kw) for i, kw in enumerate(specs)]
            for q in reqs:
                eng2.add_request(q)
            while eng2.has_work():
                eng2.step()
            return reqs
        a = run([dict(temperature=1.2, seed=5), dict(temperature=0.9, top_k=1, seed=9), dict(logprobs=2)])
        b = run([dict(temperature=1.2, seed=5), dict(temperature=1.2, seed=6), dict()])
        assert a[0].output_ids == b[0].output_ids, "a seeded request must replay its tokens"
        assert a[0].output_ids != b[1].output_ids, "another seed, another stream"
        assert a[2].output_ids == b[2].output_ids, "greedy rows do not depend on their neighbours' sampling"
        # top_k = 1 is the argmax unless the two best bf16 logits tie (both are then kept, as in vLLM)
        for i, (x, y) in enumerate(zip(a[1].output_ids, a[2].output_ids)):
            if x != y:
                top = a[2].logprob_data[i][1]
                assert top[0][1] == top[1][1], f"top_k=1 left the argmax at step {i} without a tie"
                break
        assert all(0 <= t < cfg.vocab for t in a[0].output_ids)
        # logit_bias / penalties are applied shard by shard on every rank: a +100 bias forces its token on whichever rank
        # owns it (one id per vocabulary shard), and a penalised request still runs next to an untouched one
        c = run([dict(logit_bias={5: 100.0}), dict(logit_bias={cfg.vocab - 3: 100.0}, repetition_penalty=1.4), dict()])
        assert c[0].output_ids == [5] * 8 and c[1].output_ids == [cfg.vocab - 3] * 8
        assert c[2].output_ids == b[2].output_ids
        for tok, (lp, top) in zip(a[2].output_ids, a[2].logprob_data):
            assert top[0][0] == tok and abs(top[0][1] - lp) < 1e-6 and lp <= 0
    finally:
        TP.shutdown(leader, procs)
    assert all(p.exitcode == 0 for p in procs)
