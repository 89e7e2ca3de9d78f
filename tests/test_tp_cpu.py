"""CPU, world_size 2 over gloo: the tensor-parallel host logic — Megatron sharding of the HF weights
(weights.shard_weights), the all-reduce placement (after O and after down) and the vocab-parallel
greedy pick (max, index with ties -> lowest index) — reproduces the single-rank oracle."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import pkg
from oracle import decode_oracle as O


def _tp_forward(rank, world, cfg_dict, seed, prompt, port, out_q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    W = pkg("weights")
    C = pkg("config")
    cfg = C.ModelConfig(**cfg_dict)
    ocfg = O.ModelConfig(**cfg_dict)
    w = O.make_weights(ocfg, seed=seed, norm_jitter=0.1)
    sh = {}
    for layer, kind, t in W.shard_weights(w, cfg, rank, world):
        sh[(layer, kind)] = t
    hq, hkv, I = cfg.n_q_heads // world, cfg.n_kv_heads // world, cfg.intermediate // world
    T = len(prompt)
    cs = O.cos_sin_cache(ocfg, 64)
    pos = torch.arange(T)
    cos, sin = cs[pos, :64], cs[pos, 64:]

    def allreduce(x):                       # bf16 sum over ranks, as NCCL does on the device
        y = x.float()
        dist.all_reduce(y)
        return y.to(torch.bfloat16)

    x = sh[(0, W.EMBED)][torch.tensor(prompt)]
    res = None
    for l in range(cfg.n_layers):
        if res is None:
            res, y = x, O.rms_norm(x, sh[(l, W.LN1)], cfg.rms_eps)
        else:
            y, res = O.add_rms_norm(x, res, sh[(l, W.LN1)], cfg.rms_eps)
        qkv = O.linear(y, sh[(l, W.QKV)], sh.get((l, W.QKV_BIAS)))
        q = O.rope_apply(qkv[:, :hq * 128].view(T, hq, 128), cos, sin)
        k = O.rope_apply(qkv[:, hq * 128:(hq + hkv) * 128].view(T, hkv, 128), cos, sin)
        v = qkv[:, (hq + hkv) * 128:].view(T, hkv, 128)
        a = torch.stack([O.attend(q[t], k, v, t + 1, 128 ** -0.5) for t in range(T)]).view(T, hq * 128)
        x = allreduce(O.linear(a, sh[(l, W.O)]))
        y, res = O.add_rms_norm(x, res, sh[(l, W.LN2)], cfg.rms_eps)
        gu_w = sh.get((l, W.GATE_UP_INTERLEAVED))
        if gu_w is not None:                # undo the 64-row interleave: [slab][gate64|up64]
            g3 = gu_w.view(I // 64, 2, 64, cfg.hidden)
            gu_w = torch.cat([g3[:, 0].reshape(I, -1), g3[:, 1].reshape(I, -1)], 0)
        else:
            gu_w = sh[(l, W.GATE_UP)]
        x = allreduce(O.linear(O.silu_mul(O.linear(y, gu_w)), sh[(l, W.DOWN)]))
    y, _ = O.add_rms_norm(x, res, sh[(0, W.FINAL_NORM)], cfg.rms_eps)
    logits = O.linear(y, sh[(0, W.LM_HEAD)]).float()                 # [T, V/world]
    # vocab-parallel greedy: (max, global index) per rank, gather, highest value / lowest index wins
    mx, idx = logits.max(-1)
    idx = idx + rank * logits.shape[1]
    pairs = [torch.zeros(T, 2, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(pairs, torch.stack([mx.double(), idx.double()], -1))
    best_v = torch.full((T,), -float("inf"), dtype=torch.float64)
    best_i = torch.full((T,), 2 ** 40, dtype=torch.float64)
    for pr in pairs:
        take = (pr[:, 0] > best_v) | ((pr[:, 0] == best_v) & (pr[:, 1] < best_i))
        best_v = torch.where(take, pr[:, 0], best_v)
        best_i = torch.where(take, pr[:, 1], best_i)
    gathered = [torch.zeros_like(logits) for _ in range(world)]
    dist.all_gather(gathered, logits)
    if rank == 0:
        out_q.put((torch.cat(gathered, -1), best_i.long()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("name", ["tiny-g8"])
def test_tp2_sharding_allreduce_and_vocab_parallel_pick_match_oracle(name):
    cfg = O.CONFIGS[name]
    prompt = O.make_prompt(cfg, 5, 12)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_tp_forward, args=(r, 2, cfg.to_dict(), 1234, prompt, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    logits, picked = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    w = O.make_weights(cfg, seed=1234, norm_jitter=0.1)
    om = O.OracleModel(cfg, w, capacity=64)
    want = om.forward(om.new_seq(), prompt).float()
    assert (torch.softmax(logits, -1) - torch.softmax(want, -1)).abs().max() <= 2e-3
    assert picked.tolist() == logits.argmax(-1).tolist()            # the pick == global argmax of the gathered logits
    dis = picked != want.argmax(-1)
    assert (O.top2_margin(want)[dis] < 0.08).all()


def test_shards_partition_the_weights():
    W = pkg("weights")
    C = pkg("config")
    ocfg = O.CONFIGS["tiny-g8"]
    cfg = C.ModelConfig(**ocfg.to_dict())
    w = O.make_weights(ocfg)
    full = {(l, k): t for l, k, t in W.shard_weights(w, cfg, 0, 1)}
    parts = [{(l, k): t for l, k, t in W.shard_weights(w, cfg, r, 2)} for r in range(2)]
    hq, hkv = cfg.n_q_heads // 2, cfg.n_kv_heads // 2
    for l in range(cfg.n_layers):
        qkv = full[(l, W.QKV)]
        q, k, v = qkv[:cfg.q_size], qkv[cfg.q_size:cfg.q_size + cfg.kv_size], qkv[cfg.q_size + cfg.kv_size:]
        for r in range(2):
            s = parts[r][(l, W.QKV)]
            assert torch.equal(s[:hq * 128], q[r * hq * 128:(r + 1) * hq * 128])
            assert torch.equal(s[hq * 128:(hq + hkv) * 128], k[r * hkv * 128:(r + 1) * hkv * 128])
            assert torch.equal(s[(hq + hkv) * 128:], v[r * hkv * 128:(r + 1) * hkv * 128])
        assert torch.equal(torch.cat([parts[r][(l, W.O)] for r in range(2)], 1), full[(l, W.O)])
        assert torch.equal(torch.cat([parts[r][(l, W.DOWN)] for r in range(2)], 1), full[(l, W.DOWN)])
    assert torch.equal(torch.cat([parts[r][(0, W.LM_HEAD)] for r in range(2)], 0), full[(0, W.LM_HEAD)])
    with pytest.raises(ValueError):
        list(W.shard_weights(w, cfg, 0, 3))


class _ProtoRunner:
    """Stand-in for a ModelRunner on the CPU: enough of its buffers for tp_exec's broadcast protocol."""
    def __init__(self, log):
        self.n_blocks, self.block_size, self.max_tokens, self.max_model_len, self.cfg = 8, 16, 32, 64, None
        self._h_meta = torch.zeros(4096, dtype=torch.int32)
        self._h_sample = torch.zeros(64 * 32, dtype=torch.uint8)
        self.log = log
        self.h2d_bytes_last = 0

    def pack(self, batch):
        self._h_meta[:5] = torch.tensor(batch, dtype=torch.int32)
        return 1, 1, 1, 1, 5

    def step_packed(self, T, R, S, bt_stride, max_ctx, n_ints, sync=True):
        self.log.append(("step", self._h_meta[:n_ints].tolist(), max_ctx))
        return [0]

    def pack_sampling(self, temperature, top_k, top_p, seeds, draws, min_p=None):
        n = len(temperature)
        a = self._h_sample.numpy()
        a[:32 * n] = 0
        a[0:4 * n].view("float32")[:] = temperature
        a[16 * n:24 * n].view("uint64")[:] = seeds
        return n

    def sample_packed(self, n, sync=True):
        a = self._h_sample.numpy()
        self.log.append(("sample", a[0:4 * n].view("float32").tolist(), a[16 * n:24 * n].view("uint64").tolist()))
        return [0] * n


def _proto_rank(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    X = pkg("tp_exec")
    log = []
    r = _ProtoRunner(log)
    if rank == 0:
        lead = X.TPLeader(r)
        batch = type("B", (), {"max_ctx": 33})()
        r.pack = lambda b: (r._h_meta.__setitem__(slice(0, 5), torch.tensor([1, 2, 3, 4, 5], dtype=torch.int32)), (1, 1, 1, 1, 5))[1]
        lead.step(batch)
        lead.sample_last([0.5, 0.0], [0, 0], [1.0, 1.0], [77, 78], [0, 0])
        lead.step(batch)
        lead.stop()
    else:
        X.worker_loop(r)
    q.put((rank, log))
    dist.barrier()
    dist.destroy_process_group()


def test_leader_worker_protocol_carries_steps_and_sampling():
    """world_size 2 over gloo: every step and every sampling call of rank 0 reaches the worker with the same
    payload (the device-side collectives need both ranks to make the same calls in the same order)."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() + 7) % 2000
    procs = [ctx.Process(target=_proto_rank, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    logs = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert logs[0] == logs[1]
    kinds = [e[0] for e in logs[1]]
    assert kinds == ["step", "sample", "step"]
    assert logs[1][0][1] == [1, 2, 3, 4, 5] and logs[1][0][2] == 33
    assert logs[1][1][1] == [0.5, 0.0] and logs[1][1][2] == [77, 78]
