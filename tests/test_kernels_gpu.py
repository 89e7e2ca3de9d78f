"""GPU parity, kernel by kernel, through the C-ABI, against oracle/decode_oracle.py.
Tolerances: bit-exact for integer/index/copy work (embed, KV append, argmax) and for the bf16
elementwise arithmetic of RoPE; for fp32-accumulated ops the only freedom is summation order, so
the bar is <= 1 bf16 ulp on a small fraction of elements (stated per test)."""
import numpy as np
import os

import pytest
import torch

from conftest import pkg
from oracle import decode_oracle as O

pytestmark = pytest.mark.gpu
BF = torch.bfloat16


@pytest.fixture(scope="module")
def lib():
    return pkg("_abi").lib()


@pytest.fixture(scope="module")
def abi():
    return pkg("_abi")


_keep = []


def dev(t):
    """Upload and keep alive until the test ends (a temporary would be freed — and its memory
    reused — before the asynchronous kernel reads it)."""
    d = t.cuda().contiguous()
    _keep.append(d)
    return d


@pytest.fixture(autouse=True)
def _release_uploads():
    yield
    torch.cuda.synchronize()
    _keep.clear()


def ulp_close(got, want, max_ulp=1, max_frac=0.01, atol_small=2 ** -9):
    """bf16 tensors: every element within max_ulp bf16 ulps of `want` (or atol_small near zero),
    at most max_frac of them different at all."""
    g, w = got.float().cpu(), want.float().cpu()
    diff = (g - w).abs()
    ulp = torch.maximum(w.abs(), torch.tensor(2.0 ** -126)).log2().floor().exp2() * 2.0 ** -7
    bad = diff > torch.maximum(max_ulp * ulp, torch.tensor(atol_small))
    assert not bad.any(), f"{int(bad.sum())} elements beyond {max_ulp} ulp; max diff {diff.max():.4g}"
    frac = (diff > 0).float().mean().item()
    _fractions.append((os.environ.get("PYTEST_CURRENT_TEST", "?").split(" ")[0], frac, max_frac, float(diff.max())))
    assert frac <= max_frac, f"{frac:.4f} of elements differ (allowed {max_frac})"
    return frac


_fractions = []


@pytest.fixture(scope="module", autouse=True)
def _dump_fractions():
    """Measured differing fractions of every ulp_close call, next to their bounds (gpurun_out/ulp_fractions.txt when
    that directory exists): a regression inside a bound shows up in the file before it breaks a test."""
    yield
    d = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if os.path.isdir(d) and _fractions:
        with open(os.path.join(d, "ulp_fractions.txt"), "w") as f:
            for name, frac, bound, mx in _fractions:
                f.write(f"{frac:.5f} (bound {bound}) max|d| {mx:.3e}  {name}\n")


def sync_check(abi, rc):
    abi.check(rc)
    torch.cuda.synchronize()


# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("T,n_out,k,bias", [(1, 512, 512, False), (1, 6144, 4096, False), (2, 1536, 640, True),
                                             (3, 1000, 1024, True), (4, 4096, 14336, False), (7, 264, 2048, False),
                                             (1, 2, 8, False), (9, 128256 // 8, 512, False)])
def test_linear_gemv(lib, abi, T, n_out, k, bias):
    g = torch.Generator().manual_seed(T * 1000 + n_out + k)
    x = torch.randn(T, k, generator=g).to(BF)
    w = (0.05 * torch.randn(n_out, k, generator=g)).to(BF)
    b = torch.randn(n_out, generator=g).to(BF) if bias else None
    want = O.linear(x, w, b)
    xd, wd, bd = dev(x), dev(w), (dev(b) if bias else None)
    y = torch.zeros(T, n_out, dtype=BF, device="cuda")
    sync_check(abi, lib.b200_linear_bf16(xd.data_ptr(), k, wd.data_ptr(), abi.ptr(bd), y.data_ptr(), n_out, T, n_out, k,
                                         1, None, 0, None))
    ulp_close(y, want, max_ulp=1, max_frac=0.02)


@pytest.mark.parametrize("T,n_out,k,bias", [(16, 128, 64, False), (5, 256, 128, True), (64, 6144, 4096, False),
                                             (17, 264, 512, True), (100, 1000, 1024, True), (64, 4096, 14336, False),
                                             (256, 512, 4096, False), (300, 384, 2048, True), (8, 128, 200, False),
                                             (1, 1024, 4096, False), (33, 5120 + 2048, 5120, True)])
def test_linear_tcgen05(lib, abi, T, n_out, k, bias):
    """The tcgen05/TMEM weight-streaming GEMM (impl=2): ragged T / n_out / k (TMA zero-fills the
    out-of-bounds part of a box), split-K and direct epilogues, twice in a row on the same
    workspace (the slab tickets must be left at zero)."""
    g = torch.Generator().manual_seed(T * 7 + n_out + k)
    x = torch.randn(T, k, generator=g).to(BF)
    w = (0.05 * torch.randn(n_out, k, generator=g)).to(BF)
    b = torch.randn(n_out, generator=g).to(BF) if bias else None
    want = O.linear(x, w, b)
    xd, wd, bd = dev(x), dev(w), (dev(b) if bias else None)
    wsb = lib.b200_linear_workspace_bytes(T, n_out, k)
    ws = torch.zeros(wsb, dtype=torch.uint8, device="cuda")
    for _ in range(2):
        y = torch.full((T, n_out), float("nan"), dtype=BF, device="cuda")
        sync_check(abi, lib.b200_linear_bf16(xd.data_ptr(), k, wd.data_ptr(), abi.ptr(bd), y.data_ptr(), n_out, T, n_out, k,
                                             2, ws.data_ptr(), wsb, None))
        ulp_close(y, want, max_ulp=1, max_frac=0.02)
    assert not ws[:65536].any(), "split-K tickets must be zero again after the launch"


def test_linear_global_split_k_path_in_a_fresh_process():
    """B200_TC_CLUSTER=0 (read once per process) keeps the split-K partials in the global workspace with a last-CTA
    ticket instead of reducing through DSMEM inside a cluster: the A/B path of DESIGN.md §4, kept correct by this run of
    the linear test's split shapes in a child process."""
    import subprocess, sys
    env = dict(os.environ, B200_TC_CLUSTER="0")
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", os.path.abspath(__file__), "-k",
                        "test_linear_tcgen05 and (64-6144-4096 or 256-512-4096 or 64-4096-14336 or 1-1024-4096)"],
                       env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and " passed" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_linear_tcgen05_matches_gemv_bitwise_up_to_order(lib, abi):
    """Both linear paths implement R2 (fp32 accumulate, one rounding): they may differ only by
    summation order."""
    T, n_out, k = 4, 4096, 4096
    g = torch.Generator().manual_seed(11)
    x, w = dev(torch.randn(T, k, generator=g).to(BF)), dev((0.05 * torch.randn(n_out, k, generator=g)).to(BF))
    y1 = torch.empty(T, n_out, dtype=BF, device="cuda")
    y2 = torch.empty_like(y1)
    wsb = lib.b200_linear_workspace_bytes(T, n_out, k)
    ws = torch.zeros(wsb, dtype=torch.uint8, device="cuda")
    sync_check(abi, lib.b200_linear_bf16(x.data_ptr(), k, w.data_ptr(), None, y1.data_ptr(), n_out, T, n_out, k, 1, None, 0, None))
    sync_check(abi, lib.b200_linear_bf16(x.data_ptr(), k, w.data_ptr(), None, y2.data_ptr(), n_out, T, n_out, k, 2, ws.data_ptr(), wsb, None))
    ulp_close(y2, y1.cpu(), max_ulp=1, max_frac=0.02)


@pytest.mark.parametrize("T,I,h", [(3, 1024, 512), (64, 14336, 4096), (40, 1792, 1024), (300, 256, 512)])
def test_linear_fused_silu_mul_is_bit_identical_to_unfused(lib, abi, T, I, h):
    """gate_up on interleaved slabs with SiLU*mul in the epilogue == tcgen05 linear + b200_silu_mul."""
    g = torch.Generator().manual_seed(I + T)
    x = dev(torch.randn(T, h, generator=g).to(BF))
    gate = (0.05 * torch.randn(I, h, generator=g)).to(BF)
    up = (0.05 * torch.randn(I, h, generator=g)).to(BF)
    w_plain = dev(torch.cat([gate, up], 0))
    w_il = dev(torch.stack([gate.reshape(I // 64, 64, h), up.reshape(I // 64, 64, h)], dim=1).reshape(2 * I, h))
    wsb = lib.b200_linear_workspace_bytes(T, 2 * I, h)
    ws = torch.zeros(wsb, dtype=torch.uint8, device="cuda")
    gu = torch.empty(T, 2 * I, dtype=BF, device="cuda")
    ref = torch.empty(T, I, dtype=BF, device="cuda")
    got = torch.full((T, I), float("nan"), dtype=BF, device="cuda")
    sync_check(abi, lib.b200_linear_bf16(x.data_ptr(), h, w_plain.data_ptr(), None, gu.data_ptr(), 2 * I, T, 2 * I, h, 2,
                                         ws.data_ptr(), wsb, None))
    sync_check(abi, lib.b200_silu_mul(gu.data_ptr(), ref.data_ptr(), T, I, None))
    sync_check(abi, lib.b200_linear_fused_bf16(x.data_ptr(), h, w_il.data_ptr(), None, got.data_ptr(), I, T, 2 * I, h, 1,
                                               None, None, None, None, None, 0, 0, 0, ws.data_ptr(), wsb, None))
    assert torch.equal(got, ref)
    # and the activation math against the oracle on the same linear output (the linear itself is
    # checked against the oracle in test_linear_tcgen05; through silu a 1-ulp flip of g is amplified)
    ulp_close(got, O.silu_mul(gu.cpu()), max_ulp=1, max_frac=0.005)


@pytest.mark.parametrize("T,hq,hkv,h,bias", [(5, 4, 1, 512, False), (64, 32, 8, 4096, False), (33, 5, 1, 640, True),
                                              (200, 16, 2, 1024, False)])
def test_linear_fused_rope_kv_is_bit_identical_to_unfused(lib, abi, T, hq, hkv, h, bias):
    """qkv projection with RoPE + paged KV append in the epilogue == tcgen05 linear + b200_rope_kv_append."""
    g = torch.Generator().manual_seed(T + hq)
    n_out, bs, nb = (hq + 2 * hkv) * 128, 128, 6
    x = dev(torch.randn(T, h, generator=g).to(BF))
    w = dev((0.05 * torch.randn(n_out, h, generator=g)).to(BF))
    b = dev(torch.randn(n_out, generator=g).to(BF)) if bias else None
    pos = torch.randint(0, 700, (T,), generator=g).to(torch.int32)
    slots_all = torch.randperm(nb * bs, generator=g)[:T].to(torch.int32)
    slots_all[T // 2] = -1
    posd, slotd = dev(pos), dev(slots_all)
    cs = dev(O.cos_sin_cache(O.CONFIGS["tiny"], 1024))
    wsb = lib.b200_linear_workspace_bytes(T, n_out, h)
    ws = torch.zeros(wsb, dtype=torch.uint8, device="cuda")
    out = {}
    for fused in (False, True):
        y = torch.zeros(T, n_out, dtype=BF, device="cuda")
        kc = torch.zeros(nb, hkv, bs, 128, dtype=BF, device="cuda")
        vc = torch.zeros_like(kc)
        if fused:
            sync_check(abi, lib.b200_linear_fused_bf16(x.data_ptr(), h, w.data_ptr(), abi.ptr(b), y.data_ptr(), n_out, T, n_out, h,
                                                       2, posd.data_ptr(), slotd.data_ptr(), cs.data_ptr(), kc.data_ptr(),
                                                       vc.data_ptr(), hq, hkv, bs, ws.data_ptr(), wsb, None))
        else:
            sync_check(abi, lib.b200_linear_bf16(x.data_ptr(), h, w.data_ptr(), abi.ptr(b), y.data_ptr(), n_out, T, n_out, h, 2,
                                                 ws.data_ptr(), wsb, None))
            sync_check(abi, lib.b200_rope_kv_append(y.data_ptr(), posd.data_ptr(), slotd.data_ptr(), cs.data_ptr(),
                                                    kc.data_ptr(), vc.data_ptr(), T, hq, hkv, bs, None))
        out[fused] = (y[:, :hq * 128].clone(), kc, vc)
    for a, bb in zip(out[True], out[False]):
        assert torch.equal(a, bb)
    assert out[True][1].any() and out[True][2].any()


@pytest.mark.parametrize("T,h,k", [(1, 4096, 4096), (5, 512, 1024), (16, 4096, 14336), (33, 640, 1536), (64, 4096, 4096),
                                   (64, 1024, 2048)])
def test_folded_norm_producer_is_bit_identical_to_linear_then_add_norm(lib, abi, T, h, k):
    """O / down projection that adds into the fp32 residual stream (b200_linear_residual_bf16) followed by
    b200_rmsnorm_stream == tcgen05 linear followed by the add+RMSNorm kernel: normed rows and residual bit for bit,
    twice in a row (the stream is updated in place)."""
    g = torch.Generator().manual_seed(T + h + k)
    V = 97
    table = dev(torch.randn(V, h, generator=g).to(BF))
    ids = dev(torch.randint(0, V, (T,), generator=g).to(torch.int32))
    w = dev((0.05 * torch.randn(h, k, generator=g)).to(BF))
    nw = dev((1 + 0.1 * torch.randn(h, generator=g)).to(BF))
    wsb = lib.b200_linear_workspace_bytes(T, h, k)
    ws = torch.zeros(wsb, dtype=torch.uint8, device="cuda")
    sres = torch.full((T, h), float("nan"), dtype=torch.float32, device="cuda")
    ssq = torch.full((T, h // 128), float("nan"), dtype=torch.float32, device="cuda")
    res = torch.empty(T, h, dtype=BF, device="cuda")
    sync_check(abi, lib.b200_embed_stream(ids.data_ptr(), table.data_ptr(), sres.data_ptr(), ssq.data_ptr(), T, h, V, None))
    sync_check(abi, lib.b200_embed(ids.data_ptr(), table.data_ptr(), res.data_ptr(), T, h, V, None))
    assert torch.equal(sres, res.float())
    for it in range(2):
        x = dev(torch.randn(T, k, generator=g).to(BF))
        y = torch.empty(T, h, dtype=BF, device="cuda")
        want = torch.empty(T, h, dtype=BF, device="cuda")
        got = torch.full((T, h), float("nan"), dtype=BF, device="cuda")
        sync_check(abi, lib.b200_linear_bf16(x.data_ptr(), k, w.data_ptr(), None, y.data_ptr(), h, T, h, k, 2, ws.data_ptr(), wsb, None))
        sync_check(abi, lib.b200_rmsnorm(y.data_ptr(), res.data_ptr(), nw.data_ptr(), want.data_ptr(), T, h, 1e-5, None))
        sync_check(abi, lib.b200_linear_residual_bf16(x.data_ptr(), k, w.data_ptr(), None, sres.data_ptr(), ssq.data_ptr(), T, h, k,
                                                      ws.data_ptr(), wsb, None))
        sync_check(abi, lib.b200_rmsnorm_stream(sres.data_ptr(), ssq.data_ptr(), None, nw.data_ptr(), got.data_ptr(), T, h, 1e-5, None))
        assert torch.equal(sres.to(BF), res), f"residual differs (pass {it})"
        assert torch.equal(got, want), f"normed rows differ (pass {it})"
    assert not ws[:65536].any()


@pytest.mark.parametrize("T,h,n_out,mode", [(1, 4096, 6144, 2), (1, 4096, 28672, 1), (3, 512, 1024, 0), (16, 4096, 6144, 2),
                                            (17, 640, 3072, 1), (40, 1024, 2560, 2), (64, 4096, 28672, 1), (64, 4096, 6144, 2),
                                            (9, 4096, 128256 // 4, 0)])
def test_folded_norm_consumer_is_bit_identical_to_norm_then_linear(lib, abi, T, h, n_out, mode):
    """QKV / gate_up / lm_head linears that normalise their input rows from the fp32 stream inside the kernel
    (b200_linear_normed_bf16; plain, SiLU*mul and RoPE+KV epilogues) == b200_rmsnorm_stream followed by the same
    linear over the bf16 rows: outputs and cache bit for bit.  Mode 0 also goes through the row indirection."""
    g = torch.Generator().manual_seed(T + h + n_out + mode)
    V = 211
    table = dev(torch.randn(V, h, generator=g).to(BF))
    ids = dev(torch.randint(0, V, (T,), generator=g).to(torch.int32))
    w = dev((0.05 * torch.randn(n_out, h, generator=g)).to(BF))
    nw = dev((1 + 0.1 * torch.randn(h, generator=g)).to(BF))
    sres = torch.empty(T, h, dtype=torch.float32, device="cuda")
    ssq = torch.empty(T, h // 128, dtype=torch.float32, device="cuda")
    sync_check(abi, lib.b200_embed_stream(ids.data_ptr(), table.data_ptr(), sres.data_ptr(), ssq.data_ptr(), T, h, V, None))
    sres.mul_(1.37)            # not exactly bf16 any more: s * rinv must be taken on the fp32 value
    ssq.mul_(1.37 * 1.37)
    wsb = lib.b200_linear_workspace_bytes(T, n_out, h)
    ws = torch.zeros(wsb, dtype=torch.uint8, device="cuda")
    rows = None
    R = T
    if mode == 0 and T > 1:
        rows = dev(torch.randperm(T, generator=g)[: max(T // 2, 1)].to(torch.int32))
        R = rows.numel()
    normed = torch.empty(R, h, dtype=BF, device="cuda")
    sync_check(abi, lib.b200_rmsnorm_stream(sres.data_ptr(), ssq.data_ptr(), abi.ptr(rows), nw.data_ptr(), normed.data_ptr(), R, h, 1e-5, None))
    hq, hkv, bs, nb = 0, 0, 128, 4
    pos = slots = cs = None
    ldy = n_out
    if mode == 1:
        ldy = n_out // 2
    if mode == 2:
        hkv = {48: 8, 20: 2}.get(n_out // 128, 1)
        hq = n_out // 128 - 2 * hkv
        pos = dev(torch.randint(0, 700, (T,), generator=g).to(torch.int32))
        sl = torch.randperm(nb * bs, generator=g)[:T].to(torch.int32)
        if T > 2:
            sl[T // 2] = -1
        slots = dev(sl)
        cs = dev(O.cos_sin_cache(O.CONFIGS["tiny"], 1024))
    out = {}
    for fused in (False, True):
        y = torch.full((R, ldy), float("nan"), dtype=BF, device="cuda")
        kc = torch.zeros(nb, max(hkv, 1), bs, 128, dtype=BF, device="cuda")
        vc = torch.zeros_like(kc)
        tail = (abi.ptr(pos), abi.ptr(slots), abi.ptr(cs), kc.data_ptr(), vc.data_ptr(), hq, hkv, bs, ws.data_ptr(), wsb, None)
        if fused:
            sync_check(abi, lib.b200_linear_normed_bf16(sres.data_ptr(), ssq.data_ptr(), nw.data_ptr(), 1e-5, abi.ptr(rows),
                                                        w.data_ptr(), None, y.data_ptr(), ldy, R, n_out, h, mode, *tail))
        elif mode == 0:
            sync_check(abi, lib.b200_linear_bf16(normed.data_ptr(), h, w.data_ptr(), None, y.data_ptr(), ldy, R, n_out, h, 2,
                                                 ws.data_ptr(), wsb, None))
        else:
            sync_check(abi, lib.b200_linear_fused_bf16(normed.data_ptr(), h, w.data_ptr(), None, y.data_ptr(), ldy, R, n_out, h,
                                                       mode, *tail))
        out[fused] = (y[:, :hq * 128].clone() if mode == 2 else y, kc, vc)
    for a, bb in zip(out[True], out[False]):
        assert torch.equal(a, bb)
    assert not ws[:65536].any()
    # and the normalised rows themselves against the oracle's R1 on the same fp32 stream
    src = sres.cpu() if rows is None else sres.cpu()[rows.cpu().long()]
    var = src.pow(2).mean(-1, keepdim=True)
    want = ((src * torch.rsqrt(var + 1e-5)).to(BF).float() * nw.cpu().float()).to(BF)
    ulp_close(normed, want, max_ulp=2, max_frac=0.01)


def test_linear_is_linear_at_full_width(lib, abi):
    """Size-independent property at the 8B down-proj shape: W(a*x) == a*W(x) for a power of two
    (exact in bf16/fp32), and column-sliced leading dimension works."""
    k, n_out = 14336, 4096
    g = torch.Generator().manual_seed(1)
    w = dev((0.02 * torch.randn(n_out, k, generator=g)).to(BF))
    x = dev(torch.randn(2, k + 64, generator=g).to(BF))        # ldx = k + 64
    y1 = torch.empty(2, n_out, dtype=BF, device="cuda")
    y2 = torch.empty_like(y1)
    sync_check(abi, lib.b200_linear_bf16(x.data_ptr(), k + 64, w.data_ptr(), None, y1.data_ptr(), n_out, 2, n_out, k, 1, None, 0, None))
    x4 = (x.float() * 4).to(BF)
    sync_check(abi, lib.b200_linear_bf16(x4.data_ptr(), k + 64, w.data_ptr(), None, y2.data_ptr(), n_out, 2, n_out, k, 1, None, 0, None))
    assert torch.equal((y1.float() * 4).to(BF), y2)
    ulp_close(y1, O.linear(x[:, :k].cpu(), w.cpu()), max_frac=0.02)


@pytest.mark.parametrize("T,h", [(1, 512), (5, 4096), (64, 5120), (3, 8192), (2, 640)])
@pytest.mark.parametrize("with_res", [False, True])
def test_rmsnorm(lib, abi, T, h, with_res):
    g = torch.Generator().manual_seed(h + T)
    x = torch.randn(T, h, generator=g).to(BF)
    r = torch.randn(T, h, generator=g).to(BF)
    w = (1 + 0.1 * torch.randn(h, generator=g)).to(BF)
    xd, rd, wd = dev(x), dev(r), dev(w)
    y = torch.empty(T, h, dtype=BF, device="cuda")
    if with_res:
        want, want_res = O.add_rms_norm(x, r, w, 1e-5)
        sync_check(abi, lib.b200_rmsnorm(xd.data_ptr(), rd.data_ptr(), wd.data_ptr(), y.data_ptr(), T, h, 1e-5, None))
        assert torch.equal(rd.cpu(), want_res), "residual must be bit-exact (one fp32 add, one rounding)"
    else:
        want = O.rms_norm(x, w, 1e-5)
        sync_check(abi, lib.b200_rmsnorm(xd.data_ptr(), None, wd.data_ptr(), y.data_ptr(), T, h, 1e-5, None))
    # the variance's summation order moves rsqrt by an fp32 ulp: rare 1-ulp bf16 flips at each rounding
    ulp_close(y, want, max_ulp=2, max_frac=0.01)


def test_embed_exact(lib, abi):
    tab = torch.randn(1000, 512).to(BF)
    ids = torch.tensor([0, 999, 5, 5, 17], dtype=torch.int32)
    out = torch.empty(5, 512, dtype=BF, device="cuda")
    sync_check(abi, lib.b200_embed(dev(ids).data_ptr(), dev(tab).data_ptr(), out.data_ptr(), 5, 512, 1000, None))
    assert torch.equal(out.cpu(), tab[ids.long()])


@pytest.mark.parametrize("cfgname", ["tiny", "tiny-qwen", "tiny-g8"])
def test_rope_kv_append_exact(lib, abi, cfgname):
    cfg = O.CONFIGS[cfgname]
    hq, hkv, d, bs, nb = cfg.n_q_heads, cfg.n_kv_heads, 128, 128, 5
    T = 7
    g = torch.Generator().manual_seed(3)
    qkv = torch.randn(T, (hq + 2 * hkv) * d, generator=g).to(BF)
    pos = torch.tensor([0, 1, 127, 128, 129, 300, 511], dtype=torch.int32)
    blocks = [3, 0, 4, 1]
    slots = torch.tensor([blocks[p // bs] * bs + p % bs for p in pos.tolist()], dtype=torch.int32)
    slots[5] = -1                                              # padding row: nothing written
    cs = O.cos_sin_cache(cfg, 1024)
    kc = torch.zeros(nb, hkv, bs, d, dtype=BF, device="cuda")
    vc = torch.zeros_like(kc)
    qd = dev(qkv)
    sync_check(abi, lib.b200_rope_kv_append(qd.data_ptr(), dev(pos).data_ptr(), dev(slots).data_ptr(), dev(cs).data_ptr(),
                                            kc.data_ptr(), vc.data_ptr(), T, hq, hkv, bs, None))
    q = qkv[:, :hq * d].view(T, hq, d)
    k = qkv[:, hq * d:(hq + hkv) * d].view(T, hkv, d)
    v = qkv[:, (hq + hkv) * d:].view(T, hkv, d)
    cos, sin = cs[pos.long(), :64], cs[pos.long(), 64:]
    qr, kr = O.rope_apply(q, cos, sin), O.rope_apply(k, cos, sin)
    assert torch.equal(qd.cpu()[:, :hq * d].view(T, hq, d), qr), "RoPE(q) must be bit-exact (bf16 arithmetic)"
    want_k = torch.zeros(nb, hkv, bs, d, dtype=BF)
    want_v = torch.zeros_like(want_k)
    for t in range(T):
        s = int(slots[t])
        if s >= 0:
            want_k[s // bs, :, s % bs] = kr[t]
            want_v[s // bs, :, s % bs] = v[t]
    assert torch.equal(kc.cpu(), want_k) and torch.equal(vc.cpu(), want_v)


@pytest.mark.parametrize("T,I", [(1, 1024), (3, 14336), (64, 1536)])
def test_silu_mul(lib, abi, T, I):
    gu = (2 * torch.randn(T, 2 * I, generator=torch.Generator().manual_seed(I))).to(BF)
    out = torch.empty(T, I, dtype=BF, device="cuda")
    sync_check(abi, lib.b200_silu_mul(dev(gu).data_ptr(), out.data_ptr(), T, I, None))
    ulp_close(out, O.silu_mul(gu), max_ulp=1, max_frac=0.005)   # expf vs the host libm: rare 1-ulp flips


def test_argmax_exact_with_ties(lib, abi):
    V = 128256
    lg = torch.randn(6, V, generator=torch.Generator().manual_seed(9)).to(BF)
    lg[1, 77] = 9.0; lg[1, 100000] = 9.0                       # tie -> lowest index
    lg[2, V - 1] = 50.0                                        # last element
    lg[3, :] = -1.0                                            # all equal -> 0
    lg[4, 0] = 60.0
    ids = torch.empty(6, dtype=torch.int32, device="cuda")
    mx = torch.empty(6, dtype=torch.float32, device="cuda")
    sync_check(abi, lib.b200_argmax(dev(lg).data_ptr(), V, 6, V, 0, ids.data_ptr(), mx.data_ptr(), None))
    assert ids.cpu().tolist() == O.greedy(lg).tolist()
    assert torch.equal(mx.cpu(), lg.float().max(-1).values)
    # ragged vocab (not a multiple of 8 handled by the tail loop) and an index offset
    lg2 = lg[:, :1003].contiguous()
    ld = 1008
    pad = torch.full((6, ld), -100.0, dtype=BF); pad[:, :1003] = lg2
    sync_check(abi, lib.b200_argmax(dev(pad).data_ptr(), ld, 6, 1003, 5000, ids.data_ptr(), None, None))
    assert ids.cpu().tolist() == (O.greedy(lg2) + 5000).tolist()


# ---------------------------------------------------------------------------------------------
def run_sample(lib, abi, logits, temperature, top_k, top_p, seeds, draws, min_p=None):
    R, V = logits.shape
    out = torch.empty(R, dtype=torch.int32, device="cuda")
    t = dev(torch.tensor(temperature, dtype=torch.float32))
    k = dev(torch.tensor(top_k, dtype=torch.int32))
    p = dev(torch.tensor(top_p, dtype=torch.float32))
    mp = dev(torch.tensor(min_p, dtype=torch.float32)) if min_p is not None else None
    sd = dev(torch.tensor(seeds, dtype=torch.int64))
    dr = dev(torch.tensor(draws, dtype=torch.int64))
    sync_check(abi, lib.b200_sample(dev(logits).data_ptr(), V, R, V, t.data_ptr(), k.data_ptr(), p.data_ptr(),
                                    mp.data_ptr() if mp is not None else None, sd.data_ptr(), dr.data_ptr(),
                                    out.data_ptr(), None))
    return out.cpu()


def test_sample_greedy_rows_and_top_k_1_are_argmax(lib, abi):
    g = torch.Generator().manual_seed(3)
    V = 5003
    lg = torch.randn(6, V, generator=g).to(BF)
    lg[1, 77] = lg[1, 4000] = lg[1].float().max() + 1      # tie: lowest index wins
    got = run_sample(lib, abi, lg, [0.0, 0.0, 1.0, 0.8, 0.0, 2.0], [0, 0, 1, 1, 50, 1], [1.0, 0.5, 1.0, 0.3, 0.9, 1.0],
                     [1, 2, 3, 4, 5, 6], [0, 0, 0, 9, 0, 1])
    assert got.tolist() == O.greedy(lg).tolist()
    assert got[1] == 77


def test_sample_is_a_function_of_seed_and_draw(lib, abi):
    g = torch.Generator().manual_seed(4)
    V, R = 32000, 16
    lg = (torch.randn(1, V, generator=g) * 2).to(BF).expand(R, V).contiguous()
    a = run_sample(lib, abi, lg, [1.0] * R, [0] * R, [1.0] * R, [42] * R, list(range(R)))
    b = run_sample(lib, abi, lg, [1.0] * R, [0] * R, [1.0] * R, [42] * R, list(range(R)))
    c = run_sample(lib, abi, lg, [1.0] * R, [0] * R, [1.0] * R, [43] * R, list(range(R)))
    assert a.tolist() == b.tolist()                         # same (seed, draw) -> same token
    assert len(set(a.tolist())) > R // 2                    # different draws differ
    assert a.tolist() != c.tolist()                         # another seed, another stream


@pytest.mark.parametrize("T,k,p", [(1.0, 0, 0.9), (0.7, 20, 1.0), (1.3, 40, 0.8), (0.5, 0, 0.3)])
def test_sample_support_and_frequencies_match_the_oracle(lib, abi, T, k, p):
    """4096 draws from one row: every token drawn is one the oracle keeps (ties at the cut value may add
    tokens of exactly that value), and the empirical frequencies follow the oracle's probabilities."""
    g = torch.Generator().manual_seed(int(T * 10) + k)
    V, R = 300, 4096
    row = (torch.randn(V, generator=g) * 2.5).to(BF)
    probs = O.sampling_probs(row, T, k, p)
    got = run_sample(lib, abi, row[None].expand(R, V).contiguous(), [T] * R, [k] * R, [p] * R, [7] * R, list(range(R)))
    kept = probs > 0
    cut = row.float()[kept].min()
    allowed = kept | (row.float() == cut)
    assert bool(allowed[got.long()].all()), "a dropped token was sampled"
    # renormalise over what the kernel keeps (identical to the oracle unless there is a tie at the cut)
    ref = (row.float() / T).masked_fill(~allowed, float("-inf")).softmax(-1)
    freq = torch.bincount(got.long(), minlength=V).float() / R
    sigma = (ref * (1 - ref) / R).sqrt()
    assert bool(((freq - ref).abs() <= 4.5 * sigma + 2e-3).all()), float(((freq - ref).abs() - 4.5 * sigma).max())


@pytest.mark.parametrize("seed", [21, 22, 23, 24, 25, 26, 27])
def test_sample_min_p_against_vllm_masks(lib, abi, seed):
    """min_p on the cases whose keep-masks were captured from vLLM's own MinPLogitsProcessor (+ its top-k/top-p).
    vLLM cuts in probability space (p_i < min_p * p_max), the kernel in logit space (l_i < max + T log min_p): the same
    cut except for a token whose p_i / p_max is within 1e-4 (relative) of min_p -- that BAND, and ties at the final cut
    value, are the only tokens allowed to differ, and how many there are is part of the assertion message."""
    import numpy as np
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "sampling_masks.npz"))
    row = torch.from_numpy(z[f"mp_logits_{seed}"].copy()).view(BF)
    T, mp, k, p = (float(v) for v in z[f"mp_params_{seed}"])
    k = int(k)
    keep = torch.from_numpy(z[f"mp_keep_{seed}"])
    ratio = torch.from_numpy(z[f"mp_ratio_{seed}"])
    probs = O.sampling_probs(row, T, k, p, min_p=mp)
    assert torch.equal(probs > 0, keep)                       # the reference used below IS vLLM's kept set
    band = (ratio / mp - 1).abs() < 1e-4
    ties = row.float() == row.float()[keep].min()
    allowed = keep | band | ties
    V, R = row.numel(), 4096
    got = run_sample(lib, abi, row[None].expand(R, V).contiguous(), [T] * R, [k] * R, [p] * R, [11] * R, list(range(R)),
                     min_p=[mp] * R)
    n_amb = int((band | ties).sum() - (keep & ties & ~band).sum())      # tokens the two formulations may treat differently
    assert bool(allowed[got.long()].all()), f"a token vLLM drops was sampled (band+tie tokens: {n_amb})"
    freq = torch.bincount(got.long(), minlength=V).float() / R
    sigma = (probs * (1 - probs) / R).sqrt()
    exact = ~(band | (ties & ~keep))                           # where the reference probability is unambiguous
    if n_amb == 0:
        exact = torch.ones_like(exact)
    dev_ = ((freq - probs).abs() - (4.5 * sigma + 2e-3))[exact]
    assert bool((dev_ <= 0).all()), f"frequency off by {float(dev_.max()):.4f} beyond 4.5 sigma (band+tie tokens: {n_amb})"
    if mp >= 1.0:
        assert set(got.tolist()) <= set(torch.nonzero(row.float() == row.float().max()).flatten().tolist())


def pack_adjust(rows):
    """rows: list of dicts(prompt, output, rep, freq, pres, bias) -> the CSR arrays of b200_logit_adjust (one entry per
    distinct token of a row's prompt, output and bias table)."""
    params, offsets, ids, counts, seen, bias = [], [0], [], [], [], []
    for r in rows:
        cnt, sn = {}, set(int(t) for t in r["prompt"])
        for t in r["output"]:
            cnt[int(t)] = cnt.get(int(t), 0) + 1
            sn.add(int(t))
        toks = sorted(sn | set(int(k) for k in r["bias"]))
        for t in toks:
            ids.append(t)
            counts.append(cnt.get(t, 0))
            seen.append(1 if t in sn else 0)
            bias.append(float(r["bias"].get(t, 0.0)))
        params += [r["rep"], r["freq"], r["pres"]]
        offsets.append(len(ids))
    i32 = lambda a: torch.tensor(a, dtype=torch.int32)
    return (torch.tensor(params, dtype=torch.float32), i32(offsets), i32(ids), i32(counts), i32(seen),
            torch.tensor(bias, dtype=torch.float32))


def test_logit_adjust_against_vllm_penalties(lib, abi):
    """b200_logit_adjust on the rows captured from vLLM's apply_penalties + logit_bias (tests/golden/penalties.npz): every
    touched logit equals bf16(vLLM's fp32 value) — the one rounding the kernel adds — and no other logit changes; the
    argmax of the adjusted row is vLLM's unless the rounding creates a tie.  All cases as rows of one launch (ragged)."""
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "penalties.npz"))
    seeds = sorted(int(k.split("_")[1]) for k in z.files if k.startswith("logits_"))
    V = max(len(z[f"logits_{sd}"]) for sd in seeds)
    rows, logits = [], torch.full((len(seeds), V), -30.0, dtype=BF)
    for i, sd in enumerate(seeds):
        l = torch.from_numpy(z[f"logits_{sd}"].copy()).view(BF)
        logits[i, : len(l)] = l
        rep, freq, pres = (float(v) for v in z[f"params_{sd}"])
        rows.append(dict(prompt=z[f"prompt_{sd}"].tolist(), output=z[f"output_{sd}"].tolist(), rep=rep, freq=freq, pres=pres,
                         bias={int(a): float(b) for a, b in zip(z[f"bias_ids_{sd}"], z[f"bias_{sd}"])}))
    arrs = [dev(a) for a in pack_adjust(rows)]
    ld = dev(logits.clone())
    sync_check(abi, lib.b200_logit_adjust(ld.data_ptr(), V, len(seeds), V, *[a.data_ptr() for a in arrs], None))
    got = ld.cpu()
    for i, sd in enumerate(seeds):
        want = torch.from_numpy(z[f"want_{sd}"])
        n = len(want)
        assert torch.equal(got[i, :n], want.to(BF)), f"case {sd}"
        assert torch.equal(got[i, n:], logits[i, n:])
        top = got[i, :n].float().max()
        assert got[i, :n].float()[int(want.argmax())] == top
    # a second application is NOT idempotent (penalties compound): the engine applies them once per step
    sync_check(abi, lib.b200_logit_adjust(ld.data_ptr(), V, len(seeds), V, *[a.data_ptr() for a in arrs], None))
    assert not torch.equal(ld.cpu(), got)


@pytest.mark.parametrize("V,n_top", [(5003, 5), (32000, 20), (300, 0), (64, 32)])
def test_logprobs_match_log_softmax(lib, abi, V, n_top):
    """b200_logprobs: log softmax of the raw logits for the chosen token and the n_top most likely tokens (value
    descending, ties by lowest index) against torch.log_softmax in fp32."""
    g = torch.Generator().manual_seed(V + n_top)
    R = 7
    lg = (torch.randn(R, V, generator=g) * 3).to(BF)
    lg[2, 10] = lg[2, 3] = lg[2].float().max() + 2            # a tie at the top
    lg[3] = 0.5                                               # fully degenerate row: every token ties
    tok = torch.randint(0, V, (R,), generator=g, dtype=torch.int32)
    out_lp = torch.empty(R, dtype=torch.float32, device="cuda")
    out_ids = torch.empty(R, max(n_top, 1), dtype=torch.int32, device="cuda")
    out_tlp = torch.empty(R, max(n_top, 1), dtype=torch.float32, device="cuda")
    sync_check(abi, lib.b200_logprobs(dev(lg).data_ptr(), V, R, V, dev(tok).data_ptr(), n_top, out_lp.data_ptr(),
                                      out_ids.data_ptr() if n_top else None, out_tlp.data_ptr() if n_top else None, None))
    ref = torch.log_softmax(lg.float(), dim=-1)
    assert torch.allclose(out_lp.cpu(), ref[torch.arange(R), tok.long()], atol=2e-5, rtol=1e-5)
    if n_top:
        k = min(n_top, V)
        # expected order: value descending, index ascending among equals
        order = torch.stack([torch.tensor(sorted(range(V), key=lambda i: (-float(lg[r, i]), i))[:k]) for r in range(R)])
        assert torch.equal(out_ids.cpu()[:, :k].long(), order)
        assert torch.allclose(out_tlp.cpu()[:, :k], torch.gather(ref, 1, order), atol=2e-5, rtol=1e-5)


def test_ctx_sample_last_follows_the_step(lib, abi):
    """b200_ctx_sample_last over the logits of the last step: temperature 0 reproduces the step's greedy tokens."""
    import numpy as np
    R_ = pkg("runner")
    cfg = pkg("config").get_config("tiny")
    r = R_.ModelRunner(cfg, max_tokens=16, max_model_len=256, kv_cache_gib=0.05)
    r.init_random_weights()
    T = 4
    ids = torch.tensor([5, 9, 11, 3], dtype=torch.int32, device="cuda")
    pos = torch.zeros(T, dtype=torch.int32, device="cuda")
    slots = torch.arange(T, dtype=torch.int32, device="cuda") * r.block_size
    bt = torch.arange(T, dtype=torch.int32, device="cuda").view(T, 1)
    out = r.step_device(ids, pos, slots, bt, pos + 1, 1)   # runs on the runner's stream
    torch.cuda.synchronize()
    greedy = out.cpu().tolist()
    same = r.sample_last([0.0] * T, [0] * T, [1.0] * T, [1] * T, [0] * T)
    assert same.tolist() == greedy
    hot = r.sample_last([5.0] * T, [0] * T, [1.0] * T, [1, 2, 3, 4], [0] * T)
    assert all(0 <= t < cfg.vocab for t in hot.tolist())
    lp, ids, tlp = r.logprobs_last(greedy, 3)
    assert ids[:, 0].tolist() == greedy                       # the greedy token is the most likely one
    assert np.allclose(lp, tlp[:, 0]) and (tlp[:, 0] >= tlp[:, 1]).all() and (lp <= 0).all()
    r.close()


# ---------------------------------------------------------------------------------------------
def run_attn(lib, abi, q, K, V, ctx_lens, hq, hkv, max_ctx_hint=None, seed=0, bs=128):
    """q [T,hq,128]; K,V per row: list of [n,hkv,128] -> scatter into a shuffled paged cache.
    bs=128 takes the tcgen05 kernel, bs=64 the CUDA-core kernel."""
    T = q.shape[0]
    nblk = [(n + bs - 1) // bs for n in ctx_lens]
    total = sum(nblk) + 3
    perm = torch.randperm(total, generator=torch.Generator().manual_seed(seed)).tolist()
    kc = torch.full((total, hkv, bs, 128), float("nan"), dtype=BF)   # unwritten slots are NaN on purpose
    vc = torch.full_like(kc, float("nan"))
    stride = max(nblk)
    bt = torch.zeros(T, stride, dtype=torch.int32)
    it = iter(perm)
    for t in range(T):
        for j in range(nblk[t]):
            b = next(it)
            bt[t, j] = b
            n = min(bs, ctx_lens[t] - j * bs)
            kc[b, :, :n] = K[t][j * bs:j * bs + n].permute(1, 0, 2)
            vc[b, :, :n] = V[t][j * bs:j * bs + n].permute(1, 0, 2)
    max_ctx = max_ctx_hint or max(ctx_lens)
    ws_bytes = lib.b200_paged_attn_workspace_bytes(T, hq, max_ctx, bs)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    out = torch.empty(T, hq * 128, dtype=BF, device="cuda")
    qd = dev(q.reshape(T, hq * 128))
    sync_check(abi, lib.b200_paged_attn_decode(qd.data_ptr(), hq * 128, dev(kc).data_ptr(), dev(vc).data_ptr(),
                                               dev(bt).data_ptr(), stride, None, dev(torch.tensor(ctx_lens, dtype=torch.int32)).data_ptr(),
                                               out.data_ptr(), hq * 128, T, hq, hkv, bs, max_ctx, 128 ** -0.5,
                                               ws.data_ptr(), ws_bytes, None))
    return out.cpu().view(T, hq, 128)


@pytest.mark.parametrize("bs", [128, 64])
@pytest.mark.parametrize("hq,hkv", [(4, 1), (5, 1), (16, 2), (8, 8), (4, 2)])
@pytest.mark.parametrize("ctx_lens", [[1], [63, 64, 65], [128, 129, 1], [1000, 17, 4097]])
def test_paged_attention_vs_oracle(lib, abi, hq, hkv, ctx_lens, bs):
    g = torch.Generator().manual_seed(hq * 100 + hkv + len(ctx_lens))
    T = len(ctx_lens)
    q = torch.randn(T, hq, 128, generator=g).to(BF)
    K = [torch.randn(n, hkv, 128, generator=g).to(BF) for n in ctx_lens]
    V = [torch.randn(n, hkv, 128, generator=g).to(BF) for n in ctx_lens]
    got = run_attn(lib, abi, q, K, V, ctx_lens, hq, hkv, bs=bs)
    want = torch.stack([O.attend(q[t], K[t], V[t], ctx_lens[t], 128 ** -0.5) for t in range(T)])
    assert torch.isfinite(got.float()).all()
    ulp_close(got, want, max_ulp=2, max_frac=0.02, atol_small=2 ** -8)   # measured <= 0.004 (gpurun_out/ulp_fractions.txt)


@pytest.mark.parametrize("hq,hkv", [(4, 1), (5, 1), (16, 2), (8, 8), (2, 1)])
def test_paged_attention_prefill_rows_packed(lib, abi, hq, hkv):
    """b200_paged_attn_prefill: rows of one sequence share a pass over its cache with a per-row causal mask.
    Mixed step: 3 decode rows of different sequences, a 37-row prefill chunk of sequence 3 starting at
    position 200 (so groups straddle the decode/prefill boundary and the chunk's ragged end), then one more
    decode row.  Checked against the oracle row by row and against the row-at-a-time kernel."""
    g = torch.Generator().manual_seed(hq + 31 * hkv)
    bs = 128
    seq_ctx = [150, 1, 300, 237, 64]                    # tokens cached per sequence after this step
    rows = [(0, 150), (1, 1), (2, 300)] + [(3, 200 + i + 1) for i in range(37)] + [(4, 64)]   # (seq, ctx_len)
    T = len(rows)
    K = [torch.randn(n, hkv, 128, generator=g).to(BF) for n in seq_ctx]
    V = [torch.randn(n, hkv, 128, generator=g).to(BF) for n in seq_ctx]
    nblk = [(n + bs - 1) // bs for n in seq_ctx]
    total = sum(nblk) + 2
    perm = iter(torch.randperm(total, generator=g).tolist())
    kc = torch.full((total, hkv, bs, 128), float("nan"), dtype=BF)
    vc = torch.full_like(kc, float("nan"))
    bt = torch.zeros(len(seq_ctx), max(nblk), dtype=torch.int32)
    for sq, n in enumerate(seq_ctx):
        for j in range(nblk[sq]):
            b = next(perm)
            bt[sq, j] = b
            m = min(bs, n - j * bs)
            kc[b, :, :m] = K[sq][j * bs:j * bs + m].permute(1, 0, 2)
            vc[b, :, :m] = V[sq][j * bs:j * bs + m].permute(1, 0, 2)
    q = torch.randn(T, hq, 128, generator=g).to(BF)
    r2s = torch.tensor([r[0] for r in rows], dtype=torch.int32)
    cl = torch.tensor([r[1] for r in rows], dtype=torch.int32)
    max_ctx = int(cl.max())
    wsb = lib.b200_paged_attn_workspace_bytes(T, hq, max_ctx, bs)
    ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    qd, kcd, vcd, btd, r2sd, cld = dev(q.reshape(T, hq * 128)), dev(kc), dev(vc), dev(bt), dev(r2s), dev(cl)
    outs = []
    for fn in (lib.b200_paged_attn_prefill, lib.b200_paged_attn_decode):
        out = torch.full((T, hq * 128), float("nan"), dtype=BF, device="cuda")
        sync_check(abi, fn(qd.data_ptr(), hq * 128, kcd.data_ptr(), vcd.data_ptr(), btd.data_ptr(), bt.shape[1], r2sd.data_ptr(),
                           cld.data_ptr(), out.data_ptr(), hq * 128, T, hq, hkv, bs, max_ctx, 128 ** -0.5, ws.data_ptr(), wsb, None))
        outs.append(out.cpu().view(T, hq, 128))
    want = torch.stack([O.attend(q[t], K[sq], V[sq], n, 128 ** -0.5) for t, (sq, n) in enumerate(rows)])
    assert torch.isfinite(outs[0].float()).all()
    ulp_close(outs[0], want, max_ulp=2, max_frac=0.02, atol_small=2 ** -8)      # measured <= 0.005
    ulp_close(outs[0], outs[1], max_ulp=2, max_frac=0.02, atol_small=2 ** -8)


@pytest.mark.parametrize("hq,hkv", [(8, 2), (5, 1), (8, 1)])
def test_prefill_attention_chunk_with_max_jumps(lib, abi, hq, hkv):
    """attn_prefill.cu on a 200-row chunk of one sequence that starts at position 600 (items of 128/G rows,
    ragged last item, contexts crossing tile boundaries), between two decode rows.  The keys of tokens
    >= 384 are 6x larger, so row maxima jump by far more than 2^8 in the middle of a pass: the lazy
    rescale (fold O_t, rescale, fresh accumulation) must give the oracle's result, not just the
    common no-jump path."""
    g = torch.Generator().manual_seed(7 * hq + hkv)
    bs, start, nrows = 128, 600, 200
    seq_ctx = [90, start + nrows, 513]
    rows = [(0, 90)] + [(1, start + i + 1) for i in range(nrows)] + [(2, 513)]
    T = len(rows)
    K = [torch.randn(n, hkv, 128, generator=g) for n in seq_ctx]
    K[1][384:] *= 6.0
    K = [k.to(BF) for k in K]
    V = [torch.randn(n, hkv, 128, generator=g).to(BF) for n in seq_ctx]
    nblk = [(n + bs - 1) // bs for n in seq_ctx]
    total = sum(nblk) + 1
    perm = iter(torch.randperm(total, generator=g).tolist())
    kc = torch.full((total, hkv, bs, 128), float("nan"), dtype=BF)
    vc = torch.full_like(kc, float("nan"))
    bt = torch.zeros(len(seq_ctx), max(nblk), dtype=torch.int32)
    for sq, n in enumerate(seq_ctx):
        for j in range(nblk[sq]):
            b = next(perm)
            bt[sq, j] = b
            m = min(bs, n - j * bs)
            kc[b, :, :m] = K[sq][j * bs:j * bs + m].permute(1, 0, 2)
            vc[b, :, :m] = V[sq][j * bs:j * bs + m].permute(1, 0, 2)
    q = torch.randn(T, hq, 128, generator=g).to(BF)
    r2s = torch.tensor([r[0] for r in rows], dtype=torch.int32)
    cl = torch.tensor([r[1] for r in rows], dtype=torch.int32)
    max_ctx = int(cl.max())
    wsb = lib.b200_paged_attn_workspace_bytes(T, hq, max_ctx, bs)
    ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    out = torch.full((T, hq * 128), float("nan"), dtype=BF, device="cuda")
    sync_check(abi, lib.b200_paged_attn_prefill(dev(q.reshape(T, hq * 128)).data_ptr(), hq * 128, dev(kc).data_ptr(),
                                                dev(vc).data_ptr(), dev(bt).data_ptr(), bt.shape[1], dev(r2s).data_ptr(),
                                                dev(cl).data_ptr(), out.data_ptr(), hq * 128, T, hq, hkv, bs, max_ctx,
                                                128 ** -0.5, ws.data_ptr(), wsb, None))
    got = out.cpu().view(T, hq, 128)
    want = torch.stack([O.attend(q[t], K[sq], V[sq], n, 128 ** -0.5) for t, (sq, n) in enumerate(rows)])
    assert torch.isfinite(got.float()).all()
    ulp_close(got, want, max_ulp=2, max_frac=0.02, atol_small=2 ** -8)   # measured <= 0.004 (gpurun_out/ulp_fractions.txt)


def test_paged_attention_split_invariance_long_context(lib, abi):
    """Config-3 scale (ctx 32768, G=4): the split-KV result must not depend on the split plan —
    compare the plan for max_ctx=32768 with a larger hint (different split count) and the oracle."""
    hq, hkv, n = 4, 1, 32768
    g = torch.Generator().manual_seed(5)
    q = torch.randn(1, hq, 128, generator=g).to(BF)
    K = [torch.randn(n, hkv, 128, generator=g).to(BF)]
    V = [torch.randn(n, hkv, 128, generator=g).to(BF)]
    a = run_attn(lib, abi, q, K, V, [n], hq, hkv)
    b = run_attn(lib, abi, q, K, V, [n], hq, hkv, max_ctx_hint=65536, seed=1)
    want = O.attend(q[0], K[0], V[0], n, 128 ** -0.5)[None]
    ulp_close(a, want, max_ulp=2, max_frac=0.02, atol_small=2 ** -9)
    ulp_close(b, a, max_ulp=1, max_frac=0.05, atol_small=2 ** -9)


def test_empty_inputs_are_noops(lib, abi):
    assert lib.b200_silu_mul(1, 1, 0, 1024, None) == 0
    assert lib.b200_rmsnorm(1, None, 1, 1, 0, 512, 1e-5, None) == 0
    assert lib.b200_embed(1, 1, 1, 0, 512, 10, None) == 0
    assert lib.b200_argmax(1, 8, 0, 8, 0, 1, None, None) == 0
    torch.cuda.synchronize()


def test_unsupported_shapes_error(lib, abi):
    x = torch.zeros(1, 12, dtype=BF, device="cuda")
    assert lib.b200_linear_bf16(x.data_ptr(), 12, x.data_ptr(), None, x.data_ptr(), 1, 1, 1, 12, 1, None, 0, None) == -3
    assert b"multiples of 8" in lib.b200_last_error()
    assert lib.b200_rmsnorm(x.data_ptr(), None, x.data_ptr(), x.data_ptr(), 1, 16384, 1e-5, None) == -3
