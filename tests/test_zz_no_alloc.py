"""GPU: b200_paged_attn_decode_static — the attention entry point for callers that cannot allocate.

Some users of this library run with a device-memory plan that is fixed at start-up (or behind a
binding that has no way to call cudaMalloc), so they cannot hand b200_paged_attn_decode the scratch
buffer it asks for.  The static variant keeps that scratch in an array inside the library.  This
test drives it through the C-ABI exactly as such a caller would — inputs and outputs only — and
checks that it returns what the workspace-taking entry point returns, on two calls in a row that
share the array (the second with different shapes, so leftovers from the first would show), and
that a shape too large for the array is refused with a message instead of overrunning it.

Collected last on purpose (file name): a run that stops at the first failure has then already run
every other test.  It holds no state from other tests and passes alone in a fresh process too.
"""
import pytest
import torch

from conftest import pkg

pytestmark = pytest.mark.gpu
BF = torch.bfloat16


def make_case(T, hq, hkv, ctx_lens, seed, bs=128):
    g = torch.Generator().manual_seed(seed)
    nblk = [(n + bs - 1) // bs for n in ctx_lens]
    total = sum(nblk) + 2
    kc = torch.randn(total, hkv, bs, 128, generator=g).to(BF).cuda()
    vc = torch.randn(total, hkv, bs, 128, generator=g).to(BF).cuda()
    stride = max(nblk)
    bt = torch.zeros(T, stride, dtype=torch.int32)
    perm = iter(torch.randperm(total, generator=g).tolist())
    for t in range(T):
        for j in range(nblk[t]):
            bt[t, j] = next(perm)
    q = torch.randn(T, hq * 128, generator=g).to(BF).cuda()
    return dict(q=q, kc=kc, vc=vc, bt=bt.cuda(), cl=torch.tensor(ctx_lens, dtype=torch.int32).cuda(), stride=stride,
                T=T, hq=hq, hkv=hkv, max_ctx=max(ctx_lens), bs=bs)


def run_both(lib, abi, c):
    T, hq, hkv = c["T"], c["hq"], c["hkv"]
    bs = c["bs"]
    ws_bytes = lib.b200_paged_attn_workspace_bytes(T, hq, c["max_ctx"], bs)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    ref = torch.empty(T, hq * 128, dtype=BF, device="cuda")
    got = torch.full_like(ref, float("nan"))
    abi.check(lib.b200_paged_attn_decode(c["q"].data_ptr(), hq * 128, c["kc"].data_ptr(), c["vc"].data_ptr(),
                                         c["bt"].data_ptr(), c["stride"], None, c["cl"].data_ptr(), ref.data_ptr(),
                                         hq * 128, T, hq, hkv, bs, c["max_ctx"], 128 ** -0.5, ws.data_ptr(), ws_bytes, None))
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    abi.check(lib.b200_paged_attn_decode_static(c["q"].data_ptr(), hq * 128, c["kc"].data_ptr(), c["vc"].data_ptr(),
                                                c["bt"].data_ptr(), c["stride"], None, c["cl"].data_ptr(), got.data_ptr(),
                                                hq * 128, T, hq, hkv, bs, c["max_ctx"], 128 ** -0.5, None))
    torch.cuda.synchronize()
    # the property in the entry point's name: the device's free memory is the same after the call as before it
    # (the caller of this entry point has no way to give memory back, so the library must not take any)
    assert torch.cuda.mem_get_info()[0] == free_before, "the static-scratch entry point changed the device's free memory"
    return got, ref


def test_static_scratch_attention_matches_workspace_attention_twice_in_a_row():
    abi = pkg("_abi")
    lib = abi.lib()
    assert lib.b200_paged_attn_static_scratch_bytes() == 2 << 20
    # first call: split-KV plan (few rows, long context) -> the built-in scratch is really used
    a = make_case(T=2, hq=8, hkv=2, ctx_lens=[4097, 900], seed=1)
    # (the very first launch of a kernel loads its module and may grow the context once: warm up, then measure)
    lib.b200_paged_attn_decode_static(a["q"].data_ptr(), 8 * 128, a["kc"].data_ptr(), a["vc"].data_ptr(), a["bt"].data_ptr(),
                                      a["stride"], None, a["cl"].data_ptr(), torch.empty(2, 8 * 128, dtype=BF, device="cuda").data_ptr(),
                                      8 * 128, 2, 8, 2, 128, a["max_ctx"], 128 ** -0.5, None)
    torch.cuda.synchronize()
    got, ref = run_both(lib, abi, a)
    assert torch.isfinite(got.float()).all()
    # same kernels, same split plan, same summation order: the tolerance of the attention tests
    # (<= 2 bf16 ulp) is met with room to spare — the outputs are expected to be identical
    assert (got.float() - ref.float()).abs().max() <= 2 ** -7 * ref.float().abs().max()
    assert torch.equal(got, ref)
    # second call straight after, other shapes and a different split count: leftovers must not leak
    b = make_case(T=5, hq=20, hkv=4, ctx_lens=[129, 64, 2000, 1, 777], seed=2)
    got2, ref2 = run_both(lib, abi, b)
    assert torch.isfinite(got2.float()).all()
    assert torch.equal(got2, ref2)
    # and the first case again gives the first answer again
    got3, _ = run_both(lib, abi, a)
    assert torch.equal(got3, got)
    # 64-token cache blocks take the library's other attention kernel: same contract
    c = make_case(T=3, hq=8, hkv=2, ctx_lens=[700, 65, 2049], seed=4, bs=64)
    got4, ref4 = run_both(lib, abi, c)
    assert torch.isfinite(got4.float()).all() and torch.equal(got4, ref4)


def test_static_scratch_attention_refuses_shapes_beyond_its_array():
    abi = pkg("_abi")
    lib = abi.lib()
    # 64-token blocks -> the CUDA-core kernel's split plan: 295 rows * 8 heads * 2 splits > 2 MiB
    c = make_case(T=295, hq=8, hkv=1, ctx_lens=[300] * 295, seed=3, bs=64)
    out = torch.zeros(295, 8 * 128, dtype=BF, device="cuda")
    rc = lib.b200_paged_attn_decode_static(c["q"].data_ptr(), 8 * 128, c["kc"].data_ptr(), c["vc"].data_ptr(),
                                           c["bt"].data_ptr(), c["stride"], None, c["cl"].data_ptr(), out.data_ptr(),
                                           8 * 128, 295, 8, 1, 64, 300, 128 ** -0.5, None)
    assert rc == -3
    msg = lib.b200_last_error().decode()
    assert "built-in array holds" in msg and "b200_paged_attn_decode" in msg
    torch.cuda.synchronize()
    assert not out.any(), "a refused call must not have launched anything"
