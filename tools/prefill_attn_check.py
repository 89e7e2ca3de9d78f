"""Prefill attention (attn_prefill.cu) against the row-at-a-time decode kernel on a 2048-row chunk whose
sequence already holds 1000 tokens, plus 5 decode rows of other sequences; prints max abs difference and time."""
import importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
abi = importlib.import_module("enterprise-inference_b200._abi")
lib = abi.lib()
BF = torch.bfloat16
torch.manual_seed(0)
ok = True
for (hq, hkv) in ((32, 8), (40, 8), (8, 1)):
    bs, T0, start = 128, 2048, 1000
    seq_ctx = [start + T0, 700, 129, 1, 5000, 128]
    rows = [(1, 700), (2, 129)] + [(0, start + i + 1) for i in range(T0)] + [(3, 1), (4, 5000), (5, 128)]
    T = len(rows)
    nblk = [(n + bs - 1) // bs for n in seq_ctx]
    total = sum(nblk) + 2
    kc = torch.randn(total, hkv, bs, 128, device="cuda", dtype=BF)
    vc = torch.randn(total, hkv, bs, 128, device="cuda", dtype=BF)
    perm = torch.randperm(total, device="cuda").to(torch.int32)
    bt = torch.zeros(len(seq_ctx), max(nblk), dtype=torch.int32, device="cuda")
    o = 0
    for sq, nb in enumerate(nblk):
        bt[sq, :nb] = perm[o:o + nb]
        o += nb
    q = torch.randn(T, hq * 128, device="cuda", dtype=BF)
    r2s = torch.tensor([r[0] for r in rows], dtype=torch.int32, device="cuda")
    cl = torch.tensor([r[1] for r in rows], dtype=torch.int32, device="cuda")
    max_ctx = int(cl.max())
    wsb = lib.b200_paged_attn_workspace_bytes(T, hq, max_ctx, bs)
    ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    outs, times = [], []
    for fn in (lib.b200_paged_attn_prefill, lib.b200_paged_attn_decode):
        out = torch.full((T, hq * 128), float("nan"), dtype=BF, device="cuda")
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = fn(q.data_ptr(), hq * 128, kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), bt.shape[1], r2s.data_ptr(), cl.data_ptr(),
                    out.data_ptr(), hq * 128, T, hq, hkv, bs, max_ctx, 128 ** -0.5, ws.data_ptr(), wsb, None)
            e1.record()
            torch.cuda.synchronize()
            assert rc == 0, lib.b200_last_error()
        outs.append(out.float())
        times.append(e0.elapsed_time(e1) * 1e3)
    d = (outs[0] - outs[1]).abs()
    bad = (~torch.isfinite(outs[0])).sum().item()
    worst = d.max().item()
    good = bad == 0 and worst < 2e-2
    ok = ok and good
    print(f"hq={hq} hkv={hkv} T={T}: prefill kernel {times[0]:.1f} us, row-at-a-time {times[1]:.1f} us, "
          f"max |diff| {worst:.5f}, mean {d.mean().item():.6f}, non-finite {bad}")
print("PREFILL_ATTN_CHECK", "PASS" if ok else "FAIL")
