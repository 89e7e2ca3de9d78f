"""Item timeline of CTA 0 of the tcgen05 attention (B200_ATTN_TRACE=1)."""
import ctypes as C, importlib, os, sys
os.environ["B200_ATTN_TRACE"] = "1"
os.environ.setdefault("B200_LIB", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "enterprise-inference_b200", "lib", "libb200decode_trace.so"))
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
abi = importlib.import_module("enterprise-inference_b200._abi")
lib = abi.lib()
BF = torch.bfloat16
for (B, ctx, hq, hkv) in ((64, 640, 32, 8), (8, 32768, 32, 8)):
    bs = 128
    nb_seq = (ctx + bs - 1) // bs
    n_blocks = B * nb_seq
    kc = torch.randn(n_blocks, hkv, bs, 128, device="cuda", dtype=BF)
    vc = torch.randn(n_blocks, hkv, bs, 128, device="cuda", dtype=BF)
    bt = torch.randperm(n_blocks, device="cuda").to(torch.int32).view(B, nb_seq).contiguous()
    q = torch.randn(B, hq * 128, device="cuda", dtype=BF)
    out = torch.empty_like(q)
    cl = torch.full((B,), ctx, dtype=torch.int32, device="cuda")
    wsb = lib.b200_paged_attn_workspace_bytes(B, hq, ctx, bs)
    ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for rep in range(3):
        flush.zero_()
        torch.cuda.synchronize()
        lib.b200_paged_attn_decode(q.data_ptr(), hq * 128, kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), nb_seq, None, cl.data_ptr(),
                                   out.data_ptr(), hq * 128, B, hq, hkv, bs, ctx, 128 ** -0.5, ws.data_ptr(), wsb, None)
        torch.cuda.synchronize()
    buf = (C.c_ulonglong * 64)()
    lib.b200_debug_attn_trace(buf)
    t0 = buf[0]
    vals = [(buf[i] - t0) / 1e3 if buf[i] >= t0 else None for i in range(20)]
    print(f"B={B} ctx={ctx}: start->pdl {vals[1]:.2f} us;", "items (start, first S ready, end):",
          [tuple(None if v is None else round(v, 2) for v in vals[2 + 3 * k: 5 + 3 * k]) for k in range(5)])
    tl = [(buf[i] - t0) / 1e3 if buf[i] >= t0 else None for i in range(20, 60)]
    print("  tiles (S landed, max exchanged, P stored, P handed over):",
          [tuple(None if v is None else round(v, 2) for v in tl[4 * k: 4 * k + 4]) for k in range(10)])

# ---- prefill rows: one sequence, T rows with context 1..T (row groups of 4 rows share a pass) ----
for T in (2048,):
    bs, hq, hkv = 128, 32, 8
    nb = T // bs
    kc = torch.randn(nb, hkv, bs, 128, device="cuda", dtype=BF)
    vc = torch.randn(nb, hkv, bs, 128, device="cuda", dtype=BF)
    bt = torch.arange(nb, device="cuda", dtype=torch.int32).view(1, nb).contiguous()
    q = torch.randn(T, hq * 128, device="cuda", dtype=BF)
    out = torch.empty_like(q)
    cl = torch.arange(1, T + 1, dtype=torch.int32, device="cuda")
    r2s = torch.zeros(T, dtype=torch.int32, device="cuda")
    wsb = lib.b200_paged_attn_workspace_bytes(T, hq, T, bs)
    ws = torch.empty(max(wsb, 16), dtype=torch.uint8, device="cuda")
    for rep in range(3):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        lib.b200_paged_attn_prefill(q.data_ptr(), hq * 128, kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), nb, r2s.data_ptr(),
                                    cl.data_ptr(), out.data_ptr(), hq * 128, T, hq, hkv, bs, T, 128 ** -0.5, ws.data_ptr(), wsb, None)
        e1.record()
        torch.cuda.synchronize()
    buf = (C.c_ulonglong * 64)()
    lib.b200_debug_attn_trace(buf)
    t0 = buf[0]
    vals = [(buf[i] - t0) / 1e3 if buf[i] >= t0 else None for i in range(20)]
    print(f"prefill T={T}: event {e0.elapsed_time(e1) * 1e3:.1f} us; items (start, first S ready, end):",
          [tuple(None if v is None else round(v, 2) for v in vals[2 + 3 * k: 5 + 3 * k]) for k in range(5)])
    tl = [(buf[i] - t0) / 1e3 if buf[i] >= t0 else None for i in range(20, 60)]
    print("  tiles (S landed, max exchanged, P stored, P handed over):",
          [tuple(None if v is None else round(v, 2) for v in tl[4 * k: 4 * k + 4]) for k in range(10)])
