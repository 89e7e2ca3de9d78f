"""Decode attention alone (64 rows, 8 kv heads, G = 4) over contexts that end inside a tile: us per launch after an L2 flush.
    python tools/attn_ctx_sweep.py"""
import importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
abi = importlib.import_module("enterprise-inference_b200._abi")
lib = abi.lib()
BF = torch.bfloat16
B, hq, hkv, bs = 64, 32, 8, 128
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for ctx in (512, 520, 544, 576, 608, 640):
    nb_seq = (ctx + bs - 1) // bs
    n_blocks = B * nb_seq
    kc = torch.randn(n_blocks, hkv, bs, 128, device="cuda", dtype=BF)
    vc = torch.randn(n_blocks, hkv, bs, 128, device="cuda", dtype=BF)
    bt = torch.randperm(n_blocks, device="cuda").to(torch.int32).view(B, nb_seq).contiguous()
    q = torch.randn(B, hq * 128, device="cuda", dtype=BF)
    out = torch.empty_like(q)
    cl = torch.full((B,), ctx, dtype=torch.int32, device="cuda")
    wsb = lib.b200_paged_attn_workspace_bytes(B, hq, 1024, bs)
    ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    ts = []
    for it in range(12):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        abi.check(lib.b200_paged_attn_decode(q.data_ptr(), hq * 128, kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), nb_seq, None, cl.data_ptr(),
                                             out.data_ptr(), hq * 128, B, hq, hkv, bs, 1024, 128 ** -0.5, ws.data_ptr(), wsb,
                                             torch.cuda.current_stream().cuda_stream))
        e1.record()
        torch.cuda.synchronize()
        if it >= 2:
            ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    print(f"ctx {ctx}: median {ts[len(ts) // 2]:.1f} us, min {ts[0]:.1f} us; algorithmic {B * ctx * hkv * 512 / 1e6:.0f} MB")
