"""One launch of the tcgen05 paged attention per shape, for `ncu --set full -k regex:paged_attn_tc`.

    ncu --set full --clock-control none --import-source on -k regex:paged_attn_tc -o gpurun_out/attn \
        python tools/attn_ncu.py
"""
import importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
abi = importlib.import_module("enterprise-inference_b200._abi")
lib = abi.lib()
BF = torch.bfloat16
for (B, ctx, hq, hkv) in ((8, 32768, 32, 8), (64, 640, 32, 8)):
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
    ws = torch.empty(max(wsb, 16), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    rc = lib.b200_paged_attn_decode(q.data_ptr(), hq * 128, kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), nb_seq, None,
                                    cl.data_ptr(), out.data_ptr(), hq * 128, B, hq, hkv, bs, ctx, 128 ** -0.5,
                                    ws.data_ptr(), wsb, None)
    torch.cuda.synchronize()
    print("shape", B, ctx, "rc", rc, "algorithmic bytes", B * ctx * hkv * 128 * 2 * 2)
    del kc, vc
