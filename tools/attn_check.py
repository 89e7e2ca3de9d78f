"""Quick check of the tcgen05 attention kernel vs an fp32 torch reference on the GPU (run under timeout)."""
import importlib, os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
abi = importlib.import_module("enterprise-inference_b200._abi")
lib = abi.lib()
BF = torch.bfloat16
ok = True
for (T, hq, hkv, ctx) in [(1, 4, 1, 128), (1, 4, 1, 100), (2, 8, 2, 300), (3, 32, 8, 1000), (1, 5, 1, 4097), (4, 16, 2, 640), (8, 32, 8, 2048), (1, 8, 8, 5)]:
    g = torch.Generator(device="cuda").manual_seed(ctx + T)
    bs = 128
    nb = (ctx + bs - 1) // bs
    n_blocks = T * nb + 1
    kc = torch.randn(n_blocks, hkv, bs, 128, device="cuda", generator=g).to(BF)
    vc = torch.randn(n_blocks, hkv, bs, 128, device="cuda", generator=g).to(BF)
    bt = torch.randperm(n_blocks - 1, device="cuda")[: T * nb].to(torch.int32).view(T, nb).contiguous()
    q = torch.randn(T, hq * 128, device="cuda", generator=g).to(BF)
    out = torch.full((T, hq * 128), float("nan"), device="cuda", dtype=BF)
    cl = torch.full((T,), ctx, dtype=torch.int32, device="cuda")
    wsb = lib.b200_paged_attn_workspace_bytes(T, hq, ctx, bs)
    ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    rc = lib.b200_paged_attn_decode(q.data_ptr(), hq * 128, kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), nb, None, cl.data_ptr(),
                                    out.data_ptr(), hq * 128, T, hq, hkv, bs, ctx, 128 ** -0.5, ws.data_ptr(), wsb, None)
    if rc:
        print("error", lib.b200_last_error()); ok = False; continue
    torch.cuda.synchronize()
    G = hq // hkv
    worst = 0.0
    for t in range(T):
        K = kc[bt[t].long()].permute(1, 0, 2, 3).reshape(hkv, nb * bs, 128)[:, :ctx].float()
        V = vc[bt[t].long()].permute(1, 0, 2, 3).reshape(hkv, nb * bs, 128)[:, :ctx].float()
        qq = q[t].view(hkv, G, 128).float()
        s = torch.einsum("hgd,hnd->hgn", qq, K) * 128 ** -0.5
        o = torch.einsum("hgn,hnd->hgd", torch.softmax(s, -1), V).reshape(hq * 128)
        worst = max(worst, float((out[t].float() - o).abs().max()))
    nan = int(torch.isnan(out.float()).sum())
    print(json.dumps({"T": T, "hq": hq, "hkv": hkv, "ctx": ctx, "max_abs_err": worst, "nan": nan}))
    ok &= (nan == 0 and worst < 2e-2)
print("ATTN_CHECK", "PASS" if ok else "FAIL")
