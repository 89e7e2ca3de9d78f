"""Quick numerical check of the tcgen05 linear (impl=2) against an fp32 torch matmul on the GPU.
Run under `timeout` on the GPU box: a descriptor mistake can hang a kernel."""
import importlib, os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
abi = importlib.import_module("enterprise-inference_b200._abi")
lib = abi.lib()
BF = torch.bfloat16
cases = [(16, 128, 64), (5, 128, 128), (16, 256, 512), (64, 6144, 4096), (17, 264, 512), (100, 1000, 1024),
         (64, 4096, 14336), (256, 512, 4096), (300, 384, 2048), (8, 128, 200), (64, 28672, 4096), (1, 4096, 4096),
         (2048, 1024, 512)]
if len(sys.argv) > 1:
    cases = cases[:int(sys.argv[1])]
ok = True
for (T, n, k) in cases:
    g = torch.Generator(device="cuda").manual_seed(T + n + k)
    x = torch.randn(T, k, device="cuda", generator=g).to(BF)
    w = (torch.randn(n, k, device="cuda", generator=g) * 0.05).to(BF)
    b = torch.randn(n, device="cuda", generator=g).to(BF)
    y = torch.full((T, n), float("nan"), device="cuda", dtype=BF)
    wsb = lib.b200_linear_workspace_bytes(T, n, k)
    ws = torch.zeros(wsb, dtype=torch.uint8, device="cuda")
    for rep in range(2):   # second run checks that the tickets were left at zero
        rc = lib.b200_linear_bf16(x.data_ptr(), k, w.data_ptr(), b.data_ptr(), y.data_ptr(), n, T, n, k, 2, ws.data_ptr(), wsb, None)
        if rc:
            print(json.dumps({"T": T, "n": n, "k": k, "error": lib.b200_last_error().decode()})); ok = False; break
        torch.cuda.synchronize()
        want = (x.float() @ w.float().t() + b.float()).to(BF)
        d = (y.float() - want.float()).abs()
        ulp = want.float().abs().clamp_min(2.0 ** -20).log2().floor().exp2() * 2.0 ** -7
        bad = int((d > torch.maximum(ulp, torch.tensor(2.0 ** -9, device="cuda"))).sum())
        nan = int(torch.isnan(y.float()).sum())
        print(json.dumps({"T": T, "n": n, "k": k, "rep": rep, "max_abs": float(d.nan_to_num(1e9).max()), "beyond_1ulp": bad, "nan": nan,
                          "frac_diff": float((d > 0).float().mean())}))
        ok &= (bad == 0 and nan == 0)
        y.fill_(float("nan"))
print("TC_CHECK", "PASS" if ok else "FAIL")
sys.exit(0 if ok else 1)
