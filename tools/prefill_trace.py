"""Tile timeline of CTA 0 of the prefill attention (trace build: python enterprise-inference_b200/build.py --trace)."""
import ctypes as C, importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ.setdefault("B200_LIB", os.path.join(ROOT, "enterprise-inference_b200", "lib", "libb200decode_trace.so"))
import torch
sys.path.insert(0, ROOT)
abi = importlib.import_module("enterprise-inference_b200._abi")
lib = abi.lib()
BF = torch.bfloat16
T, bs, hq, hkv = 2048, 128, 32, 8
nb = T // bs
kc = torch.randn(nb, hkv, bs, 128, device="cuda", dtype=BF)
vc = torch.randn(nb, hkv, bs, 128, device="cuda", dtype=BF)
bt = torch.arange(nb, device="cuda", dtype=torch.int32).view(1, nb).contiguous()
q = torch.randn(T, hq * 128, device="cuda", dtype=BF)
out = torch.empty_like(q)
cl = torch.arange(1, T + 1, dtype=torch.int32, device="cuda")
r2s = torch.zeros(T, dtype=torch.int32, device="cuda")
wsb = lib.b200_paged_attn_workspace_bytes(T, hq, T, bs)
ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
for rep in range(3):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    lib.b200_paged_attn_prefill(q.data_ptr(), hq * 128, kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), nb, r2s.data_ptr(),
                                cl.data_ptr(), out.data_ptr(), hq * 128, T, hq, hkv, bs, T, 128 ** -0.5, ws.data_ptr(), wsb, None)
    e1.record()
    torch.cuda.synchronize()
print(f"prefill attention T={T}: {e0.elapsed_time(e1) * 1e3:.1f} us")
buf = (C.c_ulonglong * 256)()
lib.b200_debug_prefill_trace(buf)
t0 = buf[0]
names = ["begin", "S landed", "pass1", "PV(n-1) done", "fold", "P handed"]
for n in range(40):
    v = [buf[8 + 6 * n + k] for k in range(6)]
    if v[0] < t0:
        break
    print(f"tile {n:2d}: " + "  ".join(f"{names[k]} {(v[k] - t0) / 1e3:7.2f}" if v[k] >= t0 else f"{names[k]}    -   " for k in range(6)))
