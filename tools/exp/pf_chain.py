"""Experiment: QKV linear (optionally prefetching the O weights into L2) -> ~10 us of unrelated work -> O linear."""
import ctypes as C, importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
abi = importlib.import_module("enterprise-inference_b200._abi")
lib = abi.lib()
lib.b200_debug_prefetch_hint.argtypes = [C.c_void_p, C.c_int64]
BF = torch.bfloat16
T, h = 1, 4096
x = torch.randn(T, h, device="cuda").to(BF)
Wq = [torch.randn(6144, h, device="cuda").to(BF) for _ in range(4)]
Wo = [torch.randn(4096, h, device="cuda").to(BF) for _ in range(4)]
yq = torch.empty(T, 6144, device="cuda", dtype=BF); yo = torch.empty(T, 4096, device="cuda", dtype=BF)
wsb = max(lib.b200_linear_workspace_bytes(T, 6144, h), lib.b200_linear_workspace_bytes(T, 4096, h))
ws = torch.zeros(wsb, dtype=torch.uint8, device="cuda")
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
s = torch.cuda.current_stream().cuda_stream
for pf_mb in (0, 8, 16, 33):
    for filler_cycles in (0, 20000, 40000):
        best = 1e9
        for rep in range(8):
            i = rep % 4
            flush.zero_(); torch.cuda.synchronize()
            e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            e0.record()
            if pf_mb:
                lib.b200_debug_prefetch_hint(Wo[i].data_ptr(), min(pf_mb << 20, 4096 * h * 2))
            lib.b200_linear_bf16(x.data_ptr(), h, Wq[i].data_ptr(), None, yq.data_ptr(), 6144, T, 6144, h, 2, ws.data_ptr(), wsb, s)
            if filler_cycles:
                torch.cuda._sleep(filler_cycles)
            e1.record()
            lib.b200_linear_bf16(x.data_ptr(), h, Wo[i].data_ptr(), None, yo.data_ptr(), 4096, T, 4096, h, 2, ws.data_ptr(), wsb, s)
            e2.record(); torch.cuda.synchronize()
            tot = e0.elapsed_time(e2) * 1e3
            if tot < best:
                best, a, b = tot, e0.elapsed_time(e1) * 1e3, e1.elapsed_time(e2) * 1e3
        print(f"prefetch {pf_mb:2d} MB, filler {filler_cycles:5d} cyc: total {best:6.1f} us (qkv+filler {a:5.1f}, o {b:5.1f})")
