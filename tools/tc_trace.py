"""Phase timeline of CTA 0 of the tcgen05 linear (B200_TC_TRACE=1): where does a short kernel spend its time?"""
import ctypes as C, importlib, os, sys
os.environ["B200_TC_TRACE"] = "1"
os.environ.setdefault("B200_LIB", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "enterprise-inference_b200", "lib", "libb200decode_trace.so"))
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
abi = importlib.import_module("enterprise-inference_b200._abi")
lib = abi.lib()
BF = torch.bfloat16
names = ["cta_start", "W_requested", "pdl_wait_done", "first_stage_landed", "last_mma_issued", "accum_complete",
         "tile_staged", "before_cluster", "finalize_done", "cta_end", "cluster_ready", "own_share_done"]
for (nm, T, n, k) in [("o T=64", 64, 4096, 4096), ("o T=1", 1, 4096, 4096), ("qkv T=64", 64, 6144, 4096), ("down T=64", 64, 4096, 14336),
                      ("gate_up T=64", 64, 28672, 4096), ("qkv T=1", 1, 6144, 4096), ("down T=1", 1, 4096, 14336),
                      ("gate_up T=1", 1, 28672, 4096)]:
    x = torch.randn(T, k, device="cuda").to(BF)
    ws_ = [torch.randn(n, k, device="cuda").to(BF) * 0.02 for _ in range(3)]
    y = torch.empty(T, n, device="cuda", dtype=BF)
    wsb = lib.b200_linear_workspace_bytes(T, n, k)
    ws = torch.zeros(wsb, dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for rep in range(3):
        flush.zero_()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        lib.b200_linear_bf16(x.data_ptr(), k, ws_[rep].data_ptr(), None, y.data_ptr(), n, T, n, k, 2, ws.data_ptr(), wsb, None)
        e1.record()
        torch.cuda.synchronize()
    buf = (C.c_ulonglong * 16)()
    lib.b200_debug_linear_trace(buf)
    t0 = buf[0]
    print(nm, f"event {e0.elapsed_time(e1)*1e3:.1f} us |", " ".join(f"{names[i]}={(buf[i]-t0)/1e3:.2f}" for i in (1, 2, 3, 4, 5, 6, 7, 10, 11, 8, 9)))
