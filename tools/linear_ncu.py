"""The three split-K linears of an 8B decode step at T = 64 (QKV 6144x4096, O 4096x4096, down 4096x14336), launched in
turn over rotating weights (> L2), for `ncu -k regex:linear_tc_kernel -s 6 -c 3` (one capture each) and for CUDA-event
timing without ncu.   python tools/linear_ncu.py [T]"""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
abi = importlib.import_module("enterprise-inference_b200._abi")
lib = abi.lib()
BF = torch.bfloat16
T = int(sys.argv[1]) if len(sys.argv) > 1 else 64
shapes = [("qkv", 6144, 4096), ("o", 4096, 4096), ("down", 4096, 14336)]
if len(sys.argv) > 2:      # python tools/linear_ncu.py 64 o   -> that shape only (ncu -k regex:linear_tc_kernel -s 4 -c 1)
    shapes = [sh for sh in shapes if sh[0] == sys.argv[2]]
copies = 6
W = {n: [torch.randn(no, k, device="cuda", dtype=BF) * 0.02 for _ in range(copies)] for n, no, k in shapes}
X = {n: torch.randn(T, k, device="cuda", dtype=BF) for n, no, k in shapes}
Y = {n: torch.empty(T, no, device="cuda", dtype=BF) for n, no, k in shapes}
wsb = max(lib.b200_linear_workspace_bytes(T, no, k) for _, no, k in shapes)
ws = torch.zeros(wsb, dtype=torch.uint8, device="cuda")
s = torch.cuda.current_stream().cuda_stream


def one(n, no, k, i):
    abi.check(lib.b200_linear_bf16(X[n].data_ptr(), k, W[n][i % copies].data_ptr(), None, Y[n].data_ptr(), no, T, no, k, 2,
                                   ws.data_ptr(), wsb, s))


for it in range(4):
    for n, no, k in shapes:
        one(n, no, k, it)
torch.cuda.synchronize()
for n, no, k in shapes:
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for it in range(24):
        one(n, no, k, it)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 24 * 1e3
    print(f"{n} T={T} [{no}x{k}] {us:.2f} us  {no * k * 2 / us / 1e3:.0f} GB/s")
