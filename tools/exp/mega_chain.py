"""The persistent chain kernel (csrc/mega.cu) against the per-kernel linears: correctness of every phase and the time of a
32-layer chain of the four 8B projections (QKV, O, gate_up, down shapes; 14 GB of distinct weights) at T rows.
    python tools/mega_chain.py [T] [layers]"""
import ctypes as C
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
abi = importlib.import_module("enterprise-inference_b200._abi")
lib = abi.lib()
BF = torch.bfloat16
T = int(sys.argv[1]) if len(sys.argv) > 1 else 1
L = int(sys.argv[2]) if len(sys.argv) > 2 else 32
shapes = [(6144, 4096), (4096, 4096), (28672, 4096), (4096, 14336)]
g = torch.Generator(device="cuda").manual_seed(1)
W = [[torch.empty(n, k, device="cuda", dtype=BF).normal_(0, 0.7 / k ** 0.5, generator=g) for (n, k) in shapes] for _ in range(L)]
flat = [w for layer in W for w in layer]
n = len(flat)
x0 = torch.randn(T, 4096, device="cuda", dtype=BF, generator=g)
Y = [torch.full((T, w.shape[0]), float("nan"), device="cuda", dtype=BF) for w in flat]
fn = lib.b200_debug_mega_chain
fn.restype = C.c_int
fn.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]
wp = (C.c_void_p * n)(*[w.data_ptr() for w in flat])
yp = (C.c_void_p * n)(*[y.data_ptr() for y in Y])
no = (C.c_int32 * n)(*[w.shape[0] for w in flat])
kk = (C.c_int32 * n)(*[w.shape[1] for w in flat])
stream = torch.cuda.current_stream().cuda_stream


def mega():
    abi.check(fn(n, wp, no, kk, x0.data_ptr(), yp, T, stream))


mega()
torch.cuda.synchronize()
# every phase against an fp32 reference on the chain's own inputs
worst, nbad = 0.0, 0
for i, w in enumerate(flat[:16] + flat[-4:]):
    idx = i if i < 16 else n - 4 + (i - 16)
    x = x0 if idx == 0 else Y[idx - 1][:, : w.shape[1]]
    ref = (x.float() @ w.float().t()).to(BF).float()
    got = Y[idx].float()
    assert torch.isfinite(got).all(), f"phase {idx}: non-finite output"
    d = (got - ref).abs()
    ulp = ref.abs().clamp_min(2.0 ** -20).log2().floor().exp2() * 2.0 ** -7
    bad = (d > torch.maximum(ulp, torch.tensor(2.0 ** -9, device="cuda")))
    nbad += int(bad.sum())
    worst = max(worst, float((d / ulp).max()))
print(f"T={T}: {n} phases, checked 20: {nbad} elements beyond 1 ulp, worst {worst:.2f} ulp")

# time: the chain kernel vs the same chain as per-kernel launches (PDL, CUDA graph)
wsb = max(lib.b200_linear_workspace_bytes(T, w.shape[0], w.shape[1]) for w in flat[:4])
ws = torch.zeros(wsb, dtype=torch.uint8, device="cuda")
Y2 = [torch.empty_like(y) for y in Y]


def per_kernel(s):
    for i, w in enumerate(flat):
        x = x0 if i == 0 else Y2[i - 1]
        ldx = 4096 if i == 0 else Y2[i - 1].shape[1]
        abi.check(lib.b200_linear_bf16(x.data_ptr(), ldx, w.data_ptr(), None, Y2[i].data_ptr(), w.shape[0], T, w.shape[0], w.shape[1], 2,
                                       ws.data_ptr(), wsb, s))


def timeit(f, iters=5):
    f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        f()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


abi.check(lib.b200_set_launch_pdl(1))
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    per_kernel(side.cuda_stream)
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=side):
        per_kernel(side.cuda_stream)
abi.check(lib.b200_set_launch_pdl(0))
t_old = timeit(gr.replay)
t_new = timeit(mega)
nbytes = sum(w.numel() * 2 for w in flat)
same = sum(int(torch.equal(a, b)) for a, b in zip(Y, Y2))
print(f"T={T}: per-kernel graph {t_old:.3f} ms ({nbytes / t_old / 1e6:.0f} GB/s), chain kernel {t_new:.3f} ms "
      f"({nbytes / t_new / 1e6:.0f} GB/s); {same}/{n} phase outputs bit-identical between the two")
