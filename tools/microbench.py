"""Per-kernel timings on the GPU box (CUDA events, rotating buffers larger than L2).
   python tools/microbench.py [linear|attn|step] ...   -> JSON lines on stdout"""
import ctypes as C
import importlib
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "enterprise-inference_b200"
abi = importlib.import_module(PKG + "._abi")
lib = abi.lib()
BF = torch.bfloat16
PEAK = 6542.1
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass


def time_ms(fn, iters=20, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def bench_linear(impls=(1,), Ts=(1, 2, 4)):
    shapes = {"qkv": (6144, 4096), "o": (4096, 4096), "gate_up": (28672, 4096), "down": (4096, 14336),
              "lm_head": (128256, 4096)}
    for name, (n, k) in shapes.items():
        copies = max(2, int(300e6 // (n * k * 2)) + 1)
        ws = [torch.randn(n, k, device="cuda", dtype=BF) * 0.02 for _ in range(copies)]
        for T in Ts:
            x = torch.randn(T, k, device="cuda", dtype=BF)
            y = torch.empty(T, n, device="cuda", dtype=BF)
            wsb = lib.b200_linear_workspace_bytes(T, n, k)
            wsp = torch.zeros(max(wsb, 16), dtype=torch.uint8, device="cuda")
            for impl in impls:
                i = [0]

                def fn():
                    w = ws[i[0] % copies]
                    i[0] += 1
                    rc = lib.b200_linear_bf16(x.data_ptr(), k, w.data_ptr(), None, y.data_ptr(), n, T, n, k, impl,
                                              wsp.data_ptr(), wsb, torch.cuda.current_stream().cuda_stream)
                    if rc:
                        raise RuntimeError(lib.b200_last_error())
                try:
                    ms = time_ms(fn, iters=4 * copies)
                except RuntimeError as e:
                    print(json.dumps({"kernel": "linear", "shape": name, "T": T, "impl": impl, "error": str(e)[:100]}))
                    continue
                gbs = n * k * 2 / ms / 1e6
                print(json.dumps({"kernel": "linear", "shape": name, "n": n, "k": k, "T": T, "impl": impl,
                                  "us": round(ms * 1e3, 2), "GBps": round(gbs, 1), "frac_of_measured": round(gbs / PEAK, 3),
                                  "TFLOPs": round(2.0 * n * k * T / ms / 1e9, 1)}))
        del ws
        torch.cuda.empty_cache()


def bench_attn():
    for (B, ctx, hq, hkv) in ((8, 32768, 32, 8), (64, 640, 32, 8), (1, 192, 32, 8), (1, 8192, 32, 8), (1, 2048, 8, 1)):
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
        kv_bytes = B * ctx * hkv * 128 * 2 * 2

        def fn():
            if kv_bytes < 200e6:
                flush.zero_()
            rc = lib.b200_paged_attn_decode(q.data_ptr(), hq * 128, kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), nb_seq,
                                            None, cl.data_ptr(), out.data_ptr(), hq * 128, B, hq, hkv, bs, ctx, 128 ** -0.5,
                                            ws.data_ptr(), wsb, torch.cuda.current_stream().cuda_stream)
            if rc:
                raise RuntimeError(lib.b200_last_error())
        ms = time_ms(fn, iters=10)
        if kv_bytes < 200e6:
            ms -= time_ms(lambda: flush.zero_(), iters=10)
        gbs = kv_bytes / ms / 1e6
        print(json.dumps({"kernel": "paged_attn", "B": B, "ctx": ctx, "hq": hq, "hkv": hkv, "us": round(ms * 1e3, 1),
                          "GBps": round(gbs, 1), "frac_of_measured": round(gbs / PEAK, 3)}))


def bench_step(model="llama-3.1-8b", Bs=(1, 4, 64), ctx=192):
    Cfg = importlib.import_module(PKG + ".config")
    R = importlib.import_module(PKG + ".runner")
    cfg = Cfg.get_config(model)
    r = R.ModelRunner(cfg, max_tokens=max(Bs), max_model_len=8192, kv_cache_gib=8)
    r.init_random_weights()
    for B in Bs:
        nb = (ctx + 127) // 128
        bt = torch.arange(B * nb, dtype=torch.int32, device="cuda").view(B, nb)
        ids = torch.randint(0, cfg.vocab, (B,), dtype=torch.int32, device="cuda")
        pos = torch.full((B,), ctx - 1, dtype=torch.int32, device="cuda")
        slots = (bt[:, (ctx - 1) // 128] * 128 + (ctx - 1) % 128).to(torch.int32)
        cl = pos + 1
        out = torch.empty(B, dtype=torch.int32, device="cuda")
        for graphs in (0, 1):
            for pdl in ((0, 1) if graphs else (0,)):
                os.environ["B200_PDL"] = str(pdl)
                r.set_graphs(bool(graphs))

                def fn():
                    r.step_device(ids, pos, slots, bt, cl, ctx, out_tokens=out, stream=torch.cuda.current_stream())
                ms = time_ms(fn, iters=20, warmup=3)
                bytes_step = cfg.weight_bytes() + B * ctx * cfg.kv_bytes_per_token()
                print(json.dumps({"kernel": "decode_step", "model": model, "B": B, "ctx": ctx, "graphs": graphs, "pdl": pdl,
                                  "ms": round(ms, 4), "tok_s": round(B / ms * 1e3, 1), "launches": r.last_step_launches(),
                                  "GBps": round(bytes_step / ms / 1e6, 1), "frac_of_measured": round(bytes_step / ms / 1e6 / PEAK, 3)}))
    r.close()


if __name__ == "__main__":
    what = sys.argv[1:] or ["linear", "attn", "step"]
    if "linear" in what:
        bench_linear()
    if "linear_tc" in what:
        bench_linear(impls=(1, 2), Ts=(1, 4, 16, 64, 128))
    if "linear_big" in what:
        bench_linear(impls=(2,), Ts=(256, 512, 2048))
    if "attn" in what:
        bench_attn()
    if "step" in what:
        bench_step()


def bench_prefill(model="llama-3.1-8b", T=2048, n_chunks=2):
    """One sequence prefilled in chunks of T rows (the --max-num-batched-tokens budget): rows/s."""
    Cfg = importlib.import_module(PKG + ".config")
    R = importlib.import_module(PKG + ".runner")
    cfg = Cfg.get_config(model)
    r = R.ModelRunner(cfg, max_tokens=T, max_model_len=8192, kv_cache_gib=4)
    r.init_random_weights()
    nb = (T * n_chunks + 127) // 128
    bt = torch.arange(nb, dtype=torch.int32, device="cuda").view(1, nb)
    for c in range(n_chunks):
        pos = torch.arange(c * T, (c + 1) * T, dtype=torch.int32, device="cuda")
        ids = torch.randint(0, cfg.vocab, (T,), dtype=torch.int32, device="cuda")
        rows = torch.tensor([T - 1], dtype=torch.int32, device="cuda")
        z = torch.zeros(T, dtype=torch.int32, device="cuda")

        def fn():
            r.step_device(ids, pos, pos, bt, pos + 1, (c + 1) * T, row_to_seq=z, logit_rows=rows, stream=torch.cuda.current_stream())
        ms = time_ms(fn, iters=3, warmup=1)
        flops = 2 * cfg.decode_params() * T
        print(json.dumps({"kernel": "prefill_chunk", "model": model, "T": T, "ctx_end": (c + 1) * T, "ms": round(ms, 2),
                          "rows_per_s": round(T / ms * 1e3), "TFLOPs": round(flops / ms / 1e9, 1), "launches": r.last_step_launches()}))
    r.close()


if __name__ == "__main__" and "prefill" in sys.argv[1:]:
    bench_prefill()
