"""tests/golden/sampling_masks.npz: keep-masks of vLLM's own top-k/top-p function (the pure-torch path of
[V22] vllm/v1/sample/ops/topk_topp_sampler.py:328, the same family as the reference image's sampler) on seeded
bf16 logits rows, to pin oracle.sampling_probs.  Needs the `vllm` package of this image; run on CPU:

    python tools/make_sampling_golden.py
"""
import os, sys
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vllm.v1.sample.ops.topk_topp_sampler import apply_top_k_top_p_pytorch  # noqa: E402

CASES = [  # (seed, vocab, logit scale, temperature, top_k, top_p)
    (1, 1000, 3.0, 1.0, 0, 0.9), (2, 1000, 3.0, 0.7, 50, 1.0), (3, 4096, 2.0, 1.3, 40, 0.95),
    (4, 257, 6.0, 0.5, 5, 0.5), (5, 32000, 4.0, 1.0, 0, 0.8), (6, 512, 1.0, 2.0, 511, 0.99),
    (7, 2048, 5.0, 0.9, 1, 1.0), (8, 2048, 5.0, 0.9, 0, 0.01),
]
out = {}
for seed, V, sc, T, k, p in CASES:
    g = torch.Generator().manual_seed(seed)
    logits = (torch.randn(V, generator=g) * sc).to(torch.bfloat16)
    x = (logits.float() / T)[None].clone()
    kk = torch.tensor([k]) if k else None
    pp = torch.tensor([p]) if p < 1.0 else None
    masked = apply_top_k_top_p_pytorch(x, kk, pp)[0]
    out[f"logits_{seed}"] = logits.view(torch.int16).numpy()
    out[f"keep_{seed}"] = torch.isfinite(masked).numpy()
    out[f"params_{seed}"] = np.array([T, k, p], dtype=np.float64)
# ---- min_p: vLLM's own MinPLogitsProcessor.apply ([V22] vllm/v1/sample/logits_processor/builtin.py:101), then the
# top-k/top-p function above, in the order of Sampler.sample (temperature -> argmax-invariant processors -> top-k/p).
# The processor keeps its per-request state in tensors that update_state() fills; apply() reads only min_p_count
# and min_p ([batch, 1] fp32), so those two are set directly and the real apply() runs.
from vllm.config import DeviceConfig, VllmConfig, set_current_vllm_config  # noqa: E402
from vllm.v1.sample.logits_processor.builtin import MinPLogitsProcessor  # noqa: E402

MINP_CASES = [  # (seed, vocab, logit scale, temperature, min_p, top_k, top_p)
    (21, 1000, 3.0, 1.0, 0.05, 0, 1.0), (22, 4096, 2.0, 0.7, 0.2, 0, 1.0), (23, 300, 2.5, 1.3, 0.01, 0, 1.0),
    (24, 2048, 4.0, 1.0, 0.1, 40, 1.0), (25, 2048, 4.0, 0.9, 0.02, 0, 0.9), (26, 512, 1.0, 2.0, 0.5, 20, 0.95),
    (27, 1000, 3.0, 1.0, 1.0, 0, 1.0),
]
with set_current_vllm_config(VllmConfig(device_config=DeviceConfig(device="cpu"))) as _:
    cfg_ = VllmConfig(device_config=DeviceConfig(device="cpu"))
    for seed, V, sc, T, mp, k, p in MINP_CASES:
        g = torch.Generator().manual_seed(seed)
        logits = (torch.randn(V, generator=g) * sc).to(torch.bfloat16)
        x = (logits.float() / T)[None].clone()
        proc = MinPLogitsProcessor(cfg_, torch.device("cpu"), False)
        proc.min_p_count = 1
        proc.min_p = torch.tensor([[mp]], dtype=torch.float32)
        probs = torch.softmax(x, dim=-1)[0]
        ratio = (probs / probs.max()).clone()            # what apply() compares with min_p, kept for the boundary band
        x = proc.apply(x)
        kk = torch.tensor([k]) if k else None
        pp = torch.tensor([p]) if p < 1.0 else None
        masked = apply_top_k_top_p_pytorch(x, kk, pp)[0]
        out[f"mp_logits_{seed}"] = logits.view(torch.int16).numpy()
        out[f"mp_keep_{seed}"] = torch.isfinite(masked).numpy()
        out[f"mp_ratio_{seed}"] = ratio.numpy()
        out[f"mp_params_{seed}"] = np.array([T, mp, k, p], dtype=np.float64)

np.savez_compressed(os.path.join(ROOT, "tests", "golden", "sampling_masks.npz"), **out)
print("wrote", len(CASES), "top-k/top-p cases and", len(MINP_CASES), "min_p cases")
