"""tests/golden/penalties.npz: logit rows processed by vLLM's OWN penalty code — apply_penalties
([V22] vllm/model_executor/layers/utils.py:51-89, with the pure-torch repetition penalty of vllm/_custom_ops.py:360-373,
which is what runs on host tensors) after a logit_bias add as LogitBiasLogitsProcessor.apply does it
(vllm/v1/sample/logits_processor/builtin.py:161-164) — in the order of Sampler.apply_logits_processors
(vllm/v1/sample/sampler.py:366-416).  Pins oracle.adjust_logits.  Needs the `vllm` package of this image; run on CPU:

    python tools/make_penalty_golden.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vllm.model_executor.layers.utils import apply_penalties  # noqa: E402

CASES = [  # (seed, vocab, n_prompt, n_output, repetition, frequency, presence, n_bias)
    (1, 1000, 12, 20, 1.3, 0.0, 0.0, 0), (2, 1000, 5, 40, 1.0, 0.7, 0.0, 0), (3, 4096, 30, 10, 1.0, 0.0, 1.5, 0),
    (4, 512, 8, 64, 1.15, 0.3, -0.4, 0), (5, 2048, 1, 1, 0.8, -0.5, 0.6, 0), (6, 32000, 200, 300, 1.2, 0.2, 0.2, 7),
    (7, 257, 4, 0, 1.0, 0.0, 0.0, 5), (8, 1000, 0, 16, 2.0, 1.0, 1.0, 3),
]
out = {}
for seed, V, n_p, n_o, rep, freq, pres, n_b in CASES:
    g = torch.Generator().manual_seed(seed)
    logits = (torch.randn(V, generator=g) * 3.0).to(torch.bfloat16)
    hot = torch.randint(0, V, (max(V // 20, 4),), generator=g)          # tokens repeat, as real text does
    prompt = hot[torch.randint(0, len(hot), (n_p,), generator=g)]
    output = hot[torch.randint(0, len(hot), (n_o,), generator=g)]
    bias_ids = torch.randperm(V, generator=g)[:n_b]
    if n_b:
        bias_ids[0] = hot[0]                                             # one biased token is also a penalised one
    bias = (torch.randn(n_b, generator=g) * 4.0).float()
    x = logits.float()[None].clone()
    if n_b:
        x[0, bias_ids] += bias                                           # LogitBiasLogitsProcessor.apply
    pt = prompt[None].clone() if n_p else torch.full((1, 1), V, dtype=torch.long)   # vocab_size = vLLM's padding id
    ot = output[None].clone() if n_o else torch.full((1, 1), V, dtype=torch.long)
    y = apply_penalties(x, pt, ot, torch.tensor([pres]), torch.tensor([freq]), torch.tensor([rep]))[0]
    out[f"logits_{seed}"] = logits.view(torch.int16).numpy()
    out[f"prompt_{seed}"] = prompt.numpy().astype(np.int32)
    out[f"output_{seed}"] = output.numpy().astype(np.int32)
    out[f"params_{seed}"] = np.array([rep, freq, pres], dtype=np.float64)
    out[f"bias_ids_{seed}"] = bias_ids.numpy().astype(np.int32)
    out[f"bias_{seed}"] = bias.numpy()
    out[f"want_{seed}"] = y.numpy().astype(np.float32)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "penalties.npz"), **out)
print("wrote tests/golden/penalties.npz:", len(CASES), "cases")
