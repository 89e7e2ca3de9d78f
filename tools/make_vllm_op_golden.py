"""tests/golden/vllm_native_ops.npz: outputs of vLLM's OWN native-path operators on seeded bf16 inputs, to pin
the oracle's restatement of the rounding points R1-R3 and R5 (oracle/decode_oracle.py) at the operator level.

The reference deploys the vLLM-CPU image (v0.10.2); that image cannot run here, but the `vllm` package of this
container (0.22, CUDA build) carries the same pure-PyTorch `forward_native` implementations the CPU platform
falls back to, and they run on the host:
  RMSNorm.forward_native (with and without the fused residual add)   vllm/model_executor/layers/layernorm.py
  SiluAndMul.forward_native                                           vllm/model_executor/layers/activation.py
  Llama3RotaryEmbedding / RotaryEmbedding: cos_sin_cache + forward_native   vllm/model_executor/layers/rotary_embedding/
Run on CPU (needs the image's vllm):   python tools/make_vllm_op_golden.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from vllm.config import DeviceConfig, VllmConfig, set_current_vllm_config  # noqa: E402

BF = torch.bfloat16


def bits(t: torch.Tensor) -> np.ndarray:
    return t.contiguous().view(torch.int16).numpy()


out = {}
with set_current_vllm_config(VllmConfig(device_config=DeviceConfig(device="cpu"))):
    from vllm.model_executor.layers.activation import SiluAndMul
    from vllm.model_executor.layers.layernorm import RMSNorm
    from vllm.model_executor.layers.rotary_embedding import get_rope

    g = torch.Generator().manual_seed(20250101)
    # ---- R1/R2: RMSNorm and fused add + RMSNorm, hidden sizes of the tiny and the 8B configs
    for h in (256, 4096):
        x = (torch.randn(5, h, generator=g) * 1.7).to(BF)
        res = (torch.randn(5, h, generator=g) * 3.0).to(BF)
        w = (1.0 + 0.1 * torch.randn(h, generator=g)).to(BF)
        n = RMSNorm(h, eps=1e-5)
        n.weight.data = w.clone()
        y = n.forward_native(x.clone())
        y2, r2 = n.forward_native(x.clone(), res.clone())
        out.update({f"rms_x_{h}": bits(x), f"rms_res_{h}": bits(res), f"rms_w_{h}": bits(w), f"rms_y_{h}": bits(y),
                    f"addrms_y_{h}": bits(y2), f"addrms_res_{h}": bits(r2)})
    # ---- R5: SiLU * mul
    gu = (torch.randn(7, 2 * 1536, generator=g) * 2.5).to(BF)
    out["silu_gu"] = bits(gu)
    out["silu_y"] = bits(SiluAndMul().forward_native(gu.clone()))
    # ---- R3: rotary embedding (NeoX style), llama3-scaled and plain
    ropes = {
        "llama3": (500000.0, {"rope_type": "llama3", "rope_theta": 500000.0, "factor": 8.0, "low_freq_factor": 1.0,
                              "high_freq_factor": 4.0, "original_max_position_embeddings": 8192}),
        "plain": (1000000.0, {"rope_type": "default", "rope_theta": 1000000.0}),
    }
    pos = torch.tensor([0, 1, 2, 17, 127, 128, 1000, 4095, 8191])
    q = (torch.randn(len(pos), 4 * 128, generator=g) * 1.3).to(BF)
    k = (torch.randn(len(pos), 2 * 128, generator=g) * 1.3).to(BF)
    out["rope_pos"] = pos.numpy()
    cache_rows = torch.tensor(sorted(set(range(0, 8192, 37)) | set(pos.tolist())))
    out["rope_cache_rows"] = cache_rows.numpy()
    out["rope_q"], out["rope_k"] = bits(q), bits(k)
    for name, (theta, params) in ropes.items():
        rope = get_rope(128, 8192, True, params, BF)
        cache = rope.cos_sin_cache
        assert cache.dtype == BF and tuple(cache.shape) == (8192, 128)
        q2, k2 = rope.forward_native(pos, q.clone(), k.clone())
        out[f"rope_cache_{name}"] = bits(cache[cache_rows])     # a strided sample of the 8192 rows keeps the file small
        out[f"rope_q_{name}"], out[f"rope_k_{name}"] = bits(q2), bits(k2)

np.savez_compressed(os.path.join(ROOT, "tests", "golden", "vllm_native_ops.npz"), **out)
print("wrote", len(out), "arrays")
