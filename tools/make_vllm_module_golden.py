"""tests/golden/vllm_modules.npz: outputs of vLLM's OWN model modules run on the host — LlamaMLP (merged gate_up
projection -> SiluAndMul -> down projection: [V22] vllm/model_executor/models/llama.py:77-117) and the RMSNorm module with
its fused residual add (vllm/model_executor/layers/layernorm.py) — on seeded bf16 weights and inputs.  One level above the
operator pins of tools/make_vllm_op_golden.py: the module wiring (weight packing order of gate|up, which op rounds
where) is vLLM's, not a restatement.  The CUDA wheel has no CPU platform, but these modules run eagerly on host tensors
once a single-rank gloo "tensor-parallel group" exists and TorchDynamo is off.  LlamaAttention is not captured: its
attention backend is compiled code this wheel lacks (R4 stays a restatement, narrowed against torch SDPA).

    TORCHDYNAMO_DISABLE=1 python tools/make_vllm_module_golden.py
"""
import os
import sys

os.environ.setdefault("TORCHDYNAMO_DISABLE", "1")
os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT="29997", RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
import numpy as np  # noqa: E402
import torch  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
from vllm.config import DeviceConfig, VllmConfig, set_current_vllm_config  # noqa: E402
from vllm.distributed import init_distributed_environment, initialize_model_parallel  # noqa: E402

out = {}
bits = lambda t: t.contiguous().view(torch.int16).numpy()
with set_current_vllm_config(VllmConfig(device_config=DeviceConfig(device="cpu"))):
    init_distributed_environment(world_size=1, rank=0, distributed_init_method="tcp://127.0.0.1:29997", local_rank=0, backend="gloo")
    initialize_model_parallel(1, 1)
    from vllm.model_executor.layers.layernorm import RMSNorm
    from vllm.model_executor.models.llama import LlamaMLP
    for tag, h, inter, T in (("a", 512, 1024, 5), ("b", 1024, 2048, 17), ("c", 640, 1536, 3)):
        g = torch.Generator().manual_seed(h + T)
        mlp = LlamaMLP(hidden_size=h, intermediate_size=inter, hidden_act="silu", quant_config=None, bias=False, prefix="mlp").to(torch.bfloat16)
        for p in mlp.parameters():
            p.data.copy_((0.02 * torch.randn(p.shape, generator=g)).to(torch.bfloat16))
        x = torch.randn(T, h, generator=g).to(torch.bfloat16)
        # the fixture keeps OUTPUTS only: the test regenerates weights and inputs from the same seeded stream
        # (parameter order: gate_up_proj.weight [2I, h], down_proj.weight [h, I]), checked by a digest
        assert [tuple(p.shape) for p in mlp.parameters()] == [(2 * inter, h), (h, inter)]
        out[f"mlp_{tag}_digest"] = np.array(int(mlp.gate_up_proj.weight.data.view(torch.int16).long().sum() + x.view(torch.int16).long().sum()))
        out[f"mlp_{tag}_y"] = bits(mlp(x))
        n = RMSNorm(h, eps=1e-5).to(torch.bfloat16)
        n.weight.data.copy_((1 + 0.1 * torch.randn(h, generator=g)).to(torch.bfloat16))
        res = torch.randn(T, h, generator=g).to(torch.bfloat16)
        y, r = n(x.clone(), res.clone())
        out[f"norm_{tag}_y"], out[f"norm_{tag}_r"] = bits(y), bits(r)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "vllm_modules.npz"), **out)
print("wrote tests/golden/vllm_modules.npz")
