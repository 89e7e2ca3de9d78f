"""Host-side weight plumbing: HF names -> fused, per-rank shards (pure torch, no CUDA needed).

Packing follows [V22] vllm/model_executor/models/llama.py:437-444 (q|k|v -> qkv_proj, gate|up ->
gate_up_proj); sharding is Megatron-style as in SURVEY.md §8e: QKV and gate_up column-parallel (by
head / by intermediate feature), O and down row-parallel, lm_head vocab-parallel, embedding and norms
replicated.  Weights come from the HF cache the chart mounts at /data
(/root/reference/core/helm-charts/vllm/templates/configmap.yaml:18-21).
"""
from __future__ import annotations

import glob
import json
import os
from typing import Dict, Iterator, Tuple

import torch

from .config import ModelConfig

# kinds of include/b200_decode.h (b200_weight_kind)
EMBED, LN1, QKV, QKV_BIAS, O, LN2, GATE_UP, DOWN, FINAL_NORM, LM_HEAD, GATE_UP_INTERLEAVED = range(11)


def interleave_gate_up(gate: torch.Tensor, up: torch.Tensor) -> torch.Tensor:
    """[I,h],[I,h] -> [2I,h] in 128-row slabs: 64 gate rows then the 64 matching up rows."""
    I, h = gate.shape
    return torch.stack([gate.reshape(I // 64, 64, h), up.reshape(I // 64, 64, h)], dim=1).reshape(2 * I, h)


def shard_weights(w: Dict[str, torch.Tensor], cfg: ModelConfig, rank: int = 0, world: int = 1
                  ) -> Iterator[Tuple[int, int, torch.Tensor]]:
    """Yield (layer, kind, tensor) for one TP rank, in the order the runner uploads them."""
    d = cfg.head_dim
    hq, hkv = cfg.n_q_heads // world, cfg.n_kv_heads // world
    inter, vloc = cfg.intermediate // world, cfg.vocab // world
    if cfg.n_q_heads % world or cfg.n_kv_heads % world or cfg.intermediate % world or cfg.vocab % world:
        raise ValueError(f"model does not divide over tensor_parallel_size={world}")
    qs = slice(rank * hq * d, (rank + 1) * hq * d)
    ks = slice(rank * hkv * d, (rank + 1) * hkv * d)
    is_ = slice(rank * inter, (rank + 1) * inter)
    vs = slice(rank * vloc, (rank + 1) * vloc)
    yield 0, EMBED, w["model.embed_tokens.weight"]
    for l in range(cfg.n_layers):
        p = f"model.layers.{l}."
        yield l, LN1, w[p + "input_layernorm.weight"]
        yield l, QKV, torch.cat([w[p + "self_attn.q_proj.weight"][qs], w[p + "self_attn.k_proj.weight"][ks],
                                 w[p + "self_attn.v_proj.weight"][ks]], 0)
        if cfg.qkv_bias:
            yield l, QKV_BIAS, torch.cat([w[p + "self_attn.q_proj.bias"][qs], w[p + "self_attn.k_proj.bias"][ks],
                                          w[p + "self_attn.v_proj.bias"][ks]], 0)
        yield l, O, w[p + "self_attn.o_proj.weight"][:, qs]
        yield l, LN2, w[p + "post_attention_layernorm.weight"]
        gate, up = w[p + "mlp.gate_proj.weight"][is_], w[p + "mlp.up_proj.weight"][is_]
        if inter % 64 == 0:
            yield l, GATE_UP_INTERLEAVED, interleave_gate_up(gate, up)
        else:
            yield l, GATE_UP, torch.cat([gate, up], 0)
        yield l, DOWN, w[p + "mlp.down_proj.weight"][:, is_]
    yield 0, FINAL_NORM, w["model.norm.weight"]
    yield 0, LM_HEAD, w.get("lm_head.weight", w["model.embed_tokens.weight"])[vs]


def load_safetensors_dir(path: str) -> Dict[str, torch.Tensor]:
    """All tensors of an HF checkpoint directory (model.safetensors or sharded + index)."""
    from safetensors import safe_open
    files = sorted(glob.glob(os.path.join(path, "*.safetensors")))
    if not files:
        raise FileNotFoundError(f"no *.safetensors under {path}")
    idx = os.path.join(path, "model.safetensors.index.json")
    if os.path.exists(idx):
        with open(idx) as f:
            files = sorted({os.path.join(path, v) for v in json.load(f)["weight_map"].values()})
    out = {}
    for fn in files:
        with safe_open(fn, framework="pt", device="cpu") as f:
            for k in f.keys():
                out[k] = f.get_tensor(k)
    return out


def resolve_hf_snapshot(model_id: str, hf_home: str = None) -> str:
    """Path of a cached snapshot of `model_id` under HF_HOME (no network access is attempted)."""
    if os.path.isdir(model_id):
        return model_id
    hf_home = hf_home or os.environ.get("HF_HOME", os.path.expanduser("~/.cache/huggingface"))
    base = os.path.join(hf_home, "hub", "models--" + model_id.replace("/", "--"), "snapshots")
    snaps = sorted(glob.glob(os.path.join(base, "*")))
    if not snaps:
        raise FileNotFoundError(f"{model_id} is not in the HF cache at {hf_home} (HF_HOME); the server does not download")
    return snaps[-1]
