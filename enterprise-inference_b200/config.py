"""Model shapes (SURVEY.md §8 shape table) and the RoPE cache, product side.

The reference names models only by HF id (/root/reference/core/helm-charts/vllm/xeon-values.yaml:61,124;
gaudi-values.yaml:165); the shapes are the public HF configs of those ids.
"""
from __future__ import annotations

import json
import math
import os
from dataclasses import asdict, dataclass
from typing import Dict, Optional

import torch


@dataclass
class ModelConfig:
    name: str = "tiny"
    hidden: int = 512
    n_layers: int = 2
    n_q_heads: int = 4
    n_kv_heads: int = 1
    head_dim: int = 128
    intermediate: int = 1024
    vocab: int = 1024
    qkv_bias: bool = False
    rope_theta: float = 500000.0
    rope_llama3: bool = True
    rms_eps: float = 1e-5
    max_pos: int = 32768
    # llama3 rope scaling (config.json rope_scaling): Llama-3.1 = 8 / 1 / 4 / 8192, Llama-3.2 uses factor 32
    rope_factor: float = 8.0
    rope_low_freq: float = 1.0
    rope_high_freq: float = 4.0
    rope_orig_max_pos: int = 8192

    def to_dict(self):
        return asdict(self)

    @property
    def q_size(self):
        return self.n_q_heads * self.head_dim

    @property
    def kv_size(self):
        return self.n_kv_heads * self.head_dim

    def decode_params(self) -> int:
        h, L, I, V = self.hidden, self.n_layers, self.intermediate, self.vocab
        return L * (h * (self.q_size + 2 * self.kv_size) + self.q_size * h + 3 * h * I) + V * h

    def weight_bytes(self) -> int:
        return 2 * self.decode_params()

    def kv_bytes_per_token(self) -> int:
        return 2 * self.n_kv_heads * self.head_dim * 2 * self.n_layers


MODELS: Dict[str, ModelConfig] = {
    "tiny": ModelConfig(),
    # ONE rank's share of Llama-3.1-70B at tensor-parallel 8 as a model of its own (8 of 64 q heads, 1 of 8 kv heads, 1/8 of
    # the MLP and of the vocabulary, full hidden size): the kernels of a TP8 step minus its collectives, for single-GPU
    # profiling (ncu must not wrap a multi-rank command)
    "llama-3.1-70b-tp8-shard": ModelConfig(name="llama-3.1-70b-tp8-shard", hidden=8192, n_layers=80, n_q_heads=8, n_kv_heads=1,
                                           intermediate=3584, vocab=16032, max_pos=131072),
    # the 70B head ratio (G = 8) over 8 kv heads on a small body: shards over 2, 4 and 8 tensor-parallel ranks
    "tiny-g8x8": ModelConfig(name="tiny-g8x8", hidden=1024, n_layers=2, n_q_heads=64, n_kv_heads=8, intermediate=4096,
                             vocab=4096),
    "meta-llama/Llama-3.1-8B-Instruct": ModelConfig(
        name="meta-llama/Llama-3.1-8B-Instruct", hidden=4096, n_layers=32, n_q_heads=32, n_kv_heads=8,
        intermediate=14336, vocab=128256, max_pos=131072),
    "deepseek-ai/DeepSeek-R1-Distill-Qwen-32B": ModelConfig(
        name="deepseek-ai/DeepSeek-R1-Distill-Qwen-32B", hidden=5120, n_layers=64, n_q_heads=40, n_kv_heads=8,
        intermediate=27648, vocab=152064, qkv_bias=True, rope_theta=1e6, rope_llama3=False, max_pos=131072),
    "meta-llama/Llama-3.1-70B-Instruct": ModelConfig(
        name="meta-llama/Llama-3.1-70B-Instruct", hidden=8192, n_layers=80, n_q_heads=64, n_kv_heads=8,
        intermediate=28672, vocab=128256, max_pos=131072),
}
ALIASES = {"llama-3.1-8b": "meta-llama/Llama-3.1-8B-Instruct",
           "qwen2-32b": "deepseek-ai/DeepSeek-R1-Distill-Qwen-32B",
           "llama-3.1-70b": "meta-llama/Llama-3.1-70B-Instruct"}


def get_config(name: str) -> ModelConfig:
    return MODELS[ALIASES.get(name, name)]


def from_hf_config(path: str) -> ModelConfig:
    """Read an HF config.json (the /data HF_HOME cache of the chart, templates/configmap.yaml:18-21)."""
    with open(os.path.join(path, "config.json")) as f:
        c = json.load(f)
    rs = c.get("rope_scaling") or c.get("rope_parameters") or {}
    # what the kernels implement, and nothing else: a config outside it must fail here, not produce wrong logits
    mt = c.get("model_type", "llama")
    if mt not in ("llama", "qwen2"):
        raise ValueError(f"unsupported model_type {mt!r} (llama and qwen2 decoder stacks only)")
    rope_type = rs.get("rope_type", rs.get("type")) or "default"
    if rope_type not in ("default", "llama3"):
        raise ValueError(f"unsupported rope scaling type {rope_type!r} (plain and llama3 RoPE only)")
    if c.get("use_sliding_window") and c.get("sliding_window"):
        raise ValueError("sliding-window attention is not supported")
    head_dim = c.get("head_dim") or c["hidden_size"] // c["num_attention_heads"]
    if head_dim != 128:
        raise ValueError(f"head_dim {head_dim} is not supported (the attention kernels are built for 128)")
    extra = {}
    if rope_type == "llama3":
        extra = dict(rope_factor=float(rs.get("factor", 8.0)), rope_low_freq=float(rs.get("low_freq_factor", 1.0)),
                     rope_high_freq=float(rs.get("high_freq_factor", 4.0)),
                     rope_orig_max_pos=int(rs.get("original_max_position_embeddings", 8192)))
    return ModelConfig(**extra, 
        name=c.get("_name_or_path", path), hidden=c["hidden_size"], n_layers=c["num_hidden_layers"],
        n_q_heads=c["num_attention_heads"], n_kv_heads=c.get("num_key_value_heads", c["num_attention_heads"]),
        head_dim=head_dim,
        intermediate=c["intermediate_size"], vocab=c["vocab_size"],
        qkv_bias=bool(c.get("attention_bias", False)) or c.get("model_type") == "qwen2",
        rope_theta=float(c.get("rope_theta", rs.get("rope_theta", 10000.0))),
        rope_llama3=(rope_type == "llama3"),
        rms_eps=float(c.get("rms_norm_eps", 1e-5)), max_pos=int(c.get("max_position_embeddings", 8192)))


def rope_cos_sin(cfg: ModelConfig, max_pos: Optional[int] = None) -> torch.Tensor:
    """[max_pos, head_dim] bf16 cache, cos | sin, computed on the CPU in fp32 then rounded
    (stored-bf16 is a rounding point: [V22] rotary_embedding/base.py:58-63; llama3 scaling
    llama3_rope.py:33-54)."""
    d = cfg.head_dim
    inv = 1.0 / (cfg.rope_theta ** (torch.arange(0, d, 2, dtype=torch.float32) / d))
    if cfg.rope_llama3:
        factor, low, high, orig = cfg.rope_factor, cfg.rope_low_freq, cfg.rope_high_freq, cfg.rope_orig_max_pos
        wavelen = 2 * math.pi / inv
        smooth = (orig / wavelen - low) / (high - low)
        mid = (1 - smooth) * inv / factor + smooth * inv
        inv = torch.where(wavelen < orig / high, inv, torch.where(wavelen > orig / low, inv / factor, mid))
    t = torch.arange(max_pos or cfg.max_pos, dtype=torch.float32)
    fr = torch.outer(t, inv)
    return torch.cat((fr.cos(), fr.sin()), dim=-1).to(torch.bfloat16)
