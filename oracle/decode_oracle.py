"""CPU oracle for the decode hot path (TEST INFRASTRUCTURE — never the product path).

PARITY UNPINNED end to end; operators pinned.  (R1-R3 and R5 below equal vLLM's own native-path operators bit
for bit, and the sampler's top-k/top-p masks equal vLLM's: tests/golden/vllm_native_ops.npz and
sampling_masks.npz, captured from the `vllm` package by tools/make_vllm_op_golden.py / make_sampling_golden.py.
No output of the v0.10.2 CPU image itself is available, so whole-model parity stays unpinned.)  The reference (psurabh/Enterprise-Inference v1.3.0) holds no source
for this path: its chart only names the third-party image
``public.ecr.aws/q9t5s3a7/vllm-cpu-release-repo:v0.10.2``
(/root/reference/core/helm-charts/vllm/values.yaml:23-28,
 /root/reference/core/helm-charts/vllm/xeon-values.yaml:29-32) and its tests/docs pin
no outputs (only ``temperature: 0`` smoke prompts whose completions are not recorded,
/root/reference/docs/single-node-deployment.md:80-87).  The algorithm restated here is
the published vLLM "native" forward for Llama / Qwen2 (vLLM v0.10.2 CPU build is the
pinned dependency; the readable copy consulted is the vLLM 0.22 wheel in this image,
cited as [V22]).  Because no golden vector exists in the reference, this oracle is
cross-checked against HuggingFace ``transformers`` 5.5.0 ``LlamaForCausalLM`` /
``Qwen2ForCausalLM`` (CPU, bf16) by ``oracle/make_golden.py`` — a stand-in, not the
reference — and its outputs are frozen under ``tests/golden/``.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` legs may import this module.

Rounding points (each is a decision; the CUDA path makes the same ones):
  R1  residual stream is bf16.  add+RMSNorm: s = fp32(x)+fp32(res); res' = bf16(s);
      var from the UNROUNDED s; n = bf16(s*rsqrt(var+eps)); y = bf16(fp32(n)*fp32(w))
      [V22] vllm/ir/ops/layernorm.py:10-21,40-58
  R2  linear: exact bf16*bf16 products, fp32 accumulation, (+fp32 bias), ONE rounding
      to bf16.  [V22] vllm/model_executor/layers/utils.py:277-299 (CPU linear dispatch)
  R3  RoPE: NeoX halves, cos/sin cache STORED bf16, arithmetic in bf16 — every
      elementwise op rounds: o1 = bf16(bf16(x1*c) - bf16(x2*s)), o2 = bf16(bf16(x2*c)
      + bf16(x1*s)).  [V22] rotary_embedding/base.py:58-63,69-92; common.py:160-183;
      llama3 frequency scaling: llama3_rope.py:33-54
  R4  attention (q_len 1 per query row, causal over the paged cache): scores fp32 =
      (q . k) * head_dim**-0.5, softmax fp32, P kept fp32, out = bf16(sum P*v).
      [V22] vllm/v1/attention/backends/cpu_attn.py:377-396; scale llama.py:160-161
  R5  SiLU*mul in bf16: a = bf16(fp32(bf16(silu(fp32 g))) * fp32(u)).
      [V22] vllm/model_executor/layers/activation.py:137-141
  R6  logits are bf16; sampler casts to fp32 and takes argmax, ties -> lowest index.
      [V22] vllm/v1/sample/sampler.py:90-91,235-236
"""
from __future__ import annotations

import hashlib
import math
from dataclasses import dataclass, field, asdict
from typing import Dict, List, Optional

import torch

BF16 = torch.bfloat16


@dataclass
class ModelConfig:
    """Shape of one decoder-only Llama/Qwen2 model (SURVEY.md §8 shape table)."""
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
    rope_llama3: bool = True          # llama3 scaling: factor 8, low 1, high 4, orig 8192
    rms_eps: float = 1e-5
    max_pos: int = 32768
    rope_factor: float = 8.0          # llama3 scaling parameters ([V22] rotary_embedding/llama3_rope.py:19-31)
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
        """P_dec of SURVEY.md §8: parameters streamed once per decode step."""
        h, L, I, V = self.hidden, self.n_layers, self.intermediate, self.vocab
        return L * (h * (self.q_size + 2 * self.kv_size) + self.q_size * h + 3 * h * I) + V * h

    def kv_bytes_per_token(self) -> int:
        return 2 * self.n_kv_heads * self.head_dim * 2 * self.n_layers


CONFIGS: Dict[str, ModelConfig] = {
    "tiny": ModelConfig(),
    # G=5 and bias: the Qwen2-32B head ratio on a small body
    "tiny-qwen": ModelConfig(name="tiny-qwen", hidden=640, n_layers=2, n_q_heads=5, n_kv_heads=1,
                             intermediate=1536, vocab=2048, qkv_bias=True, rope_theta=1e6,
                             rope_llama3=False),
    # G=8 (70B head ratio), 2 kv heads so TP=2 can shard it
    "tiny-g8": ModelConfig(name="tiny-g8", hidden=1024, n_layers=2, n_q_heads=16, n_kv_heads=2,
                           intermediate=2048, vocab=4096),
    # G=8 over 8 kv heads: shards over 2, 4 and 8 tensor-parallel ranks
    "tiny-g8x8": ModelConfig(name="tiny-g8x8", hidden=1024, n_layers=2, n_q_heads=64, n_kv_heads=8,
                             intermediate=4096, vocab=4096),
    "llama-3.1-8b": ModelConfig(name="llama-3.1-8b", hidden=4096, n_layers=32, n_q_heads=32,
                                n_kv_heads=8, intermediate=14336, vocab=128256, max_pos=131072),
    "qwen2-32b": ModelConfig(name="qwen2-32b", hidden=5120, n_layers=64, n_q_heads=40, n_kv_heads=8,
                             intermediate=27648, vocab=152064, qkv_bias=True, rope_theta=1e6,
                             rope_llama3=False, max_pos=131072),
    "llama-3.1-70b": ModelConfig(name="llama-3.1-70b", hidden=8192, n_layers=80, n_q_heads=64,
                                 n_kv_heads=8, intermediate=28672, vocab=128256, max_pos=131072),
}


def truncated(cfg: ModelConfig, n_layers: int, vocab: Optional[int] = None) -> ModelConfig:
    """Shape-faithful, layer-truncated copy (full h/I/heads) for full-width parity."""
    d = cfg.to_dict()
    d["n_layers"] = n_layers
    d["name"] = f"{cfg.name}-L{n_layers}"
    if vocab:
        d["vocab"] = vocab
    return ModelConfig(**d)


# --------------------------------------------------------------------------------------
# weights: HF names ([V22] vllm/model_executor/models/llama.py:437-444 packing map)
# --------------------------------------------------------------------------------------
def weight_names(cfg: ModelConfig) -> List[str]:
    names = ["model.embed_tokens.weight"]
    for i in range(cfg.n_layers):
        p = f"model.layers.{i}."
        names += [p + "input_layernorm.weight",
                  p + "self_attn.q_proj.weight", p + "self_attn.k_proj.weight",
                  p + "self_attn.v_proj.weight"]
        if cfg.qkv_bias:
            names += [p + "self_attn.q_proj.bias", p + "self_attn.k_proj.bias",
                      p + "self_attn.v_proj.bias"]
        names += [p + "self_attn.o_proj.weight", p + "post_attention_layernorm.weight",
                  p + "mlp.gate_proj.weight", p + "mlp.up_proj.weight", p + "mlp.down_proj.weight"]
    names += ["model.norm.weight", "lm_head.weight"]
    return names


def weight_shape(cfg: ModelConfig, name: str):
    h, I = cfg.hidden, cfg.intermediate
    if name.endswith("embed_tokens.weight") or name == "lm_head.weight":
        return (cfg.vocab, h)
    if name.endswith("layernorm.weight") or name == "model.norm.weight":
        return (h,)
    if "q_proj" in name:
        return (cfg.q_size, h) if name.endswith("weight") else (cfg.q_size,)
    if "k_proj" in name or "v_proj" in name:
        return (cfg.kv_size, h) if name.endswith("weight") else (cfg.kv_size,)
    if "o_proj" in name:
        return (h, cfg.q_size)
    if "gate_proj" in name or "up_proj" in name:
        return (I, h)
    if "down_proj" in name:
        return (h, I)
    raise KeyError(name)


def make_weights(cfg: ModelConfig, seed: int = 1234, norm_jitter: float = 0.0,
                 std: float = 0.02) -> Dict[str, torch.Tensor]:
    """Seeded N(0, std^2) bf16 weights, one torch.Generator stream in name order
    (SURVEY.md §8d).  Norm weights are 1 (+ N(0,norm_jitter^2) when asked, so that the
    bf16 weight multiply of R1 is exercised)."""
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    out = {}
    for n in weight_names(cfg):
        shp = weight_shape(cfg, n)
        if len(shp) == 1 and not n.endswith("bias"):
            w = torch.ones(shp, dtype=torch.float32)
            if norm_jitter:
                w = w + norm_jitter * torch.randn(shp, generator=g, dtype=torch.float32)
        elif n.endswith("bias"):
            w = 0.1 * torch.randn(shp, generator=g, dtype=torch.float32)
        else:
            w = std * torch.randn(shp, generator=g, dtype=torch.float32)
        out[n] = w.to(BF16)
    return out


def weights_digest(w: Dict[str, torch.Tensor]) -> str:
    """sha256 over the raw bf16 bytes in name order — detects RNG drift between boxes."""
    hsh = hashlib.sha256()
    for n in sorted(w):
        hsh.update(n.encode())
        hsh.update(w[n].contiguous().view(torch.int16).numpy().tobytes())
    return hsh.hexdigest()


def make_prompt(cfg: ModelConfig, request_index: int, length: int) -> List[int]:
    """Uniform token ids, seed = 1000 + request index."""
    g = torch.Generator(device="cpu")
    g.manual_seed(1000 + request_index)
    return torch.randint(0, cfg.vocab, (length,), generator=g).tolist()


# --------------------------------------------------------------------------------------
# the ops (R1..R6)
# --------------------------------------------------------------------------------------
def rope_inv_freq(cfg: ModelConfig) -> torch.Tensor:
    d = cfg.head_dim
    inv = 1.0 / (cfg.rope_theta ** (torch.arange(0, d, 2, dtype=torch.float32) / d))
    if cfg.rope_llama3:
        # [V22] rotary_embedding/llama3_rope.py:33-54
        factor, low, high, orig = cfg.rope_factor, cfg.rope_low_freq, cfg.rope_high_freq, cfg.rope_orig_max_pos
        low_wl, high_wl = orig / low, orig / high
        wl = 2 * math.pi / inv
        smooth = (orig / wl - low) / (high - low)
        inv = torch.where(wl < high_wl, inv,
                          torch.where(wl > low_wl, inv / factor,
                                      (1 - smooth) * inv / factor + smooth * inv))
    return inv


def cos_sin_cache(cfg: ModelConfig, max_pos: Optional[int] = None) -> torch.Tensor:
    """[max_pos, head_dim] bf16 = cos(0..d/2) | sin(0..d/2).  R3: stored in bf16."""
    t = torch.arange(max_pos or cfg.max_pos, dtype=torch.float32)
    fr = torch.einsum("i,j->ij", t, rope_inv_freq(cfg))
    return torch.cat((fr.cos(), fr.sin()), dim=-1).to(BF16)


def rms_norm(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    xf = x.float()
    var = xf.pow(2).mean(dim=-1, keepdim=True)
    n = (xf * torch.rsqrt(var + eps)).to(BF16)
    return (n.float() * w.float()).to(BF16)


def add_rms_norm(x: torch.Tensor, res: torch.Tensor, w: torch.Tensor, eps: float):
    s = x.float() + res.float()
    new_res = s.to(BF16)
    var = s.pow(2).mean(dim=-1, keepdim=True)
    n = (s * torch.rsqrt(var + eps)).to(BF16)
    return (n.float() * w.float()).to(BF16), new_res


def linear(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor] = None) -> torch.Tensor:
    y = x.float() @ w.float().t()
    if b is not None:
        y = y + b.float()
    return y.to(BF16)


def rope_apply(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    """x [T, heads, d] bf16; cos/sin [T, d/2] bf16.  All arithmetic in bf16 (R3)."""
    c, s = cos.unsqueeze(1), sin.unsqueeze(1)
    x1, x2 = x[..., : x.shape[-1] // 2], x[..., x.shape[-1] // 2:]
    o1 = x1 * c - x2 * s          # torch bf16 ops: each rounds to bf16
    o2 = x2 * c + x1 * s
    return torch.cat((o1, o2), dim=-1)


def attend(q: torch.Tensor, K: torch.Tensor, V: torch.Tensor, n_ctx: int, scale: float) -> torch.Tensor:
    """q [Hq, d] bf16; K,V [>=n_ctx, Hkv, d] bf16 -> [Hq, d] bf16 (R4)."""
    Hq, d = q.shape
    Hkv = K.shape[1]
    G = Hq // Hkv
    k = K[:n_ctx].float().permute(1, 0, 2)                 # [Hkv, n, d]
    v = V[:n_ctx].float().permute(1, 0, 2)
    qf = q.float().view(Hkv, G, d)
    s = torch.einsum("hgd,hnd->hgn", qf, k) * scale
    p = torch.softmax(s, dim=-1)
    o = torch.einsum("hgn,hnd->hgd", p, v)
    return o.reshape(Hq, d).to(BF16)


def silu_mul(gu: torch.Tensor) -> torch.Tensor:
    d = gu.shape[-1] // 2
    g, u = gu[..., :d], gu[..., d:]
    sg = torch.nn.functional.silu(g.float()).to(BF16)
    return (sg.float() * u.float()).to(BF16)


def greedy(logits_bf16: torch.Tensor) -> torch.Tensor:
    return logits_bf16.float().argmax(dim=-1)


def sampling_probs(logits_bf16: torch.Tensor, temperature: float, top_k: int = 0, top_p: float = 1.0,
                   min_p: float = 0.0) -> torch.Tensor:
    """The distribution vLLM's sampler draws from for one logits row [V] (temperature > 0): logits / T, top-k
    mask, top-p mask on the renormalised rest, softmax -- a restatement of
    [V22] vllm/v1/sample/sampler.py:235-256 (temperature, then the top-k/top-p sampler) and
    vllm/v1/sample/ops/topk_topp_sampler.py:328-370 (apply_top_k_top_p_pytorch: ascending sort, values below the
    k-th largest -> -inf, cumulative probability <= 1 - p -> -inf, the largest always kept).
    Returns fp32 probabilities [V] (0 for dropped tokens).  Pinned against the vLLM functions themselves
    (apply_top_k_top_p_pytorch, MinPLogitsProcessor.apply) by tests/golden/sampling_masks.npz
    (tools/make_sampling_golden.py)."""
    x = logits_bf16.float() / float(temperature)
    if min_p and min_p > 0.0:
        # [V22] vllm/v1/sample/logits_processor/builtin.py:101 (MinPLogitsProcessor.apply), run by Sampler.sample after
        # the temperature and before top-k/top-p: in PROBABILITY space, strict "<" (a token exactly at the cut stays)
        pr = torch.nn.functional.softmax(x, dim=-1)
        x = x.masked_fill(pr < torch.amax(pr, dim=-1, keepdim=True) * torch.tensor(min_p, dtype=torch.float32), float("-inf"))
    xs, idx = x.sort(descending=False)
    V = x.numel()
    if top_k and 0 < top_k < V:
        kth = xs[V - top_k]
        xs = xs.masked_fill(xs < kth, float("-inf"))
    if top_p < 1.0:
        ps = xs.softmax(dim=-1)
        cum = torch.cumsum(ps, dim=-1)
        mask = cum <= 1.0 - top_p
        mask[-1] = False
        xs = xs.masked_fill(mask, float("-inf"))
    out = torch.empty_like(x)
    out[idx] = xs
    return out.softmax(dim=-1)


def adjust_logits(logits_bf16: torch.Tensor, prompt_ids, output_ids, repetition_penalty: float = 1.0,
                  frequency_penalty: float = 0.0, presence_penalty: float = 0.0, logit_bias=None) -> torch.Tensor:
    """One logits row [V] through logit_bias and the penalties, fp32, in the image's order and arithmetic:
    [V22] vllm/v1/sample/sampler.py:366-416 (logit_bias is a non-argmax-invariant processor and runs BEFORE the
    penalties), vllm/v1/sample/logits_processor/builtin.py:161-164 (logits += bias),
    vllm/model_executor/layers/utils.py:51-89 (apply_penalties) and vllm/_custom_ops.py:360-373 (repetition: tokens
    of the prompt OR the output are scaled by 1/rep where the logit is positive, by rep otherwise — a multiplication by
    the fp32 reciprocal, not a division; then - freq * count in the OUTPUT, - pres * (count > 0)).
    Pinned against those functions themselves by tests/golden/penalties.npz (tools/make_penalty_golden.py)."""
    x = logits_bf16.float().clone()
    V = x.numel()
    if logit_bias:
        for tok, b in logit_bias.items():
            x[int(tok)] += torch.tensor(float(b), dtype=torch.float32)
    counts = torch.zeros(V, dtype=torch.float32)
    for t in output_ids:
        counts[int(t)] += 1
    seen = counts > 0
    for t in prompt_ids:
        seen[int(t)] = True
    rep = torch.tensor(float(repetition_penalty), dtype=torch.float32)
    pen = torch.where(seen, rep, torch.tensor(1.0))
    x = x * torch.where(x > 0, 1.0 / pen, pen)
    x = x - torch.tensor(float(frequency_penalty), dtype=torch.float32) * counts
    x = x - torch.tensor(float(presence_penalty), dtype=torch.float32) * (counts > 0).float()
    return x


# --------------------------------------------------------------------------------------
# the model
# --------------------------------------------------------------------------------------
@dataclass
class SeqState:
    """Contiguous (un-paged) KV of one sequence: K[l], V[l] are [capacity, Hkv, d]."""
    K: List[torch.Tensor] = field(default_factory=list)
    V: List[torch.Tensor] = field(default_factory=list)
    n: int = 0


class OracleModel:
    """Token-at-a-time restatement of [V22] vllm/model_executor/models/llama.py:223-233
    (attention), :316-333 (decoder layer), :395-434 (model), :577-582 (logits) and
    models/qwen2.py:156-171 (qkv bias)."""

    def __init__(self, cfg: ModelConfig, weights: Dict[str, torch.Tensor], capacity: int = 1024):
        self.cfg = cfg
        self.w = weights
        self.capacity = capacity
        self.cs = cos_sin_cache(cfg, max(capacity, 16))
        self.scale = cfg.head_dim ** -0.5
        self.layers = []
        for i in range(cfg.n_layers):
            p = f"model.layers.{i}."
            qkv_w = torch.cat([weights[p + f"self_attn.{x}_proj.weight"] for x in "qkv"], 0)
            qkv_b = (torch.cat([weights[p + f"self_attn.{x}_proj.bias"] for x in "qkv"], 0)
                     if cfg.qkv_bias else None)
            gu_w = torch.cat([weights[p + "mlp.gate_proj.weight"], weights[p + "mlp.up_proj.weight"]], 0)
            self.layers.append(dict(
                ln1=weights[p + "input_layernorm.weight"], qkv_w=qkv_w, qkv_b=qkv_b,
                o_w=weights[p + "self_attn.o_proj.weight"],
                ln2=weights[p + "post_attention_layernorm.weight"], gu_w=gu_w,
                d_w=weights[p + "mlp.down_proj.weight"]))

    def new_seq(self) -> SeqState:
        c = self.cfg
        st = SeqState()
        for _ in range(c.n_layers):
            st.K.append(torch.zeros(self.capacity, c.n_kv_heads, c.head_dim, dtype=BF16))
            st.V.append(torch.zeros(self.capacity, c.n_kv_heads, c.head_dim, dtype=BF16))
        return st

    def forward(self, st: SeqState, token_ids: List[int], return_hidden: bool = False, fast_linear: bool = False):
        """Append T tokens to the sequence (causal) and return bf16 logits [T, V].
        T>1 is a (chunked) prefill, T==1 a decode step: the arithmetic per token is the same.
        fast_linear: the projections as bf16 F.linear (oneDNN: fp32 accumulation in its own order, one rounding — R2
        all the same); a second correct implementation, used as a noise yardstick by the full-depth test."""
        c = self.cfg
        linear = (lambda x, w, b=None: torch.nn.functional.linear(x, w, b)) if fast_linear else globals()["linear"]
        T = len(token_ids)
        pos = torch.arange(st.n, st.n + T)
        assert st.n + T <= self.capacity
        cos = self.cs[pos, : c.head_dim // 2]
        sin = self.cs[pos, c.head_dim // 2:]
        x = self.w["model.embed_tokens.weight"][torch.tensor(token_ids)]
        res = None
        hidden = []
        for l, L in enumerate(self.layers):
            if res is None:
                res = x
                y = rms_norm(x, L["ln1"], c.rms_eps)
            else:
                y, res = add_rms_norm(x, res, L["ln1"], c.rms_eps)
            qkv = linear(y, L["qkv_w"], L["qkv_b"])
            q = qkv[:, : c.q_size].view(T, c.n_q_heads, c.head_dim)
            k = qkv[:, c.q_size: c.q_size + c.kv_size].view(T, c.n_kv_heads, c.head_dim)
            v = qkv[:, c.q_size + c.kv_size:].view(T, c.n_kv_heads, c.head_dim)
            q = rope_apply(q, cos, sin)
            k = rope_apply(k, cos, sin)
            st.K[l][st.n: st.n + T] = k
            st.V[l][st.n: st.n + T] = v
            a = torch.stack([attend(q[t], st.K[l], st.V[l], st.n + t + 1, self.scale)
                             for t in range(T)]).view(T, c.q_size)
            x = linear(a, L["o_w"])
            y, res = add_rms_norm(x, res, L["ln2"], c.rms_eps)
            gu = linear(y, L["gu_w"])
            x = linear(silu_mul(gu), L["d_w"])
            if return_hidden:
                hidden.append((x.clone(), res.clone()))
        y, _ = add_rms_norm(x, res, self.w["model.norm.weight"], c.rms_eps)
        logits = linear(y, self.w["lm_head.weight"])
        st.n += T
        return (logits, hidden) if return_hidden else logits

    def step_batch(self, states: List[SeqState], token_ids: List[int], fast_linear: bool = False, kv_batch=None):
        """One decode step for B sequences (one new token each), linears batched over the rows —
        the shape of work the image's CPU model runner does per scheduler iteration.  With
        fast_linear the projections run as bf16 torch F.linear (oneDNN: fp32 accumulate, one
        rounding — still R2) instead of an explicit fp32 matmul; used by the timed CPU baseline.
        kv_batch = (K, V), each [B, capacity, Hkv, d] with states[b].K[l] / .V[l] = row b of it for every layer and all
        sequences at the same length (the timed baseline's synthetic cache): the attention of a layer is then ONE
        batched bf16 scaled_dot_product_attention call (what the image's CPU attention backend does per step)
        instead of a Python loop over sequences — timing path only, not a parity path (P stays in the SDPA
        kernel's own precision instead of R4's fp32)."""
        c = self.cfg
        B = len(states)
        batched = kv_batch is not None and len({st.n for st in states}) == 1
        lin = (lambda x, w, b=None: torch.nn.functional.linear(x, w, b)) if fast_linear else linear
        pos = torch.tensor([st.n for st in states])
        cos, sin = self.cs[pos, : c.head_dim // 2], self.cs[pos, c.head_dim // 2:]
        x = self.w["model.embed_tokens.weight"][torch.tensor(token_ids)]
        res = None
        for l, L in enumerate(self.layers):
            if res is None:
                res = x
                y = rms_norm(x, L["ln1"], c.rms_eps)
            else:
                y, res = add_rms_norm(x, res, L["ln1"], c.rms_eps)
            qkv = lin(y, L["qkv_w"], L["qkv_b"])
            q = rope_apply(qkv[:, : c.q_size].view(B, c.n_q_heads, c.head_dim), cos, sin)
            k = rope_apply(qkv[:, c.q_size: c.q_size + c.kv_size].view(B, c.n_kv_heads, c.head_dim), cos, sin)
            v = qkv[:, c.q_size + c.kv_size:].view(B, c.n_kv_heads, c.head_dim)
            if batched:
                n = states[0].n
                kv_batch[0][:, n] = k
                kv_batch[1][:, n] = v
                att = torch.nn.functional.scaled_dot_product_attention(
                    q.unsqueeze(2), kv_batch[0][:, : n + 1].permute(0, 2, 1, 3), kv_batch[1][:, : n + 1].permute(0, 2, 1, 3),
                    scale=self.scale, enable_gqa=True)
                x = lin(att.reshape(B, c.q_size), L["o_w"])
            else:
                outs = []
                for b, st in enumerate(states):
                    st.K[l][st.n] = k[b]
                    st.V[l][st.n] = v[b]
                    outs.append(attend(q[b], st.K[l], st.V[l], st.n + 1, self.scale))
                x = lin(torch.stack(outs).view(B, c.q_size), L["o_w"])
            y, res = add_rms_norm(x, res, L["ln2"], c.rms_eps)
            x = lin(silu_mul(lin(y, L["gu_w"])), L["d_w"])
        y, _ = add_rms_norm(x, res, self.w["model.norm.weight"], c.rms_eps)
        logits = lin(y, self.w["lm_head.weight"])
        for st in states:
            st.n += 1
        return logits

    def generate(self, prompt: List[int], n_new: int, keep_logits: bool = False):
        """Greedy, ignore_eos: exactly n_new tokens."""
        st = self.new_seq()
        logits = self.forward(st, prompt)[-1:]
        toks, all_logits = [], []
        for _ in range(n_new):
            t = int(greedy(logits)[0])
            toks.append(t)
            if keep_logits:
                all_logits.append(logits[0].float().clone())
            if len(toks) == n_new:
                break
            logits = self.forward(st, [t])
        return (toks, torch.stack(all_logits)) if keep_logits else toks


def top2_margin(logits_f32: torch.Tensor) -> torch.Tensor:
    """Oracle top-1/top-2 logit margin per row — used by the margin-aware greedy check."""
    v, _ = torch.topk(logits_f32, 2, dim=-1)
    return v[..., 0] - v[..., 1]
