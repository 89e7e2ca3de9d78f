"""Generate tests/golden/*.npz: frozen oracle outputs + the HuggingFace cross-check.

Run here (CPU):  python oracle/make_golden.py
The reference has no golden vectors for this path ("parity unpinned");
these fixtures freeze (a) this repo's oracle and (b) HF transformers 5.5.0 on the same
seeded weights, as the closest executable stand-in for the vLLM-CPU image.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle.decode_oracle import (CONFIGS, OracleModel, make_prompt, make_weights,  # noqa: E402
                                  weights_digest)

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def hf_model(cfg, w):
    import transformers
    common = dict(hidden_size=cfg.hidden, num_hidden_layers=cfg.n_layers,
                  num_attention_heads=cfg.n_q_heads, num_key_value_heads=cfg.n_kv_heads,
                  head_dim=cfg.head_dim, intermediate_size=cfg.intermediate, vocab_size=cfg.vocab,
                  rms_norm_eps=cfg.rms_eps, max_position_embeddings=cfg.max_pos,
                  tie_word_embeddings=False, torch_dtype="bfloat16", attn_implementation="sdpa")
    if cfg.qkv_bias:
        hc = transformers.Qwen2Config(rope_theta=cfg.rope_theta, **common)
        m = transformers.Qwen2ForCausalLM(hc)
    else:
        rs = (dict(rope_type="llama3", factor=8.0, low_freq_factor=1.0, high_freq_factor=4.0,
                   original_max_position_embeddings=8192, rope_theta=cfg.rope_theta)
              if cfg.rope_llama3 else None)
        try:
            hc = transformers.LlamaConfig(rope_theta=cfg.rope_theta, rope_scaling=rs,
                                          attention_bias=False, **common)
        except TypeError:
            hc = transformers.LlamaConfig(rope_parameters=rs, attention_bias=False, **common)
        m = transformers.LlamaForCausalLM(hc)
    m = m.to(torch.bfloat16)
    # .to(bf16) also rounds the inv_freq buffer; rebuild the rotary module in fp32
    m.model.rotary_emb = type(m.model.rotary_emb)(config=hc)
    missing, unexpected = m.load_state_dict(w, strict=False)
    assert not [k for k in missing if "rotary" not in k], missing
    assert not unexpected, unexpected
    return m.eval()


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.manual_seed(0)
    only = sys.argv[1:]   # python oracle/make_golden.py [name ...]: regenerate just these fixtures
    for name, n_prompt, n_new in (("tiny", 24, 48), ("tiny-qwen", 17, 32), ("tiny-g8", 33, 32), ("tiny-g8x8", 40, 24)):
        if only and name not in only:
            continue
        cfg = CONFIGS[name]
        w = make_weights(cfg, seed=1234, norm_jitter=0.1)
        om = OracleModel(cfg, w, capacity=256)
        prompt = make_prompt(cfg, 0, n_prompt)
        toks, logits = om.generate(prompt, n_new, keep_logits=True)
        # teacher-forced logits over prompt+generated (what the GPU path is compared on)
        st = om.new_seq()
        tf_logits = om.forward(st, prompt + toks[:-1]).float()

        with torch.no_grad():
            hm = hf_model(cfg, w)
            ids = torch.tensor([prompt + toks[:-1]])
            hf_logits = hm(ids).logits[0].float()
        p_o = torch.softmax(tf_logits, -1)
        p_h = torch.softmax(hf_logits, -1)
        err = (p_o - p_h).abs().max().item()
        agree = (tf_logits.argmax(-1) == hf_logits.argmax(-1)).float().mean().item()
        print(f"{name}: oracle-vs-HF softmax max abs err {err:.3e}; greedy agreement {agree:.3f}; "
              f"logit max abs diff {(tf_logits - hf_logits).abs().max():.3e}")
        assert err <= 2e-2, "oracle disagrees with the HF stand-in beyond the north_star tolerance"
        np.savez_compressed(
            os.path.join(OUT, f"{name}.npz"),
            cfg=np.array(repr(cfg.to_dict())), seed=1234, norm_jitter=0.1,
            digest=np.array(weights_digest(w)),
            prompt=np.array(prompt, dtype=np.int32), tokens=np.array(toks, dtype=np.int32),
            gen_logits=logits.numpy().astype(np.float32),
            tf_logits=tf_logits.numpy().astype(np.float32),
            hf_logits=hf_logits.numpy().astype(np.float32),
            hf_softmax_err=err, hf_greedy_agreement=agree)


if __name__ == "__main__":
    main()
