"""Writes a small HF-style checkpoint into an HF_HOME hub layout: sharded safetensors + index, config.json with a
llama3 rope_scaling block, a real `tokenizers` tokenizer with a chat template, generation_config.json — what the chart
mounts at /data (/root/reference/core/helm-charts/vllm/templates/configmap.yaml:18-21) for `--model <hf id>`."""
import json
import os

import torch

from oracle import decode_oracle as O

MODEL_ID = "acme/tiny-llama-3.2"
CHAT_TEMPLATE = ("{% for m in messages %}<|{{ m['role'] }}|> {{ m['content'] }} <|end|> {% endfor %}"
                 "{% if add_generation_prompt %}<|assistant|>{% endif %}")
WORDS = ("what is deep learning ? the answer a model of layers trained on data . hello world tell me about "
         "tensor cores and memory bandwidth please explain").split()


def oracle_cfg(vocab: int) -> O.ModelConfig:
    return O.ModelConfig(name=MODEL_ID, hidden=512, n_layers=2, n_q_heads=4, n_kv_heads=1, intermediate=1024, vocab=vocab,
                         rope_theta=500000.0, rope_llama3=True, rope_factor=32.0, rope_low_freq=1.0, rope_high_freq=4.0,
                         rope_orig_max_pos=8192, max_pos=4096)


def write_checkpoint(hf_home: str, seed: int = 31):
    """Returns (snapshot_dir, oracle config, weights dict)."""
    from safetensors.torch import save_file
    from tokenizers import Tokenizer, models, pre_tokenizers
    from transformers import PreTrainedTokenizerFast
    snap = os.path.join(hf_home, "hub", "models--" + MODEL_ID.replace("/", "--"), "snapshots", "0123abcd")
    os.makedirs(snap, exist_ok=True)
    specials = ["<unk>", "<|begin|>", "<|end|>", "<|user|>", "<|assistant|>", "<|system|>"]
    vocab = {t: i for i, t in enumerate(specials + sorted(set(WORDS)))}
    while len(vocab) % 8:
        vocab[f"<pad{len(vocab)}>"] = len(vocab)
    tk = Tokenizer(models.WordLevel(vocab=vocab, unk_token="<unk>"))
    tk.pre_tokenizer = pre_tokenizers.WhitespaceSplit()
    fast = PreTrainedTokenizerFast(tokenizer_object=tk, bos_token="<|begin|>", eos_token="<|end|>", unk_token="<unk>")
    fast.chat_template = CHAT_TEMPLATE
    fast.save_pretrained(snap)
    cfg = oracle_cfg(len(vocab))
    w = O.make_weights(cfg, seed=seed, norm_jitter=0.1)
    names = sorted(w)
    half = len(names) // 2
    shards = {"model-00001-of-00002.safetensors": names[:half], "model-00002-of-00002.safetensors": names[half:]}
    for fn, ks in shards.items():
        save_file({k: w[k].contiguous() for k in ks}, os.path.join(snap, fn))
    with open(os.path.join(snap, "model.safetensors.index.json"), "w") as f:
        json.dump({"metadata": {}, "weight_map": {k: fn for fn, ks in shards.items() for k in ks}}, f)
    with open(os.path.join(snap, "config.json"), "w") as f:
        json.dump({"model_type": "llama", "architectures": ["LlamaForCausalLM"], "hidden_size": cfg.hidden,
                   "num_hidden_layers": cfg.n_layers, "num_attention_heads": cfg.n_q_heads,
                   "num_key_value_heads": cfg.n_kv_heads, "head_dim": 128, "intermediate_size": cfg.intermediate,
                   "vocab_size": cfg.vocab, "rms_norm_eps": cfg.rms_eps, "rope_theta": cfg.rope_theta,
                   "max_position_embeddings": cfg.max_pos, "torch_dtype": "bfloat16", "_name_or_path": MODEL_ID,
                   "rope_scaling": {"rope_type": "llama3", "factor": 32.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
                                    "original_max_position_embeddings": 8192}}, f)
    with open(os.path.join(snap, "generation_config.json"), "w") as f:
        json.dump({"do_sample": True, "temperature": 0.6, "top_p": 0.9}, f)
    return snap, cfg, w
