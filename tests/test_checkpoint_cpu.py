"""CPU: the real-checkpoint path of the server — `--model <hf id>` resolved in the HF_HOME cache the
chart mounts at /data (/root/reference/core/helm-charts/vllm/templates/configmap.yaml:18-21), sharded safetensors, HF
tokenizer + chat template (xeon-values.yaml:87-88), generation_config.json defaults — booted through build_engine /
create_app with a fake runner in place of the device."""
import json
import types

import numpy as np
import pytest
import torch
from fastapi.testclient import TestClient

from ckpt_util import MODEL_ID, write_checkpoint
from conftest import pkg
from oracle import decode_oracle as O

S, W, C = pkg("server"), pkg("weights"), pkg("config")


class FakeRunner:
    """Stands in for ModelRunner: records what the server hands it; next token = (sum of context) % vocab."""
    instances = []

    def __init__(self, cfg, **kw):
        self.cfg, self.kw = cfg, kw
        self.n_blocks, self.block_size = 32, kw["block_size"]
        self.max_tokens, self.max_model_len = kw["max_tokens"], kw["max_model_len"]
        self.kv, self.loaded = {}, None
        FakeRunner.instances.append(self)

    def load_weights(self, w):
        self.loaded = list(W.shard_weights(w, self.cfg, self.kw["tp_rank"], self.kw["tp_world"]))

    def step(self, b):
        bs = self.block_size
        for i in range(b.n_tokens):
            self.kv[int(b.slots[i])] = int(b.ids[i])
        out = []
        for r in b.logit_rows:
            bt = b.block_tables[int(b.row_to_seq[r])]
            out.append(sum(self.kv[int(bt[p // bs]) * bs + p % bs] for p in range(int(b.ctx_lens[r]))) % self.cfg.vocab)
        return np.array(out, dtype=np.int32)


@pytest.fixture()
def ckpt(tmp_path, monkeypatch):
    snap, cfg, w = write_checkpoint(str(tmp_path))
    monkeypatch.setenv("HF_HOME", str(tmp_path))
    return snap, cfg, w


def args_for(model, **kw):
    a = S.make_arg_parser().parse_args(["--model", model, "--max-model-len", "256", "--max-num-batched-tokens", "64",
                                        "--max-num-seqs", "4", "--block-size", "128", "--enable_chunked_prefill"])
    for k, v in kw.items():
        setattr(a, k, v)
    return a


def test_snapshot_config_weights_and_tokenizer_load_from_hf_home(ckpt):
    snap, ocfg, w = ckpt
    assert W.resolve_hf_snapshot(MODEL_ID) == snap
    cfg = C.from_hf_config(snap)
    want = ocfg.to_dict()
    assert {k: v for k, v in cfg.to_dict().items() if k != "name"} == {k: v for k, v in want.items() if k != "name"}
    assert cfg.rope_factor == 32.0, "Llama-3.2's rope factor must come from config.json, not from a constant"
    assert torch.equal(C.rope_cos_sin(cfg, 2048), O.cos_sin_cache(ocfg, 2048))
    assert not torch.equal(C.rope_cos_sin(cfg, 2048), C.rope_cos_sin(C.ModelConfig(**{**cfg.to_dict(), "rope_factor": 8.0}), 2048))
    loaded = W.load_safetensors_dir(snap)
    assert sorted(loaded) == sorted(w) and all(torch.equal(loaded[k], w[k]) for k in w)
    for world in (1, 2):                      # hq = 4 shards over 2 ranks only with hkv >= 2: expect the clear error
        if cfg.n_kv_heads % world:
            with pytest.raises(ValueError, match="tensor_parallel_size"):
                list(W.shard_weights(loaded, cfg, 0, world))
    tok = S.load_tokenizer("auto", snap, cfg.vocab)
    ids = tok.encode("what is deep learning ?", add_special_tokens=False)
    assert tok.decode(ids) == "what is deep learning ?"
    text = tok.apply_chat_template([{"role": "user", "content": "hello world"}], add_generation_prompt=True, tokenize=False)
    assert text == "<|user|> hello world <|end|> <|assistant|>"
    assert S.load_default_sampling(snap) == {"temperature": 0.6, "top_p": 0.9}
    with pytest.raises(FileNotFoundError, match="HF cache"):
        W.resolve_hf_snapshot("acme/not-downloaded")


@pytest.mark.parametrize("bad,msg", [({"model_type": "mistral"}, "model_type"), ({"rope_scaling": {"rope_type": "yarn", "factor": 4}}, "rope"),
                                     ({"head_dim": 64}, "head_dim"), ({"use_sliding_window": True, "sliding_window": 4096}, "sliding")])
def test_configs_the_kernels_do_not_implement_are_rejected(ckpt, bad, msg):
    snap = ckpt[0]
    c = json.load(open(f"{snap}/config.json"))
    c.update(bad)
    json.dump(c, open(f"{snap}/config.json", "w"))
    with pytest.raises(ValueError, match=msg):
        C.from_hf_config(snap)


def test_server_boots_from_a_checkpoint_directory(ckpt, monkeypatch):
    """build_engine(--model <hf id>) -> create_app: weights reach the runner fused and in upload order, the HF tokenizer
    and its chat template serve /v1/chat/completions, generation_config.json supplies the sampling defaults."""
    snap, ocfg, w = ckpt
    R = pkg("runner")
    FakeRunner.instances.clear()
    monkeypatch.setattr(R, "ModelRunner", FakeRunner)
    monkeypatch.setattr(torch.cuda, "set_device", lambda *_: None)
    args = args_for(MODEL_ID)
    engine, tok, workers = S.build_engine(args)
    assert workers == [] and len(FakeRunner.instances) == 1
    fr = FakeRunner.instances[0]
    assert fr.cfg.vocab == ocfg.vocab and fr.kw["max_tokens"] == 64 and fr.kw["block_size"] == 128
    kinds = [k for (_, k, _) in fr.loaded]
    assert kinds[0] == W.EMBED and kinds[-1] == W.LM_HEAD and kinds.count(W.QKV) == ocfg.n_layers
    qkv0 = next(t for (l, k, t) in fr.loaded if k == W.QKV and l == 0)
    p = "model.layers.0.self_attn."
    assert torch.equal(qkv0, torch.cat([w[p + "q_proj.weight"], w[p + "k_proj.weight"], w[p + "v_proj.weight"]], 0))
    app = S.create_app(engine, tok, MODEL_ID, args.max_model_len, default_sampling=S.load_default_sampling(snap))
    with TestClient(app) as c:
        body = {"model": MODEL_ID, "messages": [{"role": "user", "content": "what is deep learning ?"}], "max_tokens": 6, "temperature": 0}
        r = c.post("/v1/chat/completions", json=body)
        assert r.status_code == 200, r.text
        prompt_ids = tok.encode("<|user|> what is deep learning ? <|end|> <|assistant|>", add_special_tokens=False)
        ids = list(prompt_ids)
        for _ in range(6):
            ids.append(sum(ids) % ocfg.vocab)
        out = r.json()
        assert out["usage"]["prompt_tokens"] == len(prompt_ids)
        # generation stops early if the fake model emits the tokenizer's eos
        n = out["usage"]["completion_tokens"]
        assert out["choices"][0]["message"]["content"] == tok.decode(ids[len(prompt_ids):len(prompt_ids) + n], skip_special_tokens=True)
        r = c.post("/tokenize", json={"prompt": "hello world"})
        assert r.json()["tokens"][-2:] == tok.encode("hello world", add_special_tokens=False)
