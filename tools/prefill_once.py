"""One 2048-row prefill chunk (Llama-3.1-8B shapes, ctx 0..2048) for an ncu launch list:
   ncu --metrics gpu__time_duration.sum --clock-control none -k regex:b200 --csv --log-file gpurun_out/prefill.csv python tools/prefill_once.py
"""
import importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "enterprise-inference_b200"
Cfg = importlib.import_module(PKG + ".config")
R = importlib.import_module(PKG + ".runner")
os.environ.setdefault("B200_GRAPHS", "0")
T = 2048
cfg = Cfg.get_config("llama-3.1-8b")
r = R.ModelRunner(cfg, max_tokens=T, max_model_len=8192, kv_cache_gib=2)
r.init_random_weights()
nb = T // 128
bt = torch.arange(nb, dtype=torch.int32, device="cuda").view(1, nb)
pos = torch.arange(T, dtype=torch.int32, device="cuda")
ids = torch.randint(0, cfg.vocab, (T,), dtype=torch.int32, device="cuda")
rows = torch.tensor([T - 1], dtype=torch.int32, device="cuda")
z = torch.zeros(T, dtype=torch.int32, device="cuda")
torch.cuda.synchronize()
r.step_device(ids, pos, pos, bt, pos + 1, T, row_to_seq=z, logit_rows=rows, stream=torch.cuda.current_stream())
torch.cuda.synchronize()
print("launches", r.last_step_launches())
r.close()
