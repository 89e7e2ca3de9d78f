"""CPU: bench.py's contract pieces that need no GPU — both arms take `config` from one function (the driver's
`same_config`), the reference arm prints ONE JSON line with the keys the contract names, and the mean context of the
timed steps follows the request-recycling rule."""
import importlib.util
import json
import os
import subprocess
import sys

from conftest import ROOT


def load_bench():
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_mean_context_and_config_are_functions_of_the_command_line():
    b = load_bench()
    assert b.timed_mean_ctx("llama8b-bs64", 20, 3) == 512 + 3 + (1 + 20) / 2          # steps 3..22 run at ctx 516..535
    assert b.timed_mean_ctx("llama8b-bs1", 256, 8) == sum(128 + (n % 128) + 1 for n in range(8, 264)) / 256   # recycled after 128 tokens
    a, c = b.config_for("llama8b-bs64", 20, 3, 1, False), b.config_for("llama8b-bs64", 20, 3, 1, False)
    assert a == c and a["mean_ctx"] == 525.5 and a["parallelism"] == "replicas x1" and "workload" in a
    assert b.config_for("llama70b-tp", 32, 4, 8, True)["parallelism"] == "tp8"
    assert set(b.OUT_LEN) == set(b.WORKLOADS)


def test_reference_arm_prints_one_json_line_with_the_contract_keys():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "tiny", "--steps", "3",
                        "--warmup", "1"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
              "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert k in d, k
    assert d["impl"] == "reference" and d["value"] > 0 and d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    b = load_bench()
    assert d["config"] == b.config_for("tiny", 3, 3, 1, False)          # warm-up as the B200 arm clamps it (>= 3)
