# What the driver runs at round end, on one GPU: the GPU test suite, smoke(), the default bench and the reference arm.
set -x
SECONDS=0
timeout 1200 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -3
echo "gpu suite took $SECONDS s"
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
SECONDS=0; timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "default bench rc=$? took $SECONDS s"
SECONDS=0; timeout 600 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "reference arm rc=$? took $SECONDS s"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_default.json").read().strip().splitlines()[-1]); r=json.loads(open("gpurun_out/bench_ref.json").read().strip().splitlines()[-1])
print("value", d["value"], "ms", d["ms_per_step"], "e2e", d["e2e"]["value"], "roofline", d["roofline"]["frac"], d["roofline"]["whole_step"]["frac"], "clocks", d["clocks"])
print("same config:", d["config"] == r["config"], "ref", r["value"], r["ms_per_step"], "cpu_baseline", d["cpu_baseline"]["value"])
print([ (a["batch"], a["ms_per_step"], a["whole_step"]["frac"]) for a in d["also"]]); print({k:(v["ms_per_step"], v["whole_step"]["frac"]) for k,v in d["tp"].items() if isinstance(v,dict)})
PY
