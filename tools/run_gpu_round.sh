set -x
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py -x -q 2>&1 | tail -2
SECONDS=0; timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "default bench rc=$? took $SECONDS s"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_default.json").read().strip().splitlines()[-1])
print("value", d["value"], "ms", d["ms_per_step"], "e2e", d["e2e"]["value"], "roofline", d["roofline"]["frac"], d["roofline"]["whole_step"]["frac"])
print([ (a["batch"], a["ms_per_step"], a["whole_step"]["frac"]) for a in d["also"]]); print({k:(v["ms_per_step"], v["whole_step"]["frac"]) for k,v in d["tp"].items() if isinstance(v,dict)})
PY
