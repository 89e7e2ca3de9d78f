set -x
SECONDS=0; python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench took $SECONDS s"; tail -5 gpurun_out/bench_n1.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_n1.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("value","ms_per_step","e2e","launches_per_step")}, d["roofline"]["frac"], d["roofline"]["whole_step"])
print(json.dumps(d.get("also"),indent=1)); print(json.dumps(d.get("tp"),indent=1)); print(d.get("cpu_baseline",{}).get("value"))
PY
