# usage: tools/run_gpu_tp.sh N   — TP tests (N = 2) and the default bench under torchrun at N ranks
N=$1
set -x
if [ "$N" = "2" ]; then timeout 900 python -m pytest tests/test_tp_gpu.py -x -q 2>&1 | tail -5; fi
SECONDS=0
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "bench rc=$? took $SECONDS s"
python - $N <<'PY'
import json,sys
n=sys.argv[1]
d=json.loads(open(f"gpurun_out/bench_n{n}.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("value","ms_per_step","n_gpus")}); t=d["tp"]; print(t["parity"])
for k in t:
    if isinstance(t[k],dict) and "ms_per_step" in t[k]: print(k, t[k]["ms_per_step"], t[k]["whole_step"]["frac"], t[k].get("allreduce_us_per_call"), t[k]["launches_per_step"])
PY
