N=$1
set -x
run() { tag=$1; shift; env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 32 --warmup 4 --workload $WL --no-cpu-baseline > gpurun_out/tp${N}_$tag.json 2> gpurun_out/tp${N}_$tag.err; python - $N $tag <<'PY'
import json,sys
n,tag=sys.argv[1:3]
try:
    d=json.loads(open(f"gpurun_out/tp{n}_{tag}.json").read().strip().splitlines()[-1])
    print("RESULT", tag, d["ms_per_step"], "ms", d["value"], "tok/s whole", d["roofline"]["whole_step"]["frac"], "AR us", d.get("allreduce_us_per_call"), "launches", d["launches_per_step"])
except Exception as e:
    print("RESULT", tag, "FAILED", e); print(open(f"gpurun_out/tp{n}_{tag}.err").read()[-1500:])
PY
}
WL=llama70b-tp run bs64_twoshot X=1
WL=llama70b-tp run bs64_oneshot B200_AR_TWOSHOT=0
WL=llama70b-tp run bs64_nccl B200_PEER_ALLREDUCE=0
WL=llama70b-tp-bs1 run bs1 X=1
SECONDS=0
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "default bench rc=$? took $SECONDS s"
python - $N <<'PY'
import json,sys
n=sys.argv[1]
d=json.loads(open(f"gpurun_out/bench_n{n}.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("value","ms_per_step","n_gpus")}); t=d["tp"]; print(t["parity"])
for k in t:
    if isinstance(t[k],dict) and "ms_per_step" in t[k]: print(k, t[k]["ms_per_step"], t[k]["whole_step"]["frac"], t[k].get("allreduce_us_per_call"), t[k]["launches_per_step"])
PY
