#!/bin/bash
# A/B runs of bench.py under environment switches; every line lands in gpurun_out/ab_<tag>.jsonl
# usage: tools/gpu_ab.sh <tag> <workload> <steps> [ENV=VAL ...]   (one bench run)
tag=$1; wl=$2; steps=$3; shift 3
env "$@" timeout 180 python bench.py --workload "$wl" --steps "$steps" --warmup 8 --no-cpu-baseline --headline-only 2>gpurun_out/ab_${tag}.err | tail -1 > gpurun_out/ab_${tag}.json
python - "$tag" <<'PY'
import json,sys
t=sys.argv[1]
try:
    d=json.load(open(f"gpurun_out/ab_{t}.json"))
    print(t, d["ms_per_step"], "ms/step", d["value"], "tok/s", "whole", d["roofline"]["whole_step"]["frac"], "launches", d["launches_per_step"], "e2e", d["e2e"]["ms_per_step"])
except Exception as e:
    print(t, "FAILED", e); print(open(f"gpurun_out/ab_{t}.err").read()[-2000:])
PY
