set -x
timeout 400 python -m pytest tests/test_kernels_gpu.py tests/test_zz_no_alloc.py -x -q -k "attention or static" 2>&1 | tail -2
tools/gpu_ab.sh sh64 llama70b-shard8 32 X=1
tools/gpu_ab.sh sh1 llama70b-shard8-bs1 64 X=1
tools/gpu_ab.sh c32k llama8b-ctx32k 24 X=1
tools/gpu_ab.sh bs1 llama8b-bs1 64 X=1
