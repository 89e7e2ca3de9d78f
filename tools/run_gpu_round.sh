set -x
timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -k "linear" 2>&1 | tail -2
tools/gpu_ab.sh sh1_cap8 llama70b-shard8-bs1 64 X=1
tools/gpu_ab.sh sh1_cap16 llama70b-shard8-bs1 64 B200_TC_MAX_SPLIT=16
tools/gpu_ab.sh sh64_cap8 llama70b-shard8 32 X=1
tools/gpu_ab.sh sh64_cap16 llama70b-shard8 32 B200_TC_MAX_SPLIT=16
tools/gpu_ab.sh bs64 llama8b-bs64 64 X=1
