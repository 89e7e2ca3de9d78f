set -x
tools/gpu_ab.sh sh1_base llama70b-shard8-bs1 64 X=1
tools/gpu_ab.sh sh1_cap8 llama70b-shard8-bs1 64 B200_TC_MAX_SPLIT=8
tools/gpu_ab.sh sh64_base llama70b-shard8 32 X=1
tools/gpu_ab.sh sh64_cap8 llama70b-shard8 32 B200_TC_MAX_SPLIT=8
