set -x
tools/gpu_ab.sh sh64_kb16 llama70b-shard8 32 B200_TC_MIN_KB=16
tools/gpu_ab.sh sh1_kb16 llama70b-shard8-bs1 64 B200_TC_MIN_KB=16
tools/gpu_ab.sh sh64_kb8 llama70b-shard8 32 B200_TC_MIN_KB=8
