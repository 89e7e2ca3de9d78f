set -x
tools/gpu_ab.sh bs64_base llama8b-bs64 64 X=1
tools/gpu_ab.sh bs64_g148 llama8b-bs64 64 B200_ATTN_GRID=148
tools/gpu_ab.sh bs64_s2 llama8b-bs64 64 B200_ATTN_SPLITS=2
tools/gpu_ab.sh bs64_s2g148 llama8b-bs64 64 B200_ATTN_SPLITS=2 B200_ATTN_GRID=148
