set -x
nvidia-smi --query-gpu=name,memory.total --format=csv
timeout 900 python -m pytest tests/test_kernels_gpu.py -x -q -k "folded or rmsnorm or linear" 2>&1 | tail -15
timeout 900 python -m pytest tests/test_model_gpu.py -x -q 2>&1 | tail -15
tools/gpu_ab.sh bs1_default llama8b-bs1 128
tools/gpu_ab.sh bs1_nofold llama8b-bs1 128 B200_FUSE_NORM_MAX_T=0
tools/gpu_ab.sh bs1_noahead llama8b-bs1 128 B200_O_L2_AHEAD=0
tools/gpu_ab.sh bs1_neither llama8b-bs1 128 B200_O_L2_AHEAD=0 B200_FUSE_NORM_MAX_T=0
tools/gpu_ab.sh bs64_default llama8b-bs64 64
tools/gpu_ab.sh bs64_fold16 llama8b-bs64 64 B200_FUSE_NORM_MAX_T=16
