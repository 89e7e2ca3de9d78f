set -x
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py -x -q -k "adjust or penalt or sample or logprob" 2>&1 | tail -4
