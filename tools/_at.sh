set -x
L=$PWD/enterprise-inference_b200/lib
timeout 800 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py -x -q -m gpu 2>&1 | tail -2
timeout 300 python tools/microbench.py prefill 2>&1 | tail -2
B200_LIB=$L/libb200decode_old.so timeout 300 python tools/microbench.py prefill 2>&1 | tail -2
timeout 300 python tools/microbench.py prefill 2>&1 | tail -2
