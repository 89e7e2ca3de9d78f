set -x
L=$PWD/enterprise-inference_b200/lib
timeout 800 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu 2>&1 | tail -2
for v in new old new old; do
if [ $v = old ]; then E=B200_LIB=$L/libb200decode_old.so; else E=X=1; fi
tools/gpu_ab.sh ${v}64 llama8b-bs64 48 $E
tools/gpu_ab.sh ${v}1 llama8b-bs1 64 $E
tools/gpu_ab.sh ${v}32k llama8b-ctx32k 24 $E
done
