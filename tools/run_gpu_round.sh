set -x
rm -f gpurun_out/scale_report.txt
SECONDS=0
timeout 1500 python -m pytest tests/test_scale_gpu.py -q -s 2>&1 | grep -v "^    \|^$" | tail -25
echo "scale tests took $SECONDS s"
cat gpurun_out/scale_report.txt
