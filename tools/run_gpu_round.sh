set -x
rm -f gpurun_out/scale_report.txt
SECONDS=0
timeout 1500 python -m pytest tests/test_scale_gpu.py -x -q -s 2>&1 | tail -25
echo "scale tests took $SECONDS s"; SECONDS=0
timeout 1500 python -m pytest tests/ -x -q -m gpu --deselect tests/test_scale_gpu.py 2>&1 | tail -8
echo "gpu tests took $SECONDS s"
sort -r gpurun_out/ulp_fractions.txt | head -40
