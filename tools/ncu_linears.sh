set -x
for sh in o down; do
timeout 120 python tools/linear_ncu.py 64 $sh > gpurun_out/lin_$sh.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:linear_tc_kernel -s 6 -c 1 -o gpurun_out/r02_linear_t64_$sh python tools/linear_ncu.py 64 $sh > gpurun_out/ncu_lin_$sh.log 2>&1
cat gpurun_out/lin_$sh.log
done
ls -la gpurun_out/*.ncu-rep
