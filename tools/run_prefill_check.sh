timeout 300 python tools/microbench.py prefill 2>&1 | tail -4
