set -x
timeout 200 python tools/attn_ctx_sweep.py
B200_LIB=$PWD/enterprise-inference_b200/lib/libb200decode_old.so timeout 200 python tools/attn_ctx_sweep.py
