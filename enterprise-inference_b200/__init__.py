"""B200-native decode backend for the Enterprise-Inference vLLM chart (hot path only).

Layout:  csrc/ (sm_100a CUDA + C-ABI), _abi.py (ctypes binding of include/b200_decode.h),
config.py (model shapes, chart flags), runner.py (device state + one step through the C-ABI),
engine.py (block manager + continuous-batching scheduler), server.py (OpenAI surface).
PyTorch is used for device memory, streams and torch.distributed only.
"""
__version__ = "0.1.0"
