"""CPU: the C-ABI library loads and exports every symbol include/b200_decode.h declares
(no compute calls), and argument validation fails with a clear error instead of crashing."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT, pkg


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "b200_decode.h")).read()
    return sorted(set(re.findall(r"B200_API[^;(]*?\b(b200_\w+)\s*\(", src)))


def test_header_symbols_all_exported_and_bound():
    abi = pkg("_abi")
    lib = abi.lib()
    names = declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
        assert n in abi.SIGNATURES, f"{n} has no ctypes signature"
    assert sorted(abi.SIGNATURES) == names
    assert lib.b200_abi_version() == 1


def test_errors_are_reported_not_thrown():
    abi = pkg("_abi")
    lib = abi.lib()
    assert lib.b200_ctx_create(None, None) == -1
    assert b"null" in lib.b200_last_error()
    mc = abi.ModelCfg(hidden=512, n_layers=1, n_q_heads=4, n_kv_heads=1, head_dim=64, intermediate=1024,
                      vocab=1024, qkv_bias=0, rms_eps=1e-5, block_size=128, max_tokens=8,
                      max_blocks_per_seq=2, tp_rank=0, tp_world=1)
    h = C.c_void_p()
    assert lib.b200_ctx_create(C.byref(mc), C.byref(h)) == -3          # head_dim != 128
    assert b"head_dim" in lib.b200_last_error()
    mc.head_dim, mc.tp_world = 128, 3
    assert lib.b200_ctx_create(C.byref(mc), C.byref(h)) in (-1, -3)    # 4 heads / 3 ranks
    with pytest.raises(abi.B200Error):
        abi.check(lib.b200_ctx_set_graphs(None, 1))
    assert lib.b200_linear_workspace_bytes(64, 4096, 4096) > 0
    assert lib.b200_paged_attn_workspace_bytes(8, 32, 32768, 128) > 0


def test_missing_library_fails_loudly(monkeypatch):
    abi = pkg("_abi")
    monkeypatch.setattr(abi, "_lib", None)
    monkeypatch.setattr(abi, "LIB_PATH", "/nonexistent/libb200decode.so")
    with pytest.raises(ImportError, match="no CPU fallback"):
        abi.lib()


def test_sampler_uniform_is_strictly_inside_the_unit_interval():
    """The sampler draws by Gumbel-max: g = -log(-log(u)).  u == 1 would make g = +inf and that token would win
    whatever its logit (one draw in 2^24 per kept token: several garbage tokens per reply over a 128k vocabulary).
    The mapping from the 32 Philox bits must stay strictly inside (0, 1) at both end points and in between."""
    import math
    import numpy as np
    lib = pkg("_abi").lib()
    f = lib.b200_debug_uniform_from_bits
    f.restype, f.argtypes = C.c_float, [C.c_uint32]
    rng = np.random.default_rng(0)
    for bits in [0, 1, 0x1FF, 0x200, 0x7FFFFFFF, 0x80000000, 0xFFFFFE00, 0xFFFFFFFF] + rng.integers(0, 2 ** 32, 64).tolist():
        u = float(f(int(bits)))
        assert 0.0 < u < 1.0, f"bits {bits:#x} -> u = {u!r}"
        g = -math.log(-math.log(u))
        assert math.isfinite(g)
    assert f(0) == np.float32(2.0 ** -24) and f(0xFFFFFFFF) == np.float32(1.0 - 2.0 ** -24)
