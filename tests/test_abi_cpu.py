"""The C-ABI library loads on a CPU-only box and exports every symbol include/b200net.h declares."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200net.h")


def declared_symbols():
    txt = open(HEADER).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", txt)))


def test_header_declares_entry_points():
    syms = declared_symbols()
    assert len(syms) >= 35
    for must in ("b200_conv_fprop", "b200_conv_dgrad", "b200_conv_wgrad", "b200_bn_train_stats", "b200_bn_backward",
                 "b200_multi_tensor_sgd", "b200_multi_tensor_adam", "b200_gemm_bf16", "b200_bucket_pack"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    import b200net
    lib = ctypes.CDLL(b200net._C.LIB_PATH)
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, missing


def test_binding_table_matches_header():
    import b200net
    assert sorted(b200net._C.SIGNATURES) == declared_symbols()


def test_status_and_error_string_without_gpu():
    import b200net
    from b200net._C import lib, ConvDesc
    assert lib.b200_version() >= 100
    # host-only argument validation: no kernel is launched for a bad descriptor
    bad = ConvDesc(0, 8, 8, 8, 8, 3, 3, 1, 1, 0)
    assert lib.b200_conv_workspace_bytes(ctypes.byref(bad), 0) == -1
    assert "non-positive" in b200net._C.last_error()
    rc = lib.b200_conv_fprop(ctypes.byref(bad), None, None, None, None, 0, None)
    assert rc == 1
    rc = lib.b200_gemm_f32(None, None, None, 4, 4, 4, 4, 4, 4, 0, 0, None, 0, 0, None)
    assert rc == 1 and "null" in b200net._C.last_error()
    with pytest.raises(RuntimeError):
        b200net._C.check(rc, "gemm_f32")


def test_workspace_queries_are_host_only():
    import b200net
    from b200net._C import lib, ConvDesc
    # 3x3 stride-1 conv, N=32 @ 40x40x64 (class b of SURVEY 8(a')), fp32
    d = ConvDesc(32, 40, 40, 64, 64, 3, 3, 1, 1, 0)
    M, ld = 32 * 40 * 40, 576
    assert lib.b200_conv_workspace_bytes(ctypes.byref(d), 0) == M * ld * 4
    assert lib.b200_conv_workspace_bytes(ctypes.byref(d), 1) == M * ld * 4
    assert lib.b200_conv_workspace_bytes(ctypes.byref(d), 2) > M * ld * 4
    # 1x1 stride-1 runs directly on the activation matrix: no staging
    d1 = ConvDesc(32, 40, 40, 64, 256, 1, 1, 1, 0, 1)
    assert lib.b200_conv_workspace_bytes(ctypes.byref(d1), 0) == 0
    assert lib.b200_conv_workspace_bytes(ctypes.byref(d1), 1) == 0
    assert lib.b200_conv_weight_ld(7, 7, 3) == 152
    assert lib.b200_bn_workspace_bytes(1000, 64) > 0
