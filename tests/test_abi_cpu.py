"""CPU-side checks of the drop-in boundary: the library loads, exports every symbol the
headers declare, and its host-only entry points behave like the reference's.  No compute."""
import ctypes
import hashlib
import os

import numpy as np
import pytest

from tests import util


def test_library_exports_every_declared_symbol(lib_path):
    lib = ctypes.CDLL(lib_path)
    missing = [s for s in util.declared_symbols() if not hasattr(lib, s)]
    assert not missing, missing
    assert len(util.declared_symbols()) >= 38


def test_python_mirror_binds(shared_library):
    info = shared_library.rwkv_get_system_info_string()
    assert "CUDA=" in info and "CPU_FALLBACK=0" in info


@pytest.mark.parametrize("key", sorted(k for k in util.quant_sha() if not k.startswith("6v0")))
def test_quantizer_is_byte_identical_to_reference(raw_lib, tmp_path, key):
    """rwkv_quantize_model_file vs sha256 of the files the reference's quantiser wrote
    (tests/test_tiny_rwkv.c:153-167 generates them the same way)."""
    arch, rest = key.rsplit("-FP", 1)
    src, q = ("FP" + rest).split("-to-")
    out = str(tmp_path / "q.bin")
    raw_lib.lib.rwkv_set_print_errors(None, False)
    assert raw_lib.quantize(os.path.join(util.MODELS, f"tiny-rwkv-{arch}-{src}.bin"), out, q)
    assert hashlib.sha256(open(out, "rb").read()).hexdigest() == util.quant_sha()[key]


def test_quantizer_rejects_bad_arguments(raw_lib, tmp_path):
    raw_lib.lib.rwkv_set_print_errors(None, False)
    src = os.path.join(util.MODELS, "tiny-rwkv-4v0-660K-FP32.bin")
    assert not raw_lib.quantize(src, str(tmp_path / "o.bin"), "FP16")            # not a quantised format
    assert raw_lib.lib.rwkv_get_last_error(None) & 0x100                          # RWKV_ERROR_ARGS
    assert not raw_lib.quantize(str(tmp_path / "missing.bin"), str(tmp_path / "o.bin"), "Q5_1")
    err = raw_lib.lib.rwkv_get_last_error(None)
    assert (err & 0xFF00) == 0x200 and raw_lib.lib.rwkv_get_last_error(None) == 0   # FILE; read-and-clear
    q = str(tmp_path / "q.bin")
    assert raw_lib.quantize(src, q, "Q4_0")
    assert not raw_lib.quantize(q, str(tmp_path / "qq.bin"), "Q4_0")             # already quantised input


def test_init_reports_errors_like_the_reference(raw_lib, tmp_path):
    lib = raw_lib.lib
    lib.rwkv_set_print_errors(None, False)
    assert lib.rwkv_get_print_errors(None) is False
    bad = tmp_path / "bad.bin"
    bad.write_bytes(b"\0" * 64)
    assert not lib.rwkv_init_from_file(str(bad).encode(), 1, 0)
    assert lib.rwkv_get_last_error(None) != 0
    assert not lib.rwkv_init_from_file(str(tmp_path / "nope.bin").encode(), 1, 0)
    assert lib.rwkv_get_last_error(None) != 0
    lib.rwkv_free(None)                                                           # NULL is allowed (rwkv.h:203-205)
    lib.rwkv_set_print_errors(None, True)


def test_no_cpu_fallback_without_a_gpu(raw_lib):
    """Without a CUDA device the product must fail loudly, never compute on the host."""
    if raw_lib.lib.rwkv_b200_device_count() > 0:
        pytest.skip("a GPU is present")
    raw_lib.lib.rwkv_set_print_errors(None, False)
    p = os.path.join(util.MODELS, "tiny-rwkv-4v0-660K-FP32.bin")
    assert not raw_lib.lib.rwkv_init_from_file(p.encode(), 1, 0)
    err = raw_lib.lib.rwkv_get_last_error(None)
    assert err & 0x600 and (err & 0xFF) == 9                                      # CTX | UNSUPPORTED
    out = np.zeros(32, np.float32)
    raw = np.zeros(18, np.uint8)
    assert not raw_lib.lib.rwkv_b200_dequantize(2, raw.ctypes.data, 1, 32, out.ctypes.data_as(ctypes.POINTER(ctypes.c_float)))
    raw_lib.lib.rwkv_set_print_errors(None, True)


def test_k_split_rule_is_a_function_of_the_shape(lib_path):
    """csrc/kernels.h matvec_ksplit through its C-ABI hook (no device needed: 148 SMs assumed): long rows with few row blocks are
    split 2 or 4 ways down to five 1024-weight chunks per warp, tiny grids with short rows two ways, everything else not at all."""
    import ctypes
    lib = ctypes.CDLL(lib_path)
    f = lib.rwkv_b200_matvec_k_split
    f.restype = ctypes.c_int
    f.argtypes = [ctypes.c_size_t, ctypes.c_size_t]
    expect = {(5120, 20480): 4, (5120, 5120): 1, (20480, 5120): 1, (5120, 10240): 2,     # 14B shape and its 2-way shards
              (768, 3072): 2, (3072, 768): 1, (768, 768): 1,                               # 169M shape: the tiny-grid rule
              (2560, 2560): 1, (2560, 10240): 2, (2048, 7168): 1,                          # v7 2.9B / v6 1.6B shapes
              (256, 10240): 2, (160, 20480): 4, (96, 21600): 4, (67, 2048): 2, (32, 32): 1}
    got = {k: f(*k) for k in expect}
    assert got == expect
    assert all(f(*k) == v for k, v in expect.items())      # and again: no hidden state
