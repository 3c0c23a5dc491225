"""rwkv_b200_matrix_mul_no_alloc (include/rwkv_b200.h; csrc/no_alloc.cu): the quantised mat-vec with
its q8 activation scratch in a static __device__ array -- an entry point that allocates nothing.
Checked through the C ABI against the existing entry
point (rwkv_b200_matrix_mul) on two calls in a row, plus the size limit and argument errors.

Collected last on purpose (file name).
"""
import ctypes

import numpy as np
import pytest

from oracle import rwkv_oracle as O

pytestmark = pytest.mark.gpu
PF = ctypes.POINTER(ctypes.c_float)


def _bind(lib):
    lib.rwkv_b200_matrix_create.restype = ctypes.c_void_p
    lib.rwkv_b200_matrix_create.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_size_t]
    lib.rwkv_b200_matrix_mul.argtypes = [ctypes.c_void_p, PF, ctypes.c_size_t, PF]
    lib.rwkv_b200_matrix_mul.restype = ctypes.c_bool
    lib.rwkv_b200_matrix_mul_no_alloc.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
    lib.rwkv_b200_matrix_mul_no_alloc.restype = ctypes.c_bool
    lib.rwkv_b200_no_alloc_max_elements.restype = ctypes.c_size_t
    lib.rwkv_b200_matrix_free.argtypes = [ctypes.c_void_p]
    lib.rwkv_b200_matrix_free.restype = None


def _matrix(lib, t, M, K, seed):
    rng = np.random.default_rng(seed)
    w = (rng.standard_normal((M, K)) / np.sqrt(K)).astype(np.float32)
    raw = np.ascontiguousarray(O.quantize_ref(t, w.reshape(-1)))
    ptr = lib.rwkv_b200_matrix_create(t, raw.ctypes.data, M, K, 1)
    assert ptr
    return ptr


@pytest.mark.parametrize("t,M,K", [(O.TYPE_Q5_1, 1000, 2048), (O.TYPE_Q4_0, 96, 768), (O.TYPE_Q8_0, 5120, 5120)],
                         ids=["Q5_1-1000x2048", "Q4_0-96x768", "Q8_0-5120x5120"])
def test_no_alloc_matches_existing_entry_point(raw_lib, t, M, K):
    import torch
    lib = raw_lib.lib
    _bind(lib)
    lib.rwkv_set_print_errors(None, False)
    m = _matrix(lib, t, M, K, seed=M + K)
    rng = np.random.default_rng(7)
    dev = torch.device("cuda", 0)
    # two calls in a row with different inputs and column counts: the second call finds the first call's
    # blocks in the shared static scratch and must not be affected by them
    for n_vec, amp in ((3, 1.0), (1, 10.0), (8, 0.1)):
        x = (rng.standard_normal((n_vec, K)) * amp).astype(np.float32)
        y_ref = np.empty((n_vec, M), np.float32)
        assert lib.rwkv_b200_matrix_mul(m, x.ctypes.data_as(PF), n_vec, y_ref.ctypes.data_as(PF))
        x_d = torch.from_numpy(x).to(dev)
        y_d = torch.full((n_vec, M), float("nan"), dtype=torch.float32, device=dev)
        torch.cuda.synchronize()
        assert lib.rwkv_b200_matrix_mul_no_alloc(m, x_d.data_ptr(), n_vec, y_d.data_ptr()), hex(lib.rwkv_get_last_error(None))
        y = y_d.cpu().numpy()
        assert torch.equal(x_d.cpu(), torch.from_numpy(x)), "the input buffer was modified"
        # same kernels, same q8 blocks, same summation order: the tolerance of the mat-vec parity tests
        # (tests/test_gpu_codec.py: 2e-5 * scale * sqrt(K / 32)) and, in fact, the same bits
        scale = float(np.abs(y_ref).max()) + 1e-6
        assert np.abs(y - y_ref).max() <= 2e-5 * scale * np.sqrt(K / 32)
        assert np.array_equal(y, y_ref)
    lib.rwkv_b200_matrix_free(m)


def test_no_alloc_rejects_what_it_cannot_hold(raw_lib):
    import torch
    lib = raw_lib.lib
    _bind(lib)
    lib.rwkv_set_print_errors(None, False)
    limit = lib.rwkv_b200_no_alloc_max_elements()
    assert limit == 262144
    K, M = 4096, 64
    m = _matrix(lib, O.TYPE_Q4_1, M, K, seed=1)
    dev = torch.device("cuda", 0)
    n_ok, n_bad = limit // K, limit // K + 1
    x_d = torch.zeros((n_bad, K), dtype=torch.float32, device=dev)
    y_d = torch.zeros((n_bad, M), dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    assert lib.rwkv_b200_matrix_mul_no_alloc(m, x_d.data_ptr(), n_ok, y_d.data_ptr())          # exactly at the limit
    assert not lib.rwkv_b200_matrix_mul_no_alloc(m, x_d.data_ptr(), n_bad, y_d.data_ptr())     # one row past it
    err = lib.rwkv_get_last_error(None)
    assert (err & 0x100) and (err & 0xFF) == 10, hex(err)        # RWKV_ERROR_ARGS | RWKV_ERROR_SHAPE
    # host pointers are refused (nothing is staged), so is a non-quantised matrix
    xh = np.zeros((1, K), np.float32)
    assert not lib.rwkv_b200_matrix_mul_no_alloc(m, xh.ctypes.data, 1, y_d.data_ptr())
    assert lib.rwkv_get_last_error(None) & 0x100
    lib.rwkv_b200_matrix_free(m)
    w16 = np.zeros((M, K), np.float16)
    m16 = lib.rwkv_b200_matrix_create(1, w16.ctypes.data, M, K, 1)
    assert m16
    assert not lib.rwkv_b200_matrix_mul_no_alloc(m16, x_d.data_ptr(), 1, y_d.data_ptr())
    assert lib.rwkv_get_last_error(None) & 0x100
    lib.rwkv_b200_matrix_free(m16)
