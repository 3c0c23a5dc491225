"""GPU parity of the tensor-core GEMM (tcgen05 / TMEM, csrc/gemm_tc.cu) through the C ABI hook
rwkv_b200_matrix_mul_tc: FP16-dequantised weights x FP16 activations, FP32 accumulate.
Oracle: oracle.matmul_fp16 (same operand roundings, float64 accumulation), so the only
difference is the FP32 accumulation order inside the tensor core."""
import ctypes

import numpy as np
import pytest

from oracle import rwkv_oracle as O
from tests.test_gpu_codec import DevMatrix, make_matrix, PF

pytestmark = pytest.mark.gpu

TYPES = [O.TYPE_FP16, O.TYPE_Q4_0, O.TYPE_Q4_1, O.TYPE_Q5_0, O.TYPE_Q5_1, O.TYPE_Q8_0]


def mul_tc(dm, x):
    lib = dm.lib
    lib.rwkv_b200_matrix_mul_tc.argtypes = [ctypes.c_void_p, PF, ctypes.c_size_t, PF]
    lib.rwkv_b200_matrix_mul_tc.restype = ctypes.c_bool
    x = np.ascontiguousarray(x, np.float32).reshape(-1, dm.K)
    y = np.empty((x.shape[0], dm.M), np.float32)
    assert lib.rwkv_b200_matrix_mul_tc(dm.ptr, x.ctypes.data_as(PF), x.shape[0], y.ctypes.data_as(PF))
    return y


def check(dm, t, raw, M, K, x):
    y = mul_tc(dm, x)
    ref = O.matmul_fp16(t, raw, M, K, x)
    w = np.abs(O.dequantize_fp16(t, raw).reshape(M, K).astype(np.float64))
    bound = np.abs(x.astype(np.float16).astype(np.float64)) @ w.T        # sum |w x| per output
    err = np.abs(y - ref)
    # FP32 accumulation of K exact products: error <= ~K * 2^-24 * sum|w x| (loose factor 4)
    tol = 4.0 * K * 2.0 ** -24 * bound + 1e-30
    assert np.all(err <= tol), (float((err / tol).max()), float(err.max()))
    return y


@pytest.mark.parametrize("t", TYPES, ids=[O.TYPE_NAMES[t] for t in TYPES])
@pytest.mark.parametrize("M,K,T", [(128, 64, 64), (200, 256, 100), (1000, 768, 300), (96, 2048, 129), (384, 96, 64)])
def test_gemm_tc_matches_fp16_oracle(raw_lib, t, M, K, T):
    rng = np.random.default_rng(M + 3 * K + 5 * T + t)
    _, raw = make_matrix(rng, t, M, K)
    dm = DevMatrix(raw_lib.lib, t, raw, M, K)
    x = rng.standard_normal((T, K)).astype(np.float32) * rng.choice([0.1, 1.0, 10.0])
    y = check(dm, t, raw, M, K, x)
    # rows are independent: a different row count (different tile width) gives the same rows
    y2 = mul_tc(dm, x[: max(64, T // 2)])
    assert np.allclose(y2, y[: y2.shape[0]], rtol=0, atol=float(np.abs(y).max()) * 1e-5)
    dm.free()


@pytest.mark.parametrize("t", [O.TYPE_Q5_1, O.TYPE_Q8_0, O.TYPE_FP16], ids=["Q5_1", "Q8_0", "FP16"])
def test_gemm_tc_full_size(raw_lib, t):
    """BASELINE configs[2] shape (2048 x 2048, T = 1024): a one-hot column picks out one weight
    column exactly (dequantisation through the tensor path is then checked value by value), plus
    random rows against the oracle."""
    M = K = 2048
    T = 1024
    rng = np.random.default_rng(11)
    _, raw = make_matrix(rng, t, M, K)
    dm = DevMatrix(raw_lib.lib, t, raw, M, K)
    x = rng.standard_normal((T, K)).astype(np.float32)
    x[0] = 0.0; x[0, 5] = 1.0
    x[1] = 0.0; x[1, K - 1] = -2.0
    x[2] = 0.0
    y = check(dm, t, raw, M, K, x)
    w16 = O.dequantize_fp16(t, raw).reshape(M, K).astype(np.float32)
    assert np.array_equal(y[0], w16[:, 5])
    assert np.array_equal(y[1], -2.0 * w16[:, K - 1])
    assert np.all(y[2] == 0.0)
    dm.free()


@pytest.mark.parametrize("t", [O.TYPE_Q4_0, O.TYPE_Q5_1], ids=["Q4_0", "Q5_1"])
def test_gemm_tc_repeatable(raw_lib, t):
    """Stress of the producer / dequantiser / MMA hand-offs: the same product many times over, every
    repetition checked (a stale weight sub-tile shows up as a 32-row x tile-width block of wrong outputs)."""
    M, K, T = 4096, 2048, 1024
    rng = np.random.default_rng(3)
    _, raw = make_matrix(rng, t, M, K)
    dm = DevMatrix(raw_lib.lib, t, raw, M, K)
    x = rng.standard_normal((T, K)).astype(np.float32)
    ref = O.matmul_fp16(t, raw, M, K, x)
    scale = float(np.abs(ref).max())
    for rep in range(12):
        y = mul_tc(dm, x)
        bad = np.abs(y - ref) > 1e-3 * scale
        assert not bad.any(), (rep, int(bad.sum()))
    dm.free()


@pytest.mark.parametrize("t", [O.TYPE_Q5_1, O.TYPE_Q8_0, O.TYPE_FP16], ids=["Q5_1", "Q8_0", "FP16"])
@pytest.mark.parametrize("M,K,T", [(2048, 7168, 1024), (2560, 10240, 256), (160, 2048, 1024), (640, 2048, 200)])
def test_gemm_tc_few_tiles_long_k(raw_lib, t, M, K, T):
    """Shapes with few output tiles and a long K loop (ffn value, v6 time_maa_w1): the K-split launches (two CTAs
    per tile meeting in a scratch tile) against the FP16 oracle, repeated (flags must re-arm between launches)."""
    import os
    rng = np.random.default_rng(M + K + T + t)
    _, raw = make_matrix(rng, t, M, K)
    dm = DevMatrix(raw_lib.lib, t, raw, M, K)
    x = rng.standard_normal((T, K)).astype(np.float32)
    y1 = check(dm, t, raw, M, K, x)                       # whole-K tiles (default)
    os.environ["RWKV_B200_TC_KS"] = "2"                   # opt-in K split (read per launch)
    try:
        y = check(dm, t, raw, M, K, x)
        for _ in range(3):
            assert np.array_equal(mul_tc(dm, x), y)      # deterministic: own half + peer half, always in that order
    finally:
        del os.environ["RWKV_B200_TC_KS"]
    assert np.allclose(y, y1, rtol=0, atol=float(np.abs(y1).max()) * 1e-5)
    dm.free()
