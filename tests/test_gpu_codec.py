"""GPU parity of the block codecs and the dequant-fused mat-vec kernel against the oracle
(numpy restatement pinned to ggml) -- through the C ABI hooks of include/rwkv_b200.h."""
import ctypes
import os

import numpy as np
import pytest

from oracle import rwkv_oracle as O
from tests import util

pytestmark = pytest.mark.gpu
PF = ctypes.POINTER(ctypes.c_float)
QT = {"q4_0": O.TYPE_Q4_0, "q4_1": O.TYPE_Q4_1, "q5_0": O.TYPE_Q5_0, "q5_1": O.TYPE_Q5_1, "q8_0": O.TYPE_Q8_0}


def dev_dequant(lib, t, raw, rows, cols):
    out = np.empty((rows, cols), np.float32)
    raw = np.ascontiguousarray(raw)
    assert lib.rwkv_b200_dequantize(t, raw.ctypes.data, rows, cols, out.ctypes.data_as(PF))
    return out


@pytest.mark.parametrize("name", list(QT))
def test_device_dequant_bit_exact_on_ggml_vectors(raw_lib, name):
    """dequantize_row_* (ggml-quants.c:255-363) incl. d=0, max half, subnormal, all-ones qh."""
    cv = np.load(os.path.join(util.GOLDEN, "codec_vectors.npz"))
    raw = cv[f"{name}/raw"]                                   # [512 blocks][block bytes]
    y = dev_dequant(raw_lib.lib, QT[name], raw, 16, 32 * 32)  # 16 rows x 32 blocks
    assert np.array_equal(y.reshape(-1).view(np.uint32), cv[f"{name}/deq"].view(np.uint32))


@pytest.mark.parametrize("name", list(QT))
def test_device_dequant_non_finite_scales(raw_lib, name):
    """Blocks whose FP16 scale (or minimum) is +-inf or NaN: the device produces inf / NaN exactly where dequantize_row_* does (the numpy
    restatement, pinned to ggml on the finite vectors) and the same finite values elsewhere.  NaN payloads are not compared: x86
    propagates the operand's payload, the GPU returns the canonical NaN."""
    rng = np.random.default_rng(7)
    bs = {"q4_0": 18, "q4_1": 20, "q5_0": 22, "q5_1": 24, "q8_0": 34}[name]
    raw = rng.integers(0, 256, size=(64, bs), dtype=np.uint8)
    raw[:, 0:2] = (rng.standard_normal(64) * 0.05).astype(np.float16).view(np.uint8).reshape(64, 2)
    if name in ("q4_1", "q5_1"):
        raw[:, 2:4] = (rng.standard_normal(64) * 0.5).astype(np.float16).view(np.uint8).reshape(64, 2)
    for i, v in enumerate([np.inf, -np.inf, np.nan]):
        raw[i, 0:2] = np.array([v], np.float16).view(np.uint8)
        if name in ("q4_1", "q5_1"):
            raw[8 + i, 2:4] = np.array([v], np.float16).view(np.uint8)
    y = dev_dequant(raw_lib.lib, QT[name], raw, 2, 32 * 32).reshape(-1)
    with np.errstate(all="ignore"):
        ref = O.dequantize(QT[name], raw.reshape(-1))
    assert np.array_equal(np.isnan(y), np.isnan(ref)) and np.array_equal(np.isposinf(y), np.isposinf(ref)) and np.array_equal(np.isneginf(y), np.isneginf(ref))
    fin = np.isfinite(ref)
    assert np.array_equal(y[fin].view(np.uint32), ref[fin].view(np.uint32))


@pytest.mark.parametrize("fmt", util.SIX)
def test_device_dequant_bit_exact_on_model_tensors(raw_lib, fmt):
    """Every quantised tensor of the committed 6v0 files, after upload + repack."""
    f = O.read_ggmf(os.path.join(util.MODELS, f"tiny-rwkv-6v0-3m-{fmt}.bin"))
    n = 0
    for name in f.order:
        t = f.tensors[name]
        if t.dtype not in QT.values() or not name.startswith(("blocks.0.", "blocks.7.")):
            continue
        K, M = t.shape[0], int(np.prod(t.shape[1:]))
        y = dev_dequant(raw_lib.lib, t.dtype, t.data, M, K)
        assert np.array_equal(y.reshape(-1).view(np.uint32), O.dequantize(t.dtype, t.data).view(np.uint32)), name
        n += 1
    assert n >= 16


def test_device_dequant_passthrough_f16_f32(raw_lib):
    rng = np.random.default_rng(0)
    x = rng.standard_normal((5, 40)).astype(np.float32)
    assert np.array_equal(dev_dequant(raw_lib.lib, 0, x.view(np.uint8), 5, 40), x)
    h = x.astype(np.float16)
    assert np.array_equal(dev_dequant(raw_lib.lib, 1, h.view(np.uint8), 5, 40), h.astype(np.float32))


def make_matrix(rng, t, M, K):
    w = (rng.standard_normal((M, K)) / np.sqrt(K)).astype(np.float32)
    if t == O.TYPE_FP32:
        raw = w.view(np.uint8).reshape(-1)
    elif t == O.TYPE_FP16:
        raw = w.astype(np.float16).view(np.uint8).reshape(-1)
    else:
        raw = O.quantize_ref(t, w.reshape(-1))
    return O.Mat(O.GgmfTensor("w", t, (K, M), raw)), raw


class DevMatrix:
    def __init__(self, lib, t, raw, M, K, copies=1):
        lib.rwkv_b200_matrix_create.restype = ctypes.c_void_p
        lib.rwkv_b200_matrix_create.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_size_t]
        lib.rwkv_b200_matrix_mul.argtypes = [ctypes.c_void_p, PF, ctypes.c_size_t, PF]
        lib.rwkv_b200_matrix_mul.restype = ctypes.c_bool
        lib.rwkv_b200_matrix_free.argtypes = [ctypes.c_void_p]
        self.lib, self.M, self.K = lib, M, K
        raw = np.ascontiguousarray(raw)
        self.ptr = lib.rwkv_b200_matrix_create(t, raw.ctypes.data, M, K, copies)
        assert self.ptr

    def mul(self, x):
        x = np.ascontiguousarray(x, np.float32).reshape(-1, self.K)
        y = np.empty((x.shape[0], self.M), np.float32)
        assert self.lib.rwkv_b200_matrix_mul(self.ptr, x.ctypes.data_as(PF), x.shape[0], y.ctypes.data_as(PF))
        return y

    def free(self):
        self.lib.rwkv_b200_matrix_free(self.ptr)


TYPES = [O.TYPE_FP32, O.TYPE_FP16, O.TYPE_Q4_0, O.TYPE_Q4_1, O.TYPE_Q5_0, O.TYPE_Q5_1, O.TYPE_Q8_0]


@pytest.mark.parametrize("t", TYPES, ids=[O.TYPE_NAMES[t] for t in TYPES])
@pytest.mark.parametrize("M,K", [(96, 256), (1000, 768), (67, 2048), (32, 32)])
def test_matvec_matches_reference_arithmetic(raw_lib, t, M, K):
    """y = W x computed as ggml's CPU backend does (q8 activation blocks + integer block dots /
    fp16-rounded activations): only the order of the FP32 block sums may differ."""
    rng = np.random.default_rng(M * 7 + K + t)
    mat, raw = make_matrix(rng, t, M, K)
    dm = DevMatrix(raw_lib.lib, t, raw, M, K)
    for n_vec in (1, 2, 3, 8, 11):
        x = rng.standard_normal((n_vec, K)).astype(np.float32) * rng.choice([0.1, 1.0, 10.0])
        y = dm.mul(x)
        ref = np.stack([mat.matvec(x[i], "ref") for i in range(n_vec)])
        scale = np.abs(ref).max() + 1e-6
        assert np.abs(y - ref).max() <= 2e-5 * scale * np.sqrt(K / 32), (n_vec, np.abs(y - ref).max(), scale)
        if n_vec > 1:                       # column results do not depend on how many columns share the pass
            y1 = dm.mul(x[n_vec - 1])
            assert np.array_equal(y1[0], y[n_vec - 1])
    dm.free()


@pytest.mark.parametrize("t", [O.TYPE_Q4_0, O.TYPE_Q5_1, O.TYPE_Q8_0, O.TYPE_FP16], ids=["Q4_0", "Q5_1", "Q8_0", "FP16"])
def test_matvec_at_full_size(raw_lib, t):
    """BASELINE-size matrix (5120 x 5120): a few dense and one-hot vectors against the oracle, and
    linearity W(a + b) == W a + W b for FP16 weights with fp16-exact activations."""
    M = K = 5120
    rng = np.random.default_rng(5)
    mat, raw = make_matrix(rng, t, M, K)
    dm = DevMatrix(raw_lib.lib, t, raw, M, K)
    x = np.zeros((4, K), np.float32)
    x[0, 0] = 1.0
    x[1, K - 1] = -2.0
    x[2] = rng.standard_normal(K).astype(np.float32)
    x[3] = np.float32(0.0)
    y = dm.mul(x)
    assert np.all(y[3] == 0.0)
    for i in range(3):
        ref = mat.matvec(x[i], "ref")
        assert np.abs(y[i] - ref).max() <= 3e-5 * (np.abs(ref).max() + 1e-6) * np.sqrt(K / 32)
    if t == O.TYPE_FP16:
        a = rng.integers(-8, 8, K).astype(np.float32)
        b = rng.integers(-8, 8, K).astype(np.float32)
        ya, yb, yab = dm.mul(a)[0], dm.mul(b)[0], dm.mul(a + b)[0]
        assert np.allclose(yab, ya + yb, rtol=1e-5, atol=1e-4)
    dm.free()


KSPLIT_SHAPES = [(256, 10240), (160, 20480), (96, 21600)]      # few row blocks, long rows: K split 2, 4 and 4 with a ragged last chunk
_KSPLIT_CHILD = r"""
import sys, hashlib, numpy as np
sys.path.insert(0, {root!r}); sys.path.insert(0, {root!r} + "/tests")
from oracle.refload import RwkvLib
from test_gpu_codec import make_matrix, DevMatrix, KSPLIT_SHAPES
from oracle import rwkv_oracle as O
L = RwkvLib({lib!r})
for t in (O.TYPE_Q4_0, O.TYPE_Q4_1, O.TYPE_Q5_0, O.TYPE_Q5_1, O.TYPE_Q8_0):
    for M, K in KSPLIT_SHAPES:
        rng = np.random.default_rng(M + K + t)
        mat, raw = make_matrix(rng, t, M, K)
        dm = DevMatrix(L.lib, t, raw, M, K)
        x = rng.standard_normal((8, K)).astype(np.float32)
        h = hashlib.sha256()
        for n in (1, 3, 8):
            h.update(dm.mul(x[:n]).tobytes())
        print(t, M, K, h.hexdigest())
        dm.free()
"""


@pytest.mark.gpu
def test_matvec_k_split_inside_cta_and_cluster_agree(raw_lib, lib_path):
    """Long rows with few row blocks are split along K (kernels.h matvec_ksplit): inside the CTA (8 warps = 32 / ks rows x ks ranges)
    or, when the q8 block of the outputs is wanted, over a thread-block cluster.  Both add the partial sums in rank order: the
    results must agree with the oracle and with each other bit for bit (the cluster form is forced in a child process)."""
    import hashlib, subprocess, sys
    mine = {}
    for t in (O.TYPE_Q4_0, O.TYPE_Q4_1, O.TYPE_Q5_0, O.TYPE_Q5_1, O.TYPE_Q8_0):
        for M, K in KSPLIT_SHAPES:
            rng = np.random.default_rng(M + K + t)
            mat, raw = make_matrix(rng, t, M, K)
            dm = DevMatrix(raw_lib.lib, t, raw, M, K)
            x = rng.standard_normal((8, K)).astype(np.float32)
            h = hashlib.sha256()
            for n in (1, 3, 8):
                y = dm.mul(x[:n])
                h.update(y.tobytes())
                if n == 3:
                    ref = np.stack([mat.matvec(x[i], "ref") for i in range(n)])
                    assert np.abs(y - ref).max() <= 2e-5 * (np.abs(ref).max() + 1e-6) * np.sqrt(K / 32), (t, M, K)
            mine[(t, M, K)] = h.hexdigest()
            dm.free()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, RWKV_B200_KSPLIT_CLUSTER_ONLY="1")
    r = subprocess.run([sys.executable, "-c", _KSPLIT_CHILD.format(root=root, lib=lib_path)], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    theirs = {}
    for line in r.stdout.split("\n"):
        if line.strip():
            t, M, K, d = line.split()
            theirs[(int(t), int(M), int(K))] = d
    assert theirs == mine
