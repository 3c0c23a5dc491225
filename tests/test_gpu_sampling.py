"""On-device temperature / top-p sampling (csrc/sampling.cu; include/rwkv_b200.h) against token ids the REFERENCE's python/sampling.py
returned for the same seeded logits rows and uniform draws (tests/golden/sampling_golden.npz).  The device keeps sums in double where
numpy uses float32 pairwise / sequential sums, so an id may differ only where the draw (or the top-p cutoff) sits within rounding of a
step of the cumulative distribution: such rows are counted, bounded, and each is checked to be a boundary case."""
import ctypes
import os

import numpy as np
import pytest

from oracle import sampling_oracle as S
from tests import util
from tests.golden.make_golden_sampling_rows import row

pytestmark = pytest.mark.gpu
PF, PU = ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_uint32)


def dev_sample(lib, logits, T, P, u):
    lib.rwkv_b200_sample_logits.argtypes = [PF, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_float, ctypes.c_float, PF, PU]
    lib.rwkv_b200_sample_logits.restype = ctypes.c_bool
    logits = np.ascontiguousarray(logits, np.float32)
    u = np.ascontiguousarray(u, np.float32)
    out = np.empty(logits.shape[0], np.uint32)
    assert lib.rwkv_b200_sample_logits(logits.ctypes.data_as(PF), logits.shape[0], logits.shape[1], T, P, u.ctypes.data_as(PF), out.ctypes.data_as(PU))
    return out


def test_device_sampler_matches_reference_sampler(raw_lib):
    g = np.load(os.path.join(util.GOLDEN, "sampling_golden.npz"))
    n, V = int(g["n_rows"]), int(g["vocab"])
    rows = np.stack([row(i, V) for i in range(n)])
    total_bad = 0
    for c, (T, P) in enumerate(g["cases"]):
        u64 = g[f"u_{c}"]
        u = u64.astype(np.float32)
        got = dev_sample(raw_lib.lib, rows, float(T), float(P), u)
        want = g[f"tokens_{c}"]
        bad = np.nonzero(got != want)[0]
        for i in bad:                                        # float32 view of the draw, or rounding in a sum: the disagreement must be a boundary case
            alt = S.sample_logits(rows[i], float(T), float(P), float(u[i]))
            d = S.boundary_distance(rows[i], float(T), float(P), float(u64[i]))
            assert got[i] == alt or d < 1e-5, (float(T), float(P), int(i), int(got[i]), int(want[i]), d)
        total_bad += len(bad)
    assert total_bad <= 0.003 * 6 * n, total_bad


def test_device_sampler_edge_cases(raw_lib):
    V = 50277                                                # not a multiple of anything
    rng = np.random.default_rng(0)
    x = rng.standard_normal((4, V)).astype(np.float32)
    x[0, 123] = 50.0                                         # one dominant token: a mid-range draw must return it under every setting
    for T, P in ((1.0, 0.8), (0.3, 0.1), (2.0, 1.0), (0.0, 0.5)):
        out = dev_sample(raw_lib.lib, x, T, P, np.float32([0.5, 0.25, 0.0, 0.999999]))
        assert out[0] == 123 and np.all(out < V)
        for i in range(1, 4):
            if T == 0.0:
                assert out[i] == int(np.argmax(x[i]))
            else:
                assert out[i] == S.sample_logits(x[i], T, P, float(np.float32([0.5, 0.25, 0.0, 0.999999])[i])) or \
                    S.boundary_distance(x[i], T, P, float(np.float32([0.5, 0.25, 0.0, 0.999999])[i])) < 1e-5


def test_batch_eval_sample_equals_sampling_the_returned_logits(shared_library):
    from rwkv_cppy_b200 import RWKVModel, synth
    m = RWKVModel(shared_library, synth.ensure_model(shared_library, "v7-mini", "FP16-to-Q8_0"))
    B = 6
    a, b = m.create_batch(B), m.create_batch(B)
    rng = np.random.default_rng(3)
    toks = rng.integers(0, m.n_vocab, size=B).astype(np.uint32)
    for step in range(4):
        u = rng.random(B).astype(np.float32)
        logits = a.eval(toks)                                # logits to the host, sampled with the oracle
        want = np.array([S.sample_logits(logits[s], 0.9, 0.7, float(u[s])) for s in range(B)], np.uint32)
        got = b.eval_sample(toks, 0.9, 0.7, uniforms=u)      # sampled on the device, 4 bytes per stream back
        for s in range(B):
            assert got[s] == want[s] or S.boundary_distance(logits[s], 0.9, 0.7, float(u[s])) < 1e-5
        toks = want
    assert np.array_equal(a.get_state(0), b.get_state(0))
    a.free(); b.free(); m.free()
