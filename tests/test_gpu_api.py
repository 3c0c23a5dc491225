"""GPU tests of the ABI semantics the reference pins with memcmp tests
(tests/test_eval_sequence_in_chunks.c, test_logit_calculation_skipping.c,
test_context_cloning.c) plus argument errors and the additive batch interface."""
import ctypes
import os
import threading

import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu
PROMPT70 = list(b"In a shocking finding, scientist discovered a herd of dragons living i")


def fresh(raw_lib, arch="5v2", fmt="FP32"):
    raw_lib.lib.rwkv_set_print_errors(None, False)
    return raw_lib.load(util.model_path(raw_lib, arch, fmt), 2, 0)


@pytest.mark.parametrize("arch,fmt", [("5v2", "FP32"), ("4v0", "FP32-to-Q5_1"), ("6v0", "FP32-to-Q5_1"), ("7v0", "FP16")])
def test_eval_sequence_in_chunks_is_chunk_size_invariant(raw_lib, arch, fmt):
    """tests/test_eval_sequence_in_chunks.c:10-76 (5v2 FP32 there; more archs/formats here)."""
    assert len(PROMPT70) == 70
    ctx = fresh(raw_lib, arch, fmt)
    s = ctx.init_state()
    for t in PROMPT70:
        s, lg = ctx.eval(t, s, state_out=s)
    s_seq, lg_seq = ctx.eval_sequence(PROMPT70, ctx.init_state())
    for chunk in (1, 2, 8, 10, 63, 64, 70, 1000):
        s2, lg2 = ctx.eval_sequence_in_chunks(PROMPT70, chunk, ctx.init_state())
        # the chunk size is a ggml node-limit workaround: here it does not steer the computation, every size equals rwkv_eval_sequence bit for bit
        assert np.array_equal(s_seq, s2) and np.array_equal(lg_seq, lg2), chunk
    if fmt == "FP32":
        # the reference's contract (memcmp with serial evaluation, tests/test_eval_sequence_in_chunks.c:54) on the file it tests, and on every FP32 file
        assert np.array_equal(s, s_seq) and np.array_equal(lg, lg_seq)
    else:
        # 70 rows of an FP16 / quantised file run on the tensor cores (FP16 operands, like the reference's GPU offload path):
        # equal to serial evaluation within the format's own tolerance
        assert np.abs(lg - lg_seq).max() <= 2e-2 * max(1.0, float(np.abs(lg).max())), float(np.abs(lg - lg_seq).max())
        bound = util.table_tolerance(arch, fmt) * 1.05
        assert abs(float(np.sum(lg_seq - lg, dtype=np.float32))) <= max(bound, 0.05)
    s3, lg3 = ctx.eval_sequence_in_chunks(PROMPT70[:1], 4, None)        # 1-token prompt
    s4, lg4 = ctx.eval(PROMPT70[0], None)
    assert np.array_equal(s3, s4) and np.array_equal(lg3, lg4)
    ctx.free()


@pytest.mark.parametrize("arch", ["4v0", "5v1", "6v0", "7v0"])
def test_logit_calculation_skipping_keeps_state(raw_lib, arch):
    """tests/test_logit_calculation_skipping.c:14-98."""
    fmt = "FP32-to-Q5_1" if arch == "6v0" else "FP32"
    ctx = fresh(raw_lib, arch, fmt)
    a = ctx.init_state(); b = ctx.init_state()
    for t in util.LONG:
        a, _ = ctx.eval(t, a, state_out=a)
        b, none = ctx.eval(t, b, state_out=b, want_logits=False)
        assert none is None
    assert np.array_equal(a, b)
    sa, _ = ctx.eval_sequence(util.LONG, None)
    sb, _ = ctx.eval_sequence(util.LONG, None, want_logits=False)
    assert np.array_equal(sa, sb) and np.array_equal(sa, a)
    # state_out may be NULL too (rwkv.h:107-108)
    assert raw_lib.lib.rwkv_eval(ctx.ptr, 1, None, None, None)
    ctx.free()


def test_context_cloning_survives_parent_free(raw_lib):
    """tests/test_context_cloning.c:10-57."""
    ctx = fresh(raw_lib, "5v1", "FP32")
    s, lg = ctx.eval_sequence(util.PROMPT, None)
    clone = ctx.clone(2)
    ctx.free()
    s2, lg2 = clone.eval_sequence(util.PROMPT, None)
    assert np.array_equal(lg, lg2) and np.array_equal(s, s2)
    # clones evaluate concurrently from different threads (rwkv.h:93-99)
    c2 = clone.clone(1)
    out = {}

    def work(name, c):
        st = None
        for t in util.LONG:
            st, l = c.eval(t, st)
        out[name] = l
    th = [threading.Thread(target=work, args=(n, c)) for n, c in (("a", clone), ("b", c2))]
    [t.start() for t in th]; [t.join() for t in th]
    assert np.array_equal(out["a"], out["b"])
    clone.free(); c2.free()


def test_argument_errors(raw_lib):
    """rwkv_eval.inc:43,89,108,167-168: false + RWKV_ERROR_ARGS in the context's error word."""
    ctx = fresh(raw_lib, "4v0", "FP32")
    L = raw_lib.lib
    L.rwkv_set_print_errors(ctx.ptr, False)
    st = ctx.init_state(); lg = np.zeros(ctx.logits_len, np.float32)
    PF = ctypes.POINTER(ctypes.c_float)
    sp, lp = st.ctypes.data_as(PF), lg.ctypes.data_as(PF)
    assert not L.rwkv_eval(ctx.ptr, 256, sp, sp, lp)
    assert L.rwkv_get_last_error(ctx.ptr) == 0x100 and L.rwkv_get_last_error(ctx.ptr) == 0
    toks = np.array([1, 2, 999], np.uint32)
    tp = toks.ctypes.data_as(ctypes.POINTER(ctypes.c_uint32))
    assert not L.rwkv_eval_sequence(ctx.ptr, tp, 3, sp, sp, lp)
    assert L.rwkv_get_last_error(ctx.ptr) == 0x100
    assert not L.rwkv_eval_sequence(ctx.ptr, tp, 0, sp, sp, lp)
    assert L.rwkv_get_last_error(ctx.ptr) == 0x100
    assert not L.rwkv_eval_sequence_in_chunks(ctx.ptr, tp, 2, 0, sp, sp, lp)
    assert L.rwkv_get_last_error(ctx.ptr) == 0x100
    assert L.rwkv_eval_sequence(ctx.ptr, None, 5, sp, sp, lp)          # tokens NULL = prepare only (rwkv.h:135)
    assert L.rwkv_eval(ctx.ptr, 255, sp, sp, lp) and L.rwkv_get_last_error(ctx.ptr) == 0
    assert ctx.state_len == 128 * 5 * 4 and np.all(ctx.init_state().reshape(4, 5, 128)[:, 4] == np.float32(-1e30))
    ctx.free()


@pytest.mark.parametrize("arch,fmt", [("4v0", "FP16-to-Q4_0"), ("5v2", "FP16"), ("6v0", "FP16-to-Q5_0"), ("7v0", "FP32-to-Q8_0")])
def test_batch_streams_equal_individual_streams(shared_library, raw_lib, arch, fmt):
    """Additive API: B state-resident streams stepped together == B separate rwkv_eval loops, bit for bit."""
    from rwkv_cppy_b200 import RWKVModel
    m = RWKVModel(shared_library, util.model_path(raw_lib, arch, fmt))
    B, steps = 5, 6
    rng = np.random.default_rng(3)
    toks = rng.integers(0, 256, size=(steps, B)).astype(np.uint32)
    batch = m.create_batch(B)
    warm = None
    for t in util.PROMPT:
        _, warm = m.eval(t, warm, use_numpy=True)
    batch.set_state(2, warm)                                     # one stream starts from a non-trivial state
    for s in range(steps):
        logits = batch.eval(toks[s], want_logits=(s % 2 == 0))
        assert (logits is None) == (s % 2 == 1)
    logits = batch.eval(toks[0])
    for b in range(B):
        st = warm.copy() if b == 2 else None
        for s in range(steps):
            l1, st = m.eval(int(toks[s, b]), st, use_numpy=True)
        l1, st = m.eval(int(toks[0, b]), st, use_numpy=True)
        assert np.array_equal(l1, logits[b]), b
        assert np.array_equal(st, batch.get_state(b)), b
    pre = batch.get_state_prefix(64)
    assert np.array_equal(pre[1], batch.get_state(1)[:64])
    assert batch.launches_per_step(True) > 0
    ms = batch.run_timed(toks, with_logits=True)
    assert ms > 0
    batch.free(); m.free()


@pytest.mark.parametrize("B", [40, 96])
@pytest.mark.parametrize("shape,fmt", [("v7-mini", "FP16-to-Q8_0"), ("v6-mini", "FP16-to-Q5_1"), ("v4-mini", "FP16")])
def test_wide_batch_matches_individual_streams(shared_library, shape, fmt, B):
    """B >= 32 streams step together on the tensor-core path (FP16 operands, warp-per-row layer norm, coalesced
    state staging); every stream must agree with its own serial rwkv_eval loop within the FP16 tolerance."""
    from rwkv_cppy_b200 import RWKVModel, synth
    m = RWKVModel(shared_library, synth.ensure_model(shared_library, shape, fmt))
    steps = 5
    rng = np.random.default_rng(5)
    toks = rng.integers(0, 512, size=(steps, B)).astype(np.uint32)
    batch = m.create_batch(B)
    logits = None
    for s in range(steps):
        logits = batch.eval(toks[s])
    for b in (0, 1, B // 2, B - 1):
        st, l1 = None, None
        for s in range(steps):
            l1, st = m.eval(int(toks[s, b]), st, use_numpy=True)
        bs = batch.get_state(b)
        fin = np.abs(st) < 1e29
        assert np.abs(l1 - logits[b]).max() <= 3e-2 * max(1.0, float(np.abs(l1).max())), b
        assert np.abs(st[fin] - bs[fin]).max() <= 3e-2 * max(1.0, float(np.abs(st[fin]).max())), b
    batch.free(); m.free()
