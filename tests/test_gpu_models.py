"""GPU parity of the whole hot path on the reference's tiny models, through the C ABI:
the reference's own acceptance rule (tests/logit_difference_validator.inc:60-83 with the
tables of tests/test_tiny_rwkv.c:38-134) AND a direct comparison with logits / state the
unmodified reference produced for the same file (tests/golden/ref_outputs.npz)."""
import os

import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ref_out():
    return np.load(os.path.join(util.GOLDEN, "ref_outputs.npz"))


def run_serial(ctx, tokens):
    s = ctx.init_state()
    lg = None
    for t in tokens:
        s, lg = ctx.eval(t, s, state_out=s)          # state_in aliases state_out, as every reference test does
    return s, lg


@pytest.mark.parametrize("arch,fmt", util.all_cases(), ids=[f"{a}-{f}" for a, f in util.all_cases()])
def test_tiny_model(raw_lib, ref_out, arch, fmt):
    raw_lib.lib.rwkv_set_print_errors(None, False)
    ctx = raw_lib.load(util.model_path(raw_lib, arch, fmt), 2, 0)
    assert raw_lib.lib.rwkv_get_last_error(None) == 0
    assert ctx.logits_len == 256
    expected = util.expected_logits(arch)
    bound = util.table_tolerance(arch, fmt) * 1.05
    key = f"{util.ARCHS[arch]}-{fmt}"
    ref_l, ref_s, ref_l17 = ref_out[key + "/logits3"], ref_out[key + "/state3"], ref_out[key + "/logits17"]

    s, lg = run_serial(ctx, util.PROMPT)
    d_serial = float(np.sum(lg - expected, dtype=np.float32))
    s2, lg2 = ctx.eval_sequence(util.PROMPT, ctx.init_state())
    d_seq = float(np.sum(lg2 - expected, dtype=np.float32))

    # (1) against the unmodified reference on the same file.  FP32: summation order only.
    # FP16 / quantised: the q8 / fp16 activation rounding is a step function, a 1-ulp upstream
    # difference can flip one quant, which moves a logit by about one quantisation step.
    tol = 2e-5 if fmt == "FP32" else 2e-2
    lscale = max(1.0, float(np.abs(ref_l).max()))
    sscale = max(1.0, float(np.abs(ref_s[np.abs(ref_s) < 1e29]).max()))
    assert np.abs(lg - ref_l).max() <= tol * lscale, ("serial logits", float(np.abs(lg - ref_l).max()))
    assert np.abs(s - ref_s).max() <= tol * sscale, ("serial state", float(np.abs(s - ref_s).max()))
    # (2) the reference's own acceptance rule, serial and sequence mode
    assert abs(d_serial) <= bound, f"serial difference sum {d_serial} vs table {bound}"
    assert abs(d_seq) <= bound, f"sequence difference sum {d_seq} vs table {bound}"
    # (3) sequence mode is the same arithmetic per token: bit-identical to serial inside the engine
    assert np.array_equal(lg, lg2) and np.array_equal(s, s2)
    # (4) a longer prompt
    s17, lg17 = ctx.eval_sequence(util.LONG, None)
    assert np.abs(lg17 - ref_l17).max() <= (1e-4 if fmt == "FP32" else 5e-2) * max(1.0, float(np.abs(ref_l17).max()))
    ctx.free()


def test_python_wrapper_end_to_end(shared_library):
    """The mirror of the reference's RWKVModel drives the same path (numpy and torch buffers)."""
    import torch
    from rwkv_cppy_b200 import RWKVModel
    m = RWKVModel(shared_library, os.path.join(util.MODELS, "tiny-rwkv-5v2-730K-FP32.bin"))
    assert (m.n_vocab, m.n_embed, m.n_layer) == (256, 64, 12)
    logits, state = None, None
    for t in util.PROMPT:
        logits, state = m.eval(t, state, use_numpy=True)
    exp = util.expected_logits("5v2")
    assert abs(float((logits - exp).sum())) <= 0.00105
    lt, st = m.eval_sequence(util.PROMPT, None)
    assert isinstance(lt, torch.Tensor) and np.array_equal(lt.numpy(), logits) and np.array_equal(st.numpy(), state)
    lc, sc = m.eval_sequence_in_chunks(util.PROMPT, None, chunk_size=2, use_numpy=True)
    assert np.array_equal(lc, logits) and np.array_equal(sc, state)
    with pytest.raises(ValueError):
        m.eval(0, np.zeros(3, np.float32))
    with pytest.raises(ValueError):
        m.eval(256, None)
    m.free()
    with pytest.raises(ValueError):
        m.eval(0, None)


# ---- long sequences: the tolerance-mode path (tensor-core GEMM with FP16 operands + chunk-parallel
# WKV for v5 / v6 heads of 64) against token-by-token decode on the exact kernels -------------------
LONG_CASES = [("v6-mini", "FP16"), ("v6-mini", "FP16-to-Q5_1"), ("v5-mini", "FP16"), ("v5-mini", "FP16-to-Q4_0"),
              ("v7-mini", "FP16-to-Q8_0"), ("v4-mini", "FP16-to-Q4_1")]


@pytest.mark.parametrize("shape,fmt", LONG_CASES, ids=[f"{s}-{f}" for s, f in LONG_CASES])
@pytest.mark.parametrize("T", [64, 150, 523])
def test_long_sequence_matches_serial_decode(shared_library, raw_lib, shape, fmt, T):
    from rwkv_cppy_b200 import synth
    path = synth.ensure_model(shared_library, shape, fmt)
    ctx = raw_lib.load(path, 1, 0)
    rng = np.random.default_rng(T)
    tokens = [int(t) for t in rng.integers(0, 512, T)]
    s_ser, lg_ser = run_serial(ctx, tokens)
    s_seq, lg_seq = ctx.eval_sequence(tokens, ctx.init_state())
    assert np.all(np.isfinite(lg_seq)) and np.all(np.isfinite(s_seq))
    # FP16 operand rounding (2^-11 relative per operand) over K <= 1024 products and a few layers
    lscale = max(1.0, float(np.abs(lg_ser).max()))
    finite = np.abs(s_ser) < 1e29                         # v4 pp slots start at -1e30
    sscale = max(1.0, float(np.abs(s_ser[finite]).max()))
    assert np.abs(lg_seq - lg_ser).max() <= 3e-2 * lscale, float(np.abs(lg_seq - lg_ser).max() / lscale)
    assert np.abs(s_seq[finite] - s_ser[finite]).max() <= 3e-2 * sscale, float(np.abs(s_seq[finite] - s_ser[finite]).max() / sscale)
    # chunks of the sequence (state carried through the host) agree with the single call the same way
    s_c, lg_c = ctx.eval_sequence(tokens[: T // 2], ctx.init_state())
    s_c, lg_c = ctx.eval_sequence(tokens[T // 2:], s_c)
    assert np.abs(lg_c - lg_ser).max() <= 3e-2 * lscale
    ctx.free()
