"""GPU parity of the FAST paths against the UNMODIFIED reference (tests/golden/fast_outputs.npz, written by
tests/golden/make_golden_fast.py from oracle/_ref): the tensor-core sequence mat-mul (rows >= 64), the
chunk-parallel WKV5/6, every head-size-64 kernel, the wide-batch path and the BASELINE.json model shapes.

Tolerances are stated per weight format, relative to max|reference logit| (or max|reference state|):

  FP32 files                   exact kernels everywhere; only FP32 summation order differs          1e-4
  FP16 files, tensor path      same operands as the reference (FP16 weights, FP16-rounded
                               activations, FP32 accumulate), different summation order and, for
                               v5/v6, the chunked form of the recurrence                             5e-3
  quantised, tensor path       weights dequantised to FP16 (one rounding, <= 2^-11 relative) and
                               FP16 activations where the reference uses q8 activation blocks
                               (<= 1/254 relative per element): the difference is the reference's
                               own activation-quantisation noise, which grows with the coarseness
                               of nothing but the activations -- one bound for all five formats      3e-2
  quantised, exact path        the reference's arithmetic (q8 blocks, integer dots); a 1-ulp
  (decode, batches < 32)       upstream difference can flip one quant                                2e-2

The BASELINE shapes are checked at full depth and as two-layer cuts (v6-1b6-L2, v7-2b9-L2).  For scale: the UNMODIFIED
reference itself moves by 1.8e-3 (v6 1.6B) / 3.6e-3 (v7 2.9B) / 1.0e-3 (v4 14B) of max|logit| when its own input state is
perturbed by +-1 ulp ("sensitivity/*" in fast_outputs.npz, tools/ref_sensitivity.py): every q8 rounding is a step function.
These tests found a real bug in round 1's code (a layer-norm kernel reading the residual row through a const __restrict__
pointer: nvcc hoisted the loads above griddepcontrol.wait; csrc/launch.cuh RULE) that no small-shape test could see.

Observed values are appended to gpurun_out/fast_parity.txt (committed under profiles/ per round).
"""
import os

import numpy as np
import pytest

from tests import util
from tests.golden.make_golden_fast import MINI, MINI_T, PROMPT70, big_tokens, file_tag, mini_tokens, state_sample

pytestmark = pytest.mark.gpu
REPORT = os.path.join(util.ROOT, "gpurun_out", "fast_parity.txt")


@pytest.fixture(scope="module")
def fast():
    return np.load(os.path.join(util.GOLDEN, "fast_outputs.npz"))


def tol_for(fmt: str, tensor_path: bool) -> float:
    if fmt == "FP32":
        return 1e-4
    if fmt == "FP16":
        return 5e-3 if tensor_path else 2e-2
    return 3e-2 if tensor_path else 2e-2


def rel_err(a: np.ndarray, ref: np.ndarray) -> float:
    fin = np.abs(ref) < 1e29                                  # v4 pp slots start at -1e30
    scale = max(1.0, float(np.abs(ref[fin]).max()))
    return float(np.abs(a[fin] - ref[fin]).max()) / scale


def report(line: str):
    os.makedirs(os.path.dirname(REPORT), exist_ok=True)
    with open(REPORT, "a") as f:
        f.write(line + "\n")


def check(name, got, ref, tol):
    assert np.all(np.isfinite(got[np.abs(ref) < 1e29])), name
    e = rel_err(got, ref)
    report(f"{name:70s} rel_err {e:.3e}  bound {tol:.1e}")
    assert e <= tol, (name, e, tol)


def same_file(fast, key, path):
    """The synthetic file on this box must be the file the reference outputs were made from."""
    want = bytes(fast[f"files/{key}"]).decode()
    got = bytes(file_tag(path)).decode()
    assert got == want, f"synthetic model {key} differs from the one the golden outputs were generated on ({got} vs {want})"


# ---- the reference's tiny fixtures at the 70-token prompt of tests/test_eval_sequence_in_chunks.c:69 ----------
@pytest.mark.parametrize("arch,fmt", util.all_cases(), ids=[f"{a}-{f}" for a, f in util.all_cases()])
def test_tiny_models_70_tokens(raw_lib, fast, arch, fmt):
    raw_lib.lib.rwkv_set_print_errors(None, False)
    ctx = raw_lib.load(util.model_path(raw_lib, arch, fmt), 2, 0)
    key = f"tiny70/{util.ARCHS[arch]}-{fmt}"
    ref_l, ref_s = fast[key + "/logits"], fast[key + "/state"]
    # one call of 70 rows: FP16 / quantised files take the tensor-core mat-mul (and 7v0, head size 64, its own kernels)
    s, lg = ctx.eval_sequence(PROMPT70, ctx.init_state())
    check(f"{key} sequence logits", lg, ref_l, tol_for(fmt.split("-to-")[-1] if "-to-" in fmt else fmt, True))
    check(f"{key} sequence state", s, ref_s, tol_for(fmt.split("-to-")[-1] if "-to-" in fmt else fmt, True))
    # token by token on the exact kernels
    s2 = ctx.init_state()
    for t in PROMPT70:
        s2, lg2 = ctx.eval(t, s2, state_out=s2)
    check(f"{key} serial logits", lg2, ref_l, tol_for("FP32" if fmt == "FP32" else "Q", False))
    check(f"{key} serial state", s2, ref_s, tol_for("FP32" if fmt == "FP32" else "Q", False))
    ctx.free()


# ---- synthetic head-size-64 models: tensor-core mat-mul + chunk-parallel WKV (v5 / v6), v7 and v4 sequence kernels ----
@pytest.mark.parametrize("shape,fmt", MINI, ids=[f"{s}-{f}" for s, f in MINI])
def test_mini_models_long_sequences(shared_library, raw_lib, fast, shape, fmt):
    from rwkv_cppy_b200 import synth
    path = synth.ensure_model(shared_library, shape, fmt)
    same_file(fast, f"{shape}-{fmt}", path)
    ctx = raw_lib.load(path, 1, 0)
    q = fmt.split("-to-")[-1]
    for T in MINI_T:
        key = f"mini/{shape}-{fmt}/T{T}"
        s, lg = ctx.eval_sequence(mini_tokens(T), ctx.init_state())
        check(f"{key} sequence logits", lg, fast[key + "/logits"], tol_for(q, True))
        check(f"{key} sequence state", s, fast[key + "/state"], tol_for(q, True))
    # the same 150 tokens one at a time: wkv5_kernel<64> / wkv7_kernel<64> / the decode prologues against the reference
    key = f"mini/{shape}-{fmt}/T150"
    s = ctx.init_state()
    for t in mini_tokens(150):
        s, lg = ctx.eval(t, s, state_out=s)
    check(f"{key} serial logits", lg, fast[key + "/logits"], tol_for(q, False))
    check(f"{key} serial state", s, fast[key + "/state"], tol_for(q, False))
    # 40 copies of the stream through the batch API (>= 32 rows: tensor-core batch path), 64 steps
    from rwkv_cppy_b200 import RWKVModel
    m = RWKVModel(shared_library, path)
    batch = m.create_batch(40)
    toks = mini_tokens(64)
    for t in toks:
        logits = batch.eval([t] * 40)
    key = f"mini/{shape}-{fmt}/T64"
    for b in (0, 17, 39):
        check(f"{key} wide batch stream {b} logits", logits[b], fast[key + "/logits"], tol_for(q, True))
        check(f"{key} wide batch stream {b} state", batch.get_state(b), fast[key + "/state"], tol_for(q, True))
    batch.free(); m.free()
    ctx.free()


# ---- BASELINE.json shapes: model-level parity against the reference at the sizes the bench runs -------------------
def test_baseline_169m_q4_0_decode(shared_library, raw_lib, fast):
    """configs[1]: 32 decode tokens of the v4 169M-shape Q4_0 file (fused decode prologues, FP16 head)."""
    from rwkv_cppy_b200 import synth
    path = synth.ensure_model(shared_library, "v4-169m", "FP16-to-Q4_0")
    same_file(fast, "v4-169m-FP16-to-Q4_0", path)
    ctx = raw_lib.load(path, 1, 0)
    s = ctx.init_state()
    for i, t in enumerate(big_tokens(11, 32, ctx.n_vocab)):
        s, lg = ctx.eval(t, s, state_out=s)
        if i == 0:
            check("big/v4-169m Q4_0 decode logits t=1", lg, fast["big/v4-169m-FP16-to-Q4_0/logits_t1"], tol_for("Q4_0", False))
    check("big/v4-169m Q4_0 decode logits t=32", lg, fast["big/v4-169m-FP16-to-Q4_0/logits_t32"], tol_for("Q4_0", False))
    check("big/v4-169m Q4_0 decode state t=32", s, fast["big/v4-169m-FP16-to-Q4_0/state_t32"], tol_for("Q4_0", False))
    ctx.free()


def test_baseline_v6_1b6_q5_1_sequence(shared_library, raw_lib, fast):
    """configs[2]: the v6 1.6B-shape Q5_1 file: sequences of 64 and 192 tokens on the tensor-core mat-mul +
    chunk-parallel WKV6 (32 heads of 64), and 4 decode tokens on the exact kernels."""
    from rwkv_cppy_b200 import synth
    path = synth.ensure_model(shared_library, "v6-1b6", "Q5_1")
    same_file(fast, "v6-1b6-Q5_1", path)
    ctx = raw_lib.load(path, 1, 0)
    toks = big_tokens(12, 192, ctx.n_vocab)
    for T in (64, 192):
        s, lg = ctx.eval_sequence(toks[:T], ctx.init_state())
        check(f"big/v6-1b6 Q5_1 sequence T={T} logits", lg, fast[f"big/v6-1b6-Q5_1/logits_T{T}"], tol_for("Q5_1", True))
        check(f"big/v6-1b6 Q5_1 sequence T={T} state", state_sample(s), fast[f"big/v6-1b6-Q5_1/state_T{T}"], tol_for("Q5_1", True))
    s = ctx.init_state()
    for t in toks[:4]:
        s, lg = ctx.eval(t, s, state_out=s)
    check("big/v6-1b6 Q5_1 decode logits t=4", lg, fast["big/v6-1b6-Q5_1/logits_t4"], tol_for("Q5_1", False))
    ctx.free()


def test_baseline_v7_2b9_q8_0_streams(shared_library, raw_lib, fast):
    """configs[3]: the v7 2.9B-shape Q8_0 file: 2 streams x 4 tokens (exact batch path), 32 streams x 2 tokens
    (tensor-core batch path) through the batch API, and a 64-token sequence."""
    from rwkv_cppy_b200 import RWKVModel, synth
    path = synth.ensure_model(shared_library, "v7-2b9", "Q8_0")
    same_file(fast, "v7-2b9-Q8_0", path)
    m = RWKVModel(shared_library, path)
    V = m.n_vocab
    batch = m.create_batch(2)
    toks = np.stack([big_tokens(100 + st, 4, V) for st in range(2)], axis=1)       # [step][stream]
    for step in range(4):
        logits = batch.eval([int(t) for t in toks[step]])
    for st in range(2):
        check(f"big/v7-2b9 Q8_0 batch of 2, stream {st} logits t=4", logits[st], fast[f"big/v7-2b9-Q8_0/stream{st}_logits_t4"], tol_for("Q8_0", False))
        check(f"big/v7-2b9 Q8_0 batch of 2, stream {st} state t=4", state_sample(batch.get_state(st)), fast[f"big/v7-2b9-Q8_0/stream{st}_state_t4"], tol_for("Q8_0", False))
    batch.free()
    batch = m.create_batch(32)
    toks = np.stack([big_tokens(200 + st, 2, V) for st in range(32)], axis=1)
    for step in range(2):
        logits = batch.eval([int(t) for t in toks[step]])
    check("big/v7-2b9 Q8_0 batch of 32 (tensor path) logits t=2", logits[:, ::16], fast["big/v7-2b9-Q8_0/wide32_logits_t2_every16"], tol_for("Q8_0", True))
    batch.free(); m.free()
    ctx = raw_lib.load(path, 1, 0)
    s, lg = ctx.eval_sequence(big_tokens(13, 64, V), ctx.init_state())
    check("big/v7-2b9 Q8_0 sequence T=64 logits", lg, fast["big/v7-2b9-Q8_0/logits_T64"], tol_for("Q8_0", True))
    check("big/v7-2b9 Q8_0 sequence T=64 state", state_sample(s), fast["big/v7-2b9-Q8_0/state_T64"], tol_for("Q8_0", True))
    ctx.free()


@pytest.mark.parametrize("shape,fmt", [("v6-1b6-L2", "Q5_1"), ("v7-2b9-L2", "Q8_0")])
def test_baseline_shapes_two_layers(shared_library, raw_lib, fast, shape, fmt):
    """The first two layers of the v6 1.6B / v7 2.9B shapes (D = 2048 / 2560, 32 / 40 heads of 64, V = 65536): the
    big-shape kernel variants against the reference without 24-32 layers of rounding-flip amplification.  Decode
    token by token (exact kernels: format bound, observed ~1e-6 unless a quant flips) and a 64-token sequence
    (tensor-core mat-mul + chunk-parallel WKV6 / WKV7 scan)."""
    from rwkv_cppy_b200 import synth
    path = synth.ensure_model(shared_library, shape, fmt)
    same_file(fast, f"{shape}-{fmt}", path)
    ctx = raw_lib.load(path, 1, 0)
    toks = big_tokens(15, 64, ctx.n_vocab)
    s = ctx.init_state()
    for i, t in enumerate(toks[:6]):
        s, lg = ctx.eval(t, s, state_out=s)
        check(f"big/{shape} {fmt} decode logits t={i + 1}", lg, fast[f"big/{shape}-{fmt}/logits_t{i + 1}"], tol_for(fmt, False))
    check(f"big/{shape} {fmt} decode state t=6", state_sample(s), fast[f"big/{shape}-{fmt}/state_t6"], tol_for(fmt, False))
    s, lg = ctx.eval_sequence(toks, ctx.init_state())
    check(f"big/{shape} {fmt} sequence T=64 logits", lg, fast[f"big/{shape}-{fmt}/logits_T64"], tol_for(fmt, True))
    check(f"big/{shape} {fmt} sequence T=64 state", state_sample(s), fast[f"big/{shape}-{fmt}/state_T64"], tol_for(fmt, True))
    ctx.free()


def test_baseline_v4_14b_q4_1_decode(shared_library, raw_lib, fast):
    """configs[4] on one GPU: 4 decode tokens of the v4 14B-shape Q4_1 file (D = 5120: wide-row prologues, cluster K split)."""
    if "big/v4-14b-Q4_1/logits_t4" not in fast:
        pytest.skip("no 14B reference outputs in fast_outputs.npz")
    from rwkv_cppy_b200 import synth
    path = synth.ensure_model(shared_library, "v4-14b", "Q4_1")
    same_file(fast, "v4-14b-Q4_1", path)
    ctx = raw_lib.load(path, 1, 0)
    s = ctx.init_state()
    for t in big_tokens(14, 4, ctx.n_vocab):
        s, lg = ctx.eval(t, s, state_out=s)
    check("big/v4-14b Q4_1 decode logits t=4", lg, fast["big/v4-14b-Q4_1/logits_t4"], tol_for("Q4_1", False))
    check("big/v4-14b Q4_1 decode state t=4", s, fast["big/v4-14b-Q4_1/state_t4"], tol_for("Q4_1", False))
    ctx.free()
