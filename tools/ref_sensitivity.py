"""How far the UNMODIFIED reference moves when its input state is perturbed by one unit in the last place:
the yardstick for "same arithmetic, different FP32 summation order" on deep quantised models (every
q8 activation rounding is a step function; see tests/test_gpu_fast_paths.py).  Runs on CPU (build container)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle.refload import load_reference
from rwkv_cppy_b200 import synth

R = load_reference(); R.lib.rwkv_set_print_errors(None, False)
cases = [("v4-169m", "FP16-to-Q4_0"), ("v6-1b6", "Q5_1"), ("v7-2b9", "Q8_0"), ("v4-14b", "Q4_1")]
if len(sys.argv) > 1:
    cases = [c for c in cases if c[0] in sys.argv[1:]]
for shape, fmt in cases:
    c = R.load(synth.ensure_model(None, shape, fmt), min(8, os.cpu_count() or 1), 0)
    toks = [int(t) for t in np.random.default_rng(100).integers(0, c.n_vocab, 5)]
    s = c.init_state()
    for t in toks[:3]:
        s, lg = c.eval(t, s)
    s1, l1 = c.eval(toks[3], s.copy())
    for trial in range(3):
        rng = np.random.default_rng(trial)
        sp = s.copy()
        bits = sp.view(np.int32)
        bits += rng.integers(-1, 2, size=bits.shape, dtype=np.int32) * (np.abs(s) < 1e29) * (s != 0)      # +-1 ulp (or 0) on every finite non-zero entry
        s2, l2 = c.eval(toks[3], sp)
        scale = max(1.0, float(np.abs(l1).max()))
        fin = np.abs(s1) < 1e29
        print(f"{shape} {fmt}: state +-1 ulp -> logits move {np.abs(l2 - l1).max() / scale:.3e} of max|logit| (rms {np.sqrt(np.mean((l2 - l1) ** 2)):.3e}), "
              f"state moves {np.abs(s2 - s1)[fin].max() / max(1.0, np.abs(s1[fin]).max()):.3e}", flush=True)
    c.free()
