"""Diagnostic: batch vs individual streams on a tiny model, each path repeated, to tell WHICH
path is nondeterministic.  usage: batch_flake.py [arch] [fmt] [reps]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import util
from oracle.refload import RwkvLib
import rwkv_cppy_b200
from rwkv_cppy_b200 import RWKVModel
from rwkv_cppy_b200 import build as _b

arch = sys.argv[1] if len(sys.argv) > 1 else "4v0"
fmt = sys.argv[2] if len(sys.argv) > 2 else "FP16-to-Q4_0"
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 6
lib = rwkv_cppy_b200.load_rwkv_shared_library()
raw = RwkvLib(_b.build_library(verbose=False))
path = util.model_path(raw, arch, fmt)
B, steps = 5, 6
toks = np.random.default_rng(3).integers(0, 256, size=(steps, B)).astype(np.uint32)


def run_batch(m):
    batch = m.create_batch(B)
    for s in range(steps):
        batch.eval(toks[s], want_logits=(s % 2 == 0))
    logits = batch.eval(toks[0]).copy()
    states = np.stack([batch.get_state(b) for b in range(B)])
    batch.free()
    return logits, states


def run_serial(m):
    L, S = [], []
    for b in range(B):
        st = None
        for s in range(steps):
            l1, st = m.eval(int(toks[s, b]), st, use_numpy=True)
        l1, st = m.eval(int(toks[0, b]), st, use_numpy=True)
        L.append(l1.copy()); S.append(st.copy())
    return np.stack(L), np.stack(S)


def describe(tag, a, b):
    (la, sa), (lb, sb) = a, b
    for i in range(B):
        if np.array_equal(la[i], lb[i]) and np.array_equal(sa[i], sb[i]):
            continue
        D = sa.shape[1] // (5 * (sa.shape[1] // (5 * 128))) if arch == "4v0" else 0
        bad = np.nonzero(sa[i] != sb[i])[0]
        slots = sorted(set((bad // 128).tolist())) if arch == "4v0" else []
        nl = int((la[i] != lb[i]).sum())
        print(f"  {tag} stream {i}: dlogit {np.abs(la[i]-lb[i]).max():.3e} ({nl} differ) dstate "
              f"{np.abs(sa[i]-sb[i]).max():.3e} ({bad.size} differ) slots {slots}", flush=True)


ref_b = ref_s = None
for rep in range(reps):
    m = RWKVModel(lib, path)
    rb_ = run_batch(m)
    rs_ = run_serial(m)
    m.free()
    if ref_b is None:
        ref_b, ref_s = rb_, rs_
        print("rep 0: batch vs serial"); describe("b-vs-s", rb_, rs_)
        continue
    print(f"rep {rep}")
    describe("batch-vs-batch0", rb_, ref_b)
    describe("serial-vs-serial0", rs_, ref_s)
    describe("b-vs-s", rb_, rs_)
print("done")
