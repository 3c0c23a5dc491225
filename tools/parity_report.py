"""Parity report (GPU box): every tiny fixture through the product librwkv.so, compared with
the logits / state the unmodified reference produced for the same file
(tests/golden/ref_outputs.npz) and with the reference's validator rule.  Prints a table;
used to set the tolerances written in tests/test_gpu_models.py."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import util
from oracle.refload import RwkvLib
from rwkv_cppy_b200 import build


def main():
    lib = RwkvLib(build.build_library(verbose=False))
    lib.lib.rwkv_set_print_errors(None, False)
    ref = np.load(os.path.join(util.GOLDEN, "ref_outputs.npz"))
    print(f"{'fixture':28s} {'max|dlogit|':>12s} {'rel':>10s} {'max|dstate|':>12s} {'rel':>10s} {'max|dlogit17|':>13s} {'sum(l-exp)':>11s} {'bound':>9s}")
    worst = {}
    for arch, fmt in util.all_cases():
        ctx = lib.load(util.model_path(lib, arch, fmt), 2, 0)
        key = f"{util.ARCHS[arch]}-{fmt}"
        rl, rs, rl17 = ref[key + "/logits3"], ref[key + "/state3"], ref[key + "/logits17"]
        s = ctx.init_state(); lg = None
        for t in util.PROMPT:
            s, lg = ctx.eval(t, s, state_out=s)
        _, lg17 = ctx.eval_sequence(util.LONG, None)
        ls = max(1.0, float(np.abs(rl).max())); ss = max(1.0, float(np.abs(rs[np.abs(rs) < 1e29]).max())); l17s = max(1.0, float(np.abs(rl17).max()))
        dl = float(np.abs(lg - rl).max()); ds = float(np.abs(s - rs).max()); dl17 = float(np.abs(lg17 - rl17).max())
        dsum = float(np.sum(lg - util.expected_logits(arch), dtype=np.float32))
        print(f"{key:28s} {dl:12.3e} {dl/ls:10.2e} {ds:12.3e} {ds/ss:10.2e} {dl17:13.3e} {dsum:11.6f} {util.table_tolerance(arch, fmt)*1.05:9.6f}")
        cls = "FP32" if fmt == "FP32" else ("FP16" if fmt == "FP16" else "Q")
        w = worst.setdefault(cls, [0.0, 0.0, 0.0])
        w[0] = max(w[0], dl / ls); w[1] = max(w[1], ds / ss); w[2] = max(w[2], dl17 / l17s)
        ctx.free()
    print("worst relative (logits3, state3, logits17) per class:", worst)


if __name__ == "__main__":
    main()
