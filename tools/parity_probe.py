"""Per-token logit difference between the engine and the unmodified reference (oracle/_ref) on the BASELINE shapes:
tells summation-order noise amplified by q8 rounding flips (tiny at t = 1, growing with depth and tokens) from an
arithmetic difference (large at t = 1).  Runs on the GPU box: python tools/parity_probe.py [shape fmt n_tokens]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rwkv_cppy_b200
from rwkv_cppy_b200 import synth
from oracle.refload import RwkvLib, load_reference


def main():
    cases = [("v6-1b6", "Q5_1", 6), ("v7-2b9", "Q8_0", 6), ("v4-169m", "FP16-to-Q4_0", 6)]
    if len(sys.argv) >= 4:
        cases = [(sys.argv[1], sys.argv[2], int(sys.argv[3]))]
    ours = rwkv_cppy_b200.load_rwkv_shared_library()
    from rwkv_cppy_b200.build import build_library
    P = RwkvLib(build_library(verbose=False))
    R = load_reference()
    for shape, fmt, n in cases:
        path = synth.ensure_model(ours, shape, fmt)
        a, b = P.load(path, 1, 0), R.load(path, min(8, os.cpu_count() or 1), 0)
        toks = [int(t) for t in np.random.default_rng(100).integers(0, a.n_vocab, n)]
        sa, sb = a.init_state(), b.init_state()
        for i, t in enumerate(toks):
            sa, la = a.eval(t, sa)
            sb, lb = b.eval(t, sb)
            scale = max(1.0, float(np.abs(lb).max()))
            ds = np.abs(sa - sb); fin = np.abs(sb) < 1e29
            print(f"{shape} {fmt} t={i + 1}: logits rel {np.abs(la - lb).max() / scale:.3e} (max|logit| {np.abs(lb).max():.3f}), "
                  f"state rel {ds[fin].max() / max(1.0, np.abs(sb[fin]).max()):.3e}, rms logit diff {np.sqrt(np.mean((la - lb) ** 2)):.3e}", flush=True)
        # same tokens, but the engine restarts every token from the REFERENCE's state: the error of one token alone
        sb = b.init_state()
        for i, t in enumerate(toks):
            _, la = a.eval(t, sb.copy())
            sb, lb = b.eval(t, sb)
            print(f"{shape} {fmt} t={i + 1} from the reference state: logits rel {np.abs(la - lb).max() / max(1.0, float(np.abs(lb).max())):.3e}", flush=True)
        a.free(); b.free()


if __name__ == "__main__":
    main()
