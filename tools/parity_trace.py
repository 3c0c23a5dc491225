"""Localises a difference between the engine and the unmodified reference through the recurrent state: after ONE token from
the initial state, layer i's slots hold LN2(x after att_i) [ffn_xx], LN1(x entering layer i) [att_xx] and the new WKV state,
so the first slot whose error jumps names the stage that differs.  python tools/parity_trace.py <shape> <fmt> [env=val ...]"""
import os, sys
for kv in sys.argv[3:]:
    k, v = kv.split("=", 1); os.environ[k] = v
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rwkv_cppy_b200
from rwkv_cppy_b200 import synth
from rwkv_cppy_b200.build import build_library
from oracle.refload import RwkvLib, load_reference

shape, fmt = sys.argv[1], sys.argv[2]
ours = rwkv_cppy_b200.load_rwkv_shared_library()
P, R = RwkvLib(build_library(verbose=False)), load_reference()
path = synth.ensure_model(ours, shape, fmt)
a, b = P.load(path, 1, 0), R.load(path, min(8, os.cpu_count() or 1), 0)
tok = int(np.random.default_rng(100).integers(0, a.n_vocab, 1)[0])
sa, la = a.eval(tok, a.init_state())
sb, lb = b.eval(tok, b.init_state())
L, D = a.n_layer, a.n_embed
per = a.state_len // L
v4 = per == 5 * D
print(f"{shape} {fmt} {' '.join(sys.argv[3:])}: logits rel {np.abs(la - lb).max() / max(1.0, float(np.abs(lb).max())):.3e}")
A, B = sa.reshape(L, per), sb.reshape(L, per)
def rel(x, y):
    fin = np.abs(y) < 1e29
    return float(np.abs(x - y)[fin].max()) / max(1e-6, float(np.abs(y[fin]).max()))
for i in range(L):
    parts = [("att_xx", slice(D, 2 * D)), ("wkv", slice(2 * D, per)), ("ffn_xx", slice(0, D))]
    print(f"  layer {i}: " + "  ".join(f"{n} {rel(A[i, s], B[i, s]):.2e}" for n, s in parts))
