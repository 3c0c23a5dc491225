"""Repeats one tensor-core GEMM shape and reports where (rows / columns) results disagree with the FP16 oracle."""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import rwkv_oracle as O
from oracle.refload import RwkvLib
from tests.test_gpu_codec import DevMatrix, make_matrix, PF
from tests.test_gpu_gemm import mul_tc
from rwkv_cppy_b200.build import build_library

L = RwkvLib(os.environ.get("RWKV_B200_LIB") or build_library(verbose=False))
M, K, T = (int(v) for v in sys.argv[1:4])
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 10
for t in (O.TYPE_Q4_0, O.TYPE_FP16):
    rng = np.random.default_rng(7)
    _, raw = make_matrix(rng, t, M, K)
    dm = DevMatrix(L.lib, t, raw, M, K)
    x = rng.standard_normal((T, K)).astype(np.float32)
    ref = O.matmul_fp16(t, raw, M, K, x)
    scale = float(np.abs(ref).max())
    for r in range(reps):
        y = mul_tc(dm, x)
        bad = np.abs(y - ref) > 1e-3 * scale
        if bad.any():
            tt, mm = np.nonzero(bad)
            print(f"type {O.TYPE_NAMES[t]} rep {r}: {bad.sum()} bad; tokens {tt.min()}..{tt.max()} ({len(set(tt))} distinct), rows {mm.min()}..{mm.max()} ({len(set(mm))} distinct); "
                  f"token%16 {sorted(set(tt % 16))[:16]}; row//32 {sorted(set(mm // 32))[:40]}; max err {float(np.abs(y - ref).max()):.4f} scale {scale:.3f}", flush=True)
        else:
            print(f"type {O.TYPE_NAMES[t]} rep {r}: ok", flush=True)
    dm.free()
