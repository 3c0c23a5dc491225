"""Times the tensor-core GEMM hook on a few shapes (device ms per launch, TFLOP/s)."""
import ctypes, sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rwkv_cppy_b200
from oracle import rwkv_oracle as O

lib = ctypes.CDLL(os.environ["RWKV_B200_LIB"]) if os.environ.get("RWKV_B200_LIB") else rwkv_cppy_b200.load_rwkv_shared_library().library
lib.rwkv_b200_matrix_create.restype = ctypes.c_void_p
lib.rwkv_b200_matrix_create.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_size_t]
shapes = [(2048, 2048, 1024), (7168, 2048, 1024), (2048, 7168, 1024), (8192, 2048, 1024), (2560, 2560, 256), (10240, 2560, 256)]
only = int(sys.argv[1]) if len(sys.argv) > 1 else None
types = {"Q5_1": O.TYPE_Q5_1, "FP16": O.TYPE_FP16}
if os.environ.get("PROBE_TYPES"): types = {k: v for k, v in types.items() if k in os.environ["PROBE_TYPES"].split(",")}
for name, t in types.items():
    for i, (M, K, T) in enumerate(shapes):
        if only is not None and i != only: continue
        rng = np.random.default_rng(1)
        w = (rng.standard_normal((M, K)) / np.sqrt(K)).astype(np.float32)
        raw = w.astype(np.float16).view(np.uint8).reshape(-1) if t == O.TYPE_FP16 else O.quantize_ref(t, w.reshape(-1))
        raw = np.ascontiguousarray(raw)
        m = lib.rwkv_b200_matrix_create(t, raw.ctypes.data, M, K, 1)
        ms = ctypes.c_double(0)
        assert lib.rwkv_b200_matrix_bench_tc(ctypes.c_void_p(m), T, 20, ctypes.byref(ms))
        print(f"{name} M={M} K={K} T={T}: {ms.value*1e3:8.1f} us  {2.0*M*K*T/ms.value/1e9:8.1f} TFLOP/s", flush=True)
        lib.rwkv_b200_matrix_free(ctypes.c_void_p(m))
