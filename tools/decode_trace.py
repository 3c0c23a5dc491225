"""Timeline of the decode mat-vec launches (CTA 0 of each), from a -DRWKV_B200_TRACE build:
    RWKV_B200_NVCC_FLAGS=-DRWKV_B200_TRACE python -m rwkv_cppy_b200.build   (rebuild without it afterwards)
Stamps (ns, %globaltimer): 0 kernel entry, 1 before griddepcontrol.wait, 2 after it, 3 after the
prologue, 4 after the chunk loop, 5 kernel end."""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rwkv_cppy_b200
from rwkv_cppy_b200 import synth

lib = rwkv_cppy_b200.load_rwkv_shared_library()
path = synth.ensure_model(lib, "v4-169m", "FP16-to-Q4_0")
model = rwkv_cppy_b200.RWKVModel(lib, path)
batch = model.create_batch(1)
toks = np.random.default_rng(0).integers(0, model.n_vocab, size=(40, 1)).astype(np.uint32)
batch.run_timed(toks, True)
fn = lib.library.rwkv_b200_debug_trace
fn.restype = ctypes.c_int
buf = (ctypes.c_ulonglong * (8192 * 12))()
fn(buf, 8192)                                   # drop the warm-up records
ms = batch.run_timed(toks[:8], True)
n = fn(buf, 8192)
rec = np.ctypeslib.as_array(buf).reshape(-1, 12)[:n].astype(np.int64)
rec = rec[np.argsort(rec[:, 0])]
print(f"{n} records, {ms / 8 * 1e3:.1f} us/step")
per = 48
step = rec[per * 3: per * 4]
t0 = step[0, 0]
print(" pro grid | entry  wait_in wait_out  pro_done loop_done   end | prev_end->wait_out  since prev end")
prev_end = None
for r in step[:16]:
    e = r[:6] - t0
    gap = (r[2] - prev_end) if prev_end is not None else 0
    print(f"  {r[10]}  {r[11]:4d} | {e[0]:6d} {e[1]:7d} {e[2]:8d} {e[3]:9d} {e[4]:9d} {e[5]:6d} | {gap:6d}   dur(wait_out->end) {r[5]-r[2]:5d}  pro {r[3]-r[2]:5d} loop {r[4]-r[3]:5d} tail {r[5]-r[4]:5d}")
    prev_end = r[5]
d = rec[:, 5] - rec[:, 2]
for k in (0, 1, 2):
    sel = rec[:, 10] == k
    if sel.any():
        print(f"pro={k}: n={sel.sum()} wait_out->end median {np.median(d[sel]):.0f} ns; prologue {np.median((rec[:,3]-rec[:,2])[sel]):.0f}; loop {np.median((rec[:,4]-rec[:,3])[sel]):.0f}; tail {np.median((rec[:,5]-rec[:,4])[sel]):.0f}")

sel = rec[:, 10] == 1
print("LN prologue: wait_out -> stats done %.0f, mix+quantise %.0f, sync %.0f" % (
    np.median((rec[:,7]-rec[:,2])[sel]), np.median((rec[:,8]-rec[:,7])[sel]), np.median((rec[:,3]-rec[:,8])[sel])))
