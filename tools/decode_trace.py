"""Timeline of one decode step from a -DRWKV_B200_TRACE build (csrc/trace.cuh):
    RWKV_B200_BUILD_NAME=trace RWKV_B200_NVCC_FLAGS=-DRWKV_B200_TRACE python -m rwkv_cppy_b200.build
    RWKV_B200_LIB=rwkv_cppy_b200/build/variants/librwkv_trace.so python tools/decode_trace.py [shape] [format]
The first and the last CTA of every quantised mat-vec, WKV4 and tensor-parallel collective launch and every CTA of the layer-norm / mix
launches stamp %globaltimer (ns): 0 kernel entry, 1 before griddepcontrol.wait, 2 after it, 3 after the prologue / statistics, 4 after the
chunk loop / mixes, 5 kernel end (6 - 9: phases of the layer-norm statistics).  Under torchrun (WORLD_SIZE > 1) the model is loaded
tensor-parallel and rank 0 reports its own timeline."""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rwkv_cppy_b200
from rwkv_cppy_b200 import synth
from rwkv_cppy_b200.rwkv_cpp_shared_library import RWKVSharedLibrary

shape = sys.argv[1] if len(sys.argv) > 1 else "v4-169m"
fmt = sys.argv[2] if len(sys.argv) > 2 else "FP16-to-Q4_0"
lib = RWKVSharedLibrary(os.path.abspath(os.environ["RWKV_B200_LIB"])) if os.environ.get("RWKV_B200_LIB") else rwkv_cppy_b200.load_rwkv_shared_library()
world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
if world > 1:                                   # tensor parallel: run under torchrun, rank 0 reports its own timeline
    import torch, torch.distributed as dist
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib.library.rwkv_b200_set_device(local)
    if rank == 0: synth.ensure_model(None if shape == "v4-14b" else lib, shape, fmt)
    dist.barrier()
path = synth.ensure_model(None if shape == "v4-14b" else lib, shape, fmt)
if world > 1:
    ids = [lib.rwkv_b200_tp_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    model = rwkv_cppy_b200.RWKVModel(lib, path, tensor_parallel=(rank, world, ids[0]))
    if os.environ.get("RWKV_B200_TP_NCCL_ONLY") != "1":
        handles = [None] * world
        dist.all_gather_object(handles, lib.rwkv_b200_tp_ipc_handle(model._ctx))
        lib.rwkv_b200_tp_connect(model._ctx, handles)
else:
    model = rwkv_cppy_b200.RWKVModel(lib, path)
batch = model.create_batch(1)
toks = np.random.default_rng(0).integers(0, model.n_vocab, size=(16, 1)).astype(np.uint32)
batch.run_timed(toks, True)
readers = []
for name in ("rwkv_b200_debug_trace", "rwkv_b200_debug_trace_ew", "rwkv_b200_debug_trace_wkv", "rwkv_b200_debug_trace_tp"):
    fn = getattr(lib.library, name)
    fn.restype = ctypes.c_int
    readers.append(fn)
buf = (ctypes.c_ulonglong * (8192 * 12))()
for fn in readers: fn(buf, 8192)                # drop the warm-up records
STEPS = 4
ms = batch.run_timed(toks[:STEPS], True)
recs = []
for fn in readers:
    n = fn(buf, 8192)
    recs.append(np.ctypeslib.as_array(buf).reshape(-1, 12)[:n].astype(np.int64).copy())
rec = np.concatenate(recs)
rec = rec[np.argsort(rec[:, 0], kind="stable")]
print(f"{len(rec)} records, {ms / STEPS * 1e3:.1f} us/step, {model.n_layer} layers")


if world > 1:
    dist.barrier()
    if rank != 0:
        dist.destroy_process_group()
        sys.exit(0)


def kind_name(k):
    if k == 200: return "wkv4"
    if k == 300: return "tp all-reduce"
    if k == 310: return "tp ll all-reduce"
    if k == 311: return "tp ll all-reduce+mix"
    if k == 301: return "tp all-gather"
    if k == 400: return "tp ffn combine"
    if 100 <= k < 200: return f"ln_mix{k - 100}"
    return f"mv pro{k & 255} nch{(k >> 8) & 255} M{k >> 16}"


cta = rec[:, 11] >> 24
gridsz = rec[:, 11] & 0xFFFFFF
rec[:, 11] = gridsz
first = rec[cta == 0]
last = rec[(cta == gridsz - 1) & (gridsz > 1)]
ln_all = rec[(rec[:, 10] >= 100) & (rec[:, 10] < 200)]
p2p_all = rec[(rec[:, 10] == 300) | (rec[:, 10] == 310) | (rec[:, 10] == 311)]
# pair first / last CTA records of the same launch: same kind, entry stamps within the launch
per_step = len(first) // STEPS
step = first[per_step * 2: per_step * 3]
t0 = step[0, 0]
nshow = int(os.environ.get("TRACE_SHOW", "24"))
mid = len(step) // 2
print("first CTA of each launch, middle of step 2 (ns relative to the step's first record)")
print(f"{'kernel':28s} {'grid':>5s} | {'entry':>8s} {'wait_in':>8s} {'wait_out':>8s} {'pro':>8s} {'loop':>8s} {'end':>8s} | {'gap':>6s} {'dur':>6s}")
prev_end = None
for r in step[mid: mid + nshow]:
    e = r[:6] - t0
    gap = (r[2] - prev_end) if prev_end is not None else 0
    grid = r[11]
    # the last CTA's record of the same launch (entry within 100 us, same kind, same grid)
    sel = (last[:, 10] == r[10]) & (last[:, 11] == grid) & (np.abs(last[:, 2] - r[2]) < 20000)
    lend = (last[sel][:, 5].max() - t0) if sel.any() else -1
    print(f"{kind_name(int(r[10])):28s} {grid:5d} | {e[0]:8d} {e[1]:8d} {e[2]:8d} {e[3]:8d} {e[4]:8d} {e[5]:8d} | {gap:6d} {r[5]-r[2]:6d}   last CTA end {lend}")
    prev_end = max(r[5], lend + t0) if lend >= 0 else r[5]

print("\nmedians per kernel kind (ns): first CTA wait_out -> end | entry -> wait_out (time spent waiting for the predecessor) | last CTA end - first CTA end")
for k in np.unique(first[:, 10]):
    f = first[first[:, 10] == k]
    l = last[last[:, 10] == k]
    d = np.median(f[:, 5] - f[:, 2]); w = np.median(f[:, 2] - f[:, 0])
    parts = f"pro {np.median(f[:,3]-f[:,2]):.0f} loop {np.median(f[:,4]-f[:,3]):.0f} tail {np.median(f[:,5]-f[:,4]):.0f}"
    ld = np.median(l[:, 5] - l[:, 2]) if len(l) else -1
    lw = np.median(l[:, 2] - l[:, 0]) if len(l) else -1
    print(f"{kind_name(int(k)):28s} n={len(f):5d} first: dur {d:7.0f} ({parts}) waited {w:7.0f} | last CTA: dur {ld:7.0f} waited {lw:7.0f}")

# how the step divides: time from the previous launch's latest recorded end to this launch's latest recorded end (a launch whose last
# CTA leaves no record -- the cluster form of the K split -- passes its tail on to the next launch)
print("\nstep time by kernel kind: sum over launches of (this launch's latest end - previous launch's latest end), one step")
tot = {}
prev = None
for r in step:
    grid = r[11]
    sel = (last[:, 10] == r[10]) & (last[:, 11] == grid) & (np.abs(last[:, 2] - r[2]) < 20000)
    end = max(int(r[5]), int(last[sel][:, 5].max())) if sel.any() else int(r[5])
    if prev is not None:
        tot[kind_name(int(r[10]))] = tot.get(kind_name(int(r[10])), 0) + max(0, end - prev)
    prev = max(prev, end) if prev is not None else end
s = sum(tot.values())
for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    print(f"  {k:28s} {v / 1e3:8.1f} us  {100.0 * v / s:5.1f} %")
print(f"  total {s / 1e3:.1f} us (traced kernels only: the FP16 head, ln_out and the embedding are not stamped)")

# every CTA of the layer-norm launches (they all stamp): phases relative to the dependency wait returning
print("\nlayer-norm / mix launches, per CTA (ns after wait_out): row loaded | mean reduced | 2nd pass summed | variance reduced | stats done | mixes done | end")
for k in np.unique(ln_all[:, 10]):
    g = ln_all[ln_all[:, 10] == k]
    g = g[np.argsort(g[:, 0], kind="stable")]
    # split into launches: a gap of more than 10 us between entry stamps starts a new launch
    starts = np.r_[0, np.nonzero(np.diff(g[:, 0]) > 10000)[0] + 1, len(g)]
    launches = [g[a:b] for a, b in zip(starts[:-1], starts[1:])]
    mid_l = launches[len(launches) // 2]
    print(f"{kind_name(int(k))}: {len(launches)} launches of {len(mid_l)} CTAs; one launch:")
    for r in mid_l:
        rel = [int(r[i] - r[2]) for i in (6, 7, 8, 9, 3, 4, 5)]
        print("   wait_in %6d | " % int(r[2] - r[1]) + " ".join(f"{v:6d}" for v in rel))
    al = np.array([[r[i] - r[2] for i in (6, 7, 8, 9, 3, 4, 5)] for r in g])
    print("   median over all CTAs: " + " ".join(f"{np.median(al[:, i]):6.0f}" for i in range(7)) + "   max end: %.0f  (per launch median of max end: %.0f)" % (
        al[:, 6].max(), np.median([int((l[:, 5] - l[:, 2].min()).max()) for l in launches])))

if len(p2p_all):
    d = p2p_all
    print("\ntp all-reduce kernel, first / last CTA (ns after wait_out): own stores done | peers' flags seen | end   -- medians: %.0f | %.0f | %.0f  (waited for the predecessor %.0f)" % (
        np.median(d[:, 3] - d[:, 2]), np.median(d[:, 4] - d[:, 2]), np.median(d[:, 5] - d[:, 2]), np.median(d[:, 2] - d[:, 0])))
if world > 1:
    dist.destroy_process_group()
