#!/usr/bin/env python3
"""Generates tests/golden/fast_outputs.npz: outputs of the UNMODIFIED reference (oracle/_ref,
built from /root/reference) on the inputs that drive the engine's FAST paths -- tensor-core
sequence mat-mul, chunk-parallel WKV, the head-size-64 kernels, the wide-batch path, and the
BASELINE.json model shapes.  Run in the build container (the GPU box only reads the result).

  tiny70/<model>/{logits,state}         every tiny fixture after the 70-byte prompt of
                                        tests/test_eval_sequence_in_chunks.c:69 (>= 64 rows)
  mini/<shape>-<fmt>/T<n>/{logits,state}  synthetic head-size-64 models (rwkv_cppy_b200/synth.py)
                                        after n = 64 / 150 / 523 tokens
  big/<shape>-<fmt>/...                 BASELINE shapes: 169M Q4_0 decode, v6-1.6B Q5_1 sequence,
                                        v7-2.9B Q8_0 streams + sequence, 14B Q4_1 decode
  files/<name>                          sha256 of the first 64 MiB + size of each synthetic model
                                        file, so a test can tell "different file" from "different result"

Synthetic files are regenerated on the GPU box by the same deterministic writer (numpy PCG64
streams + the product's byte-identical quantiser).  `--only tiny,mini,169m,v6,v7,14b,l2` limits the
sections (existing entries of the .npz are kept).
"""
import hashlib
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle.refload import RwkvLib, load_reference  # noqa: E402
from tests import util  # noqa: E402

OUT = os.path.join(HERE, "fast_outputs.npz")
PROMPT70 = list(b"This is a port of [BlinkDL/RWKV-LM](https://github.com/BlinkDL/RWKV-LM")
MINI = [("v6-mini", "FP16"), ("v6-mini", "FP16-to-Q5_1"), ("v5-mini", "FP16"), ("v5-mini", "FP16-to-Q4_0"),
        ("v7-mini", "FP16"), ("v7-mini", "FP16-to-Q8_0"), ("v4-mini", "FP16"), ("v4-mini", "FP16-to-Q4_1")]
MINI_T = [64, 150, 523]
N_THREADS = min(8, os.cpu_count() or 1)


def file_tag(path: str) -> np.ndarray:
    h = hashlib.sha256()
    with open(path, "rb") as f:
        h.update(f.read(64 << 20))
    return np.frombuffer((h.hexdigest() + ":" + str(os.path.getsize(path))).encode(), np.uint8)


def mini_tokens(T: int):
    return [int(t) for t in np.random.default_rng(T).integers(0, 512, T)]


def big_tokens(seed: int, n: int, vocab: int):
    return [int(t) for t in np.random.default_rng(seed).integers(0, vocab, n)]


def state_sample(s: np.ndarray) -> np.ndarray:
    """Every 97th float of a large state (v7-2.9B: 21.6 MB per stream)."""
    return s[::97].copy()


def save(out):
    np.savez_compressed(OUT, **out)


def section_tiny(L, out):
    from rwkv_cppy_b200.build import build_library
    prod = RwkvLib(build_library(verbose=False))           # host-side quantiser only
    prod.lib.rwkv_set_print_errors(None, False)
    for arch, fmt in util.all_cases():
        p = util.model_path(prod, arch, fmt)
        c = L.load(p, 2, 0)
        s = c.init_state()
        for t in PROMPT70:
            s, lg = c.eval(t, s)
        key = f"tiny70/{util.ARCHS[arch]}-{fmt}"
        out[key + "/logits"] = lg.copy()
        out[key + "/state"] = s.copy()
        c.free()
        print(key, float(lg.sum()), flush=True)


def section_mini(L, ours, synth, out):
    for shape, fmt in MINI:
        p = synth.ensure_model(ours, shape, fmt)
        out[f"files/{shape}-{fmt}"] = file_tag(p)
        c = L.load(p, 2, 0)      # <= head count (4): ggml's wkv6 op returns threads >= HEADS before its barrier (ggml-cpu.c:11822) and hangs otherwise
        for T in MINI_T:
            s, lg = c.eval_sequence(mini_tokens(T), c.init_state())
            out[f"mini/{shape}-{fmt}/T{T}/logits"] = lg.copy()
            out[f"mini/{shape}-{fmt}/T{T}/state"] = s.copy()
            print("mini", shape, fmt, T, float(np.abs(lg).max()), flush=True)
        c.free()


def section_169m(L, ours, synth, out):
    # configs[1]: v4 169M Q4_0, 32 decode tokens
    p = synth.ensure_model(ours, "v4-169m", "FP16-to-Q4_0")
    out["files/v4-169m-FP16-to-Q4_0"] = file_tag(p)
    c = L.load(p, N_THREADS, 0)
    s = c.init_state()
    for i, t in enumerate(big_tokens(11, 32, c.n_vocab)):
        s, lg = c.eval(t, s)
        if i == 0:
            out["big/v4-169m-FP16-to-Q4_0/logits_t1"] = lg.copy()
    out["big/v4-169m-FP16-to-Q4_0/logits_t32"] = lg.copy()
    out["big/v4-169m-FP16-to-Q4_0/state_t32"] = s.copy()
    c.free()


def section_v6(L, ours, synth, out):
    # configs[2]: v6 1.6B Q5_1: sequences of 64 and 192 tokens (tensor-core mat-mul + chunk-parallel WKV at the real shape), 4 decode tokens
    p = synth.ensure_model(ours, "v6-1b6", "Q5_1")
    out["files/v6-1b6-Q5_1"] = file_tag(p)
    c = L.load(p, N_THREADS, 0)
    toks = big_tokens(12, 192, c.n_vocab)
    s, lg = c.eval_sequence(toks[:64], c.init_state())
    out["big/v6-1b6-Q5_1/logits_T64"] = lg.copy()
    out["big/v6-1b6-Q5_1/state_T64"] = state_sample(s)
    s, lg = c.eval_sequence(toks, c.init_state())
    out["big/v6-1b6-Q5_1/logits_T192"] = lg.copy()
    out["big/v6-1b6-Q5_1/state_T192"] = state_sample(s)
    s = c.init_state()
    for t in toks[:4]:
        s, lg = c.eval(t, s)
    out["big/v6-1b6-Q5_1/logits_t4"] = lg.copy()
    c.free()


def section_v7(L, ours, synth, out):
    # configs[3]: v7 2.9B Q8_0: 2 streams x 4 tokens (exact batch path), 32 streams x 2 tokens (tensor-core batch path), a 64-token sequence
    p = synth.ensure_model(ours, "v7-2b9", "Q8_0")
    out["files/v7-2b9-Q8_0"] = file_tag(p)
    c = L.load(p, N_THREADS, 0)
    V = c.n_vocab
    for st in range(2):
        s = c.init_state()
        for t in big_tokens(100 + st, 4, V):
            s, lg = c.eval(t, s)
        out[f"big/v7-2b9-Q8_0/stream{st}_logits_t4"] = lg.copy()
        out[f"big/v7-2b9-Q8_0/stream{st}_state_t4"] = state_sample(s)
    wide = np.empty((32, V // 16), np.float32)
    for st in range(32):
        s = c.init_state()
        for t in big_tokens(200 + st, 2, V):
            s, lg = c.eval(t, s)
        wide[st] = lg[::16]
    out["big/v7-2b9-Q8_0/wide32_logits_t2_every16"] = wide
    s, lg = c.eval_sequence(big_tokens(13, 64, V), c.init_state())
    out["big/v7-2b9-Q8_0/logits_T64"] = lg.copy()
    out["big/v7-2b9-Q8_0/state_T64"] = state_sample(s)
    c.free()


def section_l2(L, ours, synth, out):
    # two-layer cuts of the v6 1.6B / v7 2.9B shapes: 6 decode tokens + a 64-token sequence; and the reference's own movement
    # under a +-1 ulp perturbation of its input state on the full-depth files (the noise floor of "same arithmetic, other sum order")
    for shape, fmt in (("v6-1b6-L2", "Q5_1"), ("v7-2b9-L2", "Q8_0")):
        p = synth.ensure_model(ours, shape, fmt)
        out[f"files/{shape}-{fmt}"] = file_tag(p)
        c = L.load(p, N_THREADS, 0)
        toks = big_tokens(15, 64, c.n_vocab)
        s = c.init_state()
        for i, t in enumerate(toks[:6]):
            s, lg = c.eval(t, s)
            out[f"big/{shape}-{fmt}/logits_t{i + 1}"] = lg.copy()
        out[f"big/{shape}-{fmt}/state_t6"] = state_sample(s)
        s, lg = c.eval_sequence(toks, c.init_state())
        out[f"big/{shape}-{fmt}/logits_T64"] = lg.copy()
        out[f"big/{shape}-{fmt}/state_T64"] = state_sample(s)
        c.free()
        print("l2", shape, "done", flush=True)
    for shape, fmt in (("v6-1b6", "Q5_1"), ("v7-2b9", "Q8_0"), ("v4-14b", "Q4_1")):
        c = L.load(synth.ensure_model(ours, shape, fmt), N_THREADS, 0)
        toks = big_tokens(100, 4, c.n_vocab)
        s = c.init_state()
        for t in toks[:3]:
            s, lg = c.eval(t, s)
        _, l1 = c.eval(toks[3], s.copy())
        sp = s.copy()
        bits = sp.view(np.int32)
        bits += np.random.default_rng(0).integers(-1, 2, size=bits.shape, dtype=np.int32) * ((np.abs(s) < 1e29) & (s != 0))
        _, l2 = c.eval(toks[3], sp)
        out[f"sensitivity/{shape}-{fmt}"] = np.float32([np.abs(l2 - l1).max() / max(1.0, float(np.abs(l1).max())), np.sqrt(np.mean((l2 - l1) ** 2))])
        print("sensitivity", shape, out[f"sensitivity/{shape}-{fmt}"], flush=True)
        c.free()


def section_14b(L, ours, synth, out):
    # configs[4]: v4 14B Q4_1, 4 decode tokens
    p = synth.ensure_model(ours, "v4-14b", "Q4_1")
    out["files/v4-14b-Q4_1"] = file_tag(p)
    c = L.load(p, N_THREADS, 0)
    s = c.init_state()
    for t in big_tokens(14, 4, c.n_vocab):
        s, lg = c.eval(t, s)
    out["big/v4-14b-Q4_1/logits_t4"] = lg.copy()
    out["big/v4-14b-Q4_1/state_t4"] = s.copy()
    c.free()


def main():
    import rwkv_cppy_b200
    from rwkv_cppy_b200 import synth
    only = None
    if "--only" in sys.argv:
        only = set(sys.argv[sys.argv.index("--only") + 1].split(","))
    ours = rwkv_cppy_b200.load_rwkv_shared_library()        # host-side quantiser only
    ours.library.rwkv_set_print_errors(None, False)
    L = load_reference()
    L.lib.rwkv_set_print_errors(None, False)
    out = dict(np.load(OUT)) if os.path.exists(OUT) else {}
    sections = [("tiny", lambda: section_tiny(L, out)), ("mini", lambda: section_mini(L, ours, synth, out)),
                ("169m", lambda: section_169m(L, ours, synth, out)), ("v6", lambda: section_v6(L, ours, synth, out)),
                ("v7", lambda: section_v7(L, ours, synth, out)), ("14b", lambda: section_14b(L, ours, synth, out)),
                ("l2", lambda: section_l2(L, ours, synth, out))]
    for name, fn in sections:
        if only and name not in only:
            continue
        t0 = time.time()
        fn()
        save(out)                                           # after every section: a long run can be resumed with --only
        print(f"section {name} done in {time.time() - t0:.1f} s", flush=True)
    print("wrote", OUT, os.path.getsize(OUT) / 1e6, "MB")


if __name__ == "__main__":
    main()
