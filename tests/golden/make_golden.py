#!/usr/bin/env python3
"""Generates the committed fixtures under tests/golden/ (run in the build container,
where /root/reference and oracle/_ref exist; the GPU box only reads the results).

  models/tiny-rwkv-*.bin     the reference's tiny test models (tests/tiny-rwkv-*): the FP32
                             and FP16 sources of 4v0/5v1/5v2/7v0 and four pre-quantised 6v0
                             files (the 6v0 FP32/FP16 sources are missing upstream)
  expected-logits-*.bin      the reference's golden logits (tests/expected-logits-*.bin)
  ref_outputs.npz            logits + state produced by the UNMODIFIED reference
                             (oracle/_ref) on every model x format for the prompt '"in'
                             (serial) and logits for a 17-token prompt
  quant_sha256.json          sha256 of the reference's own quantised files, so the product's
                             rwkv_quantize_model_file can be checked byte-for-byte on the GPU box
  codec_vectors.npz          random blocks (incl. edge cases) + ggml dequantize_row_* outputs
                             and quantize_row_*_ref / quantize_row_q8_{0,1} outputs
"""
import ctypes, hashlib, json, os, shutil, sys, tempfile
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle.refload import load_reference  # noqa: E402

REF_TESTS = "/root/reference/tests"
ARCHS = {"4v0": "4v0-660K", "5v1": "5v1-730K", "5v2": "5v2-730K", "6v0": "6v0-3m", "7v0": "7v0-834K"}
QUANTS = ["Q4_0", "Q4_1", "Q5_0", "Q5_1", "Q8_0"]
SIX = ["FP32-to-Q5_1", "FP16-to-Q4_0", "FP32-to-Q4_1", "FP16-to-Q5_0"]
PROMPT = list(b'"in')
LONG = list(b'"in the beginning')


def main():
    L = load_reference()
    os.makedirs(os.path.join(HERE, "models"), exist_ok=True)
    for a, v in ARCHS.items():
        shutil.copy(f"{REF_TESTS}/expected-logits-{v}.bin", HERE)
        fmts = SIX + ["FP32-to-Q5_0"] if a == "6v0" else ["FP32", "FP16"]      # FP32-to-Q5_0: the file tests/test_quantization_format_compatibility.c opens as 6v0-3m-Q5_0.bin
        for f in fmts:
            shutil.copy(f"{REF_TESTS}/tiny-rwkv-{v}-{f}.bin", os.path.join(HERE, "models"))
    out, sha = {}, {}
    for a, v in ARCHS.items():
        if a == "6v0":
            fmts = SIX
        else:
            fmts = ["FP32", "FP16"] + [f"{s}-to-{q}" for s in ("FP32", "FP16") for q in QUANTS]
        for f in fmts:
            p = f"{REF_TESTS}/tiny-rwkv-{v}-{f}.bin"
            if "-to-" in f:
                sha[f"{v}-{f}"] = hashlib.sha256(open(p, "rb").read()).hexdigest()
            c = L.load(p, 2, 0)
            s = c.init_state()
            for t in PROMPT:
                s, lg = c.eval(t, s)
            out[f"{v}-{f}/logits3"] = lg.copy()
            out[f"{v}-{f}/state3"] = s.copy()
            for t in LONG[3:]:
                s, lg = c.eval(t, s)
            out[f"{v}-{f}/logits17"] = lg.copy()
            c.free()
            print("ref", v, f, float(lg.sum()))
    np.savez_compressed(os.path.join(HERE, "ref_outputs.npz"), **out)
    json.dump(sha, open(os.path.join(HERE, "quant_sha256.json"), "w"), indent=1, sort_keys=True)

    # ---- codec vectors straight from ggml (exported by the reference build) ----
    lib = L.lib
    rng = np.random.default_rng(1234)
    sizes = {"q4_0": 18, "q4_1": 20, "q5_0": 22, "q5_1": 24, "q8_0": 34}
    cv = {}
    nb = 512
    for name, bs in sizes.items():
        raw = rng.integers(0, 256, size=(nb, bs), dtype=np.uint8)
        # sane fp16 scales for most blocks; keep edge cases in the first blocks
        d = (rng.standard_normal(nb) * 0.05).astype(np.float16).view(np.uint8).reshape(nb, 2)
        raw[:, 0:2] = d
        if name in ("q4_1", "q5_1"):
            raw[:, 2:4] = (rng.standard_normal(nb) * 0.5).astype(np.float16).view(np.uint8).reshape(nb, 2)
        raw[0, 0:2] = np.array([0.0], np.float16).view(np.uint8)              # d = 0
        raw[1, 0:2] = np.array([65504.0], np.float16).view(np.uint8)          # max half
        raw[2, 0:2] = np.array([6e-8], np.float16).view(np.uint8)             # subnormal
        raw[3, 0:2] = np.array([-0.0], np.float16).view(np.uint8)
        if name in ("q5_0", "q5_1"):
            off = 2 if name == "q5_0" else 4
            raw[4, off:off + 4] = 0xFF
            raw[5, off:off + 4] = 0x00
        y = np.empty(nb * 32, np.float32)
        fn = getattr(lib, f"dequantize_row_{name}")
        fn.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64]
        fn.restype = None
        fn(raw.ctypes.data, y.ctypes.data, nb * 32)
        cv[f"{name}/raw"] = raw
        cv[f"{name}/deq"] = y
    x = (rng.standard_normal(nb * 32) * rng.choice([0.01, 1.0, 30.0], size=nb).repeat(32)).astype(np.float32)
    x[0:32] = 0.0
    x[32:64] = np.float32(1.5)            # ties
    x[64:96] = np.linspace(-1, 1, 32, dtype=np.float32)
    cv["x"] = x
    for name, bs in sizes.items():
        o = np.zeros(nb * bs, np.uint8)
        fn = getattr(lib, f"quantize_row_{name}_ref")
        fn.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64]
        fn.restype = None
        fn(x.ctypes.data, o.ctypes.data, nb * 32)
        cv[f"{name}/quant_ref"] = o
    for name, bs in (("q8_0", 34), ("q8_1", 36)):
        o = np.zeros(nb * bs, np.uint8)
        fn = getattr(lib, f"quantize_row_{name}")      # the SIMD activation quantiser that actually runs
        fn.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64]
        fn.restype = None
        fn(x.ctypes.data, o.ctypes.data, nb * 32)
        cv[f"{name}/act"] = o
    np.savez_compressed(os.path.join(HERE, "codec_vectors.npz"), **cv)
    print("done")


if __name__ == "__main__":
    main()
