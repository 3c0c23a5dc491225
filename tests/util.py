"""Shared helpers of the test-suite (CPU and GPU parts)."""
import json
import os
import re
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
MODELS = os.path.join(GOLDEN, "models")
ARCHS = {"4v0": "4v0-660K", "5v1": "5v1-730K", "5v2": "5v2-730K", "6v0": "6v0-3m", "7v0": "7v0-834K"}
QUANTS = ["Q4_0", "Q4_1", "Q5_0", "Q5_1", "Q8_0"]
SIX = ["FP32-to-Q5_1", "FP16-to-Q4_0", "FP32-to-Q4_1", "FP16-to-Q5_0"]
PROMPT = list(b'"in')
LONG = list(b'"in the beginning')
TYPE_ID = {"FP32": 0, "FP16": 1, "Q4_0": 2, "Q4_1": 3, "Q5_0": 7, "Q5_1": 8, "Q8_0": 9}

# tests/test_tiny_rwkv.c:38-134 -- |sum(logits - expected)| must be <= 1.05 * |value|
TOL_FULL = {"4v0": (0.001, -0.013652), "5v1": (0.001, -0.289921), "5v2": (0.001, 0.455912), "6v0": (0.001, -0.416620), "7v0": (0.001, 0.005766)}
TOL_Q_FP32 = {
    "4v0": (-0.160030, -0.547409, -0.170404, 0.278034, 0.076282),
    "5v1": (117.932594, -26.712271, -163.439407, -18.017435, 0.585238),
    "5v2": (35.271305, 67.015076, 25.273308, 48.068733, -9.441034),
    "6v0": (-7.588121, 21.939022, -27.332073, 3.576909, -9.539596),
    "7v0": (0.136785, 0.002614, -0.063645, -0.064663, 0.011924),
}
TOL_Q_FP16 = {
    "4v0": (0.154614, -0.539827, -0.180142, 0.294953, 0.077226),
    "5v1": (119.471931, -28.245888, -159.870956, -39.708530, -0.962695),
    "5v2": (34.135971, 65.573822, 21.588751, 29.726818, -7.242277),
    "6v0": (-7.660988, 21.797060, -27.269241, 3.405264, -9.734720),
    "7v0": (0.136678, -0.005140, -0.064447, -0.063531, 0.010921),
}


def table_tolerance(arch: str, fmt: str) -> float:
    if fmt == "FP32":
        return abs(TOL_FULL[arch][0])
    if fmt == "FP16":
        return abs(TOL_FULL[arch][1])
    src, q = fmt.split("-to-")
    return abs((TOL_Q_FP32 if src == "FP32" else TOL_Q_FP16)[arch][QUANTS.index(q)])


def all_cases():
    """(arch, fmt) for every tiny model we can materialise: FP32/FP16 sources, the five
    quantised formats from both sources (made with OUR rwkv_quantize_model_file), and the
    four pre-quantised 6v0 files."""
    out = []
    for a in ("4v0", "5v1", "5v2", "7v0"):
        out += [(a, "FP32"), (a, "FP16")]
        out += [(a, f"{s}-to-{q}") for s in ("FP32", "FP16") for q in QUANTS]
    out += [("6v0", f) for f in SIX]
    return out


_TMP = None


def model_path(raw_lib, arch: str, fmt: str) -> str:
    """Path of the model file; quantised variants of 4v0/5v1/5v2/7v0 are produced on demand
    with the product's quantiser into a temp dir."""
    global _TMP
    v = ARCHS[arch]
    p = os.path.join(MODELS, f"tiny-rwkv-{v}-{fmt}.bin")
    if os.path.exists(p):
        return p
    if _TMP is None:
        _TMP = tempfile.mkdtemp(prefix="rwkv_b200_models_")
    out = os.path.join(_TMP, f"tiny-rwkv-{v}-{fmt}.bin")
    if not os.path.exists(out):
        src, q = fmt.split("-to-")
        assert raw_lib.quantize(os.path.join(MODELS, f"tiny-rwkv-{v}-{src}.bin"), out, q)
    return out


def expected_logits(arch: str) -> np.ndarray:
    return np.fromfile(os.path.join(GOLDEN, f"expected-logits-{ARCHS[arch]}.bin"), np.float32)


def declared_symbols():
    names = []
    for h in ("rwkv.h", "rwkv_b200.h"):
        src = open(os.path.join(ROOT, "include", h)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names += re.findall(r"RWKV_API[^;(]*?\b(rwkv_\w+)\s*\(", src)
    return sorted(set(names))


def quant_sha():
    return json.load(open(os.path.join(GOLDEN, "quant_sha256.json")))
