"""Random-init model files of the named architectures, written in the reference's ggmf
format and tensor conventions (docs/FILE_FORMAT.md:9-68; python/convert_pytorch_to_ggml.py:
73-161: ggml dimension order, FP16 for 2-D tensors only, time_decay stored pre-transformed,
v6/v7 low-rank matrices transposed, v7 x_rwkvag concatenated).  Used by bench.py and the
large-shape tests: there is no network for real checkpoints, and throughput does not depend
on weight values.  Streams tensor by tensor, so a 14B file needs no 27 GB of RAM.
"""
from __future__ import annotations

import os
import struct
from dataclasses import dataclass
from typing import Optional

import numpy as np

MAGIC, VERSION = 0x67676D66, 101
FP32, FP16 = 0, 1


@dataclass
class Shape:
    name: str
    arch: int            # 4, 5 (v5.2), 6 or 7
    n_layer: int
    n_embed: int
    n_vocab: int
    ffn: int
    head_size: int = 64
    mix_lora: int = 32   # v6 time_maa_w1 rank (per mix)
    decay_lora: int = 64
    a_lora: int = 64
    v_lora: int = 32
    g_lora: int = 128


SHAPES = {
    # BASELINE.json configs[1]: RWKV v4 Pile 169M
    "v4-169m": Shape("v4-169m", 4, 12, 768, 50277, 3072),
    # configs[4]: RWKV v4 Pile 14B
    "v4-14b": Shape("v4-14b", 4, 40, 5120, 50277, 20480),
    # configs[2]: RWKV v6 World 1.6B
    "v6-1b6": Shape("v6-1b6", 6, 24, 2048, 65536, 7168, mix_lora=32, decay_lora=64),
    # configs[3]: RWKV v7 World 2.9B
    "v7-2b9": Shape("v7-2b9", 7, 32, 2560, 65536, 10240, decay_lora=96, a_lora=96, v_lora=64, g_lora=320),
    # small stand-ins for tests
    "v4-mini": Shape("v4-mini", 4, 2, 256, 512, 1024),
    "v5-mini": Shape("v5-mini", 5, 2, 256, 512, 1024),
    "v6-mini": Shape("v6-mini", 6, 2, 256, 512, 896, mix_lora=32, decay_lora=64),
    "v7-mini": Shape("v7-mini", 7, 3, 256, 512, 1024, decay_lora=64, a_lora=64, v_lora=32, g_lora=128),
}


class _Writer:
    def __init__(self, path: str, s: Shape, data_type: int, seed: int):
        self.f = open(path, "wb")
        self.f.write(struct.pack("<6I", MAGIC, VERSION, s.n_vocab, s.n_embed, s.n_layer, data_type))
        self.rng = np.random.default_rng(seed)
        self.mat_type = data_type

    def tensor(self, name: str, arr: np.ndarray, force_f32: bool = False):
        """arr in numpy (row-major) order; ggml dims are the reverse."""
        two_d = arr.ndim == 2 and not force_f32
        dtype = self.mat_type if two_d else FP32
        key = name.encode()
        self.f.write(struct.pack("<3I", arr.ndim, len(key), dtype))
        self.f.write(struct.pack(f"<{arr.ndim}I", *reversed(arr.shape)))
        self.f.write(key)
        self.f.write(np.ascontiguousarray(arr, np.float16 if dtype == FP16 else np.float32).tobytes())

    def matrix(self, name: str, rows: int, cols: int, gain: float = 1.0, force_f32: bool = False):
        """U(-1,1)/sqrt(cols) * gain, generated in row chunks."""
        key = name.encode()
        dtype = FP32 if force_f32 else self.mat_type
        self.f.write(struct.pack("<3I", 2, len(key), dtype))
        self.f.write(struct.pack("<2I", cols, rows))
        self.f.write(key)
        step = max(1, (1 << 24) // cols)
        scale = np.float32(gain / np.sqrt(cols))
        for r0 in range(0, rows, step):
            n = min(step, rows - r0)
            a = (self.rng.random((n, cols), dtype=np.float32) * np.float32(2) - np.float32(1)) * scale
            self.f.write(a.astype(np.float16 if dtype == FP16 else np.float32).tobytes())

    def ln(self, prefix: str, D: int):
        self.tensor(prefix + ".weight", 1 + 0.02 * self.rng.standard_normal(D))
        self.tensor(prefix + ".bias", 0.02 * self.rng.standard_normal(D))

    def close(self):
        self.f.close()


def write_model(path: str, shape: Shape, data_type: int = FP16, seed: int = 0) -> str:
    s = shape
    D, V, F, L = s.n_embed, s.n_vocab, s.ffn, s.n_layer
    S = s.head_size
    H = D // S
    w = _Writer(path, s, data_type, seed)
    rng = w.rng
    w.matrix("emb.weight", V, D, gain=np.sqrt(D) * 0.5)
    w.ln("blocks.0.ln0", D)
    for i in range(L):
        p = f"blocks.{i}."
        w.ln(p + "ln1", D)
        w.ln(p + "ln2", D)
        if s.arch == 4:
            for n in "kvr":
                w.tensor(p + f"att.time_mix_{n}", rng.random(D))
            w.tensor(p + "att.time_first", 0.3 * rng.standard_normal(D))
            w.tensor(p + "att.time_decay", -np.exp(rng.uniform(-6, -1, D)))          # stored as -exp(.) (convert :125-126)
        elif s.arch == 5:
            # v5.2: per-channel decay stored as exp(-exp(.)), time_faaaa raw (convert :112-123)
            for n in "kvrg":
                w.tensor(p + f"att.time_mix_{n}", rng.random(D))
            w.tensor(p + "att.time_decay", np.exp(-np.exp(rng.uniform(-6, -1, (H, S, 1)))))
            w.tensor(p + "att.time_faaaa", 0.3 * rng.standard_normal((H, S, 1)))
        elif s.arch == 6:
            for n in "xwkvrg":
                w.tensor(p + f"att.time_maa_{n}", rng.random(D))
            w.matrix(p + "att.time_maa_w1", 5 * s.mix_lora, D, gain=0.5)
            w.tensor(p + "att.time_maa_w2", rng.uniform(-1, 1, (5, D, s.mix_lora)) * 0.1 / np.sqrt(s.mix_lora))
            w.tensor(p + "att.time_decay", rng.uniform(-6, -1, (H, S, 1)))
            w.matrix(p + "att.time_decay_w1", s.decay_lora, D, gain=0.5)
            w.matrix(p + "att.time_decay_w2", D, s.decay_lora, gain=0.1)
            w.tensor(p + "att.time_faaaa", 0.3 * rng.standard_normal((H, S, 1)))
        else:
            w.tensor(p + "att.x_rwkvag", rng.random((6, 1, D)))
            w.tensor(p + "att.w0", rng.uniform(-6, 0, (1, 1, D)))
            w.matrix(p + "att.w1", s.decay_lora, D, gain=0.5)
            w.matrix(p + "att.w2", D, s.decay_lora, gain=0.1)
            w.tensor(p + "att.a0", 0.1 * rng.standard_normal((1, 1, D)))
            w.matrix(p + "att.a1", s.a_lora, D, gain=0.5)
            w.matrix(p + "att.a2", D, s.a_lora, gain=0.1)
            if i != 0:
                w.tensor(p + "att.v0", 0.1 * rng.standard_normal((1, 1, D)))
                w.matrix(p + "att.v1", s.v_lora, D, gain=0.5)
                w.matrix(p + "att.v2", D, s.v_lora, gain=0.1)
            w.matrix(p + "att.g1", s.g_lora, D, gain=0.5)
            w.matrix(p + "att.g2", D, s.g_lora, gain=1.0)
            w.tensor(p + "att.k_k", 0.85 + 0.05 * rng.standard_normal((1, 1, D)))
            w.tensor(p + "att.k_a", 1.0 + 0.05 * rng.standard_normal((1, 1, D)))
            w.tensor(p + "att.r_k", 0.1 * rng.standard_normal((H, S)), force_f32=True)
        w.matrix(p + "att.key.weight", D, D, gain=0.5 if s.arch >= 5 else 1.0)
        w.matrix(p + "att.value.weight", D, D)
        w.matrix(p + "att.receptance.weight", D, D)
        if s.arch in (5, 6):
            w.matrix(p + "att.gate.weight", D, D)
        w.matrix(p + "att.output.weight", D, D, gain=0.5)
        if s.arch >= 5:
            w.ln(p + "att.ln_x", D)
        if s.arch in (4, 5):
            w.tensor(p + "ffn.time_mix_k", rng.random(D))
            w.tensor(p + "ffn.time_mix_r", rng.random(D))
        elif s.arch == 6:
            w.tensor(p + "ffn.time_maa_k", rng.random(D))
            w.tensor(p + "ffn.time_maa_r", rng.random(D))
        else:
            w.tensor(p + "ffn.x_k", rng.random((1, 1, D)))
        w.matrix(p + "ffn.key.weight", F, D)
        w.matrix(p + "ffn.value.weight", D, F, gain=0.5)
        if s.arch != 7:
            w.matrix(p + "ffn.receptance.weight", D, D)
    w.ln("ln_out", D)
    w.matrix("head.weight", V, D)
    w.close()
    return path


def ensure_model(lib, shape_name: str, fmt: str, cache_dir: Optional[str] = None, seed: int = 0) -> str:
    """Returns the path of `<shape>-<fmt>.bin` (fmt: FP16, FP32 or FP16-to-Qx_y), creating it
    -- and quantising with the product's rwkv_quantize_model_file -- when missing."""
    cache_dir = cache_dir or os.environ.get("RWKV_B200_MODEL_CACHE", "/tmp/rwkv_b200_models")
    os.makedirs(cache_dir, exist_ok=True)
    shape = SHAPES[shape_name]
    src_fmt = fmt.split("-to-")[0]
    src = os.path.join(cache_dir, f"{shape_name}-{src_fmt}.bin")
    if not os.path.exists(src):
        tmp = src + f".tmp{os.getpid()}"
        write_model(tmp, shape, FP16 if src_fmt == "FP16" else FP32, seed)
        os.replace(tmp, src)
    if "-to-" not in fmt:
        return src
    out = os.path.join(cache_dir, f"{shape_name}-{fmt}.bin")
    if not os.path.exists(out):
        tmp = out + f".tmp{os.getpid()}"
        lib.library.rwkv_set_print_errors(None, False)
        lib.rwkv_quantize_model_file(src, tmp, fmt.split("-to-")[1])
        os.replace(tmp, out)
    return out
