"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the reference's hot path.

This is the ORACLE: a CPU restatement of rwkv.cpp's inference algorithm (ggmf file
codec, ggml block (de)quantisation, the q8 activation quantisers and block dot
products of ggml's CPU backend, and the RWKV v4 / v5.1 / v5.2 / v6 / v7 time-mix,
channel-mix and WKV recurrences).  It exists so the CUDA engine can be checked on
seeded inputs.  It is NOT part of the product: only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline leg may import it.  rwkv_cppy_b200 never does.

Pinning (see tests/test_oracle_*.py): checked against
  * the reference's golden logits tests/expected-logits-*.bin (copied to tests/golden/),
  * the unmodified reference compiled by oracle/Makefile (oracle/_ref), and
  * ggml's dequantize_row_* / quantize_row_*_ref exported by that build.

Every function cites the reference file:line it restates (paths relative to
/root/reference).  Shapes: ggml order is (ne0, ne1, ne2) with ne0 fastest; numpy
arrays here use the reversed order, so a ggml matrix (K, M) is an ndarray [M, K].
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np

F32 = np.float32

# ---------------------------------------------------------------------------------
# ggmf container (rwkv_file_format.inc:5-26,102-197; docs/FILE_FORMAT.md:9-68)
# ---------------------------------------------------------------------------------
RWKV_FILE_MAGIC = 0x67676D66
TYPE_FP32, TYPE_FP16, TYPE_Q4_0, TYPE_Q4_1 = 0, 1, 2, 3
TYPE_Q5_0, TYPE_Q5_1, TYPE_Q8_0 = 7, 8, 9
TYPE_NAMES = {0: "FP32", 1: "FP16", 2: "Q4_0", 3: "Q4_1", 7: "Q5_0", 8: "Q5_1", 9: "Q8_0"}
TYPE_FROM_NAME = {v: k for k, v in TYPE_NAMES.items()}
# (block elements, block bytes)  -- ggml-common.h:161-208
TYPE_BLOCK = {0: (1, 4), 1: (1, 2), 2: (32, 18), 3: (32, 20), 7: (32, 22), 8: (32, 24), 9: (32, 34)}


@dataclass
class GgmfTensor:
    name: str
    dtype: int
    shape: Tuple[int, ...]        # ggml order (ne0, ne1, ne2)
    data: np.ndarray              # raw uint8


@dataclass
class GgmfFile:
    version: int
    n_vocab: int
    n_embed: int
    n_layer: int
    data_type: int
    tensors: Dict[str, GgmfTensor] = field(default_factory=dict)
    order: List[str] = field(default_factory=list)


def tensor_nbytes(dtype: int, shape) -> int:
    """rwkv_utilities.inc:1-4: type_size * n / blck_size."""
    blk, nb = TYPE_BLOCK[dtype]
    n = 1
    for s in shape:
        n *= int(s)
    return nb * n // blk


def read_ggmf(path: str) -> GgmfFile:
    raw = np.fromfile(path, dtype=np.uint8)
    buf = raw.tobytes()
    magic, version, n_vocab, n_embed, n_layer, data_type = struct.unpack_from("<6I", buf, 0)
    if magic != RWKV_FILE_MAGIC:
        raise ValueError("bad magic")
    if version not in (100, 101):
        raise ValueError("bad version")
    f = GgmfFile(version, n_vocab, n_embed, n_layer, data_type)
    off = 24
    while off < len(buf):
        dim_count, key_len, dtype = struct.unpack_from("<3I", buf, off)
        off += 12
        if dim_count not in (1, 2, 3):
            raise ValueError("bad dim_count")
        shape = struct.unpack_from(f"<{dim_count}I", buf, off)
        off += 4 * dim_count
        name = buf[off:off + key_len].decode("utf-8")
        off += key_len
        n = tensor_nbytes(dtype, shape)
        f.tensors[name] = GgmfTensor(name, dtype, tuple(shape), raw[off:off + n])
        f.order.append(name)
        off += n
    return f


def write_ggmf(path: str, f: GgmfFile) -> None:
    with open(path, "wb") as out:
        out.write(struct.pack("<6I", RWKV_FILE_MAGIC, f.version, f.n_vocab, f.n_embed, f.n_layer, f.data_type))
        for name in f.order:
            t = f.tensors[name]
            key = name.encode("utf-8")
            out.write(struct.pack("<3I", len(t.shape), len(key), t.dtype))
            out.write(struct.pack(f"<{len(t.shape)}I", *t.shape))
            out.write(key)
            out.write(np.ascontiguousarray(t.data).tobytes())


# ---------------------------------------------------------------------------------
# Block codecs (ggml-common.h:161-208; ggml-quants.c:31-217 quantize_row_*_ref,
# :255-363 dequantize_row_*)
# ---------------------------------------------------------------------------------
def _fp16_field(blocks: np.ndarray, lo: int) -> np.ndarray:
    return np.ascontiguousarray(blocks[:, lo:lo + 2]).view(np.float16).reshape(-1).astype(F32)


def unpack_blocks(dtype: int, raw: np.ndarray):
    """Split raw block bytes into (q int16 [nb,32], d f32 [nb], m f32 [nb] or None).
    q is the integer the reference multiplies by d: Q4_0 nib-8, Q4_1 nib, Q5_0 q5-16,
    Q5_1 q5, Q8_0 int8.  Element j<16 = low nibble of qs[j]; j+16 = high nibble;
    5th bits: qh bit j and bit j+16 (ggml-quants.c:296-346)."""
    blk, nbytes = TYPE_BLOCK[dtype]
    b = raw.reshape(-1, nbytes)
    d = _fp16_field(b, 0)
    m = None
    if dtype == TYPE_Q8_0:
        q = b[:, 2:34].view(np.int8).astype(np.int16)
        return q, d, m
    if dtype in (TYPE_Q4_0, TYPE_Q4_1):
        qs_off = 2 if dtype == TYPE_Q4_0 else 4
        if dtype == TYPE_Q4_1:
            m = _fp16_field(b, 2)
        qs = b[:, qs_off:qs_off + 16]
        q = np.concatenate([qs & 0x0F, qs >> 4], axis=1).astype(np.int16)
        if dtype == TYPE_Q4_0:
            q -= 8
        return q, d, m
    # Q5_0 / Q5_1
    qh_off = 2 if dtype == TYPE_Q5_0 else 4
    if dtype == TYPE_Q5_1:
        m = _fp16_field(b, 2)
    qh = np.ascontiguousarray(b[:, qh_off:qh_off + 4]).view(np.uint32).reshape(-1)
    qs = b[:, qh_off + 4:qh_off + 20]
    bits = ((qh[:, None] >> np.arange(32, dtype=np.uint32)[None, :]) & 1).astype(np.int16)
    q = np.concatenate([qs & 0x0F, qs >> 4], axis=1).astype(np.int16) | (bits << 4)
    if dtype == TYPE_Q5_0:
        q -= 16
    return q, d, m


def dequantize(dtype: int, raw: np.ndarray) -> np.ndarray:
    """Bit-exact dequantize_row_q{4_0,4_1,5_0,5_1,8_0} (ggml-quants.c:255,275,296,322,349):
    y = q*d (+ m) in FP32 (q*d is exact, so one rounding regardless of FMA contraction)."""
    if dtype == TYPE_FP32:
        return raw.view(F32).copy()
    if dtype == TYPE_FP16:
        return raw.view(np.float16).astype(F32)
    q, d, m = unpack_blocks(dtype, raw)
    y = q.astype(F32) * d[:, None]
    if m is not None:
        # x0*d + m : compiled with FMA contraction on x86 (-march=native); q*d is exact in
        # FP32 (<=19 significant bits) so fused and unfused agree bit-for-bit.
        y = y + m[:, None]
    return y.reshape(-1).astype(F32)


def dequantize_fp16(dtype: int, raw: np.ndarray) -> np.ndarray:
    """Weights as the tensor-core sequence path sees them: fp16(q*d + m) with ONE rounding
    (float64 holds q*d + m exactly for every fp16 d, m pair of practical range).  This is the
    dequantise-to-FP16 step of the reference's GPU offload path (ggml-cuda convert + FP16 GEMM),
    restated; the CPU path never rounds weights to FP16."""
    if dtype == TYPE_FP16:
        return raw.view(np.float16).copy()
    q, d, m = unpack_blocks(dtype, raw)
    y = q.astype(np.float64) * d[:, None].astype(np.float64)
    if m is not None:
        y = y + m[:, None].astype(np.float64)
    return y.reshape(-1).astype(np.float16)


def matmul_fp16(dtype: int, raw: np.ndarray, M: int, K: int, x: np.ndarray) -> np.ndarray:
    """out[T][M] of the tensor-core path: FP16 weights x FP16-rounded activations, exact products,
    accumulated here in float64 (the device accumulates in FP32 in an unspecified order)."""
    w = dequantize_fp16(dtype, raw).reshape(M, K).astype(np.float64)
    xh = np.asarray(x, F32).astype(np.float16).astype(np.float64)
    return xh @ w.T


def _to_fp16_bytes(x: np.ndarray) -> np.ndarray:
    return x.astype(np.float16).view(np.uint8).reshape(-1, 2)


def quantize_ref(dtype: int, x: np.ndarray) -> np.ndarray:
    """quantize_row_q{4_0,4_1,5_0,5_1,8_0}_ref (ggml-quants.c:31,68,105,149,194); returns raw bytes."""
    x = np.ascontiguousarray(x, dtype=F32).reshape(-1, 32)
    nb = x.shape[0]
    blk, nbytes = TYPE_BLOCK[dtype]
    out = np.zeros((nb, nbytes), np.uint8)
    one = F32(1.0)
    if dtype in (TYPE_Q4_0, TYPE_Q5_0):
        ax = np.abs(x)
        idx = np.argmax(ax, axis=1)            # first max: `if (amax < fabsf(v))`
        mx = x[np.arange(nb), idx]
        mx = np.where(ax[np.arange(nb), idx] == 0, F32(0), mx)
        div = F32(-8.0) if dtype == TYPE_Q4_0 else F32(-16.0)
        d = (mx / div).astype(F32)
        with np.errstate(divide="ignore"):
            idv = np.where(d != 0, one / np.where(d != 0, d, one), F32(0)).astype(F32)
        out[:, 0:2] = _to_fp16_bytes(d)
        off = F32(8.5) if dtype == TYPE_Q4_0 else F32(16.5)
        cap = 15 if dtype == TYPE_Q4_0 else 31
        # (int8_t)(x0 + 8.5f): C truncation toward zero of a value in int8 range
        qi = np.minimum(cap, np.trunc((x * idv[:, None]).astype(F32) + off).astype(np.int32).astype(np.int8).astype(np.int32))
        qi = qi.astype(np.uint8)
    elif dtype in (TYPE_Q4_1, TYPE_Q5_1):
        mn = x.min(axis=1)
        mxv = x.max(axis=1)
        lv = F32(15.0) if dtype == TYPE_Q4_1 else F32(31.0)
        d = ((mxv - mn).astype(F32) / lv).astype(F32)
        idv = np.where(d != 0, one / np.where(d != 0, d, one), F32(0)).astype(F32)
        out[:, 0:2] = _to_fp16_bytes(d)
        out[:, 2:4] = _to_fp16_bytes(mn)
        v = (((x - mn[:, None]).astype(F32) * idv[:, None]).astype(F32) + F32(0.5)).astype(F32)
        if dtype == TYPE_Q4_1:
            qi = np.minimum(15, np.trunc(v).astype(np.int32).astype(np.int8).astype(np.int32)).astype(np.uint8)
        else:
            qi = np.trunc(v).astype(np.int32).astype(np.uint8)
    elif dtype == TYPE_Q8_0:
        amax = np.abs(x).max(axis=1)
        d = (amax / F32(127.0)).astype(F32)
        idv = np.where(d != 0, one / np.where(d != 0, d, one), F32(0)).astype(F32)
        out[:, 0:2] = _to_fp16_bytes(d)
        v = (x * idv[:, None]).astype(F32)
        q = np.where(v >= 0, np.floor(v + F32(0.5)), np.ceil(v - F32(0.5)))   # roundf: ties away
        out[:, 2:34] = q.astype(np.int8).view(np.uint8)
        return out.reshape(-1)
    else:
        raise ValueError("not a quantised type")
    lo, hi = qi[:, :16], qi[:, 16:]
    packed = (lo & 0x0F) | ((hi & 0x0F) << 4)
    if dtype == TYPE_Q4_0:
        out[:, 2:18] = packed
    elif dtype == TYPE_Q4_1:
        out[:, 4:20] = packed
    else:
        bits = np.concatenate([(lo >> 4) & 1, (hi >> 4) & 1], axis=1).astype(np.uint32)
        qh = (bits << np.arange(32, dtype=np.uint32)[None, :]).sum(axis=1).astype(np.uint32)
        qoff = 2 if dtype == TYPE_Q5_0 else 4
        out[:, qoff:qoff + 4] = qh.view(np.uint8).reshape(-1, 4)
        out[:, qoff + 4:qoff + 20] = packed
    return out.reshape(-1)


def quantize_act_q8(x: np.ndarray):
    """Activation quantiser as it actually runs on x86 (ggml-cpu-quants.c:783-850 q8_0,
    :1083-1150 q8_1): d = amax/127 stored as FP16, id = 127/amax, round-to-nearest-even,
    s = fp16(d * sum(q)) with the UNROUNDED d.  Returns (q int32 [nb,32], d f32 [nb], s f32 [nb])."""
    x = np.ascontiguousarray(x, dtype=F32).reshape(-1, 32)
    amax = np.abs(x).max(axis=1).astype(F32)
    d = (amax / F32(127.0)).astype(F32)
    with np.errstate(divide="ignore", invalid="ignore"):
        idv = np.where(amax != 0, F32(127.0) / np.where(amax != 0, amax, F32(1)), F32(0)).astype(F32)
    q = np.rint((x * idv[:, None]).astype(F32)).astype(np.int32)
    d16 = d.astype(np.float16).astype(F32)
    s16 = (d * q.sum(axis=1).astype(F32)).astype(F32).astype(np.float16).astype(F32)
    return q, d16, s16


# ---------------------------------------------------------------------------------
# mul_mat (ggml-cpu.c:7377 ggml_compute_forward_mul_mat, traits :253-311, dots
# ggml-cpu-quants.c:1741,2321,2614,2969,3343; ggml-cpu.c:1352 f32, :1463 f16)
# ---------------------------------------------------------------------------------
class Mat:
    """A ggml matrix (K, M): `matvec(x)` = ggml_mul_mat(W, x) for x of K floats.

    mode "ref"   : what the CPU reference computes -- activations quantised to Q8_0
                   (Q4_0/Q5_0/Q8_0 weights) or Q8_1 (Q4_1/Q5_1), integer block dots,
                   FP32 accumulate; F16 weights see FP16-rounded activations.
    mode "exact" : dequantised FP32 weights times FP32 activations (float64 accumulate).
    """

    def __init__(self, t: GgmfTensor):
        self.dtype = t.dtype
        self.K = int(t.shape[0])
        self.M = int(np.prod(t.shape[1:])) if len(t.shape) > 1 else 1
        self.shape = t.shape
        self.raw = t.data
        self._deq = None
        self._q = None

    def deq(self) -> np.ndarray:
        if self._deq is None:
            self._deq = dequantize(self.dtype, self.raw).reshape(self.M, self.K)
        return self._deq

    def _blocks(self):
        if self._q is None:
            q, d, m = unpack_blocks(self.dtype, self.raw)
            kb = self.K // 32
            self._q = (q.reshape(self.M, kb, 32).astype(np.int32), d.reshape(self.M, kb),
                       None if m is None else m.reshape(self.M, kb))
        return self._q

    def matvec(self, x: np.ndarray, mode: str = "ref") -> np.ndarray:
        x = np.asarray(x, dtype=F32).reshape(-1)
        assert x.shape[0] == self.K, (x.shape, self.K, self.M)
        if mode == "exact" or self.dtype == TYPE_FP32:
            if mode == "exact":
                return (self.deq().astype(np.float64) @ x.astype(np.float64)).astype(F32)
            return (self.deq() @ x).astype(F32)
        if self.dtype == TYPE_FP16:
            xh = x.astype(np.float16).astype(F32)          # ggml_fp32_to_fp16_row, RNE (F16C)
            return (self.deq() @ xh).astype(F32)           # FP16xFP16 products are exact in FP32
        qw, dw, mw = self._blocks()
        qx, dx, sx = quantize_act_q8(x)
        sumi = np.einsum("mkj,kj->mk", qw, qx, dtype=np.int64).astype(F32)   # exact ints
        acc = (dw * dx[None, :]).astype(F32) * sumi
        if mw is not None:
            acc = acc + (mw * sx[None, :]).astype(F32)
        return acc.astype(F32).sum(axis=1, dtype=F32)

    def row(self, i: int) -> np.ndarray:
        """ggml_get_rows on one row (ggml-cpu.c:8464): dequantised FP32."""
        blk, nb = TYPE_BLOCK[self.dtype]
        rb = nb * self.K // blk
        return dequantize(self.dtype, self.raw[i * rb:(i + 1) * rb])


def _vec(t: GgmfTensor) -> np.ndarray:
    return dequantize(t.dtype, t.data).reshape(-1)


# ---------------------------------------------------------------------------------
# Elementwise pieces
# ---------------------------------------------------------------------------------
def layer_norm(x, w, b, eps=1e-5):
    """rwkv_operators.inc:93 + ggml-cpu.c:6880: mean/var with double accumulators, then *w + b."""
    x = x.astype(F32)
    mean = F32(x.astype(np.float64).sum() / x.shape[0])
    v = (x - mean).astype(F32)
    var = F32((v * v).astype(F32).astype(np.float64).sum() / x.shape[0])
    scale = F32(1.0) / np.sqrt(F32(var + F32(eps)), dtype=F32)
    y = (v * scale).astype(F32)
    return ((y * w).astype(F32) + b).astype(F32)


def group_norm(x, H, eps):
    """ggml_norm over each head of S channels (rwkv_graph.inc:281-285,376-380)."""
    S = x.shape[0] // H
    xs = x.reshape(H, S).astype(F32)
    mean = (xs.astype(np.float64).sum(axis=1) / S).astype(F32)
    v = (xs - mean[:, None]).astype(F32)
    var = ((v * v).astype(F32).astype(np.float64).sum(axis=1) / S).astype(F32)
    scale = (F32(1.0) / np.sqrt((var + F32(eps)).astype(F32), dtype=F32)).astype(F32)
    return (v * scale[:, None]).astype(F32).reshape(-1)


def sigmoid(x):
    return (F32(1.0) / (F32(1.0) + np.exp(-x.astype(F32), dtype=F32))).astype(F32)


def silu(x):
    x = x.astype(F32)
    return (x / (F32(1.0) + np.exp(-x, dtype=F32))).astype(F32)


def _fma(a, b, c):
    """FP32 fused multiply-add emulated through float64 (product exact, one extra rounding)."""
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(F32)


def mix_v4(x, xp, mu):
    """rwkv_graph.inc:94-97: x*mu + (xp - xp*mu), each ggml op rounding separately."""
    return ((x * mu).astype(F32) + (xp - (xp * mu).astype(F32)).astype(F32)).astype(F32)


# ---------------------------------------------------------------------------------
# WKV recurrences
# ---------------------------------------------------------------------------------
def wkv4(k, v, aa, bb, pp, u, w):
    """rwkv_graph.inc:119-161 (w = time_decay already -exp(.) in the file)."""
    ww = (u + k).astype(F32)
    qq = np.maximum(pp, ww)
    e1 = np.exp((pp - qq).astype(F32), dtype=F32)
    e2 = np.exp((ww - qq).astype(F32), dtype=F32)
    a = ((e1 * aa).astype(F32) + (e2 * v).astype(F32)).astype(F32)
    b = ((e1 * bb).astype(F32) + e2).astype(F32)
    ww = (pp + w).astype(F32)
    qq = np.maximum(ww, k)
    e1 = np.exp((ww - qq).astype(F32), dtype=F32)
    e2 = np.exp((k - qq).astype(F32), dtype=F32)
    aa2 = ((e1 * aa).astype(F32) + (e2 * v).astype(F32)).astype(F32)
    bb2 = ((e1 * bb).astype(F32) + e2).astype(F32)
    return (a / b).astype(F32), aa2, bb2, qq.astype(F32)


def wkv6(k, v, r, u, w, S):
    """ggml_compute_forward_rwkv_wkv6_f32 (ggml-cpu.c:11803; vector body :11880-11922).
    k,v,r,w: [H,S]; u: [H,S]; S: [H,S(i=key),S(j=value)].  Returns (y [H,S], S')."""
    H, N = k.shape
    y = np.zeros((H, N), F32)
    Sn = np.empty_like(S)
    for i in range(N):
        kv = (v * k[:, i:i + 1]).astype(F32)                    # [H, j]
        temp = _fma(kv, u[:, i:i + 1], S[:, i, :])
        y = _fma(temp, r[:, i:i + 1], y)
        Sn[:, i, :] = _fma(S[:, i, :], w[:, i:i + 1], kv)
    return y, Sn


def wkv7(r, w, k, v, a, b, S):
    """rwkv_wkv_v7_impl (rwkv_operators_wkv_v7.inc:37-121). r,w,k,v,a,b: [H,S];
    S: [H, i=value row, j=key col]. Returns (y [H,S], S')."""
    H, N = r.shape
    sa = np.zeros((H, N), F32)
    for j in range(N):                                           # sequential scalar sum (:61-63)
        sa = _fma(a[:, j:j + 1], S[:, :, j], sa)
    kv = (v[:, :, None] * k[:, None, :]).astype(F32)
    Sn = _fma(sa[:, :, None], b[:, None, :], _fma(S, w[:, None, :], kv))
    y = np.zeros((H, N), F32)
    for j in range(N):
        y = _fma(Sn[:, :, j], r[:, j:j + 1], y)
    return y, Sn


# ---------------------------------------------------------------------------------
# Model (rwkv_model_loading.inc:128-285 parameter binding, :319-340 arch detection,
# :403-409 head geometry) and forward (rwkv_graph.inc:56-543, :611-720)
# ---------------------------------------------------------------------------------
class OracleModel:
    def __init__(self, path: str, mode: str = "ref"):
        f = read_ggmf(path)
        self.file = f
        self.mode = mode
        self.n_vocab, self.n_embed, self.n_layer = f.n_vocab, f.n_embed, f.n_layer
        T = f.tensors
        self.major, self.minor = 4, 0
        if "blocks.0.att.ln_x.weight" in T:
            self.major, self.minor = 5, (2 if "blocks.0.att.gate.weight" in T else 1)
        if "blocks.0.att.time_maa_x" in T:
            self.major, self.minor = 6, 0
        if "blocks.0.att.r_k" in T:
            self.major, self.minor = 7, 0
        D = self.n_embed
        self.head_count = self.head_size = 0
        if self.major == 7:
            self.head_count = T["blocks.0.att.r_k"].shape[1]
        elif self.major >= 5:
            self.head_count = T["blocks.0.att.time_decay"].shape[2]
        if self.head_count:
            self.head_size = D // self.head_count
        self.V = {}
        self.M = {}
        for name, t in T.items():
            is_mat = len(t.shape) == 2 and (name.endswith(".weight") and ("ln" not in name.split(".")[-2])
                                            or name.split(".")[-1] in ("time_maa_w1", "time_decay_w1", "time_decay_w2",
                                                                       "w1", "w2", "a1", "a2", "v1", "v2", "g1", "g2"))
            if is_mat:
                self.M[name] = Mat(t)
            else:
                self.V[name] = dequantize(t.dtype, t.data)
        if self.major == 6:
            for i in range(self.n_layer):
                t = T[f"blocks.{i}.att.time_maa_w2"]           # ggml (E, D, 5)
                E = t.shape[0]
                subs = []
                blk, nb = TYPE_BLOCK[t.dtype]
                per = nb * E * D // blk
                for n in range(5):
                    subs.append(Mat(GgmfTensor("", t.dtype, (E, D), t.data[n * per:(n + 1) * per])))
                self.M[f"blocks.{i}.att.time_maa_w2"] = subs

    # -- state -------------------------------------------------------------------
    @property
    def state_len(self) -> int:
        """rwkv.cpp:171-179."""
        per = (2 + self.head_size) if self.major >= 5 else 5
        return self.n_embed * per * self.n_layer

    def init_state(self) -> np.ndarray:
        """rwkv_eval.inc:224-241."""
        s = np.zeros(self.state_len, F32)
        if self.major < 5:
            D = self.n_embed
            s.reshape(self.n_layer, 5, D)[:, 4, :] = F32(-1e30)
        return s

    def mv(self, name, x):
        return self.M[name].matvec(x, self.mode)

    # -- blocks ------------------------------------------------------------------
    def _att_v4(self, p, x, st):
        V = self.V
        x0 = layer_norm(x, V[p + "ln1.weight"], V[p + "ln1.bias"])
        xp = st[1]
        xk = mix_v4(x0, xp, V[p + "att.time_mix_k"])
        xv = mix_v4(x0, xp, V[p + "att.time_mix_v"])
        xr = mix_v4(x0, xp, V[p + "att.time_mix_r"])
        r = sigmoid(self.mv(p + "att.receptance.weight", xr))
        k = self.mv(p + "att.key.weight", xk)
        v = self.mv(p + "att.value.weight", xv)
        wkv, aa, bb, pp = wkv4(k, v, st[2], st[3], st[4], V[p + "att.time_first"], V[p + "att.time_decay"])
        st[1], st[2], st[3], st[4] = x0, aa, bb, pp
        return self.mv(p + "att.output.weight", (r * wkv).astype(F32))

    def _ffn_v4_v5(self, p, x, st):
        V = self.V
        x0 = layer_norm(x, V[p + "ln2.weight"], V[p + "ln2.bias"])
        xp = st[0]
        xk = mix_v4(x0, xp, V[p + "ffn.time_mix_k"])
        xr = mix_v4(x0, xp, V[p + "ffn.time_mix_r"])
        st[0] = x0
        r = sigmoid(self.mv(p + "ffn.receptance.weight", xr))
        k = self.mv(p + "ffn.key.weight", xk)
        k = np.square(np.maximum(k, F32(0))).astype(F32)
        return (r * self.mv(p + "ffn.value.weight", k)).astype(F32)

    def _att_v5(self, p, x, st):
        V, H, S, D = self.V, self.head_count, self.head_size, self.n_embed
        x0 = layer_norm(x, V[p + "ln1.weight"], V[p + "ln1.bias"])
        xp = st[1]
        xk = mix_v4(x0, xp, V[p + "att.time_mix_k"])
        xv = mix_v4(x0, xp, V[p + "att.time_mix_v"])
        xr = mix_v4(x0, xp, V[p + "att.time_mix_r"])
        g = None
        if self.minor >= 2:
            xg = mix_v4(x0, xp, V[p + "att.time_mix_g"])
        st[1] = x0
        r = self.mv(p + "att.receptance.weight", xr).reshape(H, S)
        k = self.mv(p + "att.key.weight", xk).reshape(H, S)
        v = self.mv(p + "att.value.weight", xv).reshape(H, S)
        if self.minor >= 2:
            g = silu(self.mv(p + "att.gate.weight", xg))
            u = V[p + "att.time_faaaa"].reshape(H, S)
            w = V[p + "att.time_decay"].reshape(H, S)
        else:
            u = np.repeat(V[p + "att.time_first"].reshape(H, 1), S, axis=1)
            w = np.repeat(V[p + "att.time_decay"].reshape(H, 1), S, axis=1)
        y, Sn = wkv6(k, v, r, u, w, st[2].reshape(H, S, S))
        st[2] = Sn.reshape(-1)
        y = group_norm(y.reshape(-1), H, 1e-5)
        y = ((y * V[p + "att.ln_x.weight"]).astype(F32) + V[p + "att.ln_x.bias"]).astype(F32)
        if g is not None:
            y = (y * g).astype(F32)
        return self.mv(p + "att.output.weight", y)

    def _att_v6(self, p, x, st):
        V, H, S, D = self.V, self.head_count, self.head_size, self.n_embed
        x0 = layer_norm(x, V[p + "ln1.weight"], V[p + "ln1.bias"])
        sx = (st[1] - x0).astype(F32)
        st[1] = x0
        xxx = ((sx * V[p + "att.time_maa_x"]).astype(F32) + x0).astype(F32)
        t = np.tanh(self.mv(p + "att.time_maa_w1", xxx), dtype=F32)
        E = t.shape[0] // 5
        ms = [self.M[p + "att.time_maa_w2"][n].matvec(t[n * E:(n + 1) * E], self.mode) for n in range(5)]

        def mixed(m, name):
            return (((m + V[p + name]).astype(F32) * sx).astype(F32) + x0).astype(F32)
        xw, xk, xv = mixed(ms[0], "att.time_maa_w"), mixed(ms[1], "att.time_maa_k"), mixed(ms[2], "att.time_maa_v")
        xr, xg = mixed(ms[3], "att.time_maa_r"), mixed(ms[4], "att.time_maa_g")
        r = self.mv(p + "att.receptance.weight", xr).reshape(H, S)
        k = self.mv(p + "att.key.weight", xk).reshape(H, S)
        v = self.mv(p + "att.value.weight", xv).reshape(H, S)
        g = silu(self.mv(p + "att.gate.weight", xg))
        w = self.mv(p + "att.time_decay_w2", np.tanh(self.mv(p + "att.time_decay_w1", xw), dtype=F32))
        w = (w + V[p + "att.time_decay"].reshape(-1)).astype(F32)
        w = np.exp(-np.exp(w, dtype=F32), dtype=F32).reshape(H, S)
        u = V[p + "att.time_faaaa"].reshape(H, S)
        y, Sn = wkv6(k, v, r, u, w, st[2].reshape(H, S, S))
        st[2] = Sn.reshape(-1)
        y = group_norm(y.reshape(-1), H, 64e-5)
        y = ((y * V[p + "att.ln_x.weight"]).astype(F32) + V[p + "att.ln_x.bias"]).astype(F32)
        y = (y * g).astype(F32)
        return self.mv(p + "att.output.weight", y)

    def _ffn_v6(self, p, x, st):
        V = self.V
        x0 = layer_norm(x, V[p + "ln2.weight"], V[p + "ln2.bias"])
        sx = (st[0] - x0).astype(F32)
        st[0] = x0
        xk = ((sx * V[p + "ffn.time_maa_k"]).astype(F32) + x0).astype(F32)
        xr = ((sx * V[p + "ffn.time_maa_r"]).astype(F32) + x0).astype(F32)
        r = sigmoid(self.mv(p + "ffn.receptance.weight", xr))
        k = np.square(np.maximum(self.mv(p + "ffn.key.weight", xk), F32(0))).astype(F32)
        return (r * self.mv(p + "ffn.value.weight", k)).astype(F32)

    def _att_v7(self, p, x, st, i, vf):
        V, H, S, D = self.V, self.head_count, self.head_size, self.n_embed
        x0 = layer_norm(x, V[p + "ln1.weight"], V[p + "ln1.bias"])
        sx = (st[1] - x0).astype(F32)
        st[1] = x0
        mu = V[p + "att.x_rwkvag"].reshape(6, D)
        xr, xw, xk, xv, xa, xg = [((sx * mu[n]).astype(F32) + x0).astype(F32) for n in range(6)]
        r = self.mv(p + "att.receptance.weight", xr)
        g = self.mv(p + "att.g2", sigmoid(self.mv(p + "att.g1", xg)))
        a = sigmoid((self.mv(p + "att.a2", self.mv(p + "att.a1", xa)) + V[p + "att.a0"]).astype(F32))
        w = (self.mv(p + "att.w2", np.tanh(self.mv(p + "att.w1", xw), dtype=F32)) + V[p + "att.w0"]).astype(F32)
        w = np.exp((sigmoid(w) * F32(-0.606531)).astype(F32), dtype=F32)
        k = self.mv(p + "att.key.weight", xk)
        kk = (k * V[p + "att.k_k"]).astype(F32).reshape(H, S)
        # rwkv_l2norm_impl (rwkv_operators.inc:40-83)
        nrm = np.sqrt((kk * kk).astype(F32).sum(axis=1, dtype=F32), dtype=F32)
        kk = (kk * (F32(1.0) / np.maximum(nrm, F32(1e-12)))[:, None]).astype(F32).reshape(-1)
        ka = (k * V[p + "att.k_a"]).astype(F32)
        k = (k + ((a * ka).astype(F32) - ka).astype(F32)).astype(F32)
        v = self.mv(p + "att.value.weight", xv)
        if i == 0:
            vf = v
        else:
            gate = sigmoid((self.mv(p + "att.v2", self.mv(p + "att.v1", xv)) + V[p + "att.v0"]).astype(F32))
            v = (v + ((vf - v).astype(F32) * gate).astype(F32)).astype(F32)
        y, Sn = wkv7(r.reshape(H, S), w.reshape(H, S), k.reshape(H, S), v.reshape(H, S),
                     (-kk).reshape(H, S), (kk * a).astype(F32).reshape(H, S), st[2].reshape(H, S, S))
        st[2] = Sn.reshape(-1)
        y = group_norm(y.reshape(-1), H, 64e-5)
        y = ((y * V[p + "att.ln_x.weight"]).astype(F32) + V[p + "att.ln_x.bias"]).astype(F32)
        rk = ((k * r).astype(F32) * V[p + "att.r_k"].reshape(-1)).astype(F32).reshape(H, S).sum(axis=1, dtype=F32)
        y = (y + (v.reshape(H, S) * rk[:, None]).astype(F32).reshape(-1)).astype(F32)
        y = (y * g).astype(F32)
        return self.mv(p + "att.output.weight", y), vf

    def _ffn_v7(self, p, x, st):
        V = self.V
        x0 = layer_norm(x, V[p + "ln2.weight"], V[p + "ln2.bias"])
        sx = (st[0] - x0).astype(F32)
        st[0] = x0
        xk = ((sx * V[p + "ffn.x_k"]).astype(F32) + x0).astype(F32)
        k = np.square(np.maximum(self.mv(p + "ffn.key.weight", xk), F32(0))).astype(F32)
        return self.mv(p + "ffn.value.weight", k)

    # -- entry points ------------------------------------------------------------
    def eval(self, token: int, state: Optional[np.ndarray], want_logits: bool = True):
        """rwkv_eval (rwkv_eval.inc:38) over the serial graph (rwkv_graph.inc:611-720).
        Returns (new_state, logits or None)."""
        D, L = self.n_embed, self.n_layer
        if state is None:
            state = self.init_state()
        state = state.astype(F32).copy()
        V = self.V
        x = self.M["emb.weight"].row(int(token))
        x = layer_norm(x, V["blocks.0.ln0.weight"], V["blocks.0.ln0.bias"])
        per = (2 + self.head_size) if self.major >= 5 else 5
        vf = None
        for i in range(L):
            p = f"blocks.{i}."
            base = i * per * D
            if self.major >= 5:
                st = [state[base:base + D].copy(), state[base + D:base + 2 * D].copy(),
                      state[base + 2 * D:base + per * D].copy()]
            else:
                st = [state[base + n * D:base + (n + 1) * D].copy() for n in range(5)]
            if self.major == 7:
                o, vf = self._att_v7(p, x, st, i, vf)
                x = (x + o).astype(F32)
                x = (x + self._ffn_v7(p, x, st)).astype(F32)
            elif self.major == 6:
                x = (x + self._att_v6(p, x, st)).astype(F32)
                x = (x + self._ffn_v6(p, x, st)).astype(F32)
            elif self.major == 5:
                x = (x + self._att_v5(p, x, st)).astype(F32)
                x = (x + self._ffn_v4_v5(p, x, st)).astype(F32)
            else:
                x = (x + self._att_v4(p, x, st)).astype(F32)
                x = (x + self._ffn_v4_v5(p, x, st)).astype(F32)
            state[base:base + per * D] = np.concatenate(st)
        logits = None
        if want_logits:
            xo = layer_norm(x, V["ln_out.weight"], V["ln_out.bias"])
            logits = self.mv("head.weight", xo)
        return state, logits

    def eval_sequence(self, tokens, state, want_logits: bool = True):
        """rwkv_eval_sequence (rwkv_eval.inc:79): per-column arithmetic of the sequence graph
        equals the serial graph's, so this is the serial loop (logits of the last token)."""
        logits = None
        for n, t in enumerate(tokens):
            state, lg = self.eval(t, state, want_logits and n == len(tokens) - 1)
            if lg is not None:
                logits = lg
        return state, logits
