"""TEST INFRASTRUCTURE ONLY -- ctypes loader for the compiled, unmodified reference
(oracle/_ref/librwkv_ref_<isa>.so, built by oracle/Makefile from /root/reference).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference`
legs may import this module.  The product path (rwkv_cppy_b200) never does.

The binding mirrors the reference's own ctypes layer
(python/rwkv_cpp/rwkv_cpp_shared_library.py:49-107) so both libraries can be driven
through the same `RwkvLib` class: the product `librwkv.so` exports the same symbols.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
P_FLOAT = ctypes.POINTER(ctypes.c_float)
P_U32 = ctypes.POINTER(ctypes.c_uint32)


def _cpu_flags() -> set:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def ref_library_path() -> Optional[str]:
    """Pick the ISA variant this host can execute (avx512 needs F+BW+DQ+VL+VNNI+VBMI)."""
    flags = _cpu_flags()
    need512 = {"avx512f", "avx512bw", "avx512dq", "avx512vl", "avx512cd", "avx512_vnni", "avx512vbmi"}
    order = ["avx512", "avx2"] if need512 <= flags else ["avx2"]
    forced = os.environ.get("RWKV_REF_ISA")
    if forced:
        order = [forced]
    for isa in order:
        p = os.path.join(_HERE, "_ref", f"librwkv_ref_{isa}.so")
        if os.path.exists(p):
            return p
    return None


class RwkvLib:
    """Thin ctypes binding of the rwkv.h C ABI (rwkv.h:76-221) usable for BOTH the
    reference build and the product librwkv.so."""

    def __init__(self, path: str):
        self.path = path
        lib = ctypes.CDLL(path)
        self.lib = lib
        lib.rwkv_init_from_file.argtypes = [ctypes.c_char_p, ctypes.c_uint32, ctypes.c_uint32]
        lib.rwkv_init_from_file.restype = ctypes.c_void_p
        lib.rwkv_clone_context.argtypes = [ctypes.c_void_p, ctypes.c_uint32]
        lib.rwkv_clone_context.restype = ctypes.c_void_p
        lib.rwkv_eval.argtypes = [ctypes.c_void_p, ctypes.c_uint32, P_FLOAT, P_FLOAT, P_FLOAT]
        lib.rwkv_eval.restype = ctypes.c_bool
        lib.rwkv_eval_sequence.argtypes = [ctypes.c_void_p, P_U32, ctypes.c_size_t, P_FLOAT, P_FLOAT, P_FLOAT]
        lib.rwkv_eval_sequence.restype = ctypes.c_bool
        lib.rwkv_eval_sequence_in_chunks.argtypes = [
            ctypes.c_void_p, P_U32, ctypes.c_size_t, ctypes.c_size_t, P_FLOAT, P_FLOAT, P_FLOAT]
        lib.rwkv_eval_sequence_in_chunks.restype = ctypes.c_bool
        for name in ("rwkv_get_n_vocab", "rwkv_get_n_embed", "rwkv_get_n_layer",
                     "rwkv_get_state_len", "rwkv_get_logits_len"):
            getattr(lib, name).argtypes = [ctypes.c_void_p]
            getattr(lib, name).restype = ctypes.c_size_t
        lib.rwkv_init_state.argtypes = [ctypes.c_void_p, P_FLOAT]
        lib.rwkv_init_state.restype = None
        lib.rwkv_free.argtypes = [ctypes.c_void_p]
        lib.rwkv_free.restype = None
        lib.rwkv_quantize_model_file.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p]
        lib.rwkv_quantize_model_file.restype = ctypes.c_bool
        lib.rwkv_get_system_info_string.argtypes = []
        lib.rwkv_get_system_info_string.restype = ctypes.c_char_p
        lib.rwkv_set_print_errors.argtypes = [ctypes.c_void_p, ctypes.c_bool]
        lib.rwkv_set_print_errors.restype = None
        lib.rwkv_get_print_errors.argtypes = [ctypes.c_void_p]
        lib.rwkv_get_print_errors.restype = ctypes.c_bool
        lib.rwkv_get_last_error.argtypes = [ctypes.c_void_p]
        lib.rwkv_get_last_error.restype = ctypes.c_int
        if hasattr(lib, "rwkv_b200_device_count"):      # the product library also exports the additive API
            lib.rwkv_b200_device_count.restype = ctypes.c_int
            lib.rwkv_b200_dequantize.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t, P_FLOAT]
            lib.rwkv_b200_dequantize.restype = ctypes.c_bool

    def system_info(self) -> str:
        return self.lib.rwkv_get_system_info_string().decode()

    def quantize(self, src: str, dst: str, fmt: str) -> bool:
        return bool(self.lib.rwkv_quantize_model_file(src.encode(), dst.encode(), fmt.encode()))

    def load(self, path: str, n_threads: int = 2, n_gpu_layers: int = 0) -> "RwkvCtx":
        ptr = self.lib.rwkv_init_from_file(path.encode(), n_threads, n_gpu_layers)
        if not ptr:
            raise RuntimeError(f"rwkv_init_from_file failed for {path}: error 0x{self.lib.rwkv_get_last_error(None):x}")
        return RwkvCtx(self, ptr)


def _fp(a: Optional[np.ndarray]):
    if a is None:
        return None
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(P_FLOAT)


class RwkvCtx:
    def __init__(self, lib: RwkvLib, ptr):
        self.L = lib
        self.ptr = ptr
        l = lib.lib
        self.n_vocab = l.rwkv_get_n_vocab(ptr)
        self.n_embed = l.rwkv_get_n_embed(ptr)
        self.n_layer = l.rwkv_get_n_layer(ptr)
        self.state_len = l.rwkv_get_state_len(ptr)
        self.logits_len = l.rwkv_get_logits_len(ptr)

    def clone(self, n_threads: int = 2) -> "RwkvCtx":
        p = self.L.lib.rwkv_clone_context(self.ptr, n_threads)
        if not p:
            raise RuntimeError("rwkv_clone_context failed")
        return RwkvCtx(self.L, p)

    def init_state(self) -> np.ndarray:
        s = np.empty(self.state_len, np.float32)
        self.L.lib.rwkv_init_state(self.ptr, _fp(s))
        return s

    def eval(self, token: int, state_in, state_out=None, logits_out=None, want_logits=True):
        """Returns (state_out, logits_out). state_in may be None (fresh state)."""
        if state_out is None:
            state_out = np.empty(self.state_len, np.float32)
        if logits_out is None and want_logits:
            logits_out = np.empty(self.logits_len, np.float32)
        ok = self.L.lib.rwkv_eval(self.ptr, token, _fp(state_in), _fp(state_out), _fp(logits_out))
        if not ok:
            raise ValueError(f"rwkv_eval failed: 0x{self.L.lib.rwkv_get_last_error(self.ptr):x}")
        return state_out, logits_out

    def eval_sequence(self, tokens: Sequence[int], state_in, state_out=None, logits_out=None, want_logits=True):
        toks = np.ascontiguousarray(tokens, dtype=np.uint32)
        if state_out is None:
            state_out = np.empty(self.state_len, np.float32)
        if logits_out is None and want_logits:
            logits_out = np.empty(self.logits_len, np.float32)
        ok = self.L.lib.rwkv_eval_sequence(self.ptr, toks.ctypes.data_as(P_U32), len(toks),
                                           _fp(state_in), _fp(state_out), _fp(logits_out))
        if not ok:
            raise ValueError(f"rwkv_eval_sequence failed: 0x{self.L.lib.rwkv_get_last_error(self.ptr):x}")
        return state_out, logits_out

    def eval_sequence_in_chunks(self, tokens, chunk, state_in, state_out=None, logits_out=None, want_logits=True):
        toks = np.ascontiguousarray(tokens, dtype=np.uint32)
        if state_out is None:
            state_out = np.empty(self.state_len, np.float32)
        if logits_out is None and want_logits:
            logits_out = np.empty(self.logits_len, np.float32)
        ok = self.L.lib.rwkv_eval_sequence_in_chunks(self.ptr, toks.ctypes.data_as(P_U32), len(toks), chunk,
                                                     _fp(state_in), _fp(state_out), _fp(logits_out))
        if not ok:
            raise ValueError(f"rwkv_eval_sequence_in_chunks failed: 0x{self.L.lib.rwkv_get_last_error(self.ptr):x}")
        return state_out, logits_out

    def free(self):
        if self.ptr:
            self.L.lib.rwkv_free(self.ptr)
            self.ptr = None


def load_reference() -> RwkvLib:
    p = ref_library_path()
    if p is None:
        raise FileNotFoundError("oracle/_ref/librwkv_ref_*.so not built; run `make -C oracle`")
    return RwkvLib(p)
