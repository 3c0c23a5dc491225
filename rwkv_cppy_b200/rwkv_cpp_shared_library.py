"""ctypes binding of librwkv.so -- the host-side mirror of the reference's
python/rwkv_cpp/rwkv_cpp_shared_library.py (class and method names, argument meaning and
error behaviour follow :21-373 there, so code written against the reference wrapper runs
unchanged).  The library behind it is the B200 engine; there is no CPU path: a call that
computes raises when the CUDA extension or the GPU is missing.

Additive methods (`rwkv_b200_*`) bind include/rwkv_b200.h.
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional, Sequence

P_FLOAT = ctypes.POINTER(ctypes.c_float)
P_INT = ctypes.POINTER(ctypes.c_int32)
P_U32 = ctypes.POINTER(ctypes.c_uint32)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIBRARY_NAME = "librwkv.so"


class RWKVContext:
    """Opaque handle of a loaded model (rwkv_cpp_shared_library.py:21)."""

    def __init__(self, ptr) -> None:
        self.ptr = ptr


class RWKVSharedLibrary:
    """Loaded librwkv.so with typed entry points (rwkv_cpp_shared_library.py:26-373)."""

    def __init__(self, shared_library_path: str) -> None:
        self.path = shared_library_path
        # RTLD_LOCAL (ctypes' default): a process may also hold the reference's own librwkv.so (tests, the
        # bench's cpu_baseline leg); with RTLD_GLOBAL that library's internal rwkv_* calls would bind here
        self.library = lib = ctypes.CDLL(shared_library_path)
        vp, u32, sz, b = ctypes.c_void_p, ctypes.c_uint32, ctypes.c_size_t, ctypes.c_bool

        def sig(name, args, res):
            fn = getattr(lib, name)
            fn.argtypes, fn.restype = args, res

        sig("rwkv_init_from_file", [ctypes.c_char_p, u32, u32], vp)
        sig("rwkv_clone_context", [vp, u32], vp)
        sig("rwkv_gpu_offload_layers", [vp, u32], b)
        sig("rwkv_eval", [vp, ctypes.c_int32, P_FLOAT, P_FLOAT, P_FLOAT], b)
        sig("rwkv_eval_sequence", [vp, P_INT, sz, P_FLOAT, P_FLOAT, P_FLOAT], b)
        sig("rwkv_eval_sequence_in_chunks", [vp, P_INT, sz, sz, P_FLOAT, P_FLOAT, P_FLOAT], b)
        for n in ("rwkv_get_n_vocab", "rwkv_get_n_embed", "rwkv_get_n_layer", "rwkv_get_state_len", "rwkv_get_logits_len"):
            sig(n, [vp], sz)
        sig("rwkv_get_state_buffer_element_count", [vp], u32)
        sig("rwkv_get_logits_buffer_element_count", [vp], u32)
        sig("rwkv_init_state", [vp, P_FLOAT], None)
        sig("rwkv_free", [vp], None)
        sig("rwkv_quantize_model_file", [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p], b)
        sig("rwkv_get_system_info_string", [], ctypes.c_char_p)
        sig("rwkv_set_print_errors", [vp, b], None)
        sig("rwkv_get_print_errors", [vp], b)
        sig("rwkv_get_last_error", [vp], ctypes.c_int)
        # additive (include/rwkv_b200.h)
        sig("rwkv_b200_device_count", [], ctypes.c_int)
        sig("rwkv_b200_set_device", [ctypes.c_int], b)
        sig("rwkv_b200_set_l2_flush", [sz], None)
        sig("rwkv_b200_batch_create", [vp, sz], vp)
        sig("rwkv_b200_batch_free", [vp], None)
        sig("rwkv_b200_batch_size", [vp], sz)
        sig("rwkv_b200_batch_set_state", [vp, sz, P_FLOAT], b)
        sig("rwkv_b200_batch_get_state", [vp, sz, P_FLOAT], b)
        sig("rwkv_b200_batch_get_state_prefix", [vp, sz, P_FLOAT], b)
        sig("rwkv_b200_batch_eval", [vp, P_U32, P_FLOAT], b)
        sig("rwkv_b200_batch_eval_sample", [vp, P_U32, ctypes.c_float, ctypes.c_float, P_FLOAT, P_U32], b)
        sig("rwkv_b200_sample_logits", [P_FLOAT, sz, sz, ctypes.c_float, ctypes.c_float, P_FLOAT, P_U32], b)
        sig("rwkv_b200_batch_run_timed", [vp, P_U32, sz, ctypes.c_int, ctypes.POINTER(ctypes.c_double)], b)
        sig("rwkv_b200_batch_launches_per_step", [vp, ctypes.c_int], sz)
        sig("rwkv_b200_sequence_run_timed", [vp, P_U32, sz, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_double)], b)
        sig("rwkv_b200_dequantize", [ctypes.c_int, vp, sz, sz, P_FLOAT], b)
        sig("rwkv_b200_matrix_create", [ctypes.c_int, vp, sz, sz, sz], vp)
        sig("rwkv_b200_matrix_free", [vp], None)
        sig("rwkv_b200_matrix_mul", [vp, P_FLOAT, sz, P_FLOAT], b)
        sig("rwkv_b200_matrix_bench", [vp, sz, ctypes.c_int, ctypes.POINTER(ctypes.c_double)], b)
        sig("rwkv_b200_matrix_mul_tc", [vp, P_FLOAT, sz, P_FLOAT], b)
        sig("rwkv_b200_matrix_bench_tc", [vp, sz, ctypes.c_int, ctypes.POINTER(ctypes.c_double)], b)
        sig("rwkv_b200_matrix_bytes", [vp], sz)
        sig("rwkv_b200_tp_unique_id", [vp], b)
        sig("rwkv_b200_tp_init_from_file", [ctypes.c_char_p, ctypes.c_int, ctypes.c_int, vp], vp)
        sig("rwkv_b200_tp_ipc_handle", [vp, vp], b)
        sig("rwkv_b200_tp_connect", [vp, vp], b)
        self.nullptr = ctypes.cast(0, P_FLOAT)

    # ---- reference surface ---------------------------------------------------------------
    def rwkv_init_from_file(self, model_file_path: str, thread_count: int, offload_layers: int) -> RWKVContext:
        ptr = self.library.rwkv_init_from_file(model_file_path.encode("utf-8"), ctypes.c_uint32(thread_count), ctypes.c_uint32(offload_layers))
        if not ptr:
            raise ValueError(f"rwkv_init_from_file failed (error 0x{self.library.rwkv_get_last_error(None):x}), check stderr")
        return RWKVContext(ptr)

    # ---- tensor parallel (additive) -------------------------------------------------------
    def rwkv_b200_tp_unique_id(self) -> bytes:
        """128-byte NCCL unique id; create on rank 0 and broadcast it to the other ranks."""
        buf = ctypes.create_string_buffer(128)
        if not self.library.rwkv_b200_tp_unique_id(buf):
            raise ValueError("rwkv_b200_tp_unique_id failed (NCCL not available?)")
        return buf.raw

    def rwkv_b200_tp_init_from_file(self, model_file_path: str, rank: int, world_size: int, unique_id: bytes) -> RWKVContext:
        assert len(unique_id) == 128
        buf = ctypes.create_string_buffer(unique_id, 128)
        ptr = self.library.rwkv_b200_tp_init_from_file(model_file_path.encode("utf-8"), rank, world_size, buf)
        if not ptr:
            raise ValueError(f"rwkv_b200_tp_init_from_file failed (error 0x{self.library.rwkv_get_last_error(None):x}), check stderr")
        return RWKVContext(ptr)

    def rwkv_b200_tp_ipc_handle(self, ctx: RWKVContext) -> bytes:
        buf = ctypes.create_string_buffer(64)
        if not self.library.rwkv_b200_tp_ipc_handle(ctx.ptr, buf):
            raise ValueError("rwkv_b200_tp_ipc_handle failed, check stderr")
        return buf.raw

    def rwkv_b200_tp_connect(self, ctx: RWKVContext, handles: List[bytes]) -> None:
        blob = b"".join(handles)
        buf = ctypes.create_string_buffer(blob, len(blob))
        if not self.library.rwkv_b200_tp_connect(ctx.ptr, buf):
            raise ValueError("rwkv_b200_tp_connect failed, check stderr")

    def rwkv_clone_context(self, ctx: RWKVContext, thread_count: int) -> RWKVContext:
        ptr = self.library.rwkv_clone_context(ctx.ptr, ctypes.c_uint32(thread_count))
        if not ptr:
            raise ValueError("rwkv_clone_context failed, check stderr")
        return RWKVContext(ptr)

    def rwkv_gpu_offload_layers(self, ctx: RWKVContext, layer_count: int) -> bool:
        return bool(self.library.rwkv_gpu_offload_layers(ctx.ptr, ctypes.c_uint32(layer_count)))

    def rwkv_eval(self, ctx: RWKVContext, token: int, state_in_address: Optional[int], state_out_address: int, logits_out_address: int) -> None:
        ok = self.library.rwkv_eval(ctx.ptr, ctypes.c_int32(token), ctypes.cast(state_in_address or 0, P_FLOAT),
                                    ctypes.cast(state_out_address or 0, P_FLOAT), ctypes.cast(logits_out_address or 0, P_FLOAT))
        if not ok:
            raise ValueError("rwkv_eval failed, check stderr")

    def rwkv_eval_sequence(self, ctx: RWKVContext, tokens: Sequence[int], state_in_address: Optional[int],
                           state_out_address: int, logits_out_address: int) -> None:
        arr = (ctypes.c_int32 * len(tokens))(*tokens)
        ok = self.library.rwkv_eval_sequence(ctx.ptr, arr, ctypes.c_size_t(len(tokens)), ctypes.cast(state_in_address or 0, P_FLOAT),
                                             ctypes.cast(state_out_address or 0, P_FLOAT), ctypes.cast(logits_out_address or 0, P_FLOAT))
        if not ok:
            raise ValueError("rwkv_eval_sequence failed, check stderr")

    def rwkv_eval_sequence_in_chunks(self, ctx: RWKVContext, tokens: Sequence[int], chunk_size: int, state_in_address: Optional[int],
                                     state_out_address: int, logits_out_address: int) -> None:
        arr = (ctypes.c_int32 * len(tokens))(*tokens)
        ok = self.library.rwkv_eval_sequence_in_chunks(ctx.ptr, arr, ctypes.c_size_t(len(tokens)), ctypes.c_size_t(chunk_size),
                                                       ctypes.cast(state_in_address or 0, P_FLOAT), ctypes.cast(state_out_address or 0, P_FLOAT),
                                                       ctypes.cast(logits_out_address or 0, P_FLOAT))
        if not ok:
            raise ValueError("rwkv_eval_sequence_in_chunks failed, check stderr")

    def rwkv_get_n_vocab(self, ctx: RWKVContext) -> int:
        return self.library.rwkv_get_n_vocab(ctx.ptr)

    def rwkv_get_n_embed(self, ctx: RWKVContext) -> int:
        return self.library.rwkv_get_n_embed(ctx.ptr)

    def rwkv_get_n_layer(self, ctx: RWKVContext) -> int:
        return self.library.rwkv_get_n_layer(ctx.ptr)

    def rwkv_get_state_buffer_element_count(self, ctx: RWKVContext) -> int:
        return self.library.rwkv_get_state_buffer_element_count(ctx.ptr)

    def rwkv_get_logits_buffer_element_count(self, ctx: RWKVContext) -> int:
        return self.library.rwkv_get_logits_buffer_element_count(ctx.ptr)

    def rwkv_free(self, ctx: RWKVContext) -> None:
        self.library.rwkv_free(ctx.ptr)
        ctx.ptr = self.nullptr

    def rwkv_quantize_model_file(self, model_file_path_in: str, model_file_path_out: str, format_name: str) -> None:
        ok = self.library.rwkv_quantize_model_file(model_file_path_in.encode("utf-8"), model_file_path_out.encode("utf-8"), format_name.encode("utf-8"))
        if not ok:
            raise ValueError("rwkv_quantize_model_file failed, check stderr")

    def rwkv_get_system_info_string(self) -> str:
        return self.library.rwkv_get_system_info_string().decode("utf-8")


def library_path() -> str:
    return os.path.join(_HERE, LIBRARY_NAME)


def load_rwkv_shared_library() -> RWKVSharedLibrary:
    """Finds librwkv.so next to this package (building it is `__graft_entry__.build()` /
    `python -m rwkv_cppy_b200.build`).  Mirrors rwkv_cpp_shared_library.py:375: raises when
    the library is missing -- nothing here can compute without it."""
    for p in (library_path(), os.path.join(_HERE, "..", "bin", LIBRARY_NAME), os.path.join(_HERE, "..", "build", LIBRARY_NAME), os.path.join(_HERE, "..", LIBRARY_NAME)):
        if os.path.isfile(p):
            return RWKVSharedLibrary(os.path.abspath(p))
    raise ValueError(f"{LIBRARY_NAME} not found; run `python -m rwkv_cppy_b200.build` first")
