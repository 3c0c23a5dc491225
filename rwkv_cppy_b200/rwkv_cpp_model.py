"""RWKVModel -- host-side mirror of the reference's python/rwkv_cpp/rwkv_cpp_model.py:22-364
(same constructor arguments, eval / eval_sequence / eval_sequence_in_chunks / free, same
buffer validation and (logits, state) return order) over the B200 engine, plus RWKVBatch,
the additive state-resident many-stream interface (include/rwkv_b200.h).
"""
from __future__ import annotations

import ctypes
import multiprocessing
import os
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import rwkv_cpp_shared_library as _lib

try:  # torch is optional for the wrapper, exactly as in the reference
    import torch  # noqa: F401
    _HAVE_TORCH = True
except Exception:  # pragma: no cover
    _HAVE_TORCH = False


class RWKVModel:
    def __init__(self, shared_library: _lib.RWKVSharedLibrary, model_path: str,
                 thread_count: int = max(1, multiprocessing.cpu_count() // 2), gpu_layer_count: int = 0, **kwargs) -> None:
        if "gpu_layers_count" in kwargs:
            gpu_layer_count = kwargs["gpu_layers_count"]
        if not os.path.isfile(model_path):
            raise ValueError(f"{model_path} is not a file")
        if not (thread_count > 0):
            raise ValueError("Thread count must be > 0")
        if not (gpu_layer_count >= 0):
            raise ValueError("GPU layer count must be >= 0")
        self._library = shared_library
        tp = kwargs.get("tensor_parallel")      # additive: (rank, world_size, unique_id) -> one shard of the model per process / GPU
        if tp is not None:
            self._ctx = shared_library.rwkv_b200_tp_init_from_file(model_path, int(tp[0]), int(tp[1]), tp[2])
        else:
            self._ctx = shared_library.rwkv_init_from_file(model_path, thread_count, gpu_layer_count)
        self._state_buffer_element_count = shared_library.rwkv_get_state_buffer_element_count(self._ctx)
        self._logits_buffer_element_count = shared_library.rwkv_get_logits_buffer_element_count(self._ctx)
        self._valid = True

    # ---- properties --------------------------------------------------------------------
    @property
    def n_vocab(self) -> int:
        return self._library.rwkv_get_n_vocab(self._ctx)

    @property
    def n_embed(self) -> int:
        return self._library.rwkv_get_n_embed(self._ctx)

    @property
    def n_layer(self) -> int:
        return self._library.rwkv_get_n_layer(self._ctx)

    @property
    def state_len(self) -> int:
        return self._state_buffer_element_count

    # ---- evaluation ----------------------------------------------------------------------
    def _prepare(self, state_in, state_out, logits_out, use_numpy):
        if not self._valid:
            raise ValueError("Model was freed")
        use_numpy = self._detect_numpy_usage([state_in, state_out, logits_out], use_numpy)
        if state_in is not None:
            self._validate_tensor(state_in, "state_in", self._state_buffer_element_count)
            state_in_ptr = self._get_data_ptr(state_in)
        else:
            state_in_ptr = 0
        if state_out is not None:
            self._validate_tensor(state_out, "state_out", self._state_buffer_element_count)
        else:
            state_out = self._zeros_float32(self._state_buffer_element_count, use_numpy)
        if logits_out is not None:
            self._validate_tensor(logits_out, "logits_out", self._logits_buffer_element_count)
        else:
            logits_out = self._zeros_float32(self._logits_buffer_element_count, use_numpy)
        return state_in_ptr, state_out, logits_out

    def eval(self, token: int, state_in, state_out=None, logits_out=None, use_numpy: bool = False):
        """One token; returns (logits, state) (rwkv_cpp_model.py:85-148)."""
        state_in_ptr, state_out, logits_out = self._prepare(state_in, state_out, logits_out, use_numpy)
        self._library.rwkv_eval(self._ctx, token, state_in_ptr, self._get_data_ptr(state_out), self._get_data_ptr(logits_out))
        return logits_out, state_out

    def eval_sequence(self, tokens: List[int], state_in, state_out=None, logits_out=None, use_numpy: bool = False):
        """Whole sequence in one pass; logits of the last token (rwkv_cpp_model.py:150-223)."""
        state_in_ptr, state_out, logits_out = self._prepare(state_in, state_out, logits_out, use_numpy)
        self._library.rwkv_eval_sequence(self._ctx, tokens, state_in_ptr, self._get_data_ptr(state_out), self._get_data_ptr(logits_out))
        return logits_out, state_out

    def eval_sequence_in_chunks(self, tokens: List[int], state_in, state_out=None, logits_out=None, chunk_size: int = 16, use_numpy: bool = False):
        """Sequence in chunks of `chunk_size` (rwkv_cpp_model.py:225-299)."""
        state_in_ptr, state_out, logits_out = self._prepare(state_in, state_out, logits_out, use_numpy)
        self._library.rwkv_eval_sequence_in_chunks(self._ctx, tokens, chunk_size, state_in_ptr,
                                                   self._get_data_ptr(state_out), self._get_data_ptr(logits_out))
        return logits_out, state_out

    def free(self) -> None:
        if not self._valid:
            raise ValueError("Already freed")
        self._valid = False
        self._library.rwkv_free(self._ctx)

    def __del__(self) -> None:
        if getattr(self, "_valid", False):
            self.free()

    # ---- additive ------------------------------------------------------------------------
    def create_batch(self, n_streams: int) -> "RWKVBatch":
        if not self._valid:
            raise ValueError("Model was freed")
        return RWKVBatch(self, n_streams)

    # ---- helpers (rwkv_cpp_model.py:320-364) ---------------------------------------------
    @staticmethod
    def _is_pytorch_tensor(tensor) -> bool:
        return hasattr(tensor, "__module__") and tensor.__module__ == "torch"

    def _detect_numpy_usage(self, tensors, use_numpy_by_default: bool) -> bool:
        for t in tensors:
            if t is not None:
                return not self._is_pytorch_tensor(t)
        return use_numpy_by_default or not _HAVE_TORCH

    def _validate_tensor(self, tensor, name: str, size: int) -> None:
        if self._is_pytorch_tensor(tensor):
            if tensor.device != torch.device("cpu"):
                raise ValueError(f"{name} is not on CPU")
            if tensor.dtype != torch.float32:
                raise ValueError(f"{name} is not of type float32")
            if tuple(tensor.shape) != (size,):
                raise ValueError(f"{name} has invalid shape {tuple(tensor.shape)}, expected ({size})")
            if not tensor.is_contiguous():
                raise ValueError(f"{name} is not contiguous")
        else:
            if tensor.dtype != np.float32:
                raise ValueError(f"{name} is not of type float32")
            if tensor.shape != (size,):
                raise ValueError(f"{name} has invalid shape {tensor.shape}, expected ({size})")
            if not tensor.data.contiguous:
                raise ValueError(f"{name} is not contiguous")

    def _get_data_ptr(self, tensor):
        return tensor.data_ptr() if self._is_pytorch_tensor(tensor) else tensor.ctypes.data

    def _zeros_float32(self, element_count: int, use_numpy: bool):
        if use_numpy:
            return np.zeros(element_count, dtype=np.float32)
        return torch.zeros(element_count, dtype=torch.float32, device="cpu")


class RWKVBatch:
    """n independent streams (chat sessions / ReservoirRWKV reservoirs) whose states stay in
    HBM; one `eval` steps all of them with one token each (additive, include/rwkv_b200.h)."""

    def __init__(self, model: RWKVModel, n_streams: int) -> None:
        self._model = model
        self._lib = model._library.library
        self._ptr = self._lib.rwkv_b200_batch_create(model._ctx.ptr, n_streams)
        if not self._ptr:
            raise ValueError("rwkv_b200_batch_create failed, check stderr")
        self.n_streams = n_streams

    def set_state(self, stream: int, state: Optional[np.ndarray]) -> None:
        p = None if state is None else np.ascontiguousarray(state, np.float32).ctypes.data_as(_lib.P_FLOAT)
        if not self._lib.rwkv_b200_batch_set_state(self._ptr, stream, p):
            raise ValueError("rwkv_b200_batch_set_state failed")

    def get_state(self, stream: int) -> np.ndarray:
        out = np.empty(self._model.state_len, np.float32)
        if not self._lib.rwkv_b200_batch_get_state(self._ptr, stream, out.ctypes.data_as(_lib.P_FLOAT)):
            raise ValueError("rwkv_b200_batch_get_state failed")
        return out

    def get_state_prefix(self, n: int) -> np.ndarray:
        out = np.empty((self.n_streams, n), np.float32)
        if not self._lib.rwkv_b200_batch_get_state_prefix(self._ptr, n, out.ctypes.data_as(_lib.P_FLOAT)):
            raise ValueError("rwkv_b200_batch_get_state_prefix failed")
        return out

    def eval(self, tokens: Sequence[int], want_logits: bool = True) -> Optional[np.ndarray]:
        toks = np.ascontiguousarray(tokens, np.uint32)
        if toks.shape != (self.n_streams,):
            raise ValueError(f"tokens has invalid shape {toks.shape}, expected ({self.n_streams},)")
        logits = np.empty((self.n_streams, self._model.n_vocab), np.float32) if want_logits else None
        lp = logits.ctypes.data_as(_lib.P_FLOAT) if want_logits else None
        if not self._lib.rwkv_b200_batch_eval(self._ptr, toks.ctypes.data_as(_lib.P_U32), lp):
            raise ValueError("rwkv_b200_batch_eval failed, check stderr")
        return logits

    def eval_sample(self, tokens: Sequence[int], temperature: float = 1.0, top_p: float = 0.8, uniforms: Optional[np.ndarray] = None,
                    rng: Optional[np.random.Generator] = None) -> np.ndarray:
        """One step of every stream + temperature / top-p sampling on the device (python/sampling.py semantics): returns the
        sampled token ids [n_streams]; the logits never leave the GPU.  `uniforms` (one draw in [0, 1) per stream) default to rng.random()."""
        toks = np.ascontiguousarray(tokens, np.uint32)
        if toks.shape != (self.n_streams,):
            raise ValueError(f"tokens has invalid shape {toks.shape}, expected ({self.n_streams},)")
        if uniforms is None:
            uniforms = (rng or np.random.default_rng()).random(self.n_streams)
        u = np.ascontiguousarray(uniforms, np.float32)
        u = np.minimum(u, np.float32(1.0 - 2.0 ** -24))            # a double just below 1 may round up to 1.0f
        if u.shape != (self.n_streams,):
            raise ValueError(f"uniforms has invalid shape {u.shape}, expected ({self.n_streams},)")
        out = np.empty(self.n_streams, np.uint32)
        if not self._lib.rwkv_b200_batch_eval_sample(self._ptr, toks.ctypes.data_as(_lib.P_U32), float(temperature), float(top_p),
                                                     u.ctypes.data_as(_lib.P_FLOAT), out.ctypes.data_as(_lib.P_U32)):
            raise ValueError("rwkv_b200_batch_eval_sample failed, check stderr")
        return out

    def run_timed(self, tokens: np.ndarray, with_logits: bool = True) -> float:
        """tokens [n_steps][n_streams]; returns device milliseconds for all steps."""
        toks = np.ascontiguousarray(tokens, np.uint32)
        assert toks.ndim == 2 and toks.shape[1] == self.n_streams
        ms = ctypes.c_double(0.0)
        if not self._lib.rwkv_b200_batch_run_timed(self._ptr, toks.ctypes.data_as(_lib.P_U32), toks.shape[0], int(with_logits), ctypes.byref(ms)):
            raise ValueError("rwkv_b200_batch_run_timed failed, check stderr")
        return ms.value

    def launches_per_step(self, with_logits: bool = True) -> int:
        return int(self._lib.rwkv_b200_batch_launches_per_step(self._ptr, int(with_logits)))

    def free(self) -> None:
        if self._ptr:
            self._lib.rwkv_b200_batch_free(self._ptr)
            self._ptr = None

    def __del__(self) -> None:
        try:
            self.free()
        except Exception:
            pass
