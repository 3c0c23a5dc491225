"""The reference's UNMODIFIED Python wrapper (python/rwkv_cpp, imported from /root/reference: build container only) binds every
entry point it declares against the product librwkv.so: RWKVSharedLibrary's constructor sets argtypes / restype on each symbol
and raises if one is missing (python/rwkv_cpp/rwkv_cpp_shared_library.py:49-107).  Without a GPU the model constructor must fail
loudly through the reference's own error path (no CPU fallback)."""
import os
import sys

import pytest

REF_PY = "/root/reference/python"


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF_PY, "rwkv_cpp")), reason="reference sources not mounted (GPU box)")
def test_reference_wrapper_binds_product_library(lib_path):
    sys.path.insert(0, REF_PY)
    try:
        from rwkv_cpp import rwkv_cpp_shared_library as ref_lib
        from rwkv_cpp import rwkv_cpp_model as ref_model
    finally:
        sys.path.remove(REF_PY)
    lib = ref_lib.RWKVSharedLibrary(lib_path)
    info = lib.rwkv_get_system_info_string()
    assert "CPU_FALLBACK=0" in info
    for name in ("rwkv_init_from_file", "rwkv_eval", "rwkv_eval_sequence", "rwkv_eval_sequence_in_chunks", "rwkv_get_n_vocab", "rwkv_get_n_embed",
                 "rwkv_get_n_layer", "rwkv_get_state_buffer_element_count", "rwkv_get_logits_buffer_element_count", "rwkv_free",
                 "rwkv_quantize_model_file", "rwkv_get_system_info_string"):
        assert hasattr(lib.library, name), name
    import torch
    if not torch.cuda.is_available():
        from tests import util
        with pytest.raises(ValueError):                           # the reference wrapper turns a NULL context into ValueError
            ref_model.RWKVModel(lib, os.path.join(util.MODELS, "tiny-rwkv-4v0-660K-FP32.bin"))
