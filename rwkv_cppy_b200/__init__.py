"""rwkv_cppy_b200 -- B200-native engine behind the rwkv.h C ABI (librwkv.so) and the
host-side mirror of the reference's python/rwkv_cpp wrapper.

    from rwkv_cppy_b200 import RWKVModel, load_rwkv_shared_library
    model = RWKVModel(load_rwkv_shared_library(), "model.bin")
    logits, state = model.eval(token, None)

Nothing here computes without the CUDA library and a GPU: there is no CPU path.
"""
from .rwkv_cpp_shared_library import RWKVContext, RWKVSharedLibrary, load_rwkv_shared_library, library_path  # noqa: F401
from .rwkv_cpp_model import RWKVModel, RWKVBatch  # noqa: F401
