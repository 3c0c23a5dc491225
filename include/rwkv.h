/* rwkv.h -- C ABI of the B200-native RWKV engine (librwkv.so).
 *
 * This is the DROP-IN BOUNDARY.  Every entry point below has the name, argument list,
 * argument meaning and error behaviour of the reference's public header (cited per
 * function as rwkv.h:<line> / implementation file:line, paths relative to the reference
 * tree), so that the reference's ctypes layer (python/rwkv_cpp/rwkv_cpp_shared_library.py:49-107),
 * its C tests (tests/ *.c) and esn.cpp link against this library unchanged.
 *
 * Everything behind these symbols is new: a C++ host over hand-written sm_100a CUDA
 * kernels.  There is no CPU fallback -- every call that computes needs a CUDA device and
 * fails (false / NULL + error flags) when there is none.
 *
 * Thread model (same contract as the reference, rwkv.h:64-67,93-103): a context may be
 * used from any thread, one call at a time per context; for parallel inference clone the
 * context.  Clones share the weights in HBM (atomic refcount) and own their stream,
 * captured launch graphs and activation workspace.
 */
#ifndef RWKV_H
#define RWKV_H

#include <stddef.h>
#include <stdint.h>
#include <stdbool.h>

#if defined(_WIN32) && !defined(__MINGW32__)
#    define RWKV_API __declspec(dllexport)
#else
#    define RWKV_API __attribute__((visibility("default")))
#endif

/* ggmf container constants (rwkv.h:22-29; rwkv_file_format.inc:115-139). */
#define RWKV_FILE_MAGIC       0x67676d66
#define RWKV_FILE_VERSION_0   100
#define RWKV_FILE_VERSION_1   101
#define RWKV_FILE_VERSION_MIN RWKV_FILE_VERSION_0
#define RWKV_FILE_VERSION_MAX RWKV_FILE_VERSION_1
#define RWKV_FILE_VERSION     RWKV_FILE_VERSION_MAX

#ifdef __cplusplus
extern "C" {
#endif

/* Error word = (category << 8) | detail; several may be OR-ed together (rwkv.h:38-62). */
enum rwkv_error_flags {
    RWKV_ERROR_NONE = 0,

    /* categories */
    RWKV_ERROR_ARGS         = 1 << 8,
    RWKV_ERROR_FILE         = 2 << 8,
    RWKV_ERROR_MODEL        = 3 << 8,
    RWKV_ERROR_MODEL_PARAMS = 4 << 8,
    RWKV_ERROR_GRAPH        = 5 << 8,   /* here: a CUDA launch / graph failure */
    RWKV_ERROR_CTX          = 6 << 8,

    /* details */
    RWKV_ERROR_ALLOC         = 1,
    RWKV_ERROR_FILE_OPEN     = 2,
    RWKV_ERROR_FILE_STAT     = 3,
    RWKV_ERROR_FILE_READ     = 4,
    RWKV_ERROR_FILE_WRITE    = 5,
    RWKV_ERROR_FILE_MAGIC    = 6,
    RWKV_ERROR_FILE_VERSION  = 7,
    RWKV_ERROR_DATA_TYPE     = 8,
    RWKV_ERROR_UNSUPPORTED   = 9,
    RWKV_ERROR_SHAPE         = 10,
    RWKV_ERROR_DIMENSION     = 11,
    RWKV_ERROR_KEY           = 12,
    RWKV_ERROR_DATA          = 13,
    RWKV_ERROR_PARAM_MISSING = 14
};

struct rwkv_context;   /* opaque */

/* rwkv.h:76 / rwkv.cpp:218.  ctx == NULL sets the calling thread's global flag. */
RWKV_API void rwkv_set_print_errors(struct rwkv_context * ctx, const bool print_errors);
/* rwkv.h:80 / rwkv.cpp:224. */
RWKV_API bool rwkv_get_print_errors(const struct rwkv_context * ctx);
/* rwkv.h:84 / rwkv.cpp:229.  Read-and-clear; ctx == NULL reads the thread-local word
 * that init / quantize failures are recorded in. */
RWKV_API enum rwkv_error_flags rwkv_get_last_error(struct rwkv_context * ctx);

/* rwkv.h:91 / rwkv.cpp:71.  Parses the ggmf file, uploads every tensor to HBM (quantised
 * blocks repacked into plane layout) and prepares the single-token launch graph.
 * n_threads is accepted and ignored.  n_gpu_layers is accepted; any value places the
 * whole model on the GPU.  Returns NULL on failure. */
RWKV_API struct rwkv_context * rwkv_init_from_file(const char * model_file_path, const uint32_t n_threads, const uint32_t n_gpu_layers);

/* rwkv.h:99 / rwkv.cpp:123.  New context over the same weights; survives the parent. */
RWKV_API struct rwkv_context * rwkv_clone_context(struct rwkv_context * ctx, const uint32_t n_threads);

/* rwkv.h:109 / rwkv_eval.inc:38.  One token.  state_in NULL = fresh state; state_in may
 * alias state_out; state_out / logits_out may be NULL; logits_out NULL skips the head and
 * leaves the state bit-identical.  Buffers are caller-owned HOST float32 of
 * rwkv_get_state_len / rwkv_get_logits_len elements.  token >= n_vocab -> false + ARGS. */
RWKV_API bool rwkv_eval(struct rwkv_context * ctx, const uint32_t token, const float * state_in, float * state_out, float * logits_out);

/* rwkv.h:140 / rwkv_eval.inc:79.  Whole sequence, logits of the last token.  tokens NULL =
 * prepare only (returns true).  sequence_len 0 -> false + ARGS.  No node limit here. */
RWKV_API bool rwkv_eval_sequence(struct rwkv_context * ctx, const uint32_t * tokens, const size_t sequence_len, const float * state_in, float * state_out, float * logits_out);

/* rwkv.h:165 / rwkv_eval.inc:158.  Host loop over chunks; state stays in HBM between
 * chunks; logits only for the final chunk.  chunk_size 0 -> false + ARGS. */
RWKV_API bool rwkv_eval_sequence_in_chunks(struct rwkv_context * ctx, const uint32_t * tokens, const size_t sequence_len, const size_t chunk_size, const float * state_in, float * state_out, float * logits_out);

/* rwkv.h:177-201 / rwkv.cpp:155-185. */
RWKV_API size_t rwkv_get_n_vocab(const struct rwkv_context * ctx);
RWKV_API size_t rwkv_get_n_embed(const struct rwkv_context * ctx);
RWKV_API size_t rwkv_get_n_layer(const struct rwkv_context * ctx);
RWKV_API size_t rwkv_get_state_len(const struct rwkv_context * ctx);
RWKV_API size_t rwkv_get_logits_len(const struct rwkv_context * ctx);

/* Legacy counters the Python wrapper binds (rwkv.cpp:145-153;
 * rwkv_cpp_shared_library.py:91-95). */
RWKV_API uint32_t rwkv_get_state_buffer_element_count(const struct rwkv_context * ctx);
RWKV_API uint32_t rwkv_get_logits_buffer_element_count(const struct rwkv_context * ctx);

/* rwkv.h:201 / rwkv_eval.inc:224.  Zeros; v4 `pp` slots = -1e30. */
RWKV_API void rwkv_init_state(const struct rwkv_context * ctx, float * state);

/* rwkv.h:205 / rwkv.cpp:187.  NULL is allowed. */
RWKV_API void rwkv_free(struct rwkv_context * ctx);

/* rwkv.h:218 / rwkv_quantize.inc:16.  FP32/FP16 ggmf -> Q4_0|Q4_1|Q5_0|Q5_1|Q8_0 ggmf,
 * blocks byte-identical to ggml's quantize_row_*_ref.  Host-only. */
RWKV_API bool rwkv_quantize_model_file(const char * model_file_path_in, const char * model_file_path_out, const char * format_name);

/* rwkv.h:221 / rwkv.cpp:239.  Static storage. */
RWKV_API const char * rwkv_get_system_info_string(void);

/* Named by older callers and docs (docs/cuBLAS_on_Windows.md:50-51); absent from the
 * reference revision, kept as a compatibility entry: the model is always fully on the
 * GPU, so this only validates ctx and returns true. */
RWKV_API bool rwkv_gpu_offload_layers(struct rwkv_context * ctx, const uint32_t n_layers);

#ifdef __cplusplus
}
#endif

#endif /* RWKV_H */
