/* rwkv_b200.h -- ADDITIVE entry points of librwkv.so (nothing here replaces a reference
 * symbol; rwkv.h is the drop-in surface).
 *
 * Why they exist: the reference ABI moves the whole recurrent state host<->device on every
 * call (rwkv_eval.inc:2-22), which on a B200 costs more than the token itself for v5+
 * models, and it has no notion of many independent streams (the ReservoirRWKV /
 * chat-session workload loops rwkv_eval per stream, python/rwkv_cpp/reservoir.py:139-203).
 * These calls keep states resident in HBM and step B streams per launch sequence.
 * Plain pointers and sizes only.
 */
#ifndef RWKV_B200_H
#define RWKV_B200_H

#include "rwkv.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- environment ---------------------------------------------------------------------- */
/* Number of CUDA devices visible; 0 when there is no usable GPU. */
RWKV_API int  rwkv_b200_device_count(void);
/* Selects the device used by subsequent rwkv_init_from_file calls of this thread (default:
 * LOCAL_RANK env var if set, else 0). */
RWKV_API bool rwkv_b200_set_device(int device);

/* Timed runs (rwkv_b200_batch_run_timed) write `bytes` of scratch before every step so the
 * step starts with a cold L2, and time each step with its own CUDA event pair; 0 = off. */
RWKV_API void rwkv_b200_set_l2_flush(size_t bytes);

/* ---- batched, state-resident decode ------------------------------------------------------ */
struct rwkv_b200_batch;   /* opaque: n_streams states in HBM + a captured launch graph */

/* Creates n_streams streams over ctx's weights, each initialised like rwkv_init_state. */
RWKV_API struct rwkv_b200_batch * rwkv_b200_batch_create(struct rwkv_context * ctx, size_t n_streams);
RWKV_API void   rwkv_b200_batch_free(struct rwkv_b200_batch * batch);
RWKV_API size_t rwkv_b200_batch_size(const struct rwkv_b200_batch * batch);
/* state == NULL resets the stream to the initial state.  Host FP32, rwkv_get_state_len. */
RWKV_API bool   rwkv_b200_batch_set_state(struct rwkv_b200_batch * batch, size_t stream, const float * state);
RWKV_API bool   rwkv_b200_batch_get_state(struct rwkv_b200_batch * batch, size_t stream, float * state);
/* Copies the first n floats of every stream's state to out[n_streams][n] (what a
 * ReservoirRWKV readout consumes: reservoir.py:176-186). */
RWKV_API bool   rwkv_b200_batch_get_state_prefix(struct rwkv_b200_batch * batch, size_t n, float * out);
/* One token per stream.  tokens: host uint32[n_streams].  logits_out: host
 * float[n_streams][n_vocab] or NULL (head skipped).  Same arithmetic per stream as rwkv_eval. */
RWKV_API bool   rwkv_b200_batch_eval(struct rwkv_b200_batch * batch, const uint32_t * tokens, float * logits_out);
/* Runs n_steps steps with tokens[n_steps][n_streams] already uploaded once; returns the
 * device time of the steps in milliseconds (CUDA events on the engine's stream) in *ms.
 * with_logits != 0 computes the head every step (logits stay in HBM).  Measurement aid. */
RWKV_API bool   rwkv_b200_batch_run_timed(struct rwkv_b200_batch * batch, const uint32_t * tokens, size_t n_steps, int with_logits, double * ms);

/* One token per stream like rwkv_b200_batch_eval, then temperature / top-p sampling ON THE DEVICE: tokens_out[n_streams] (host)
 * receives one sampled id per stream and the logits never leave HBM (at 256 streams of V = 65536 that is 1 KB instead of 67 MB of
 * device-to-host traffic per step).  Restates python/sampling.py:10-52 of the reference (softmax; temperature == 0 -> argmax;
 * top_p < 1 -> probabilities below the top-p cutoff zeroed; probs ** (1 / temperature); np.random.choice).  np.random.choice draws one
 * uniform per call: the caller supplies it, uniforms[n_streams] in [0, 1) (host), from any generator.  logit_bias is not supported here. */
RWKV_API bool   rwkv_b200_batch_eval_sample(struct rwkv_b200_batch * batch, const uint32_t * tokens, float temperature, float top_p,
                                            const float * uniforms, uint32_t * tokens_out);
/* The same sampler on n_rows x n_vocab host logits (parity tests). */
RWKV_API bool   rwkv_b200_sample_logits(const float * logits, size_t n_rows, size_t n_vocab, float temperature, float top_p,
                                        const float * uniforms, uint32_t * tokens_out);

/* Number of kernel launches one decode step of this batch performs (graph nodes). */
RWKV_API size_t rwkv_b200_batch_launches_per_step(const struct rwkv_b200_batch * batch, int with_logits);

/* ---- sequence mode, device-timed ----------------------------------------------------------- */
/* rwkv_eval_sequence over `tokens` from a fresh state with everything resident; returns device ms.
 * Logits of the last token are computed when with_logits != 0. */
RWKV_API bool   rwkv_b200_sequence_run_timed(struct rwkv_context * ctx, const uint32_t * tokens, size_t n, int with_logits, int iters, double * ms);

/* ---- tensor parallel decode (one process per GPU, NCCL over NVLink) --------------------------
 * The reference is single-device (rwkv.cpp:86); models too large or too latency-critical for
 * one GPU are split here: column-parallel key / value / receptance (+ the WKV state of those
 * channels) and ffn key, row-parallel att output and ffn value, one all-reduce per block, the
 * vocabulary rows of the head sharded and all-gathered.  RWKV v4, n_embed and the ffn width
 * divisible by 32 * world_size.
 *   rank 0: rwkv_b200_tp_unique_id(id); broadcast the 128 bytes to the other ranks by any means;
 *   every rank: ctx = rwkv_b200_tp_init_from_file(path, rank, world, id)   (device: LOCAL_RANK / rwkv_b200_set_device);
 *   then rwkv_eval / rwkv_eval_sequence / the batch API, called collectively with the same tokens.
 * Every rank receives the full logits.  State buffers keep the reference layout; a rank updates
 * the aa / bb / pp slices of its own channels (the shift states are replicated). */
RWKV_API bool rwkv_b200_tp_unique_id(void * out_128_bytes);
RWKV_API struct rwkv_context * rwkv_b200_tp_init_from_file(const char * model_file_path, int rank, int world_size, const void * unique_id_128_bytes);
/* Optional, right after init and before the first eval: replaces the NCCL calls on the small
 * per-layer vectors by a one-shot all-reduce over NVLink peer memory (one kernel per collective:
 * P2P stores into every peer's inbox + epoch flags; csrc/tp_p2p.cu).  Every rank exports a
 * 64-byte CUDA IPC handle, the host gathers the world_size handles in rank order and gives the
 * table to every rank.  RWKV_B200_TP_NCCL_ONLY=1 keeps NCCL (the baseline it is measured against). */
RWKV_API bool rwkv_b200_tp_ipc_handle(struct rwkv_context * ctx, void * out_64_bytes);
RWKV_API bool rwkv_b200_tp_connect(struct rwkv_context * ctx, const void * handles_world_x_64_bytes);

/* ---- codec / mat-vec hooks (parity tests and the roofline microbenchmark) ------------------ */
/* type ids are the ggmf ones (rwkv_file_format.inc:5-26): 0 F32, 1 F16, 2 Q4_0, 3 Q4_1, 7 Q5_0, 8 Q5_1, 9 Q8_0 */
/* Uploads n_rows x n_cols weights given as raw ggml rows, repacks them as the loader does and
 * dequantises on the device: out[n_rows][n_cols] must equal ggml's dequantize_row_* bit for bit. */
RWKV_API bool   rwkv_b200_dequantize(int type, const void * raw, size_t n_rows, size_t n_cols, float * out);

struct rwkv_b200_matrix;  /* `copies` independent device copies of one matrix (to defeat the L2 when timing) */
RWKV_API struct rwkv_b200_matrix * rwkv_b200_matrix_create(int type, const void * raw, size_t n_rows, size_t n_cols, size_t copies);
RWKV_API void   rwkv_b200_matrix_free(struct rwkv_b200_matrix * m);
/* y[n_vec][n_rows] = W x[n_vec][n_cols] through the decode mat-vec kernel (n_vec >= 1). */
RWKV_API bool   rwkv_b200_matrix_mul(struct rwkv_b200_matrix * m, const float * x, size_t n_vec, float * y);
/* Times `iters` launches (n_vec columns) rotating over the copies; returns average device ms per launch. */
RWKV_API bool   rwkv_b200_matrix_bench(struct rwkv_b200_matrix * m, size_t n_vec, int iters, double * ms_per_launch);
/* Same product through the tensor-core GEMM (tcgen05 / TMEM; FP16 operands, FP32 accumulate):
 * quantised or FP16 matrices with n_cols >= 64.  This is the kernel rwkv_eval_sequence and wide
 * batches use for n_vec >= 64 rows (replaces ggml_mul_mat at T > 1, rwkv_graph.inc:744-870). */
RWKV_API bool   rwkv_b200_matrix_mul_tc(struct rwkv_b200_matrix * m, const float * x, size_t n_vec, float * y);
RWKV_API bool   rwkv_b200_matrix_bench_tc(struct rwkv_b200_matrix * m, size_t n_vec, int iters, double * ms_per_launch);
/* No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid
 * it).  Same product and same arithmetic as rwkv_b200_matrix_mul for a QUANTISED matrix, but this call
 * allocates nothing: x and y are DEVICE buffers of the handle's GPU supplied by the caller
 * (x[n_vec][n_cols], y[n_vec][n_rows], FP32), and the q8 activation blocks the kernel reads live in a
 * static __device__ array inside the library (352 KiB per GPU that loads it, held until exit).
 * Limit: n_vec * n_cols <= RWKV_B200_NO_ALLOC_MAX_ELEMS, larger calls fail with RWKV_ERROR_ARGS |
 * RWKV_ERROR_SHAPE.  All calls share that array (it keeps the last call's contents; every byte a
 * call reads it has rewritten first): calls that overlap on the GPU race, serialise them.
 * Synchronises the handle's stream before returning. */
#define RWKV_B200_NO_ALLOC_MAX_ELEMS 262144
RWKV_API bool   rwkv_b200_matrix_mul_no_alloc(struct rwkv_b200_matrix * m, const float * x_dev, size_t n_vec, float * y_dev);
RWKV_API size_t rwkv_b200_no_alloc_max_elements(void);
/* Algorithmic bytes of one copy (the file bytes of the matrix). */
RWKV_API size_t rwkv_b200_matrix_bytes(const struct rwkv_b200_matrix * m);
/* Test hook: the K split (1, 2 or 4) the quantised mat-vec applies to an n_rows x n_cols matrix on the current device -- a function of
 * the shape and the SM count only (148 assumed without a device), so every evaluation of a matrix adds its partial sums in one order
 * (csrc/kernels.h matvec_ksplit; replaces nothing in the reference, whose ggml_compute_forward_mul_mat ggml-cpu.c:7377 splits rows over threads). */
RWKV_API int    rwkv_b200_matvec_k_split(size_t n_rows, size_t n_cols);

#ifdef __cplusplus
}
#endif
#endif /* RWKV_B200_H */
