// Host-callable launchers of the sm_100a kernels + the plain-data descriptors they take.
// Every launcher enqueues on the given stream and returns cudaGetLastError().
#pragma once
#include "common.h"
#include <cstdlib>

namespace rb {

// An activation matrix [T][K] in the forms the mat-vec kernels consume:
//   f    FP32 values (always present)
//   q    int8 per-32 quantised values  } the q8 activation blocks of ggml's CPU backend
//   d    fp16-rounded scale per block   } (quantize_row_q8_0 / q8_1 as compiled for x86,
//   s    fp16-rounded d*sum(q) (q8_1)   }  ggml-cpu-quants.c:783-850,1083-1150), split into
//   isum sum(q) per block (int)         }  planes.  q == nullptr when K % 32 != 0.
//   h    FP16 copy [T][K] for the tensor-core MATMUL (optional; filled on demand)
struct Act {
    float *  f = nullptr;
    int8_t * q = nullptr;
    float *  d = nullptr;
    float *  s = nullptr;
    int *    isum = nullptr;
    uint16_t * h = nullptr;
    int      K = 0;          // row length (row stride of f and q; K/32 for d, s, isum)
};

enum Epilogue : int {
    EPI_STORE = 0,       // out = y
    EPI_SIGMOID,         // out = 1/(1+exp(-y))
    EPI_RELU_SQ,         // out = max(y,0)^2
    EPI_SILU,            // out = y/(1+exp(-y))
    EPI_TANH,            // out = tanh(y)
    EPI_ADD_TO,          // out = out + y                       (residual)
    EPI_MULADD_TO,       // out = out + aux[t][m] * y           (ffn: x += r * (Wv k))
    EPI_BIAS_SIGMOID,    // out = sigmoid(y + vec[m])           (v7 a, v-gate)
    EPI_V6_DECAY,        // out = exp(-exp(y + vec[m]))         (rwkv_graph.inc:349-356)
    EPI_V7_DECAY,        // out = exp(-0.606531 * sigmoid(y + vec[m]))   (rwkv_graph.inc:418-426)
    EPI_ADD_AUX,         // out = aux[t][m] + y                 (tensor parallel: rank 0 folds the residual into its partial sum)
    EPI_TP_PUSH,         // y          -> slot [tp_off + m] of this rank in every peer's low-latency inbox (fused all-reduce, tp.h)
    EPI_TP_PUSH_SIGMOID, // sigmoid(y) -> same                  (the ffn receptance slice travels with the partial sums)
};

struct MatvecJob {
    DevMat        W;
    Act           x;                  // input rows, T of them (x.K == W.K)
    int           x_off = 0;          // element offset into each input row (slices of a wider row)
    float *       out = nullptr;      // [T][ldo]
    int           ldo = 0;
    int           epi = EPI_STORE;
    const float * aux = nullptr;      // [T][ld_aux] (EPI_MULADD_TO)
    int           ld_aux = 0;
    const float * vec = nullptr;      // [M] per-row constant (bias / decay base)
    Act           out_act;            // optional quantised copy of `out` (out_act.q may be null)
    const float * pro_mu = nullptr;   // PRO_LNMIX launches: this job's token-shift mix vector [K]
    int           tp_off = 0;         // EPI_TP_PUSH*: entry offset inside this rank's inbox slot
    int           ks = 1;             // K split of this job when the launch splits inside the CTA (MatvecBatch::ks_warps; filled by launch_matvec)
};

// Optional prologue of a single-row quantised mat-vec launch (decode of models up to LN_SMALL_D
// wide, where the step is a chain of dependent launches, not bytes): every CTA rebuilds the
// input vector of its job in shared memory instead of reading q8 blocks a separate kernel wrote.
//   PRO_LNMIX  x -> layer norm -> token-shift mix with the job's pro_mu  (what ln_mix_kernel does; CTA 0
//              stores the new shift state)
//   PRO_WKV4   k, v, r, state -> r * wkv  (what wkv4_kernel does; CTA 0 stores the new state)
// Both reproduce the standalone kernels bit for bit (same per-element arithmetic, same
// reduction order, same q8 blocks).
constexpr int LN_SMALL_D = 2048;      // rows up to here keep their prologue operands in registers across the dependency wait
constexpr int LN_PRO_MAX_D = 8192;    // widest row a prologue (and the 256-thread layer-norm shape) handles
enum PrologueKind : int { PRO_NONE = 0, PRO_LNMIX = 1, PRO_WKV4 = 2 };

// Tensor-parallel decode, all-reduce fused into the mat-vecs (tp.h "low-latency inbox"): the
// epilogue of a row-split mat-vec stores {value, flag} pairs straight into every peer's inbox
// (EPI_TP_PUSH), and the prologue of the next mat-vec waits for the flags of its site and rebuilds
// the replicated residual row itself -- no collective launch.
constexpr int TP_LL_WORLD = 8;
struct TpFused {
    unsigned char * ll[TP_LL_WORLD] = {};   // low-latency inbox of every rank (ll[rank] is local memory), null = off
    int rank = 0, world = 1;
    const unsigned int * step = nullptr;    // pass counter (local; bumped by the first kernel of every pass)
    unsigned int * err = nullptr;           // host-mapped error word: set when a peer did not deliver within ~2 s
    unsigned int slot_entries = 0;          // entries per source slot
    int site_push = -1;                     // site id the EPI_TP_PUSH* jobs of this launch publish (flag = step << 8 | site)
};
struct TpWait {
    int site = -1;                          // >= 0: the prologue first gathers x = base (+ | + r *) sum of the ranks' partials of that site
    int gated = 0;                          // 1: x = base + r * (p0 + p1 + ...) with r from the slots' second section (ffn); 0: x = ((base + p0) + p1) + ...
    int n = 0, n_slice = 0;                 // D and D / world
    float * x_out = nullptr;                // CTA 0 stores the gathered row here (the other buffer of the residual ping-pong)
};

struct MatvecPrologue {
    int kind = PRO_NONE, D = 0, mix_kind = 0;
    const float * x = nullptr, * ln_w = nullptr, * ln_b = nullptr;          // PRO_LNMIX
    const float * k = nullptr, * v = nullptr, * r = nullptr, * u = nullptr, * decay = nullptr;   // PRO_WKV4
    const float * state_in = nullptr; float * state_out = nullptr;          // shift slot (LNMIX) / aa slot (WKV4; bb, pp follow at +state_D, +2*state_D)
    int state_D = 0;                                                         // PRO_WKV4: floats between the aa / bb / pp slots (0: D)
    TpWait wait;                                                             // PRO_LNMIX under fused tensor parallelism
};

constexpr int MATVEC_MAX_JOBS = 8;
struct MatvecBatch {
    MatvecPrologue pro;
    TpFused tp;                             // fused tensor-parallel collectives (ll[0] == nullptr: none)
    MatvecJob job[MATVEC_MAX_JOBS];
    int     cta_begin[MATVEC_MAX_JOBS + 1];   // prefix sum of CTAs per job (filled by launch_matvec)
    int     n_jobs = 0;
    int     T = 1;
    int     ks = 1;                         // K split (cluster size along x) of the quantised kernel; all jobs of a batch share it
    // K split inside the CTA instead (filled by launch_matvec): the 8 warps of a CTA cover 32 / ks rows, warp w takes K range w % ks of
    // row group w / ks, and the partial sums meet in the CTA's own shared memory in the same rank order -- the same bits as the cluster
    // form without a cluster launch or DSMEM, and every job of the launch may have its own split.  Used whenever no split job has to
    // emit the q8 block of its 32 outputs (that needs 32 rows per CTA: the cluster form).
    int     ks_warps = 0;
    // L2 prefetch of the weights of the NEXT dependent launch: a decode step is a chain of dependent launches, and between two of them
    // (tail of one, dependency wait + activation fetch of the next) HBM idles; every warp of this launch asks the L2 for its slice of the
    // next matrices before it waits, so the memory system keeps streaming through the hand-off.
    const void * pf_ptr[3] = {};
    unsigned     pf_bytes[3] = {};
    int          n_pf = 0;
};
constexpr unsigned MATVEC_PF_MAX_BYTES = 48u << 20;      // per matrix: the L2 also holds the stream of the running launch
// K split of a quantised (M x K) mat-vec: a function of the shape only, so that every evaluation
// of the same matrix adds its partial sums in the same order.  Splits while the row blocks alone
// would leave most SMs idle and every CTA keeps at least five 1024-weight chunks (below that the
// fixed cost of a CTA outweighs the extra parallelism: measured on 5120x5120 vs 5120x20480).
inline bool matvec_ksplit_disabled() {
    static const bool off = [] { const char * e = getenv("RWKV_B200_NO_KSPLIT"); return e && e[0] == '1'; }();
    return off;
}
inline int matvec_ksplit(uint32_t M, uint32_t K) {
    static const int min_chunks = [] { const char * e = getenv("RWKV_B200_KSPLIT_MIN_CHUNKS"); const int v = e ? atoi(e) : 0; return v > 0 ? v : 5; }();   // tuning switch
    const int row_blocks = (int) ((M + 31) / 32), nch = (int) ((K / 32 + TILE_BLOCKS - 1) / TILE_BLOCKS);
    int ks = 1;
    if (matvec_ksplit_disabled()) return 1;
    while (row_blocks * ks < 3 * device_sm_count() && ks < 4 && nch / (ks * 2) >= min_chunks) ks *= 2;
    // Tiny grids (narrow models: up to 32 row blocks, i.e. a launch that leaves three quarters of the SMs idle whatever happens) with
    // short rows of 2 - 4 chunks: two ways, one or two chunks per warp, so the chain after the dependency wait is one tile shorter
    // (169M-shape ffn value 768 x 3072: 237.8 -> 229.8 us per token).  Wider shapes measured slower with splits this fine.
    if (ks == 1 && row_blocks <= 32 && nch >= 2 && nch <= 4) ks = 2;
    return ks;
}
// a split job can be split inside the CTA unless it emits the q8 block of its outputs (see MatvecBatch::ks_warps)
inline bool matvec_split_in_cta_enabled() {           // RWKV_B200_KSPLIT_CLUSTER_ONLY=1: every split launch takes the cluster form (timing experiments)
    static const bool on = [] { const char * e = getenv("RWKV_B200_KSPLIT_CLUSTER_ONLY"); return !(e && e[0] == '1'); }();
    return on;
}
inline bool matvec_split_in_cta(const MatvecJob & j) { return matvec_split_in_cta_enabled() && j.out_act.f == nullptr; }

// All jobs of a batch must have the same weight type.
cudaError_t launch_matvec(const MatvecBatch & b, cudaStream_t st);
// once per device, outside stream capture (opts the bulk-async kernel in to > 48 KB shared memory)
cudaError_t matvec_async_init();
cudaError_t wkv_init();            // same for the WKV scan kernels
cudaError_t elementwise_init();    // and the sequence layer-norm kernel

// ---- tensor-core MATMUL (sequence mode / wide batches; matmul_tc.cu) ----------------------------
// out[T][M] = epi(X[T][K] * W[M][K]^T) with FP16 operands (quantised weights dequantised to FP16
// on the fly) and FP32 accumulation on tcgen05 / TMEM.  Same epilogues as the mat-vec jobs.
constexpr int MATMUL_MAX_JOBS = 6;
struct MatmulJob {
    DevMat           W;                 // quantised (needs W.tc) or T_F16
    const uint16_t * x16 = nullptr;     // FP16 activations, T rows of W.K values, row stride ldx
    int              ldx = 0;
    float *          out = nullptr; int ldo = 0;
    int              epi = EPI_STORE;
    const float *    aux = nullptr; int ld_aux = 0;
    const float *    vec = nullptr;
    uint16_t *       out_h = nullptr; int ld_h = 0;   // optional FP16 copy of the result (input of a chained MATMUL)
};
bool matmul_tc_enabled();                                // driver entry point found and RWKV_B200_NO_TC != 1
// smallest row count that takes the tensor path: 64 tokens of a sequence, 32 streams of a batch (a batch step re-reads every
// weight whatever its width, so the tensor path pays earlier: 32 v7-2.9B streams 7.74 -> 5.99 ms); RWKV_B200_TC_MIN_ROWS sets both
int  matmul_tc_min_rows(bool batch = false);
bool matmul_tc_supported(const DevMat & W, int T, bool batch = false);
cudaError_t matmul_tc_init();                            // once per device, outside stream capture
// Scratch of the K-split launches (per context: one MATMUL at a time per stream): raw partial tiles + one flag per
// (tile, epilogue warp), zero-initialised.  Without it every launch keeps whole K loops per CTA.
struct MatmulScratch { float * buf = nullptr; unsigned int * flags = nullptr; size_t floats = 0; };
constexpr size_t MATMUL_SCRATCH_FLOATS = (size_t) 74 * 256 * 128, MATMUL_SCRATCH_FLAGS = 74 * 8;
cudaError_t launch_matmul_tc(const MatmulJob * jobs, int n_jobs, int T, cudaStream_t st, const MatmulScratch * scratch = nullptr);   // all jobs share one weight type
cudaError_t launch_to_half(const float * x, uint16_t * out, size_t n, cudaStream_t st);  // n % 4 == 0

// ---- codec -----------------------------------------------------------------------------
// raw ggml rows (device copy of the file bytes, M x K) -> warp-tile layout (tiled_bytes(type, M, K) bytes at `out`)
cudaError_t launch_repack(uint32_t type, const uint8_t * raw, size_t M, size_t K, void * out, cudaStream_t st);
// raw ggml rows -> the tensor-core chunk layout (DevMat::tc)
cudaError_t launch_repack_tc(uint32_t type, const uint8_t * raw, size_t M, size_t K, void * out, cudaStream_t st);
// device layout -> FP32 [M][K]; bit-exact dequantize_row_* (ggml-quants.c:255-363)
cudaError_t launch_dequant(const DevMat & W, float * out, cudaStream_t st);
// device layout -> FP16 [M][K] (operand of the tensor-core sequence MATMUL)
cudaError_t launch_dequant_f16(const DevMat & W, void * out_half, cudaStream_t st);

// ---- embedding + ln0 --------------------------------------------------------------------
// step != nullptr: thread 0 also bumps the tensor-parallel pass counter
cudaError_t launch_embed_ln0(const DevMat & emb, const int * tokens, int T, const float * w, const float * b, float * x, cudaStream_t st, unsigned int * step = nullptr);

// ---- layer norm + token shift + mixes ---------------------------------------------------
enum MixKind : int { MIX_LERP = 0 /* v4/v5: x*mu + (xp - xp*mu) */, MIX_SX = 1 /* v6/v7: x + (xp-x)*mu */ };
constexpr int MIX_MAX_OUT = 6;
struct LnMixArgs {
    const float * x = nullptr;       // [R][D] residual stream
    const float * ln_w = nullptr, * ln_b = nullptr;
    // token shift source: row r uses state_in[(batch ? r : 0)*state_stride + slot ..] when r == 0 or batch,
    // else LN(x[r-1]).  The new shift value LN(x[r]) is stored for the last row (sequence) / every row (batch).
    const float * state_in = nullptr; float * state_out = nullptr;
    size_t state_stride = 0;         // floats between consecutive streams' states
    int    batch = 0;                // 0: rows are consecutive tokens of one stream; 1: rows are independent streams
    int    R = 1, D = 0;
    int    kind = MIX_LERP, n_out = 0;
    const float * mu[MIX_MAX_OUT] = {};
    Act    out[MIX_MAX_OUT];
    float * x0 = nullptr;            // optional [R][D] copy of LN(x)
    float * sx = nullptr;            // optional [R][D] xp - LN(x)
    // fused tensor parallelism (R == 1): the row is first gathered from the peers' partial sums (kernels.h TpWait; `x` is the base row)
    TpFused tp;
    TpWait  wait;
    int     dbg_skip_stats = 0;      // timing experiments only
};
cudaError_t launch_ln_mix(const LnMixArgs & a, cudaStream_t st);

// plain LN of selected rows -> Act (ln_out before the head). row_idx == nullptr: rows 0..R-1.
cudaError_t launch_ln_act(const float * x, int ldx, int R, int D, const float * w, const float * b, Act out, cudaStream_t st);

// v6: x_i = x0 + sx * (maa_i + m_i), i = w,k,v,r,g  (rwkv_graph.inc:312-336)
struct V6Mix2Args { const float * x0, * sx; const float * m[5]; const float * maa[5]; Act out[5]; int R, D; };
cudaError_t launch_v6_mix2(const V6Mix2Args & a, cudaStream_t st);

// Same mixes with the low-rank second stage folded in (long sequences on the tensor-core path):
//   m_i[t][e] = sum_j lora[t][i*32 + j] * W2_i[e][j]   (rwkv_graph.inc:312-324; rank 32 per mix),  out_i = fp16(x0 + sx * (maa_i + m_i))
// so the five [R][D] FP32 products are never written to memory.  W2_i: FP32 [D][32] row-major.
struct V6MixLoraArgs { const float * x0, * sx, * lora; int ld_lora; const float * w2[5]; const float * maa[5]; uint16_t * out_h[5]; int R, D; };
cudaError_t launch_v6_mix_lora(const V6MixLoraArgs & a, cudaStream_t st);

// standalone activation quantiser (when a producer could not fuse it)
cudaError_t launch_quantize_act(Act a, int R, cudaStream_t st);

// ---- WKV recurrences --------------------------------------------------------------------
struct WkvArgs {
    int R = 1, D = 0, H = 0, S = 0, batch = 0;
    int state_D = 0, chan0 = 0;           // v4 tensor parallel: the rows hold D local channels [chan0, chan0 + D) of state slots state_D wide (0 = D)
    size_t state_stride = 0;
    const float * state_in = nullptr; float * state_out = nullptr;   // pointers at this layer's att state slot(s)
    const float * r = nullptr, * k = nullptr, * v = nullptr, * w = nullptr, * g = nullptr, * a = nullptr;   // [R][D]
    const float * u = nullptr;            // time_first / time_faaaa  ([D] or [H] when per-head scalar)
    const float * decay = nullptr;        // static decay ([D] or [H]); v6 uses per-token w instead
    int   per_head_scalar = 0;            // v5.1: u, decay are one value per head
    const float * lnx_w = nullptr, * lnx_b = nullptr;
    float gn_eps = 1e-5f;
    // v7 extras
    const float * k_k = nullptr, * k_a = nullptr, * r_k = nullptr, * vgate = nullptr;
    float * v_first = nullptr;            // [R][D]; written when layer == 0, read otherwise
    int   layer = 0;
    Act   y;                              // output rows
    float * yraw = nullptr;               // [R][D] scratch: raw scan outputs of the sequence kernels (before the head norm)
    // chunk-parallel form (wkv_chunk.cu): scratch for ceil(R/16) chunks, [H][chunks][S][S] and [H][chunks][S];
    // both set => long sequences of head-size-64 models take it (tolerance mode, like the tensor-core MATMUL)
    float * chunk_ds = nullptr, * chunk_pc = nullptr;
};
constexpr int WKV_SEQ_MIN_ROWS = 8;       // sequences at least this long use the scan kernels (same arithmetic, different work split)
cudaError_t launch_wkv4(const WkvArgs & a, cudaStream_t st);   // state slots: aa, bb, pp at state + {0,1,2}*D
cudaError_t launch_wkv5(const WkvArgs & a, cudaStream_t st);   // v5.1 / v5.2 / v6 (w != nullptr => per-token decay)
cudaError_t launch_wkv7(const WkvArgs & a, cudaStream_t st);
bool wkv7_chunked_supported(const WkvArgs & a);                // head size 64, one stream, >= 64 rows, scratch present ([H][ceil(R/32)][64][64] floats x 2)
constexpr int WKV7_CHUNK_TOKENS = 32;
bool wkv5_chunked_supported(const WkvArgs & a);                // head size 64, one stream, scratch present, RWKV_B200_NO_WKV_CHUNK != 1
cudaError_t launch_wkv5_chunked(const WkvArgs & a, cudaStream_t st);
constexpr int WKV_CHUNK_TOKENS = 16;

// ---- on-device sampling (sampling.cu; python/sampling.py:10-52 of the reference) ---------------------
// one token id per row of logits [rows][V]; probs = scratch [rows][V]; u = one uniform in [0, 1) per row (device)
cudaError_t launch_sample(const float * logits, float * probs, int rows, int V, float temperature, float top_p, const float * u, uint32_t * out, cudaStream_t st);

// tensor parallel ffn tail: x[t][m] += recv[t][D + m] * recv[t][m]  (recv = all-reduced [partial Wv k | r shards], 2*D per row)
cudaError_t launch_tp_ffn_combine(float * x, const float * recv, int R, int D, cudaStream_t st);
// fused tensor parallelism: materialise the residual row of a site outside a mat-vec prologue (before ln_out)
cudaError_t launch_tp_gather_row(const TpFused & tp, const TpWait & w, const float * base, cudaStream_t st);

} // namespace rb
