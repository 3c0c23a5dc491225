// Quantised mat-vec, bulk-async staged: the HBM-bound decode kernel.
//
// Same arithmetic as the CPU reference (q8 activation blocks, exact integer block dots,
// FP32 scale and accumulate -- ggml-cpu-quants.c:1741,2321,2614,2969,3343; activations
// ggml-cpu-quants.c:783-850,1083-1150); what is new is the data movement:
//   * weights live in the warp-tile layout (common.h): the next 4 rows x 1024 weights of a
//     warp are ONE contiguous run, fetched with one cp.async.bulk (1-D TMA copy, SASS UBLKCP)
//     that completes on an mbarrier;
//   * every warp owns a private ring of GA_STAGES tiles in shared memory and keeps
//     GA_STAGES-1 copies in flight while it computes, so the bytes in flight per SM
//     (16-24 warps x 3 tiles x 2.3-4.3 KB) are set by shared memory, not by registers;
//   * lanes read lane-consecutive 16-byte nibble groups / 4-byte qh / 2-4-byte scales from
//     the tile (conflict-free), unpack with mask/shift, and use dp4a against the q8
//     activation block (L1-resident, loaded once per lane for the 4 rows).
// One CTA = 8 warps = 32 consecutive rows, so the epilogue can also emit the q8 activation
// block of its 32 outputs for a chained mat-vec.  Per output the FP32 summation order is
// lane-strided over K + butterfly: it depends on K only, never on T or the grid.
#include "matvec_common.cuh"
#include "launch.cuh"
#include "ptx.cuh"
#include "tp_ll.cuh"
#include "trace.cuh"
#include <cstdlib>
#include <algorithm>

namespace rb {

constexpr int GA_WARPS = 8, GA_R = TILE_ROWS, GA_KC = TILE_BLOCKS;
constexpr int GA_KSPLIT_MAX = 4;

template <int TYPE> struct QFmt;
template <> struct QFmt<T_Q4_0> { static constexpr int QB = 16, HB = 0, DB = 2, STAGES = 4; };
template <> struct QFmt<T_Q4_1> { static constexpr int QB = 16, HB = 0, DB = 4, STAGES = 4; };
template <> struct QFmt<T_Q5_0> { static constexpr int QB = 16, HB = 4, DB = 2, STAGES = 4; };
template <> struct QFmt<T_Q5_1> { static constexpr int QB = 16, HB = 4, DB = 4, STAGES = 4; };
template <> struct QFmt<T_Q8_0> { static constexpr int QB = 32, HB = 0, DB = 2, STAGES = 3; };

template <int TYPE> __host__ __device__ constexpr int ga_stage_bytes() { return GA_R * GA_KC * (QFmt<TYPE>::QB + QFmt<TYPE>::HB + QFmt<TYPE>::DB); }
template <int TYPE, int S> __host__ __device__ constexpr int ga_smem_bytes() { return GA_WARPS * S * (ga_stage_bytes<TYPE>() + 8); }

// Per-lane cursor into the q8 activation rows (block `lane` of column col0 of the current chunk).  The pointers are
// loop-carried: an iteration costs one add per plane, nothing is re-derived from the kernel parameters.
template <int TYPE, int NCOL>
struct XCursor {
    static constexpr bool HAS_S = (TYPE == T_Q4_1 || TYPE == T_Q5_1), HAS_I = (TYPE == T_Q4_0 || TYPE == T_Q5_0);
    const int8_t * q; const float * d; const float * s; const int * isum;
    int K, nbrow, tmax;
    // positioned at chunk `chunk` of the row
    __device__ __forceinline__ void init(const Act & x, int x_off, int col0, int T, int lane, int chunk) {
        K = x.K; nbrow = x.K >> 5; tmax = T - 1 - col0;
        q = x.q + (size_t) col0 * K + x_off + lane * 32 + (size_t) chunk * (GA_KC * 32);
        const size_t b0 = (size_t) col0 * nbrow + (x_off >> 5) + lane + (size_t) chunk * GA_KC;
        d = x.d + b0; s = x.s + b0; isum = x.isum + b0;
        pin();
    }
    // keeps the pointers in registers (the compiler otherwise rebuilds them from constant memory every iteration)
    __device__ __forceinline__ void pin() {
        asm volatile("" : "+l"(q), "+l"(d));
        if (HAS_S) asm volatile("" : "+l"(s));
        if (HAS_I) asm volatile("" : "+l"(isum));
    }
    __device__ __forceinline__ void next() { q += GA_KC * 32; d += GA_KC; s += GA_KC; isum += GA_KC; pin(); }
    // Fire-and-forget loads of this lane's share of the next `n` chunks: pulls the whole activation
    // row into L1 in one round trip.  Under a saturated L2 a dependent load per chunk costs a
    // microsecond each; this way the later load() calls hit L1 (or merge with the miss in flight).
    // nb_left = blocks of the row from the current chunk on
    __device__ __forceinline__ void warm(int n, int nb_left, int lane) const {
#pragma unroll
        for (int cc = 0; cc < NCOL; cc++) {
            const int dt = NCOL == 1 ? 0 : min(cc, tmax);
#pragma unroll 1
            for (int c = 0; c < n; c++) {
                if (c * GA_KC + lane < nb_left) {
                    uint32_t t0, t1;
                    asm volatile("ld.global.ca.u32 %0, [%1];" : "=r"(t0) : "l"(q + (size_t) dt * K + c * (GA_KC * 32)));
                    asm volatile("ld.global.ca.u32 %0, [%1];" : "=r"(t1) : "l"(d + dt * nbrow + c * GA_KC));
                    if (HAS_S) asm volatile("ld.global.ca.u32 %0, [%1];" : "=r"(t1) : "l"(s + dt * nbrow + c * GA_KC));
                    if (HAS_I) asm volatile("ld.global.ca.u32 %0, [%1];" : "=r"(t1) : "l"(isum + dt * nbrow + c * GA_KC));
                }
            }
        }
    }
    __device__ __forceinline__ void load(XBlock<NCOL> & xb) const {             // this lane's block of the current chunk
#pragma unroll
        for (int cc = 0; cc < NCOL; cc++) {
            const int dt = NCOL == 1 ? 0 : min(cc, tmax);
            const int8_t * qp = q + (size_t) dt * K;
            const uint4 lo = *reinterpret_cast<const uint4 *>(qp);
            const uint4 hi = *reinterpret_cast<const uint4 *>(qp + 16);
            xb.w[cc][0] = lo.x; xb.w[cc][1] = lo.y; xb.w[cc][2] = lo.z; xb.w[cc][3] = lo.w;
            xb.w[cc][4] = hi.x; xb.w[cc][5] = hi.y; xb.w[cc][6] = hi.z; xb.w[cc][7] = hi.w;
            const int bo = dt * nbrow;
            xb.d[cc] = d[bo];
            if (HAS_S) xb.s[cc] = s[bo];
            if (HAS_I) xb.isum[cc] = isum[bo];
        }
    }
};

// ---- decode prologues (kernels.h MatvecPrologue) ---------------------------------------------------
// The input vector of the launch as q8 blocks in shared memory (what the chunk loop otherwise
// reads from the Act planes in global memory).
template <bool ON, int PIT = 8>
struct ProXT {
    static constexpr int N = ON ? PIT * 256 : 32;            // kernels without a prologue carry a stub
    __align__(16) int8_t q[N];
    float d[N / 32], s[N / 32];
    int   isum[N / 32];
};
constexpr int PRO_NT = GA_WARPS * 32;                        // thread `tid` holds elements it * 256 + tid, it < ITEMS

// values of elements (it0 + g)*256 + tid, g < G -> q8 blocks ((it0 + g)*8 + warp) in shared memory; the blocks of a
// group are reduced together so their shuffle chains overlap (same values as act_store_block)
template <int G, typename PX>
__device__ __forceinline__ void pro_store_group(PX & px, const float (&v)[G], int it0, int D) {
    const int tid = threadIdx.x;
    float amax[G]; int q[G], sum[G];
#pragma unroll
    for (int g = 0; g < G; g++) amax[g] = (it0 + g) * PRO_NT + tid < D ? fabsf(v[g]) : 0.0f;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1)
#pragma unroll
        for (int g = 0; g < G; g++) amax[g] = fmaxf(amax[g], __shfl_xor_sync(FULL, amax[g], off));
#pragma unroll
    for (int g = 0; g < G; g++) {
        const float id = amax[g] != 0.0f ? __fdiv_rn(127.0f, amax[g]) : 0.0f;
        q[g] = (it0 + g) * PRO_NT + tid < D ? __float2int_rn(__fmul_rn(v[g], id)) : 0;
        sum[g] = q[g];
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1)
#pragma unroll
        for (int g = 0; g < G; g++) sum[g] += __shfl_xor_sync(FULL, sum[g], off);
#pragma unroll
    for (int g = 0; g < G; g++) {
        const int e = (it0 + g) * PRO_NT + tid;
        if (e - (tid & 31) < D) {                            // warps past the row own no block (uniform)
            px.q[e] = (int8_t) q[g];
            if ((tid & 31) == 0) {
                const float d = __fdiv_rn(amax[g], 127.0f);
                px.d[e >> 5] = round_f16(d); px.s[e >> 5] = round_f16(__fmul_rn(d, (float) sum[g])); px.isum[e >> 5] = sum[g];
            }
        }
    }
}

// part 1 (before pdl_wait): constants and the read-only input state.  Rows wider than LN_SMALL_D fetch
// them after the wait instead (they come from L2 and the registers are needed for the row itself).
template <int ITEMS> struct ProLnRegs { float xs[ITEMS <= 8 ? ITEMS : 1], w[ITEMS <= 8 ? ITEMS : 1], b[ITEMS <= 8 ? ITEMS : 1], mu[ITEMS <= 8 ? ITEMS : 1]; };
template <int ITEMS>
__device__ __forceinline__ void pro_lnmix_pre(const MatvecPrologue & P, const float * __restrict__ mu, ProLnRegs<ITEMS> & r) {
    if constexpr (ITEMS <= 8) {
#pragma unroll
        for (int it = 0; it < ITEMS; it++) {
            const int e = it * PRO_NT + threadIdx.x;
            const bool valid = e < P.D;
            r.xs[it] = valid ? P.state_in[e] : 0.0f;
            r.w[it] = valid ? P.ln_w[e] : 0.0f; r.b[it] = valid ? P.ln_b[e] : 0.0f;
            r.mu[it] = valid ? mu[e] : 0.0f;
        }
    }
}
// part 2 (after pdl_wait): the row itself; same arithmetic as ln_mix_kernel<., ., 256> (elementwise.cu)
template <int ITEMS, typename PX>
__device__ __forceinline__ void pro_lnmix_post(const MatvecPrologue & P, const TpFused & tpf, const float * __restrict__ mu, const ProLnRegs<ITEMS> & r, PX & px, double (*red)[32] TRACE_ARG) {
    float x[1][ITEMS];
    if (P.wait.site >= 0) {                                  // fused tensor parallelism: the row is still in flight from the peers
        const uint2 * inbox = tp_inbox(tpf, P.wait.site);
        const unsigned want = tp_flag(tpf, P.wait.site);
#pragma unroll
        for (int it = 0; it < ITEMS; it++) {
            const int e = it * PRO_NT + threadIdx.x;
            x[0][it] = 0.0f;
            if (e < P.D) {
                x[0][it] = tp_gather_element(tpf, P.wait, inbox, want, P.x[e], e);
                if (blockIdx.x == 0) P.wait.x_out[e] = x[0][it];
            }
        }
    } else {
#pragma unroll
        for (int it = 0; it < ITEMS; it++) {
            const int e = it * PRO_NT + threadIdx.x;
            x[0][it] = e < P.D ? P.x[e] : 0.0f;
        }
    }
    float scale[1];
    ln_rows_stats<ITEMS, 1>(x, P.D, PRO_NT, red, scale, 1e-5f TRACE_PASS);
    constexpr int G = 4;
#pragma unroll
    for (int it0 = 0; it0 < ITEMS; it0 += G) {
        float v[G];
#pragma unroll
        for (int g = 0; g < G; g++) {
            const int it = it0 + g, e = it * PRO_NT + threadIdx.x;
            float xs, w, b, m;
            if constexpr (ITEMS <= 8) { xs = r.xs[it]; w = r.w[it]; b = r.b[it]; m = r.mu[it]; }
            else {
                const bool valid = e < P.D;
                xs = valid ? P.state_in[e] : 0.0f; w = valid ? P.ln_w[e] : 0.0f; b = valid ? P.ln_b[e] : 0.0f; m = valid ? mu[e] : 0.0f;
            }
            const float x0 = __fadd_rn(__fmul_rn(__fmul_rn(x[0][it], scale[0]), w), b);
            if (P.mix_kind == MIX_LERP) v[g] = __fadd_rn(__fmul_rn(x0, m), __fsub_rn(xs, __fmul_rn(xs, m)));
            else                        v[g] = __fadd_rn(__fmul_rn(__fsub_rn(xs, x0), m), x0);
            if (e < P.D && blockIdx.x == 0) P.state_out[e] = x0;
        }
        pro_store_group<G>(px, v, it0, P.D);
    }
}

template <int ITEMS> struct ProWkvRegs { float aa[ITEMS <= 8 ? ITEMS : 1], bb[ITEMS <= 8 ? ITEMS : 1], pp[ITEMS <= 8 ? ITEMS : 1], u[ITEMS <= 8 ? ITEMS : 1], w[ITEMS <= 8 ? ITEMS : 1]; };
template <int ITEMS>
__device__ __forceinline__ void pro_wkv4_pre(const MatvecPrologue & P, ProWkvRegs<ITEMS> & r) {
    if constexpr (ITEMS <= 8) {
        const int sd = P.state_D ? P.state_D : P.D;
#pragma unroll
        for (int it = 0; it < ITEMS; it++) {
            const int e = it * PRO_NT + threadIdx.x;
            const bool valid = e < P.D;
            r.aa[it] = valid ? P.state_in[e] : 0.0f; r.bb[it] = valid ? P.state_in[sd + e] : 0.0f; r.pp[it] = valid ? P.state_in[2 * sd + e] : 0.0f;
            r.u[it] = valid ? P.u[e] : 0.0f; r.w[it] = valid ? P.decay[e] : 0.0f;
        }
    }
}
// same arithmetic as wkv4_kernel (wkv.cu; rwkv_operators.inc / rwkv_graph.inc v4 att)
template <int ITEMS, typename PX>
__device__ __forceinline__ void pro_wkv4_post(const MatvecPrologue & P, const ProWkvRegs<ITEMS> & g, PX & px) {
    constexpr int G = 4;
    const int sd = P.state_D ? P.state_D : P.D;
#pragma unroll
    for (int it0 = 0; it0 < ITEMS; it0 += G) {
        float k[G], v[G], r[G], out[G], aa[G], bb[G], pp[G], u[G], w[G];
#pragma unroll
        for (int q = 0; q < G; q++) {
            const int it = it0 + q, e = it * PRO_NT + threadIdx.x;
            const bool valid = e < P.D;
            k[q] = valid ? P.k[e] : 0.0f; v[q] = valid ? P.v[e] : 0.0f; r[q] = valid ? P.r[e] : 0.0f;
            if constexpr (ITEMS <= 8) { aa[q] = g.aa[it]; bb[q] = g.bb[it]; pp[q] = g.pp[it]; u[q] = g.u[it]; w[q] = g.w[it]; }
            else {
                aa[q] = valid ? P.state_in[e] : 0.0f; bb[q] = valid ? P.state_in[sd + e] : 0.0f; pp[q] = valid ? P.state_in[2 * sd + e] : 0.0f;
                u[q] = valid ? P.u[e] : 0.0f; w[q] = valid ? P.decay[e] : 0.0f;
            }
        }
#pragma unroll
        for (int q = 0; q < G; q++) {
            const int e = (it0 + q) * PRO_NT + threadIdx.x;
            float ww = __fadd_rn(u[q], k[q]);
            float qq = fmaxf(pp[q], ww);
            float e1 = expf(__fsub_rn(pp[q], qq)), e2 = expf(__fsub_rn(ww, qq));
            const float num = __fadd_rn(__fmul_rn(e1, aa[q]), __fmul_rn(e2, v[q]));
            const float den = __fadd_rn(__fmul_rn(e1, bb[q]), e2);
            ww = __fadd_rn(pp[q], w[q]);
            qq = fmaxf(ww, k[q]);
            e1 = expf(__fsub_rn(ww, qq)); e2 = expf(__fsub_rn(k[q], qq));
            const float na = __fadd_rn(__fmul_rn(e1, aa[q]), __fmul_rn(e2, v[q]));
            const float nb = __fadd_rn(__fmul_rn(e1, bb[q]), e2);
            out[q] = e < P.D ? __fmul_rn(r[q], __fdiv_rn(num, den)) : 0.0f;
            if (e < P.D && blockIdx.x == 0) { P.state_out[e] = na; P.state_out[sd + e] = nb; P.state_out[2 * sd + e] = qq; }
        }
        pro_store_group<G>(px, out, it0, P.D);
    }
}

static_assert(sizeof(MatvecBatch) <= 4096, "kernel parameter space");

template <int TYPE, int NCOL, int S, int PRO = PRO_NONE, int PIT = 4>
__global__ void __launch_bounds__(GA_WARPS * 32, NCOL == 1 ? ((QFmt<TYPE>::QB == 32 || (PRO != PRO_NONE && PIT >= 16)) ? 2 : 3) : (NCOL == 2 ? 2 : 1))
matvec_q_kernel(const __grid_constant__ MatvecBatch B) {
    static_assert(PRO == PRO_NONE || NCOL == 1, "prologues are single-row");
    using F = QFmt<TYPE>;
    constexpr int BB = F::QB + F::HB + F::DB;              // bytes per block
    constexpr int STAGE = ga_stage_bytes<TYPE>();
    constexpr bool HAS_M = (TYPE == T_Q4_1 || TYPE == T_Q5_1);
    constexpr bool PREFETCH_X = (NCOL == 1);               // next chunk's activations one iteration ahead
    extern __shared__ __align__(128) uint8_t ga_smem[];
    __shared__ float tile[NCOL][GA_WARPS * GA_R];
    __shared__ ProXT<PRO != PRO_NONE, PIT> pro_xs;
    __shared__ double pro_red[PRO == PRO_LNMIX ? 2 : 1][32];

    TRACE_DECL; TRACE(0);
    pdl_trigger();
    int j = 0;
#pragma unroll 1
    while (j + 1 < B.n_jobs && (int) blockIdx.x >= B.cta_begin[j + 1]) j++;
    const MatvecJob & J = B.job[j];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int M = J.W.M, nb = J.W.K >> 5, T = B.T;
    // K split: the `ks` CTAs of a cluster share one block of 32 rows, CTA kr takes a contiguous range
    // of the row's chunks; partial sums meet in the leader's shared memory (see the tail).
    // ... or, B.ks_warps: the split happens inside the CTA -- its 8 warps cover 32 / ks rows, warp w takes K range w % ks of row
    // group w / ks; the partial sums meet in this CTA's shared memory, in the same rank order.
    const bool ksw = B.ks_warps != 0;
    const int ks = ksw ? J.ks : B.ks, local = (int) blockIdx.x - B.cta_begin[j];
    const int kr = ks > 1 ? (ksw ? warp % ks : local % ks) : 0;
    const int tile_row0 = ksw ? local * (GA_WARPS * GA_R / ks) : (ks > 1 ? local / ks : local) * GA_WARPS * GA_R;
    const int row0 = tile_row0 + (ksw ? warp / ks : warp) * GA_R;
    const int col0 = blockIdx.y * NCOL;
    const int nch_all = (nb + GA_KC - 1) / GA_KC;
    const int cbase = ks > 1 ? kr * nch_all / ks : 0;
    const int cend  = ks > 1 ? (kr + 1) * nch_all / ks : nch_all;
    const int nch = row0 < M ? cend - cbase : 0;                 // warps past the last row stay idle

    uint8_t * my = ga_smem + (size_t) warp * S * STAGE;
    const uint32_t my_s = smem_u32(my);
    const uint32_t bars = smem_u32(ga_smem + (size_t) GA_WARPS * S * STAGE) + warp * S * 8;
    const uint8_t * src0 = reinterpret_cast<const uint8_t *>(J.W.qs) + (size_t) (row0 / GA_R) * J.W.rg_stride + (size_t) cbase * STAGE;
    uint32_t last_bytes = (uint32_t) ((GA_R * (nb - (nch_all - 1) * GA_KC) * BB + 15) & ~15);   // the row's final (possibly short) chunk
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < S; s++) mbar_init(bars + s * 8, 1);
        mbar_init_fence();
        fence_proxy_async_smem();
        // weights do not depend on the preceding kernel: start streaming at once
#pragma unroll
        for (int s = 0; s < S; s++) {
            if (s < nch) {
                const uint32_t bytes = (cbase + s == nch_all - 1) ? last_bytes : (uint32_t) STAGE;
                mbar_expect_tx(bars + s * 8, bytes);
                bulk_g2s(my_s + s * STAGE, src0 + (size_t) s * STAGE, bytes, bars + s * 8);
            }
        }
    }
    if (B.n_pf > 0 && lane == 0) {                       // this warp's slice of the next launch's weights -> L2 (see MatvecBatch)
        const unsigned nw = gridDim.x * gridDim.y * GA_WARPS, wid = (blockIdx.y * gridDim.x + blockIdx.x) * GA_WARPS + warp;
        for (int p = 0; p < B.n_pf; p++) {
            const unsigned per = ((B.pf_bytes[p] + nw - 1) / nw + 15u) & ~15u, off = wid * per;
            if (off < B.pf_bytes[p]) {
                const unsigned len = min(per, B.pf_bytes[p] - off) & ~15u;
                if (len) bulk_prefetch_l2(reinterpret_cast<const uint8_t *>(B.pf_ptr[p]) + off, len);
            }
        }
    }
    XCursor<TYPE, NCOL> xc;
    if (PRO == PRO_NONE) xc.init(J.x, J.x_off, col0, T, lane, cbase);
    ProLnRegs<PIT> pln; ProWkvRegs<PIT> pwk;
    if (PRO == PRO_LNMIX) pro_lnmix_pre<PIT>(B.pro, J.pro_mu, pln);
    if (PRO == PRO_WKV4)  pro_wkv4_pre<PIT>(B.pro, pwk);
    __syncwarp();
    TRACE(1);
    pdl_wait();                                         // activations (and the residual) come from the preceding kernels
    TRACE(2);
    if constexpr (PRO != PRO_NONE) {
        if (PRO == PRO_LNMIX) pro_lnmix_post<PIT>(B.pro, B.tp, J.pro_mu, pln, pro_xs, pro_red TRACE_PASS);
        if (PRO == PRO_WKV4)  pro_wkv4_post<PIT>(B.pro, pwk, pro_xs);
        __syncthreads();
        xc.q = pro_xs.q + lane * 32; xc.d = pro_xs.d + lane; xc.s = pro_xs.s + lane; xc.isum = pro_xs.isum + lane;
        xc.K = 0; xc.nbrow = 0; xc.tmax = 0;
    }

    TRACE(3);
    // The epilogue's operands, one (row, column) per lane.  Fetched up front in the variants that run
    // the small, latency-bound grids (4-stage ring / prologue kernels); the 3-stage variants stream
    // large matrices at the register limit and fetch them in the tail instead.
    constexpr bool EPI_EARLY = (S == 4) || PRO != PRO_NONE;
    EpiOperands epre;
    if (EPI_EARLY && lane < GA_R * NCOL && row0 + lane / NCOL < M && col0 + lane % NCOL < T)
        epre = load_epilogue_operands(J, col0 + lane % NCOL, row0 + lane / NCOL);

    RowAcc<TYPE, NCOL> ra[GA_R];
#pragma unroll
    for (int r = 0; r < GA_R; r++) ra[r].clear();

    XBlock<NCOL> xb, xn;
    int left = nb - cbase * GA_KC;                       // blocks of the row from the current chunk on
    if (PRO == PRO_NONE && NCOL <= 2 && nch > 2) xc.warm(nch, left, lane);
    if (PREFETCH_X && nch > 0 && lane < left) xc.load(xb);
    // the refill cursor: source of chunk c + S, and the iteration whose refill is the row's final (short) chunk
    const uint8_t * src_next = src0 + (size_t) S * STAGE;
    const int c_last = nch_all - 1 - cbase - S;
    asm volatile("" : "+r"(last_bytes), "+l"(src_next));
    int s = 0;
    uint32_t phase = 0;
#pragma unroll 1
    for (int c = 0; c < nch; c++) {                      // c counts this CTA's chunks; cbase + c is the chunk of the row
        const int cnt = min(GA_KC, left);
        const bool act = lane < cnt;
        if (PREFETCH_X) { xc.next(); if (c + 1 < nch && lane < left - GA_KC) xc.load(xn); }
        else { if (act) xc.load(xb); xc.next(); }       // before the wait: independent of the tile
        left -= GA_KC;
        mbar_wait(bars + s * 8, phase);

        // the warp's 4 rows x (this lane's block) from the stage into registers, then hand the
        // stage back to the copy engine before doing the arithmetic
        const uint8_t * st = my + (size_t) s * STAGE;
        uint4 q0[GA_R], q1[GA_R];
        uint32_t qh[GA_R], dm[GA_R];
        if (cnt == GA_KC) {
            const uint8_t * pq = st + lane * F::QB;
            const uint8_t * ph = st + GA_R * GA_KC * F::QB + lane * 4;
            const uint8_t * pd = st + GA_R * GA_KC * (F::QB + F::HB) + lane * F::DB;
#pragma unroll
            for (int r = 0; r < GA_R; r++) {
                q0[r] = *reinterpret_cast<const uint4 *>(pq + r * GA_KC * F::QB);
                if (F::QB == 32) q1[r] = *reinterpret_cast<const uint4 *>(pq + r * GA_KC * F::QB + 16);
                if (F::HB) qh[r] = *reinterpret_cast<const uint32_t *>(ph + r * GA_KC * 4);
                if (HAS_M) dm[r] = *reinterpret_cast<const uint32_t *>(pd + r * GA_KC * 4);
                else       dm[r] = *reinterpret_cast<const uint16_t *>(pd + r * GA_KC * 2);
            }
        } else if (act) {
#pragma unroll
            for (int r = 0; r < GA_R; r++) {
                const int e = r * cnt + lane;
                q0[r] = *reinterpret_cast<const uint4 *>(st + (size_t) e * F::QB);
                if (F::QB == 32) q1[r] = *reinterpret_cast<const uint4 *>(st + (size_t) e * F::QB + 16);
                if (F::HB) qh[r] = *reinterpret_cast<const uint32_t *>(st + GA_R * cnt * F::QB + e * 4);
                const uint8_t * dp = st + GA_R * cnt * (F::QB + F::HB) + e * F::DB;
                if (HAS_M) dm[r] = *reinterpret_cast<const uint32_t *>(dp);
                else       dm[r] = *reinterpret_cast<const uint16_t *>(dp);
            }
        }
        __syncwarp();                                   // every lane has read the stage
        if (lane == 0 && c + S < nch) {
            const uint32_t bytes = c == c_last ? last_bytes : (uint32_t) STAGE;
            mbar_expect_tx(bars + s * 8, bytes);
            bulk_g2s(my_s + s * STAGE, src_next, bytes, bars + s * 8);
        }
        src_next += STAGE;
        if (act) {
#pragma unroll
            for (int r = 0; r < GA_R; r++) {
                float dw, mw = 0.0f;
                if (HAS_M) { const __half2 h2 = *reinterpret_cast<const __half2 *>(&dm[r]); dw = __low2float(h2); mw = __high2float(h2); }
                else dw = __half2float(__ushort_as_half((unsigned short) dm[r]));
                dot_block<TYPE, NCOL>(ra[r], xb, q0[r], F::QB == 32 ? q1[r] : make_uint4(0, 0, 0, 0), F::HB ? qh[r] : 0u, dw, mw);
            }
        }
        if (PREFETCH_X) xb = xn;
        if (++s == S) { s = 0; phase ^= 1; }
    }

    TRACE(4);
    // ---- warp totals; lane (r*NCOL + c) finishes output (row0 + r, col0 + c) ----------------------
    float mine = 0.0f;
#pragma unroll
    for (int r = 0; r < GA_R; r++)
#pragma unroll
        for (int c = 0; c < NCOL; c++) {
            float v = warp_sum(ra[r].acc[c]);
            if (HAS_M) v = __fadd_rn(v, warp_sum(ra[r].summ[c]));
            if (lane == r * NCOL + c) mine = v;
        }
    __shared__ float part[GA_KSPLIT_MAX][GA_WARPS * GA_R * NCOL];
    if (ks > 1 && ksw) {
        // partial sums of the ks warps of a row group -> its first warp, summed there in rank order (no q8 epilogue in this form:
        // the other warps are done)
        float * pw = &part[0][0];
        if (lane < GA_R * NCOL) pw[warp * GA_R * NCOL + lane] = mine;
        if (!EPI_EARLY && kr == 0 && lane < GA_R * NCOL && row0 + lane / NCOL < M && col0 + lane % NCOL < T)
            epre = load_epilogue_operands(J, col0 + lane % NCOL, row0 + lane / NCOL);      // in flight across the barrier
        __syncthreads();
        if (kr != 0) return;
        if (lane < GA_R * NCOL) {
            float tsum = mine;
            for (int k = 1; k < ks; k++) tsum = __fadd_rn(tsum, pw[(warp + k) * GA_R * NCOL + lane]);
            mine = tsum;
        }
    } else if (ks > 1) {
        // partial sums of the cluster -> leader (DSMEM), summed there in rank order
        const int slot = warp * GA_R * NCOL + min(lane, GA_R * NCOL - 1);
        uint32_t remote;
        asm volatile("mapa.shared::cluster.u32 %0, %1, 0;" : "=r"(remote) : "r"(smem_u32(&part[kr][slot])));
        if (lane < GA_R * NCOL) asm volatile("st.shared::cluster.f32 [%0], %1;" :: "r"(remote), "f"(mine) : "memory");
        asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
        // the leader's epilogue operands travel while the cluster gathers (they were an L2 round trip after the barrier)
        if (!EPI_EARLY && kr == 0 && lane < GA_R * NCOL && row0 + lane / NCOL < M && col0 + lane % NCOL < T)
            epre = load_epilogue_operands(J, col0 + lane % NCOL, row0 + lane / NCOL);
        asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
        if (kr != 0) return;
        if (lane < GA_R * NCOL) {
            float tsum = part[0][slot];
            for (int k = 1; k < ks; k++) tsum = __fadd_rn(tsum, part[k][slot]);
            mine = tsum;
        }
    }
    if (NCOL == 1 && (J.epi == EPI_TP_PUSH || J.epi == EPI_TP_PUSH_SIGMOID)) {        // uniform per CTA
        // fused all-reduce, producer side: lane = (row r, peer p) stores {value, flag} into slot `rank` of peer p's inbox
        static_assert(GA_R * TP_LL_WORLD == 32, "one lane per (row, peer)");
        float val = __shfl_sync(FULL, mine, lane & (GA_R - 1));
        if (J.epi == EPI_TP_PUSH_SIGMOID) val = sigmoidf_ref(val);
        const int m = row0 + (lane & (GA_R - 1)), peer = lane / GA_R;
        if (m < M && peer < B.tp.world) tp_push(B.tp, B.tp.site_push, tp_flag(B.tp, B.tp.site_push), peer, (size_t) J.tp_off + m, val);
        TRACE(5);
        TRACE_FLUSH(PRO | (nch_all << 8) | ((unsigned long long) M << 16), gridDim.x);
        return;
    }
    const bool want_q = (J.out_act.f != nullptr);
    if (lane < GA_R * NCOL) {
        const int r = lane / NCOL, c = lane % NCOL;
        const int m = row0 + r, t = col0 + c;
        float v = 0.0f;
        if (m < M && t < T) {
            if (!EPI_EARLY && ks == 1) epre = load_epilogue_operands(J, t, m);
            v = apply_epilogue_pre(J, mine, epre);
            if (J.out) J.out[(size_t) t * J.ldo + m] = v;
        }
        if (want_q) tile[c][warp * GA_R + r] = v;
    }
    if (want_q) {                                       // uniform per CTA
        __syncthreads();
        if (warp < NCOL && col0 + warp < T) {
            const int m = tile_row0 + lane;
            act_store_block(J.out_act, (size_t) (col0 + warp), m, m < M, tile[warp][lane]);
        }
    }
    TRACE(5);
    TRACE_FLUSH(PRO | (nch_all << 8) | ((unsigned long long) M << 16), gridDim.x);
}

} // namespace rb
TRACE_EXPORT(rwkv_b200_debug_trace)
namespace rb {

static int ga_stage_override() {
    static const int v = [] { const char * e = getenv("RWKV_B200_GA_STAGES"); return e ? atoi(e) : 0; }();
    return v;
}

template <int TYPE, int NCOL>
static cudaError_t launch_async_n(const MatvecBatch & b, int total_ctas, int ygrid, cudaStream_t st) {
    // Ring depth: 3 stages let three CTAs share an SM (24 warps hide the unpack latency) and win
    // once the grid fills them (>= 3 x 148 CTAs); smaller grids run one or two CTAs per SM and
    // want the deeper ring for bytes in flight.  Q8_0 tiles are 4.25 KB: 3 stages either way.
    // The depth never changes the arithmetic.
    // (14B shape, 160-CTA att.output launches: 3 stages measured 2 % faster per token than 4; the deep ring only pays below one CTA per SM)
    int S = (QFmt<TYPE>::QB == 32 || total_ctas * ygrid >= device_sm_count()) ? 3 : 4;
    if (ga_stage_override() == 3 || ga_stage_override() == 4) S = ga_stage_override();
    const int cl = b.ks > 1 ? b.ks : 0;
    // Measured and not kept (14B Q4_1 decode, 2.32 ms per token with the rule above): a 5-stage whole-row ring at 2 CTAs per SM
    // 3.02 ms; a 2-stage ring with 64 registers at 4 CTAs per SM (the 480-CTA att k / v / r launch then fits in one wave) 2.57 ms
    // -- the layer-norm CTAs that share the SMs with the waiting mat-vec CTAs slow down, and two tiles per warp are too few.
    if (S == 3) return launch_pdl_cluster(matvec_q_kernel<TYPE, NCOL, 3>, dim3(total_ctas, ygrid), dim3(GA_WARPS * 32), ga_smem_bytes<TYPE, 3>(), st, cl, b);
    return launch_pdl_cluster(matvec_q_kernel<TYPE, NCOL, 4>, dim3(total_ctas, ygrid), dim3(GA_WARPS * 32), ga_smem_bytes<TYPE, 4>(), st, cl, b);
}

template <int TYPE, int PRO>
static cudaError_t launch_async_pro(const MatvecBatch & b, int total, cudaStream_t st) {
    constexpr int S = QFmt<TYPE>::QB == 32 ? 3 : 4;          // a CTA's rows are a handful of chunks: the ring holds most of them while the prologue runs
    const dim3 grid(total, 1), block(GA_WARPS * 32);
    if (b.pro.D <= 4 * PRO_NT)  return launch_pdl(matvec_q_kernel<TYPE, 1, S, PRO, 4>, grid, block, ga_smem_bytes<TYPE, S>(), st, b);
    if (b.pro.D <= 8 * PRO_NT)  return launch_pdl(matvec_q_kernel<TYPE, 1, S, PRO, 8>, grid, block, ga_smem_bytes<TYPE, S>(), st, b);
    return cudaErrorInvalidValue;                            // wider rows: not instantiated (see forward.cpp can_fuse: fusion loses there)
}

template <int TYPE>
static cudaError_t launch_async_t(const MatvecBatch & b, int total, cudaStream_t st) {
    if (b.pro.kind != PRO_NONE) {
        if (b.T != 1 || b.ks != 1 || b.pro.D > LN_SMALL_D || b.pro.D % 32 != 0) return cudaErrorInvalidValue;
        if (b.pro.wait.site >= 0 && (b.pro.kind != PRO_LNMIX || !b.tp.ll[0] || !b.pro.wait.x_out)) return cudaErrorInvalidValue;
        for (int j = 0; j < b.n_jobs; j++)
            if ((int) b.job[j].W.K != b.pro.D || b.job[j].x_off != 0 || (b.pro.kind == PRO_LNMIX && !b.job[j].pro_mu)) return cudaErrorInvalidValue;
        return b.pro.kind == PRO_LNMIX ? launch_async_pro<TYPE, PRO_LNMIX>(b, total, st) : launch_async_pro<TYPE, PRO_WKV4>(b, total, st);
    }
    for (int j = 0; j < b.n_jobs; j++)
        if ((b.job[j].epi == EPI_TP_PUSH || b.job[j].epi == EPI_TP_PUSH_SIGMOID) && (b.T != 1 || !b.tp.ll[0] || b.tp.site_push < 0)) return cudaErrorInvalidValue;
    if (b.T == 1)      return launch_async_n<TYPE, 1>(b, total, 1, st);
    else if (b.T == 2) return launch_async_n<TYPE, 2>(b, total, 1, st);
    else if (b.T <= 4) return launch_async_n<TYPE, 4>(b, total, 1, st);
    else               return launch_async_n<TYPE, 8>(b, total, (b.T + 7) / 8, st);
}

// b.cta_begin already filled by launch_matvec (32 rows per CTA)
cudaError_t launch_matvec_async(const MatvecBatch & b, int total_ctas, cudaStream_t st) {
    switch (b.job[0].W.type) {
        case T_Q4_0: return launch_async_t<T_Q4_0>(b, total_ctas, st);
        case T_Q4_1: return launch_async_t<T_Q4_1>(b, total_ctas, st);
        case T_Q5_0: return launch_async_t<T_Q5_0>(b, total_ctas, st);
        case T_Q5_1: return launch_async_t<T_Q5_1>(b, total_ctas, st);
        case T_Q8_0: return launch_async_t<T_Q8_0>(b, total_ctas, st);
        default:     return cudaErrorInvalidValue;
    }
}

template <int TYPE, int NCOL, int S>
static cudaError_t set_smem_one() {
    return cudaFuncSetAttribute(matvec_q_kernel<TYPE, NCOL, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, ga_smem_bytes<TYPE, S>());
}

template <int TYPE>
static cudaError_t set_smem_attr() {
    cudaError_t e = cudaSuccess;
    if (e == cudaSuccess) e = set_smem_one<TYPE, 1, 3>();
    if (e == cudaSuccess) e = set_smem_one<TYPE, 2, 3>();
    if (e == cudaSuccess) e = set_smem_one<TYPE, 4, 3>();
    if (e == cudaSuccess) e = set_smem_one<TYPE, 8, 3>();
    if (e == cudaSuccess) e = set_smem_one<TYPE, 1, 4>();
    if (e == cudaSuccess) e = set_smem_one<TYPE, 2, 4>();
    if (e == cudaSuccess) e = set_smem_one<TYPE, 4, 4>();
    if (e == cudaSuccess) e = set_smem_one<TYPE, 8, 4>();
    constexpr int SP = QFmt<TYPE>::QB == 32 ? 3 : 4;
    if (e == cudaSuccess) e = cudaFuncSetAttribute(matvec_q_kernel<TYPE, 1, SP, PRO_LNMIX, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, ga_smem_bytes<TYPE, SP>());
    if (e == cudaSuccess) e = cudaFuncSetAttribute(matvec_q_kernel<TYPE, 1, SP, PRO_WKV4, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, ga_smem_bytes<TYPE, SP>());
    if (e == cudaSuccess) e = cudaFuncSetAttribute(matvec_q_kernel<TYPE, 1, SP, PRO_LNMIX, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, ga_smem_bytes<TYPE, SP>());
    if (e == cudaSuccess) e = cudaFuncSetAttribute(matvec_q_kernel<TYPE, 1, SP, PRO_WKV4, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, ga_smem_bytes<TYPE, SP>());
    return e;
}

// Opt in to > 48 KB dynamic shared memory; call once per device before the first launch
// (and outside stream capture).
cudaError_t matvec_async_init() {
    cudaError_t e = set_smem_attr<T_Q4_0>();
    if (e == cudaSuccess) e = set_smem_attr<T_Q4_1>();
    if (e == cudaSuccess) e = set_smem_attr<T_Q5_0>();
    if (e == cudaSuccess) e = set_smem_attr<T_Q5_1>();
    if (e == cudaSuccess) e = set_smem_attr<T_Q8_0>();
    return e;
}

} // namespace rb
