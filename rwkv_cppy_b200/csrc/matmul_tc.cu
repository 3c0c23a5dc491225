// Tensor-core MATMUL for sequence mode and wide batches: out[T][M] = epi(X[T][K] * W[M][K]^T).
//
// Replaces the same ggml_mul_mat nodes as the mat-vec kernels (rwkv_graph.inc block builders),
// for T >= 64 rows where the product is a dense contraction.  Arithmetic: weights dequantised
// to FP16 with ONE rounding (w16 = fp16(q*d + m), HFMA2 / HMUL2), activations rounded to FP16,
// products exact, FP32 accumulation in tensor memory -- the numerics of the reference's own
// GPU offload path (dequantise-to-FP16 + FP16 MATMUL), not of its CPU path; FP32 files never
// take this path.
//
// Structure (one persistent CTA per SM, 18 warps, tiles of 128 weight rows x BN tokens):
//   warp 0      producer: per 64-wide K step one cp.async.bulk of the quantised 128x64 chunk
//               (DevMat::tc layout, 4.5-8.5 KB) into a ring, and one TMA tensor load of the
//               BN x 64 FP16 activation box (128-byte swizzle) into the stage;
//   warps 10-17 dequantisers: thread = (row, block); unpack 32 weights, convert to FP16 and
//               store them into the stage's A tile in the K-major 128B-swizzled layout;
//   warp 1      one elected thread issues 4 x tcgen05.mma (M128 x N=BN x K16, kind::f16) per
//               stage, accumulator in TMEM (double buffered: 2 x BN columns), and commits the
//               stage / the accumulator to mbarriers;
//   warps 2-9   epilogue: tcgen05.ld 32 lanes x 32 columns, the same fused epilogues as the
//               mat-vec kernels, coalesced FP32 stores (+ optional FP16 copy for a chained MATMUL).
// FP16 weight matrices (FP16 files, head) skip the dequantisers: the A tile is a second TMA load.
#include "matvec_common.cuh"
#include "launch.cuh"
#include "ptx.cuh"

#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_fp16.h>
#include <cstdlib>
#include <cstring>

namespace rb {

constexpr int GT_BM = TC_ROWS, GT_BK = TC_KB;
constexpr int GT_WARPS = 18, GT_THREADS = GT_WARPS * 32;
constexpr int GT_EPI_WARPS = 8, GT_DQ_WARP0 = 2 + GT_EPI_WARPS;
constexpr int GT_SQ = 4;                                  // quantised-chunk ring depth
constexpr int GT_A_BYTES = GT_BM * GT_BK * 2;             // 16 KB

struct GtJob {
    CUtensorMap xmap;                                     // FP16 activations [T][K]
    CUtensorMap wmap;                                     // FP16 weights [M][K] (T_F16 only)
    const uint8_t * tc;
    uint32_t M, K, kblocks;
    float * out; int ldo; int epi; const float * aux; int ld_aux; const float * vec;
    __half * out_h; int ld_h;
    int tile_begin, m_tiles;
};
struct GtBatch {
    GtJob job[MATMUL_MAX_JOBS];
    int n_jobs, T, n_tiles, total_tiles;
    // K split (ks = 2): CTAs 2c and 2c+1 share output tile(s) c, c + gridDim/2, ...; the odd CTA takes the upper half of
    // K, leaves its raw accumulators in `scratch` ([tile][BN tokens][128 rows] FP32) and raises one flag per epilogue
    // warp; the even CTA adds them to its own before the epilogue op (own + peer: a fixed order) and clears the flag.
    int ks;
    float * scratch; unsigned int * flags;
};

template <int BN> __host__ __device__ constexpr int gt_stages() { return BN == 256 ? 3 : BN == 128 ? 4 : 5; }
template <int TYPE> __host__ __device__ constexpr int gt_chunk_bytes() {
    return TYPE == T_F16 ? 0 : 2 * GT_BM * (TYPE == T_Q4_0 ? 18 : TYPE == T_Q4_1 ? 20 : TYPE == T_Q5_0 ? 22 : TYPE == T_Q5_1 ? 24 : 34);
}
template <int TYPE, int BN> __host__ __device__ constexpr int gt_smem_bytes() {
    return 1024 + gt_stages<BN>() * (GT_A_BYTES + BN * GT_BK * 2) + GT_SQ * gt_chunk_bytes<TYPE>() + 256;
}

// two FP16 values 1024 + b from two bytes of w (bytes lo_idx, lo_idx + 1)
__device__ __forceinline__ __half2 bytes_to_h2_biased(uint32_t w, int pair) {
    const uint32_t r = __byte_perm(w, 0x64646464u, pair == 0 ? 0x4140 : 0x4342);
    return *reinterpret_cast<const __half2 *>(&r);
}

// 32 weights of one block -> 4 x 16 bytes of FP16 (elements 0-7, 8-15, 16-23, 24-31)
template <int TYPE>
__device__ __forceinline__ void dequant_block_f16(uint4 (&out)[4], const uint4 q0, const uint4 q1, const uint32_t qh, const uint32_t dm) {
    int wv[8];
    unpack_block<TYPE>(wv, q0, q1, qh);
    const __half2 dmh = *reinterpret_cast<const __half2 *>(&dm);
    const __half2 d2 = __half2half2(__low2half(dmh));
    const __half2 m2 = __half2half2(__high2half(dmh));
    // 1024 + q is exact in FP16 for q < 1024; subtracting the bias (and the zero point) is exact too
    const float bias = TYPE == T_Q4_0 ? 1032.0f : TYPE == T_Q5_0 ? 1040.0f : TYPE == T_Q8_0 ? 1152.0f : 1024.0f;
    const __half2 b2 = __float2half2_rn(bias);
#pragma unroll
    for (int c = 0; c < 4; c++) {
        uint32_t h[4];
#pragma unroll
        for (int i = 0; i < 2; i++) {
            uint32_t w = (uint32_t) wv[2 * c + i];
            if (TYPE == T_Q8_0) w ^= 0x80808080u;          // signed byte -> byte + 128
#pragma unroll
            for (int p = 0; p < 2; p++) {
                __half2 v = __hsub2(bytes_to_h2_biased(w, p), b2);
                if (TYPE == T_Q4_1 || TYPE == T_Q5_1) v = __hfma2(v, d2, m2);
                else v = __hmul2(v, d2);
                h[2 * i + p] = *reinterpret_cast<const uint32_t *>(&v);
            }
        }
        out[c] = make_uint4(h[0], h[1], h[2], h[3]);
    }
}

// One warp's share of a finished accumulator tile: lane = weight row m; the warp takes the
// 16-column (token) chunks `half_id`, `half_id + 2`, ... of the BN columns.  Stores are 128-byte
// coalesced (32 lanes = 32 consecutive m of one token row).
// K-split, upper half: raw accumulators -> scratch tile (row = TMEM lane of this thread)
template <int BN>
__device__ __forceinline__ void epilogue_spill(float * scr, uint32_t tcol, int lane_row, int t0, int T, int half_id) {
#pragma unroll 1
    for (int c0 = half_id * 16; c0 < BN; c0 += 32) {
        if (t0 + c0 >= T) break;                                  // warp-uniform
        uint32_t v[16];
        tmem_ld_32x16(tcol + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; i++) __stcg(scr + (size_t) (c0 + i) * GT_BM + lane_row, __uint_as_float(v[i]));
    }
}

template <int EPI, int BN>
__device__ __forceinline__ void epilogue_tile(const GtJob & J, uint32_t tcol, int m, int t0, int T, int half_id, const float * peer = nullptr, int lane_row = 0) {
    const bool row_ok = m < (int) J.M;
    float * const out = J.out; __half * const out_h = J.out_h;
    const int ldo = J.ldo, ld_h = J.ld_h;
#pragma unroll 1
    for (int c0 = half_id * 16; c0 < BN; c0 += 32) {
        if (t0 + c0 >= T) break;                                  // warp-uniform
        uint32_t v[16];
        tmem_ld_32x16(tcol + c0, v);
        tmem_ld_wait();
        const int nv = T - t0 - c0;
        if (peer) {                                                // K split: add the other half's partial sums (own + peer)
            float pp[16];
#pragma unroll
            for (int i = 0; i < 16; i++) pp[i] = __ldcg(peer + (size_t) (c0 + i) * GT_BM + lane_row);
#pragma unroll
            for (int i = 0; i < 16; i++) v[i] = __float_as_uint(__fadd_rn(__uint_as_float(v[i]), pp[i]));
        }
        if (row_ok) {
            // read-modify-write epilogues: all 16 operand loads go out before the first store (the compiler
            // may not move a load of `out` above a store to `out`, which made this 16 dependent round trips)
            constexpr bool RMW = (EPI == EPI_ADD_TO || EPI == EPI_MULADD_TO);
            constexpr bool AUX = (EPI == EPI_MULADD_TO || EPI == EPI_ADD_AUX);
            float po[16], pa[16];
            if (RMW) {
#pragma unroll
                for (int i = 0; i < 16; i++) po[i] = i < nv ? out[(size_t) (t0 + c0 + i) * ldo + m] : 0.0f;
            }
            if (AUX) {
#pragma unroll
                for (int i = 0; i < 16; i++) pa[i] = i < nv ? J.aux[(size_t) (t0 + c0 + i) * J.ld_aux + m] : 0.0f;
            }
#pragma unroll
            for (int i = 0; i < 16; i++) {
                if (i < nv) {
                    const int t = t0 + c0 + i;
                    float y = __uint_as_float(v[i]);
                    if (EPI == EPI_ADD_TO)         y = __fadd_rn(po[i], y);
                    else if (EPI == EPI_MULADD_TO) y = __fadd_rn(po[i], __fmul_rn(pa[i], y));
                    else if (EPI == EPI_ADD_AUX)   y = __fadd_rn(pa[i], y);
                    else                           y = apply_epilogue_c<EPI>(J, y, t, m);
                    if (out) out[(size_t) t * ldo + m] = y;
                    if (out_h) out_h[(size_t) t * ld_h + m] = __float2half_rn(y);
                }
            }
        }
    }
}

template <int TYPE, int BN>
__global__ void __launch_bounds__(GT_THREADS, 1)
matmul_tc_kernel(const __grid_constant__ GtBatch B) {
    constexpr bool QUANT = (TYPE != T_F16);
    constexpr int S = gt_stages<BN>();
    constexpr int B_BYTES = BN * GT_BK * 2;
    constexpr int CH = gt_chunk_bytes<TYPE>();
    constexpr int QB = TYPE == T_Q8_0 ? 32 : 16;
    constexpr int HB = (TYPE == T_Q5_0 || TYPE == T_Q5_1) ? 4 : 0;
    constexpr int DB = (TYPE == T_Q4_1 || TYPE == T_Q5_1) ? 4 : 2;
    constexpr uint32_t TMEM_COLS = 2 * BN;
    extern __shared__ uint8_t gt_smem_raw[];

    pdl_trigger();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t base = (smem_u32(gt_smem_raw) + 1023u) & ~1023u;
    uint8_t * base_p = gt_smem_raw + (base - smem_u32(gt_smem_raw));
    const uint32_t sA = base, sB = base + S * GT_A_BYTES, sQ = sB + S * B_BYTES;
    uint8_t * pA = base_p; uint8_t * pQ = base_p + S * (GT_A_BYTES + B_BYTES);
    const uint32_t bars = sQ + GT_SQ * CH;
    const uint32_t bar_full = bars, bar_empty = bars + 8 * S, bar_qfull = bars + 16 * S, bar_qempty = bar_qfull + 8 * GT_SQ,
                   bar_tfull = bar_qempty + 8 * GT_SQ, bar_tempty = bar_tfull + 16, tmem_slot = bar_tempty + 16;
    volatile uint32_t * tmem_slot_p = reinterpret_cast<volatile uint32_t *>(base_p + (tmem_slot - base));

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < S; s++) { mbar_init(bar_full + 8 * s, QUANT ? 1 + 8 : 1); mbar_init(bar_empty + 8 * s, 1); }
        for (int s = 0; s < GT_SQ; s++) { mbar_init(bar_qfull + 8 * s, 1); mbar_init(bar_qempty + 8 * s, 8); }
        for (int s = 0; s < 2; s++) { mbar_init(bar_tfull + 8 * s, 1); mbar_init(bar_tempty + 8 * s, GT_EPI_WARPS); }
        mbar_init_fence();
        fence_proxy_async_smem();
        for (int j = 0; j < B.n_jobs; j++) { tma_prefetch_desc(&B.job[j].xmap); if (!QUANT) tma_prefetch_desc(&B.job[j].wmap); }
    }
    if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_p;

    const int n_tiles = B.n_tiles;
    const int ks = B.ks, krank = ks > 1 ? (int) (blockIdx.x & 1) : 0;
    const int tile0 = ks > 1 ? (int) (blockIdx.x >> 1) : (int) blockIdx.x, tile_step = ks > 1 ? (int) (gridDim.x >> 1) : (int) gridDim.x;
    // this CTA's share of the K steps of a job
    auto kb_lo = [&](uint32_t kblocks) { return krank ? kblocks / 2 : 0u; };
    auto kb_hi = [&](uint32_t kblocks) { return (ks > 1 && !krank) ? kblocks / 2 : kblocks; };
    auto locate = [&](int tile, int & j, int & mt, int & nt) {
        j = 0;
#pragma unroll 1
        while (j + 1 < B.n_jobs && tile >= B.job[j + 1].tile_begin) j++;
        const int local = tile - B.job[j].tile_begin;
        mt = local / n_tiles; nt = local % n_tiles;
    };

    if (warp == 0) {
        // ===== producer =====
        if (lane == 0) {
            pdl_wait();                                          // the activations come from the preceding kernels
            uint32_t it = 0;
            for (int tile = tile0; tile < B.total_tiles; tile += tile_step) {
                int j, mt, nt; locate(tile, j, mt, nt);
                const GtJob & J = B.job[j];
                for (uint32_t kb = kb_lo(J.kblocks); kb < kb_hi(J.kblocks); kb++, it++) {
                    const uint32_t s = it % S, ph = (it / S) & 1;
                    if (QUANT) {
                        const uint32_t qs = it % GT_SQ, qph = (it / GT_SQ) & 1;
                        mbar_wait(bar_qempty + 8 * qs, qph ^ 1);
                        mbar_expect_tx(bar_qfull + 8 * qs, CH);
                        bulk_g2s(sQ + qs * CH, J.tc + ((size_t) mt * J.kblocks + kb) * CH, CH, bar_qfull + 8 * qs);
                    }
                    mbar_wait(bar_empty + 8 * s, ph ^ 1);
#ifdef RB_ABL_NOB
                    mbar_arrive(bar_full + 8 * s);
#else
                    mbar_expect_tx(bar_full + 8 * s, QUANT ? B_BYTES : B_BYTES + GT_A_BYTES);
                    tma_load_2d(sB + s * B_BYTES, &J.xmap, (int) (kb * GT_BK), nt * BN, bar_full + 8 * s);
#endif
                    if (!QUANT) tma_load_2d(sA + s * GT_A_BYTES, &J.wmap, (int) (kb * GT_BK), mt * GT_BM, bar_full + 8 * s);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        constexpr uint32_t idesc = umma_idesc_f16(GT_BM, BN);
        uint32_t it = 0, tl = 0;
        for (int tile = tile0; tile < B.total_tiles; tile += tile_step, tl++) {
            int j, mt, nt; locate(tile, j, mt, nt);
            const uint32_t kblocks = B.job[j].kblocks;
            const uint32_t buf = tl & 1, bph = (tl >> 1) & 1;
            mbar_wait(bar_tempty + 8 * buf, bph ^ 1);            // epilogue has drained this accumulator
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + buf * BN;
            const uint32_t kfirst = kb_lo(kblocks), klast = kb_hi(kblocks);
            for (uint32_t kb = kfirst; kb < klast; kb++, it++) {
                const uint32_t s = it % S, ph = (it / S) & 1;
                mbar_wait(bar_full + 8 * s, ph);
                tc_fence_after();
                if (lane == 0) {
                    const uint64_t ad = umma_desc_k_sw128(sA + s * GT_A_BYTES), bd = umma_desc_k_sw128(sB + s * B_BYTES);
#pragma unroll
                    for (int k = 0; k < GT_BK / 16; k++)          // +32 bytes (16 halves) along K inside the swizzle atom
#ifdef RB_ABL_NOMMA
                        if (kb == 0 && k == 0)
#endif
                        tc_mma_f16(d_tmem, ad + 2 * k, bd + 2 * k, idesc, ((kb - kfirst) | (uint32_t) k) != 0 ? 1u : 0u);
                    tc_commit(bar_empty + 8 * s);                 // stage free once these MMAs have read it
                    if (kb + 1 == klast) tc_commit(bar_tfull + 8 * buf);
                }
                __syncwarp();
            }
        }
    } else if (warp < GT_DQ_WARP0) {
        // ===== epilogue: warp -> TMEM lane quadrant (warp % 4, fixed by the hardware) x column half =====
        const int q = warp & 3, half_id = (warp - 2) >> 2;
        pdl_wait();
        uint32_t tl = 0;
        for (int tile = tile0; tile < B.total_tiles; tile += tile_step, tl++) {
            int j, mt, nt; locate(tile, j, mt, nt);
            const GtJob & J = B.job[j];
            const uint32_t buf = tl & 1, bph = (tl >> 1) & 1;
            mbar_wait(bar_tfull + 8 * buf, bph);
            tc_fence_after();
            const uint32_t tcol = tmem_base + ((uint32_t) (q * 32) << 16) + buf * BN;
            const int m = mt * GT_BM + q * 32 + lane, t0 = nt * BN;
            const float * peer = nullptr;
            if (ks > 1) {
                float * scr = B.scratch + (size_t) tile * BN * GT_BM;
                unsigned int * flag = B.flags + (size_t) tile * GT_EPI_WARPS + (warp - 2);
                if (krank) {                                      // upper half of K: spill, publish, next tile
                    epilogue_spill<BN>(scr, tcol, q * 32 + lane, t0, B.T, half_id);
                    __threadfence();
                    __syncwarp();
                    if (lane == 0) asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(flag), "r"(1u) : "memory");
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_tempty + 8 * buf);
                    continue;
                }
                if (lane == 0) {
                    unsigned int seen = 0;
                    const long long tstart = clock64();
                    do {
                        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(flag) : "memory");
                    } while (seen == 0 && clock64() - tstart < 4000000000LL);      // ~2 s: never hang the GPU on a lost partner
                    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" :: "l"(flag), "r"(0u) : "memory");     // ready for the next launch
                }
                __syncwarp();
                peer = scr;
            }
            const int lrow = q * 32 + lane;
            switch (J.epi) {
                case EPI_STORE:        epilogue_tile<EPI_STORE, BN>(J, tcol, m, t0, B.T, half_id, peer, lrow); break;
                case EPI_SIGMOID:      epilogue_tile<EPI_SIGMOID, BN>(J, tcol, m, t0, B.T, half_id, peer, lrow); break;
                case EPI_RELU_SQ:      epilogue_tile<EPI_RELU_SQ, BN>(J, tcol, m, t0, B.T, half_id, peer, lrow); break;
                case EPI_SILU:         epilogue_tile<EPI_SILU, BN>(J, tcol, m, t0, B.T, half_id, peer, lrow); break;
                case EPI_TANH:         epilogue_tile<EPI_TANH, BN>(J, tcol, m, t0, B.T, half_id, peer, lrow); break;
                case EPI_ADD_TO:       epilogue_tile<EPI_ADD_TO, BN>(J, tcol, m, t0, B.T, half_id, peer, lrow); break;
                case EPI_MULADD_TO:    epilogue_tile<EPI_MULADD_TO, BN>(J, tcol, m, t0, B.T, half_id, peer, lrow); break;
                case EPI_BIAS_SIGMOID: epilogue_tile<EPI_BIAS_SIGMOID, BN>(J, tcol, m, t0, B.T, half_id, peer, lrow); break;
                case EPI_V6_DECAY:     epilogue_tile<EPI_V6_DECAY, BN>(J, tcol, m, t0, B.T, half_id, peer, lrow); break;
                case EPI_ADD_AUX:      epilogue_tile<EPI_ADD_AUX, BN>(J, tcol, m, t0, B.T, half_id, peer, lrow); break;
                default:               epilogue_tile<EPI_V7_DECAY, BN>(J, tcol, m, t0, B.T, half_id, peer, lrow); break;
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_tempty + 8 * buf);
        }
    } else if (QUANT) {
        // ===== dequantisers: thread = (row, block) of the 128 x 64 chunk =====
        const int d = threadIdx.x - GT_DQ_WARP0 * 32;
        const int row = d & (GT_BM - 1), blk = d >> 7;
        const int e = blk * GT_BM + row;
        uint32_t it = 0;
        for (int tile = tile0; tile < B.total_tiles; tile += tile_step) {
            int j, mt, nt; locate(tile, j, mt, nt);
            const uint32_t kblocks = B.job[j].kblocks;
            for (uint32_t kb = kb_lo(kblocks); kb < kb_hi(kblocks); kb++, it++) {
                const uint32_t s = it % S, ph = (it / S) & 1, qs = it % GT_SQ, qph = (it / GT_SQ) & 1;
                mbar_wait(bar_qfull + 8 * qs, qph);
                const uint8_t * cq = pQ + qs * CH;
                const uint4 q0 = *reinterpret_cast<const uint4 *>(cq + e * QB);
                const uint4 q1 = QB == 32 ? *reinterpret_cast<const uint4 *>(cq + e * QB + 16) : make_uint4(0, 0, 0, 0);
                const uint32_t qh = HB ? *reinterpret_cast<const uint32_t *>(cq + 2 * GT_BM * QB + e * 4) : 0u;
                const uint8_t * dp = cq + 2 * GT_BM * (QB + HB) + e * DB;
                const uint32_t dm = DB == 4 ? *reinterpret_cast<const uint32_t *>(dp) : (uint32_t) *reinterpret_cast<const uint16_t *>(dp);
                // the slot goes back to the producer only once the loads have RETURNED: the arrive consumes the loaded words
                const uint32_t dep = __reduce_or_sync(0xffffffffu, (q0.x ^ q0.w) | (q1.x ^ q1.w) | qh | dm);
                if (lane == 0) mbar_arrive_after(bar_qempty + 8 * qs, dep);
                uint4 h[4];
#ifdef RB_ABL_NODQ
                h[0] = q0; h[1] = q0; h[2] = make_uint4(qh, dm, qh, dm); h[3] = q1;
#else
                dequant_block_f16<TYPE>(h, q0, q1, qh, dm);
#endif
                mbar_wait(bar_empty + 8 * s, ph ^ 1);               // the MMAs that read this stage have completed
                uint8_t * ar = pA + s * GT_A_BYTES + row * 128;
#pragma unroll
                for (int c = 0; c < 4; c++)
                    *reinterpret_cast<uint4 *>(ar + (((blk * 4 + c) ^ (row & 7)) << 4)) = h[c];
#ifndef RB_ABL_NOFENCE
                fence_proxy_async_smem();                           // generic-proxy stores -> visible to tcgen05.mma
#endif
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_full + 8 * s);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

// ---- FP32 -> FP16 activation rows ---------------------------------------------------------
__global__ void to_half_kernel(const float * x /* written by the preceding kernel: NOT __restrict__ (invariant loads may be hoisted above griddepcontrol.wait) */, __half * out, size_t n4) {
    pdl_trigger();
    pdl_wait();
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t) gridDim.x * blockDim.x) {
        const float4 f = reinterpret_cast<const float4 *>(x)[i];
        const __half2 a = __floats2half2_rn(f.x, f.y), b = __floats2half2_rn(f.z, f.w);
        uint2 o; o.x = *reinterpret_cast<const uint32_t *>(&a); o.y = *reinterpret_cast<const uint32_t *>(&b);
        reinterpret_cast<uint2 *>(out)[i] = o;
    }
}

cudaError_t launch_to_half(const float * x, uint16_t * out, size_t n, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    if (n % 4 != 0) return cudaErrorInvalidValue;
    const size_t n4 = n / 4;
    const unsigned grid = (unsigned) std::min<size_t>((n4 + 255) / 256, (size_t) device_sm_count() * 8);
    return launch_pdl(to_half_kernel, dim3(grid), dim3(256), 0, st, x, reinterpret_cast<__half *>(out), n4);
}

// ---- host side ----------------------------------------------------------------------------
static PFN_cuTensorMapEncodeTiled_v12000 encode_fn() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn = [] {
        void * p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess || qres != cudaDriverEntryPointSuccess) p = nullptr;
        return (PFN_cuTensorMapEncodeTiled_v12000) p;
    }();
    return fn;
}

// FP16 row-major [rows][ld] matrix, box = 64 columns x box_rows rows, 128-byte swizzle, zero fill out of bounds
static bool make_map(CUtensorMap * map, const void * base, uint64_t cols, uint64_t rows, uint64_t ld, uint32_t box_rows) {
    auto fn = encode_fn();
    if (!fn) return false;
    const cuuint64_t gdim[2] = {cols, rows};
    const cuuint64_t gstride[1] = {ld * 2};
    const cuuint32_t box[2] = {(cuuint32_t) GT_BK, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void *>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

bool matmul_tc_enabled() {
    static const bool on = [] { const char * e = getenv("RWKV_B200_NO_TC"); return !(e && e[0] == '1') && encode_fn() != nullptr; }();
    return on;
}

int matmul_tc_min_rows(bool batch) {
    static const int v = [] { const char * e = getenv("RWKV_B200_TC_MIN_ROWS"); const int n = e ? atoi(e) : 0; return n > 0 ? n : 0; }();
    return v > 0 ? v : (batch ? 32 : 64);
}

bool matmul_tc_supported(const DevMat & W, int T, bool batch) {
    if (!matmul_tc_enabled() || T < matmul_tc_min_rows(batch)) return false;
    if (W.K < 64 || W.K % 8 != 0) return false;
    if (W.type == T_F16) return true;
    return type_quantized(W.type) && W.tc != nullptr;
}

template <int TYPE, int BN>
static cudaError_t launch_gt(const GtBatch & b, cudaStream_t st) {
    const int nsm = device_sm_count();
    const int grid = b.ks > 1 ? 2 * std::min(b.total_tiles, nsm / 2) : std::min(b.total_tiles, nsm);
    return launch_pdl(matmul_tc_kernel<TYPE, BN>, dim3(grid), dim3(GT_THREADS), gt_smem_bytes<TYPE, BN>(), st, b);
}
template <int TYPE>
static cudaError_t launch_gt_bn(const GtBatch & b, int BN, cudaStream_t st) {
    if (BN == 256) return launch_gt<TYPE, 256>(b, st);
    if (BN == 128) return launch_gt<TYPE, 128>(b, st);
    return launch_gt<TYPE, 64>(b, st);
}

cudaError_t launch_matmul_tc(const MatmulJob * jobs, int n_jobs, int T, cudaStream_t st, const MatmulScratch * scratch) {
    if (n_jobs <= 0) return cudaSuccess;
    if (n_jobs > MATMUL_MAX_JOBS) return cudaErrorInvalidValue;
    // Tile width: wide tiles reuse the dequantised weights over more tokens, narrow tiles fill the
    // 148 SMs when there are few weight rows.  Cost model (cycles per 64-wide K step of one CTA,
    // L2-bound estimates): 800 / 520 / 330 for BN = 256 / 128 / 64, times the number of tile rounds.
    int m_tiles_total = 0;
    for (int j = 0; j < n_jobs; j++) m_tiles_total += ((int) jobs[j].W.M + GT_BM - 1) / GT_BM;
    int BN = 64, KS = 1; double best = 1e300;
    const int nsm = device_sm_count();
    const int cand[3] = {256, 128, 64}; const double cost[3] = {800.0, 520.0, 330.0};
    static const int force_bn = [] { const char * e = getenv("RWKV_B200_TC_BN"); return e ? atoi(e) : 0; }();   // tests: pin the tile width
    // K split: measured slower than whole-K tiles on every model shape of this round (the single-round launches it
    // applies to expose the epilogue, which doubles: spill + merge; ffn value 2048 x 7168 x 1024: 112 us split at 256-token
    // tiles vs 83 us at 128-token tiles), so it is opt-in: RWKV_B200_TC_KS=2 splits whenever the shape allows (read per
    // launch so that the tests can switch it).
    const int force_ks = [] { const char * e = getenv("RWKV_B200_TC_KS"); return e ? atoi(e) : 1; }();
    uint32_t kb_max = 0, kb_min = 0xffffffffu; bool kb_even = true;
    for (int j = 0; j < n_jobs; j++) {
        const uint32_t kb = (jobs[j].W.K + GT_BK - 1) / GT_BK;
        kb_max = std::max(kb_max, kb); kb_min = std::min(kb_min, kb); kb_even = kb_even && (kb % 2 == 0);
    }
    // K split over CTA pairs: two CTAs share a tile and meet in an L2-resident scratch tile (deterministic: own half + peer half).
    const bool can_split = scratch && scratch->buf && scratch->flags && kb_even && kb_min >= 16 && force_ks != 1;
    for (int c = 0; c < 3; c++) {
        if (force_bn && cand[c] != force_bn) continue;
        if (cand[c] > 64 && cand[c] / 2 >= T) continue;           // tile would be more than half empty
        const int tiles = m_tiles_total * ((T + cand[c] - 1) / cand[c]);
        for (int ks = 1; ks <= 2; ks++) {
            if (ks == 2 && !(can_split && tiles <= nsm / 2 && (size_t) tiles * cand[c] * GT_BM <= scratch->floats)) continue;
            if (ks == 1 && force_ks == 2 && can_split && tiles <= nsm / 2 && (size_t) tiles * cand[c] * GT_BM <= scratch->floats) continue;
            const double t = (double) ((tiles * ks + nsm - 1) / nsm) * cost[c] * (double) (kb_max / ks) + (ks == 2 ? 2500.0 : 0.0);
            if (t < best) { best = t; BN = cand[c]; KS = ks; }
        }
    }
    GtBatch b;
    memset(&b, 0, sizeof(b));
    b.n_jobs = n_jobs; b.T = T; b.n_tiles = (T + BN - 1) / BN;
    b.ks = KS; b.scratch = KS > 1 ? scratch->buf : nullptr; b.flags = KS > 1 ? scratch->flags : nullptr;
    const uint32_t type = jobs[0].W.type;
    int total = 0;
    for (int j = 0; j < n_jobs; j++) {
        const MatmulJob & g = jobs[j];
        if (g.W.type != type) return cudaErrorInvalidValue;
        GtJob & o = b.job[j];
        if (!make_map(&o.xmap, g.x16, g.W.K, (uint64_t) T, (uint64_t) g.ldx, (uint32_t) BN)) return cudaErrorInvalidValue;
        if (type == T_F16 && !make_map(&o.wmap, g.W.qs, g.W.K, g.W.M, g.W.K, GT_BM)) return cudaErrorInvalidValue;
        o.tc = (const uint8_t *) g.W.tc; o.M = g.W.M; o.K = g.W.K; o.kblocks = (g.W.K + GT_BK - 1) / GT_BK;
        o.out = g.out; o.ldo = g.ldo; o.epi = g.epi; o.aux = g.aux; o.ld_aux = g.ld_aux; o.vec = g.vec;
        o.out_h = reinterpret_cast<__half *>(g.out_h); o.ld_h = g.ld_h;
        o.m_tiles = ((int) g.W.M + GT_BM - 1) / GT_BM; o.tile_begin = total;
        total += o.m_tiles * b.n_tiles;
    }
    b.total_tiles = total;
    switch (type) {
        case T_F16:  return launch_gt_bn<T_F16>(b, BN, st);
        case T_Q4_0: return launch_gt_bn<T_Q4_0>(b, BN, st);
        case T_Q4_1: return launch_gt_bn<T_Q4_1>(b, BN, st);
        case T_Q5_0: return launch_gt_bn<T_Q5_0>(b, BN, st);
        case T_Q5_1: return launch_gt_bn<T_Q5_1>(b, BN, st);
        case T_Q8_0: return launch_gt_bn<T_Q8_0>(b, BN, st);
        default:     return cudaErrorInvalidValue;
    }
}

template <int TYPE>
static cudaError_t gt_attr() {
    cudaError_t e = cudaFuncSetAttribute(matmul_tc_kernel<TYPE, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, gt_smem_bytes<TYPE, 256>());
    if (e == cudaSuccess) e = cudaFuncSetAttribute(matmul_tc_kernel<TYPE, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, gt_smem_bytes<TYPE, 128>());
    if (e == cudaSuccess) e = cudaFuncSetAttribute(matmul_tc_kernel<TYPE, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, gt_smem_bytes<TYPE, 64>());
    return e;
}

// once per device, outside stream capture
cudaError_t matmul_tc_init() {
    cudaError_t e = gt_attr<T_F16>();
    if (e == cudaSuccess) e = gt_attr<T_Q4_0>();
    if (e == cudaSuccess) e = gt_attr<T_Q4_1>();
    if (e == cudaSuccess) e = gt_attr<T_Q5_0>();
    if (e == cudaSuccess) e = gt_attr<T_Q5_1>();
    if (e == cudaSuccess) e = gt_attr<T_Q8_0>();
    return e;
}

} // namespace rb
