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
#include "gemv_common.cuh"
#include "launch.cuh"

namespace rb {

constexpr int GA_WARPS = 8, GA_R = TILE_ROWS, GA_KC = TILE_BLOCKS;

template <int TYPE> struct QFmt;
template <> struct QFmt<T_Q4_0> { static constexpr int QB = 16, HB = 0, DB = 2, STAGES = 4; };
template <> struct QFmt<T_Q4_1> { static constexpr int QB = 16, HB = 0, DB = 4, STAGES = 4; };
template <> struct QFmt<T_Q5_0> { static constexpr int QB = 16, HB = 4, DB = 2, STAGES = 4; };
template <> struct QFmt<T_Q5_1> { static constexpr int QB = 16, HB = 4, DB = 4, STAGES = 4; };
template <> struct QFmt<T_Q8_0> { static constexpr int QB = 32, HB = 0, DB = 2, STAGES = 3; };

template <int TYPE> __host__ __device__ constexpr int ga_stage_bytes() { return GA_R * GA_KC * (QFmt<TYPE>::QB + QFmt<TYPE>::HB + QFmt<TYPE>::DB); }
template <int TYPE> __host__ __device__ constexpr int ga_smem_bytes() { return GA_WARPS * QFmt<TYPE>::STAGES * (ga_stage_bytes<TYPE>() + 8); }

// ---- PTX wrappers ------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra.uni WAIT_DONE;\n\t"
        "bra.uni WAIT_LOOP;\n\t"
        "WAIT_DONE:\n\t}" :: "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void * src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

template <int TYPE, int NCOL>
__global__ void __launch_bounds__(GA_WARPS * 32)
gemv_async_kernel(const __grid_constant__ GemvBatch B) {
    using F = QFmt<TYPE>;
    constexpr int S = F::STAGES;
    constexpr int BB = F::QB + F::HB + F::DB;              // bytes per block
    constexpr int STAGE = ga_stage_bytes<TYPE>();
    constexpr bool HAS_M = (TYPE == T_Q4_1 || TYPE == T_Q5_1);
    extern __shared__ __align__(128) uint8_t ga_smem[];
    __shared__ float tile[NCOL][GA_WARPS * GA_R];

    pdl_trigger();
    int j = 0;
#pragma unroll 1
    while (j + 1 < B.n_jobs && (int) blockIdx.x >= B.cta_begin[j + 1]) j++;
    const GemvJob & J = B.job[j];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int M = J.W.M, nb = J.W.K >> 5, T = B.T;
    const int tile_row0 = ((int) blockIdx.x - B.cta_begin[j]) * GA_WARPS * GA_R;
    const int row0 = tile_row0 + warp * GA_R;
    const int col0 = blockIdx.y * NCOL;
    const int nch = row0 < M ? (nb + GA_KC - 1) / GA_KC : 0;     // warps past the last row stay idle

    uint8_t * my = ga_smem + (size_t) warp * S * STAGE;
    const uint32_t bars = smem_u32(ga_smem + (size_t) GA_WARPS * S * STAGE) + warp * S * 8;
    const uint8_t * src0 = reinterpret_cast<const uint8_t *>(J.W.qs) + (size_t) (row0 / GA_R) * J.W.rg_stride;
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < S; s++) mbar_init(bars + s * 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        // weights do not depend on the preceding kernel: start streaming at once
#pragma unroll
        for (int s = 0; s < S; s++) {
            if (s < nch) {
                const int cnt = min(GA_KC, nb - s * GA_KC);
                const uint32_t bytes = (uint32_t) ((GA_R * cnt * BB + 15) & ~15);
                mbar_expect_tx(bars + s * 8, bytes);
                bulk_g2s(smem_u32(my + (size_t) s * STAGE), src0 + (size_t) s * GA_R * GA_KC * BB, bytes, bars + s * 8);
            }
        }
    }
    __syncwarp();
    pdl_wait();                                         // activations (and the residual) come from the preceding kernels

    RowAcc<TYPE, NCOL> ra[GA_R];
#pragma unroll
    for (int r = 0; r < GA_R; r++) ra[r].clear();

    for (int c = 0; c < nch; c++) {
        const int s = c % S;
        const int cnt = min(GA_KC, nb - c * GA_KC);
        const bool act = lane < cnt;
        XBlock<NCOL> xb;
        if (act) load_xblock<TYPE, NCOL>(xb, J.x, J.x_off, col0, T, c * GA_KC + lane);     // before the wait: independent of the tile
        mbar_wait(bars + s * 8, (uint32_t) ((c / S) & 1));
        const uint8_t * st = my + (size_t) s * STAGE;
        if (act) {
#pragma unroll
            for (int r = 0; r < GA_R; r++) {
                const int e = r * cnt + lane;
                const uint4 * q = reinterpret_cast<const uint4 *>(st + (size_t) e * F::QB);
                const uint4 q0 = q[0];
                const uint4 q1 = (F::QB == 32) ? q[1] : make_uint4(0, 0, 0, 0);
                uint32_t qh = 0;
                if (F::HB) qh = *reinterpret_cast<const uint32_t *>(st + GA_R * cnt * F::QB + e * 4);
                float dw, mw = 0.0f;
                const uint8_t * dp = st + GA_R * cnt * (F::QB + F::HB) + e * F::DB;
                if (HAS_M) { const __half2 dm = *reinterpret_cast<const __half2 *>(dp); dw = __low2float(dm); mw = __high2float(dm); }
                else dw = __half2float(*reinterpret_cast<const __half *>(dp));
                dot_block<TYPE, NCOL>(ra[r], xb, q0, q1, qh, dw, mw);
            }
        }
        __syncwarp();                                   // every lane has consumed the stage
        if (lane == 0 && c + S < nch) {
            const int cn = c + S, cnt2 = min(GA_KC, nb - cn * GA_KC);
            const uint32_t bytes = (uint32_t) ((GA_R * cnt2 * BB + 15) & ~15);
            mbar_expect_tx(bars + s * 8, bytes);
            bulk_g2s(smem_u32(my + (size_t) s * STAGE), src0 + (size_t) cn * GA_R * GA_KC * BB, bytes, bars + s * 8);
        }
    }

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
    const bool want_q = (J.out_act.f != nullptr);
    if (lane < GA_R * NCOL) {
        const int r = lane / NCOL, c = lane % NCOL;
        const int m = row0 + r, t = col0 + c;
        float v = 0.0f;
        if (m < M && t < T) {
            v = apply_epilogue(J, mine, t, m);
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
}

template <int TYPE, int NCOL>
static cudaError_t launch_async_n(const GemvBatch & b, int total_ctas, int ygrid, cudaStream_t st) {
    return launch_pdl(gemv_async_kernel<TYPE, NCOL>, dim3(total_ctas, ygrid), dim3(GA_WARPS * 32), ga_smem_bytes<TYPE>(), st, b);
}

template <int TYPE>
static cudaError_t launch_async_t(const GemvBatch & b, int total, cudaStream_t st) {
    if (b.T == 1)      return launch_async_n<TYPE, 1>(b, total, 1, st);
    else if (b.T == 2) return launch_async_n<TYPE, 2>(b, total, 1, st);
    else if (b.T <= 4) return launch_async_n<TYPE, 4>(b, total, 1, st);
    else               return launch_async_n<TYPE, 8>(b, total, (b.T + 7) / 8, st);
}

// b.cta_begin already filled by launch_gemv (32 rows per CTA)
cudaError_t launch_gemv_async(const GemvBatch & b, int total_ctas, cudaStream_t st) {
    switch (b.job[0].W.type) {
        case T_Q4_0: return launch_async_t<T_Q4_0>(b, total_ctas, st);
        case T_Q4_1: return launch_async_t<T_Q4_1>(b, total_ctas, st);
        case T_Q5_0: return launch_async_t<T_Q5_0>(b, total_ctas, st);
        case T_Q5_1: return launch_async_t<T_Q5_1>(b, total_ctas, st);
        case T_Q8_0: return launch_async_t<T_Q8_0>(b, total_ctas, st);
        default:     return cudaErrorInvalidValue;
    }
}

template <int TYPE>
static cudaError_t set_smem_attr() {
    cudaError_t e = cudaFuncSetAttribute(gemv_async_kernel<TYPE, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, ga_smem_bytes<TYPE>());
    if (e == cudaSuccess) e = cudaFuncSetAttribute(gemv_async_kernel<TYPE, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, ga_smem_bytes<TYPE>());
    if (e == cudaSuccess) e = cudaFuncSetAttribute(gemv_async_kernel<TYPE, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, ga_smem_bytes<TYPE>());
    if (e == cudaSuccess) e = cudaFuncSetAttribute(gemv_async_kernel<TYPE, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, ga_smem_bytes<TYPE>());
    return e;
}

// Opt in to > 48 KB dynamic shared memory; call once per device before the first launch
// (and outside stream capture).
cudaError_t gemv_async_init() {
    cudaError_t e = set_smem_attr<T_Q4_0>();
    if (e == cudaSuccess) e = set_smem_attr<T_Q4_1>();
    if (e == cudaSuccess) e = set_smem_attr<T_Q5_0>();
    if (e == cudaSuccess) e = set_smem_attr<T_Q5_1>();
    if (e == cudaSuccess) e = set_smem_attr<T_Q8_0>();
    return e;
}

} // namespace rb
