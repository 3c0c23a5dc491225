// Dequant-fused mat-vec (decode) / thin mat-mat (up to 8 columns per pass) kernel.
//
// Replaces ggml_compute_forward_mul_mat + ggml_vec_dot_q*_q8_* / _f16 / _f32
// (ggml-cpu.c:7377,7287; ggml-cpu-quants.c:1741,2321,2614,2969,3343; ggml-cpu.c:1352,1463)
// and computes what they compute:
//   quantised W : activations are q8 blocks (see Act); per block an exact integer dot
//                 sumi = sum q_w*q_x (dp4a), then acc += (d_w*d_x)*sumi [+ m_w*s_x] in FP32;
//   F16 W       : activations rounded to FP16 (ggml_fp32_to_fp16_row), products exact, FP32 sum;
//   F32 W       : FP32 FMA.
// Only the order of the FP32 block sums differs from the CPU (lane-strided + butterfly
// instead of 8 AVX lanes + hsum).
//
// Work split: one CTA = 32 consecutive rows of one job (8 warps x 4 rows); lanes stride the
// K/32 blocks of a row, so a warp request for one row is 32 consecutive 16-byte nibble
// groups = 512 B coalesced (+128 B qh, +64/128 B scales).  The activation block (32 B of
// int8 + d + s) is loaded once per lane iteration and reused for the 4 rows.  HBM-bound:
// the roofline is bytes(W)/t against the measured copy bandwidth.
// A CTA owning 32 consecutive outputs can quantise them for the next mat-vec in its
// epilogue (one q8 block per column), so chained products need no extra launch.
#include "matvec_common.cuh"
#include "launch.cuh"
#include "tp_ll.cuh"
#include <cstdlib>

namespace rb {

cudaError_t launch_matvec_async(const MatvecBatch & b, int total_ctas, cudaStream_t st);

constexpr int MATVEC_WARPS = 8;
constexpr int MATVEC_R     = 4;                       // rows per warp
constexpr int MATVEC_ROWS  = MATVEC_WARPS * MATVEC_R;     // 32 rows per CTA = one q8 block of outputs

template <int TYPE, int NCOL>
__global__ void __launch_bounds__(MATVEC_WARPS * 32)
matvec_kernel(const __grid_constant__ MatvecBatch B) {
    pdl_trigger();
    // which job / row block
    int j = 0;
#pragma unroll 1
    while (j + 1 < B.n_jobs && (int) blockIdx.x >= B.cta_begin[j + 1]) j++;
    const MatvecJob & J = B.job[j];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int M = J.W.M, K = J.W.K, T = B.T;
    const int row0 = ((int) blockIdx.x - B.cta_begin[j]) * MATVEC_ROWS + warp * MATVEC_R;
    const int col0 = blockIdx.y * NCOL;

    float res[MATVEC_R][NCOL];
    {   // pull the head of this warp's rows towards L2 while the producer of x is still running
        const size_t esz = (TYPE == T_F32) ? 4 : 2;
#pragma unroll
        for (int r = 0; r < MATVEC_R; r++) {
            const char * p = reinterpret_cast<const char *>(J.W.qs) + (size_t) min(row0 + r, M - 1) * K * esz + (size_t) lane * 128;
            if ((size_t) lane * 128 < (size_t) K * esz) asm volatile("prefetch.global.L2 [%0];" :: "l"(p));
        }
    }
    pdl_wait();

    {
        static_assert(TYPE == T_F32 || TYPE == T_F16, "quantised types use matvec_q.cu");
        constexpr int VEC = (TYPE == T_F32) ? 4 : 8;
        float acc[MATVEC_R][NCOL];
#pragma unroll
        for (int r = 0; r < MATVEC_R; r++)
#pragma unroll
            for (int c = 0; c < NCOL; c++) acc[r][c] = 0.0f;
        const int nchunk = (K % VEC == 0 && J.x_off % VEC == 0 && J.x.K % 4 == 0) ? K / VEC : 0;   // else rows are not 16-byte aligned
        for (int ch = lane; ch < nchunk; ch += 32) {
            float xv[NCOL][VEC];
#pragma unroll
            for (int c = 0; c < NCOL; c++) {
                const int t = min(col0 + c, T - 1);
                const float * xp = J.x.f + (size_t) t * J.x.K + J.x_off + (size_t) ch * VEC;
#pragma unroll
                for (int i = 0; i < VEC; i += 4) {
                    const float4 f = *reinterpret_cast<const float4 *>(xp + i);
                    xv[c][i] = f.x; xv[c][i + 1] = f.y; xv[c][i + 2] = f.z; xv[c][i + 3] = f.w;
                }
                if (TYPE == T_F16) {
#pragma unroll
                    for (int i = 0; i < VEC; i++) xv[c][i] = round_f16(xv[c][i]);
                }
            }
            uint4 wq[MATVEC_R];
#pragma unroll
            for (int r = 0; r < MATVEC_R; r++) {
                const int row = min(row0 + r, M - 1);
                if (TYPE == T_F32) wq[r] = *reinterpret_cast<const uint4 *>(reinterpret_cast<const float *>(J.W.qs) + (size_t) row * K + (size_t) ch * 4);
                else               wq[r] = *reinterpret_cast<const uint4 *>(reinterpret_cast<const __half *>(J.W.qs) + (size_t) row * K + (size_t) ch * 8);
            }
#pragma unroll
            for (int r = 0; r < MATVEC_R; r++) {
                float wf[VEC];
                if (TYPE == T_F32) {
                    wf[0] = __uint_as_float(wq[r].x); wf[1] = __uint_as_float(wq[r].y); wf[2] = __uint_as_float(wq[r].z); wf[3] = __uint_as_float(wq[r].w);
                } else {
                    const uint32_t u[4] = {wq[r].x, wq[r].y, wq[r].z, wq[r].w};
#pragma unroll
                    for (int i = 0; i < 4; i++) {
                        const __half2 h = *reinterpret_cast<const __half2 *>(&u[i]);
                        wf[2 * i] = __low2float(h); wf[2 * i + 1] = __high2float(h);
                    }
                }
#pragma unroll
                for (int c = 0; c < NCOL; c++)
#pragma unroll
                    for (int i = 0; i < VEC; i++) acc[r][c] = __fmaf_rn(wf[i], xv[c][i], acc[r][c]);
            }
        }
        // scalar tail when K is not a multiple of the vector width
        for (int k = nchunk * VEC + lane; k < K; k += 32) {
#pragma unroll
            for (int r = 0; r < MATVEC_R; r++) {
                const int row = min(row0 + r, M - 1);
                const float wv = (TYPE == T_F32) ? reinterpret_cast<const float *>(J.W.qs)[(size_t) row * K + k]
                                                 : __half2float(reinterpret_cast<const __half *>(J.W.qs)[(size_t) row * K + k]);
#pragma unroll
                for (int c = 0; c < NCOL; c++) {
                    const int t = min(col0 + c, T - 1);
                    float xs = J.x.f[(size_t) t * J.x.K + J.x_off + k];
                    if (TYPE == T_F16) xs = round_f16(xs);
                    acc[r][c] = __fmaf_rn(wv, xs, acc[r][c]);
                }
            }
        }
#pragma unroll
        for (int r = 0; r < MATVEC_R; r++)
#pragma unroll
            for (int c = 0; c < NCOL; c++) res[r][c] = warp_sum(acc[r][c]);
    }

    // ---- epilogue: lane (r*NCOL + c) finishes output (row0 + r, col0 + c) ------------------
    __shared__ float tile[NCOL][MATVEC_ROWS];
    float mine = 0.0f;
#pragma unroll
    for (int r = 0; r < MATVEC_R; r++)
#pragma unroll
        for (int c = 0; c < NCOL; c++) if (lane == r * NCOL + c) mine = res[r][c];
    if (NCOL == 1 && (J.epi == EPI_TP_PUSH || J.epi == EPI_TP_PUSH_SIGMOID)) {        // fused tensor-parallel all-reduce, producer side (see matvec_q.cu)
        static_assert(MATVEC_R * TP_LL_WORLD == 32, "one lane per (row, peer)");
        float val = __shfl_sync(FULL, mine, lane & (MATVEC_R - 1));
        if (J.epi == EPI_TP_PUSH_SIGMOID) val = sigmoidf_ref(val);
        const int m = row0 + (lane & (MATVEC_R - 1)), peer = lane / MATVEC_R;
        if (m < M && peer < B.tp.world) tp_push(B.tp, B.tp.site_push, tp_flag(B.tp, B.tp.site_push), peer, (size_t) J.tp_off + m, val);
        return;
    }
    const bool want_q = (J.out_act.f != nullptr);
    if (lane < MATVEC_R * NCOL) {
        const int r = lane / NCOL, c = lane % NCOL;
        const int m = row0 + r, t = col0 + c;
        float v = 0.0f;
        if (m < M && t < T) {
            v = apply_epilogue(J, mine, t, m);
            if (J.out) J.out[(size_t) t * J.ldo + m] = v;
        }
        if (want_q) tile[c][warp * MATVEC_R + r] = v;
    }
    if (want_q) {                                   // uniform per CTA
        __syncthreads();
        if (warp < NCOL && col0 + warp < T) {
            const int m = ((int) blockIdx.x - B.cta_begin[j]) * MATVEC_ROWS + lane;
            act_store_block(J.out_act, (size_t) (col0 + warp), m, m < M, tile[warp][lane]);
        }
    }
}

// ---- FP32 weights with a very short K (v6 time_maa_w2: K = 32) over many rows ---------------
// The mat-vec kernel re-reads such a matrix once per 8 columns; here a thread keeps its weight
// row in registers and walks 32 activation rows staged in shared memory.  The result is
// bit-identical to matvec_kernel<T_F32>: lane c of that kernel accumulates chunk c (4 weights,
// FMA chain), lanes >= K/4 hold 0, and the xor-butterfly (16, 8, 4, 2, 1) combines them; the
// same tree is evaluated here in one thread (x + 0 is exact, so the empty lanes drop out).
template <int NCH>   // K / 4: 8 or 16
__global__ void __launch_bounds__(256) smallk_kernel(const __grid_constant__ MatvecBatch B) {
    constexpr int K = NCH * 4, TT = 32;
    __shared__ __align__(16) float xs[TT][K];
    pdl_trigger();
    const MatvecJob & J = B.job[blockIdx.z];
    const int m = blockIdx.x * 256 + threadIdx.x, t0 = blockIdx.y * TT, M = J.W.M, T = B.T;
    float w[K];
    {
        const float4 * wp = reinterpret_cast<const float4 *>(reinterpret_cast<const float *>(J.W.qs) + (size_t) min(m, M - 1) * K);
#pragma unroll
        for (int c = 0; c < NCH; c++) { const float4 f = wp[c]; w[4 * c] = f.x; w[4 * c + 1] = f.y; w[4 * c + 2] = f.z; w[4 * c + 3] = f.w; }
    }
    pdl_wait();
    for (int x = threadIdx.x; x < TT * NCH; x += 256) {
        const int t = x / NCH, c = x % NCH;
        const float4 f = t0 + t < T ? *reinterpret_cast<const float4 *>(J.x.f + (size_t) (t0 + t) * J.x.K + J.x_off + c * 4) : make_float4(0, 0, 0, 0);
        *reinterpret_cast<float4 *>(&xs[t][c * 4]) = f;
    }
    __syncthreads();
    if (m >= M) return;
    const int nt = min(TT, T - t0);
    for (int t = 0; t < nt; t++) {
        float p[NCH];
#pragma unroll
        for (int c = 0; c < NCH; c++) {
            float acc = 0.0f;
#pragma unroll
            for (int i = 0; i < 4; i++) acc = __fmaf_rn(w[4 * c + i], xs[t][4 * c + i], acc);
            p[c] = acc;
        }
#pragma unroll
        for (int o = NCH / 2; o > 0; o >>= 1)
#pragma unroll
            for (int l = 0; l < o; l++) p[l] = p[l] + p[l + o];          // lane l of the butterfly step `xor o`
        const float y = apply_epilogue(J, p[0], t0 + t, m);
        J.out[(size_t) (t0 + t) * J.ldo + m] = y;
    }
}

static bool smallk_ok(const MatvecBatch & b) {
    if (b.T < 16 || b.n_jobs < 1) return false;
    const uint32_t K = b.job[0].W.K;
    if (K != 32 && K != 64) return false;
    for (int j = 0; j < b.n_jobs; j++) {
        const MatvecJob & g = b.job[j];
        if (g.W.type != T_F32 || g.W.K != K || g.W.M != b.job[0].W.M || g.out == nullptr || g.out_act.f != nullptr) return false;
        if (g.x_off % 4 != 0 || g.x.K % 4 != 0) return false;
    }
    return true;
}

static cudaError_t launch_smallk(const MatvecBatch & b, cudaStream_t st) {
    const dim3 grid((b.job[0].W.M + 255) / 256, (b.T + 31) / 32, b.n_jobs);
    if (b.job[0].W.K == 32) return launch_pdl(smallk_kernel<8>, grid, dim3(256), 0, st, b);
    return launch_pdl(smallk_kernel<16>, grid, dim3(256), 0, st, b);
}

template <int TYPE>
static cudaError_t launch_typed(const MatvecBatch & b, int total_ctas, cudaStream_t st) {
    const dim3 block(MATVEC_WARPS * 32);
    if (b.T == 1)      return launch_pdl(matvec_kernel<TYPE, 1>, dim3(total_ctas, 1), block, 0, st, b);
    else if (b.T == 2) return launch_pdl(matvec_kernel<TYPE, 2>, dim3(total_ctas, 1), block, 0, st, b);
    else if (b.T <= 4) return launch_pdl(matvec_kernel<TYPE, 4>, dim3(total_ctas, 1), block, 0, st, b);
    else               return launch_pdl(matvec_kernel<TYPE, 8>, dim3(total_ctas, (b.T + 7) / 8), block, 0, st, b);
}


cudaError_t launch_matvec(const MatvecBatch & b_in, cudaStream_t st) {
    if (b_in.n_jobs == 0) return cudaSuccess;
    MatvecBatch b = b_in;
    const uint32_t type = b.job[0].W.type;
    int total = 0;
    b.ks = 1; b.ks_warps = 0;
    if (type_quantized(type)) {
        // the split of a matrix is a function of its shape only (kernels.h); where it happens is a property of the launch
        bool in_cta = matvec_split_in_cta_enabled() && b.pro.kind == PRO_NONE, any = false;
        for (int j = 0; j < b.n_jobs; j++) {
            b.job[j].ks = matvec_ksplit(b.job[j].W.M, b.job[j].W.K);
            if (b.job[j].ks > 1) { any = true; in_cta = in_cta && matvec_split_in_cta(b.job[j]); }
        }
        if (any && in_cta) b.ks_warps = 1;
        else if (any) {                                          // cluster form: one split for the whole launch (callers group by split)
            b.ks = b.job[0].ks;
            for (int j = 1; j < b.n_jobs; j++) if (b.job[j].ks != b.ks) return cudaErrorInvalidValue;
        }
    }
    for (int j = 0; j < b.n_jobs; j++) {
        if (b.job[j].W.type != type) return cudaErrorInvalidValue;
        b.cta_begin[j] = total;
        const int rows_per_cta = b.ks_warps ? MATVEC_ROWS / b.job[j].ks : MATVEC_ROWS;
        total += (((int) b.job[j].W.M + rows_per_cta - 1) / rows_per_cta) * b.ks;
    }
    b.cta_begin[b.n_jobs] = total;
    for (int j = 0; j < b.n_jobs; j++)
        if ((b.job[j].epi == EPI_TP_PUSH || b.job[j].epi == EPI_TP_PUSH_SIGMOID) && (b.T != 1 || !b.tp.ll[0] || b.tp.site_push < 0)) return cudaErrorInvalidValue;
    if (type_quantized(type)) return launch_matvec_async(b, total, st);
    if (b.pro.kind != PRO_NONE) return cudaErrorInvalidValue;      // prologues exist in the quantised kernel only
    if (type == T_F32 && smallk_ok(b)) return launch_smallk(b, st);
    if (type == T_F32) return launch_typed<T_F32>(b, total, st);
    if (type == T_F16) return launch_typed<T_F16>(b, total, st);
    return cudaErrorInvalidValue;
}

} // namespace rb
