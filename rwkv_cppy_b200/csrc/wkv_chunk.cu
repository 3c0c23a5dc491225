// v5 / v6 WKV over long sequences, chunk-parallel form (head size 64, rows >= 64 of FP16 /
// quantised models -- the rows that also take the tensor-core GEMM, i.e. the tolerance-mode
// sequence path; everything else keeps the exact scan of wkv.cu).
//
// Reference recurrence (ggml_compute_forward_rwkv_wkv6_f32, ggml-cpu.c:11945-11980), per head,
// key index i, value index j:
//     y_t[j]    = sum_i r_t[i] * (u[i] k_t[i] v_t[j] + S_{t-1}[i][j])
//     S_t[i][j] = S_{t-1}[i][j] w_t[i] + k_t[i] v_t[j]
// is sequential in t: the scan kernel spends 180 ns per token whatever the GPU.  Unrolling the
// recurrence over a chunk of C = 16 tokens with LP_t[i] = sum_{q<t} log w_q[i]:
//     S_{t-1}  = S_in * exp(LP_t)  +  sum_{s<t} exp(LP_t - LP_{s+1}) k_s v_s^T
//     y_t[j]   = sum_i r_t[i] exp(LP_t[i]) S_in[i][j]                                  (inter)
//              + sum_{s<t} (sum_i r_t[i] k_s[i] exp(LP_t[i] - LP_{s+1}[i])) v_s[j]      (intra)
//              + (sum_i r_t[i] u[i] k_t[i]) v_t[j]                                      (bonus)
//     S_out    = S_in * exp(LP_C) + sum_s exp(LP_C - LP_{s+1}) k_s v_s^T
// Every exponent is <= 0 (s + 1 <= t), so nothing overflows however strong the decay is.
// Three launches per layer, the first and last parallel over (head, chunk):
//   wkv_chunk_state_kernel   dS_c = sum_s exp(LP_C - LP_{s+1}) k_s v_s^T and P_c = prod_q w_q (plain product)
//   wkv_chunk_scan_kernel    S_start[c] = state entering chunk c (T/16 multiply-adds per entry;
//                            overwrites dS_c in place), final state -> state_out
//   wkv_chunk_out_kernel     y of the chunk from S_start[c] (+ per-head norm, ln_x, gate fused)
// Sums are FP32 in a different order than the scan: results agree to FP32 rounding, not bit
// for bit (tests/test_gpu_models.py compares a sequence pass with token-by-token decode).
#include "device_utils.cuh"
#include "launch.cuh"

namespace rb {

namespace {

constexpr int CS = 64;            // head size
constexpr int CC = 16;            // tokens per chunk
constexpr int KP = CS + 1;        // padded row stride where a warp reads one column of 16 different rows

__device__ __forceinline__ void cpa16(void * smem, const void * gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"((uint32_t) __cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cpa_commit_wait() {
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// log of one decay factor; w = 0 (exp(-exp(x)) underflowed) behaves as a factor of 1e-38
__device__ __forceinline__ float log_decay(float w) { return logf(fmaxf(w, 1.0e-38f)); }

// decay factors of token t0 + t for this thread's 4 channels (1 past the end of the sequence)
__device__ __forceinline__ float4 load_decay4(const WkvArgs & a, int h, int t0) {
    const int t = threadIdx.x >> 4, c4 = (threadIdx.x & 15) * 4;
    if (t0 + t >= a.R) return make_float4(1.0f, 1.0f, 1.0f, 1.0f);
    if (a.w) return *reinterpret_cast<const float4 *>(a.w + (size_t) (t0 + t) * a.D + h * CS + c4);
    if (a.per_head_scalar) { const float v = a.decay[h]; return make_float4(v, v, v, v); }
    return *reinterpret_cast<const float4 *>(a.decay + h * CS + c4);
}
// lw[t][i] <- log w of token t0 + t, one float4 per thread
__device__ __forceinline__ void store_log_decay(const float4 w, float (*lw)[CS]) {
    const int t = threadIdx.x >> 4, c4 = (threadIdx.x & 15) * 4;
    *reinterpret_cast<float4 *>(&lw[t][c4]) = make_float4(log_decay(w.x), log_decay(w.y), log_decay(w.z), log_decay(w.w));
}

__device__ __forceinline__ float4 load_row4(const float * p, const WkvArgs & a, int h, int t0) {
    const int t = threadIdx.x >> 4, c4 = (threadIdx.x & 15) * 4;
    if (t0 + t >= a.R) return make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    return *reinterpret_cast<const float4 *>(p + (size_t) (t0 + t) * a.D + h * CS + c4);
}

// ---- 1: chunk-local state increment -------------------------------------------------------------
__global__ void __launch_bounds__(256) wkv_chunk_state_kernel(const __grid_constant__ WkvArgs a, float * __restrict__ ds, float * __restrict__ pc, int nchunks) {
    __shared__ __align__(16) float lw[CC][CS];
    __shared__ __align__(16) float lp[CC + 1][CS];
    __shared__ __align__(16) float kd[CC][CS];
    __shared__ __align__(16) float sv[CC][CS];
    pdl_trigger();
    const int tid = threadIdx.x, c = blockIdx.x, h = blockIdx.y, t0 = c * CC;
    pdl_wait();
    const int t = tid >> 4, c4 = (tid & 15) * 4;
    {
        const float4 w4 = load_decay4(a, h, t0);
        store_log_decay(w4, lw);
        *reinterpret_cast<float4 *>(&kd[t][c4]) = w4;      // raw factors for the chunk product (kd proper is built below)
    }
    const float4 k4 = load_row4(a.k, a, h, t0);
    *reinterpret_cast<float4 *>(&sv[t][c4]) = load_row4(a.v, a, h, t0);
    __syncthreads();
    if (tid < CS) {
        // LP by sums of logs (used in ratios inside the chunk only); the factor that carries the state
        // from chunk to chunk is the plain product, so no log / exp error compounds along the sequence
        float acc = 0.0f, prod = 1.0f;
        lp[0][tid] = 0.0f;
#pragma unroll
        for (int q = 0; q < CC; q++) { acc += lw[q][tid]; lp[q + 1][tid] = acc; prod *= kd[q][tid]; }
        pc[((size_t) h * nchunks + c) * CS + tid] = prod;
    }
    __syncthreads();
    {
        const float4 le = *reinterpret_cast<const float4 *>(&lp[CC][c4]);
        const float4 ls = *reinterpret_cast<const float4 *>(&lp[t + 1][c4]);
        *reinterpret_cast<float4 *>(&kd[t][c4]) = make_float4(k4.x * __expf(le.x - ls.x), k4.y * __expf(le.y - ls.y),
                                                               k4.z * __expf(le.z - ls.z), k4.w * __expf(le.w - ls.w));
    }
    __syncthreads();
    // dS[i][j] = sum_s kd[s][i] v[s][j]; thread = 4 x 4 tile
    const int i0 = (tid >> 4) * 4, j0 = (tid & 15) * 4;
    float acc[4][4] = {};
#pragma unroll
    for (int s = 0; s < CC; s++) {
        const float4 ka = *reinterpret_cast<const float4 *>(&kd[s][i0]);
        const float4 vb = *reinterpret_cast<const float4 *>(&sv[s][j0]);
        const float kk[4] = {ka.x, ka.y, ka.z, ka.w}, vv[4] = {vb.x, vb.y, vb.z, vb.w};
#pragma unroll
        for (int ii = 0; ii < 4; ii++)
#pragma unroll
            for (int jj = 0; jj < 4; jj++) acc[ii][jj] = fmaf(kk[ii], vv[jj], acc[ii][jj]);
    }
    float * out = ds + ((size_t) h * nchunks + c) * CS * CS;
#pragma unroll
    for (int ii = 0; ii < 4; ii++)
        *reinterpret_cast<float4 *>(out + (size_t) (i0 + ii) * CS + j0) = make_float4(acc[ii][0], acc[ii][1], acc[ii][2], acc[ii][3]);
}

// ---- 2: state entering every chunk ----------------------------------------------------------------
// grid (H, 4): CTA = 16 key rows of a head, thread = one float4 of the state.
__global__ void __launch_bounds__(256) wkv_chunk_scan_kernel(const __grid_constant__ WkvArgs a, float * __restrict__ ds, const float * __restrict__ pc, int nchunks) {
    pdl_trigger();
    const int tid = threadIdx.x, h = blockIdx.x;
    const int i = blockIdx.y * 16 + (tid >> 4), j0 = (tid & 15) * 4;
    const size_t so = (size_t) h * CS * CS + (size_t) i * CS + j0;
    float4 cur = *reinterpret_cast<const float4 *>(a.state_in + so);
    pdl_wait();
    float4 * p = reinterpret_cast<float4 *>(ds + (size_t) h * nchunks * CS * CS + (size_t) i * CS + j0);
    const float * pp = pc + (size_t) h * nchunks * CS + i;
    constexpr int U = 8;
    for (int c0 = 0; c0 < nchunks; c0 += U) {
        float4 d[U]; float f[U];
#pragma unroll
        for (int q = 0; q < U; q++)
            if (c0 + q < nchunks) { d[q] = p[(size_t) (c0 + q) * (CS * CS / 4)]; f[q] = pp[(size_t) (c0 + q) * CS]; }
#pragma unroll
        for (int q = 0; q < U; q++)
            if (c0 + q < nchunks) {
                p[(size_t) (c0 + q) * (CS * CS / 4)] = cur;
                cur = make_float4(fmaf(cur.x, f[q], d[q].x), fmaf(cur.y, f[q], d[q].y), fmaf(cur.z, f[q], d[q].z), fmaf(cur.w, f[q], d[q].w));
            }
    }
    *reinterpret_cast<float4 *>(a.state_out + so) = cur;
}

// ---- 3: outputs of a chunk ----------------------------------------------------------------------------
__global__ void __launch_bounds__(256) wkv_chunk_out_kernel(const __grid_constant__ WkvArgs a, const float * __restrict__ ds, int nchunks) {
    __shared__ __align__(16) float sst[CS * CS];          // S_start of the chunk; later the 4 partial y tiles, then y
    __shared__ __align__(16) float lw[CC][CS];            // log w, then rqT[i][t] = r_t[i] exp(LP_t[i])
    __shared__ float lp[CC + 1][KP];
    __shared__ float sr[CC][CS];
    __shared__ float sk[CC][KP];
    __shared__ __align__(16) float sv[CC][CS];
    __shared__ __align__(16) float sa[CC][CC];            // sa[s][t] = A[t][s]
    __shared__ float su[CS];
    pdl_trigger();
    const int tid = threadIdx.x, c = blockIdx.x, h = blockIdx.y, t0 = c * CC;
    const int t = tid >> 4, c4 = (tid & 15) * 4;
    if (tid < CS) su[tid] = a.per_head_scalar ? a.u[h] : a.u[h * CS + tid];
    pdl_wait();
    {
        const float * src = ds + ((size_t) h * nchunks + c) * CS * CS;
#pragma unroll
        for (int q = 0; q < 4; q++) cpa16(sst + (q * 256 + tid) * 4, src + (q * 256 + tid) * 4);
    }
    store_log_decay(load_decay4(a, h, t0), lw);
    {
        const float4 r4 = load_row4(a.r, a, h, t0), k4 = load_row4(a.k, a, h, t0);
        sr[t][c4] = r4.x; sr[t][c4 + 1] = r4.y; sr[t][c4 + 2] = r4.z; sr[t][c4 + 3] = r4.w;
        sk[t][c4] = k4.x; sk[t][c4 + 1] = k4.y; sk[t][c4 + 2] = k4.z; sk[t][c4 + 3] = k4.w;
        *reinterpret_cast<float4 *>(&sv[t][c4]) = load_row4(a.v, a, h, t0);
    }
    __syncthreads();
    if (tid < CS) {
        float acc = 0.0f;
        lp[0][tid] = 0.0f;
#pragma unroll
        for (int q = 0; q < CC; q++) { acc += lw[q][tid]; lp[q + 1][tid] = acc; }
    }
    __syncthreads();
    // A[t][s]: thread (t, s); rqT (into lw's storage: lw is dead after the prefix sums)
    {
        const int s = tid & 15;
        float acc = 0.0f;
        if (s < t) {
#pragma unroll 8
            for (int i = 0; i < CS; i++) acc = fmaf(sr[t][i] * sk[s][i], __expf(lp[t][i] - lp[s + 1][i]), acc);
        } else if (s == t) {
#pragma unroll 8
            for (int i = 0; i < CS; i++) acc = fmaf(sr[t][i] * sk[t][i], su[i], acc);
        }
        sa[s][t] = acc;
    }
    float (*rqT)[CC] = reinterpret_cast<float (*)[CC]>(&lw[0][0]);     // [CS][CC]
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const int e = q * 256 + tid, i = e >> 4, tt = e & 15;
        rqT[i][tt] = sr[tt][i] * __expf(lp[tt][i]);
    }
    cpa_commit_wait();
    __syncthreads();
    // y[t][j]: thread = 4 tokens x 4 columns, K split over 4 thread groups (16 keys + 4 intra tokens each)
    const int g = tid >> 6, tq = ((tid >> 4) & 3) * 4, j0 = (tid & 15) * 4;
    float acc[4][4] = {};
#pragma unroll 4
    for (int ii = 0; ii < 16; ii++) {
        const int i = g * 16 + ii;
        const float4 ra = *reinterpret_cast<const float4 *>(&rqT[i][tq]);
        const float4 sb = *reinterpret_cast<const float4 *>(&sst[i * CS + j0]);
        const float rr[4] = {ra.x, ra.y, ra.z, ra.w}, ss[4] = {sb.x, sb.y, sb.z, sb.w};
#pragma unroll
        for (int x = 0; x < 4; x++)
#pragma unroll
            for (int y = 0; y < 4; y++) acc[x][y] = fmaf(rr[x], ss[y], acc[x][y]);
    }
#pragma unroll
    for (int ii = 0; ii < 4; ii++) {
        const int s = g * 4 + ii;
        const float4 aa = *reinterpret_cast<const float4 *>(&sa[s][tq]);
        const float4 vb = *reinterpret_cast<const float4 *>(&sv[s][j0]);
        const float av[4] = {aa.x, aa.y, aa.z, aa.w}, vv[4] = {vb.x, vb.y, vb.z, vb.w};
#pragma unroll
        for (int x = 0; x < 4; x++)
#pragma unroll
            for (int y = 0; y < 4; y++) acc[x][y] = fmaf(av[x], vv[y], acc[x][y]);
    }
    __syncthreads();                                       // every group is done reading S_start
#pragma unroll
    for (int x = 0; x < 4; x++)
        *reinterpret_cast<float4 *>(&sst[(g * CC + tq + x) * CS + j0]) = make_float4(acc[x][0], acc[x][1], acc[x][2], acc[x][3]);
    __syncthreads();
    // thread (t, 4 columns): total, per-head norm over the 16 threads of a token, ln_x, gate
    float y[4];
    {
        const float4 p0 = *reinterpret_cast<const float4 *>(&sst[(0 * CC + t) * CS + c4]);
        const float4 p1 = *reinterpret_cast<const float4 *>(&sst[(1 * CC + t) * CS + c4]);
        const float4 p2 = *reinterpret_cast<const float4 *>(&sst[(2 * CC + t) * CS + c4]);
        const float4 p3 = *reinterpret_cast<const float4 *>(&sst[(3 * CC + t) * CS + c4]);
        y[0] = (p0.x + p1.x) + (p2.x + p3.x); y[1] = (p0.y + p1.y) + (p2.y + p3.y);
        y[2] = (p0.z + p1.z) + (p2.z + p3.z); y[3] = (p0.w + p1.w) + (p2.w + p3.w);
    }
    double s1 = ((double) y[0] + (double) y[1]) + ((double) y[2] + (double) y[3]);
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) s1 += __shfl_xor_sync(FULL, s1, o);
    const float mean = (float) (s1 / CS);
    double s2 = 0.0;
#pragma unroll
    for (int x = 0; x < 4; x++) { y[x] = __fsub_rn(y[x], mean); s2 += (double) __fmul_rn(y[x], y[x]); }
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) s2 += __shfl_xor_sync(FULL, s2, o);
    const float scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn((float) (s2 / CS), a.gn_eps)));
    const bool row_ok = t0 + t < a.R;
    {
        const int e = h * CS + c4;
        const float4 lw4 = *reinterpret_cast<const float4 *>(a.lnx_w + e), lb4 = *reinterpret_cast<const float4 *>(a.lnx_b + e);
        float4 g4 = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
        if (a.g && row_ok) g4 = *reinterpret_cast<const float4 *>(a.g + (size_t) (t0 + t) * a.D + e);
        y[0] = __fadd_rn(__fmul_rn(__fmul_rn(y[0], scale), lw4.x), lb4.x);
        y[1] = __fadd_rn(__fmul_rn(__fmul_rn(y[1], scale), lw4.y), lb4.y);
        y[2] = __fadd_rn(__fmul_rn(__fmul_rn(y[2], scale), lw4.z), lb4.z);
        y[3] = __fadd_rn(__fmul_rn(__fmul_rn(y[3], scale), lw4.w), lb4.w);
        if (a.g) { y[0] = __fmul_rn(y[0], g4.x); y[1] = __fmul_rn(y[1], g4.y); y[2] = __fmul_rn(y[2], g4.z); y[3] = __fmul_rn(y[3], g4.w); }
    }
    __syncthreads();                                       // partial tiles consumed
    *reinterpret_cast<float4 *>(&sst[t * CS + c4]) = make_float4(y[0], y[1], y[2], y[3]);
    __syncthreads();
    // rows out in whatever forms the consumer wants (FP32 / FP16 / q8 blocks): warp = 32 consecutive channels
    const int warp = tid >> 5, lane = tid & 31;
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const int blk = q * 8 + warp, tt = blk >> 1, half = blk & 1;
        if (t0 + tt < a.R) act_store_block(a.y, (size_t) (t0 + tt), h * CS + half * 32 + lane, true, sst[tt * CS + half * 32 + lane]);
    }
}

} // namespace

bool wkv5_chunked_supported(const WkvArgs & a) {
    static const bool off = [] { const char * e = getenv("RWKV_B200_NO_WKV_CHUNK"); return e && e[0] == '1'; }();
    return !off && !a.batch && a.S == CS && a.D == a.H * CS && a.chunk_ds != nullptr && a.chunk_pc != nullptr && (a.D % 4) == 0;
}

cudaError_t launch_wkv5_chunked(const WkvArgs & a, cudaStream_t st) {
    if (!wkv5_chunked_supported(a)) return cudaErrorInvalidValue;
    const int nchunks = (a.R + CC - 1) / CC;
    cudaError_t e = launch_pdl(wkv_chunk_state_kernel, dim3(nchunks, a.H), dim3(256), 0, st, a, a.chunk_ds, a.chunk_pc, nchunks);
    if (e != cudaSuccess) return e;
    e = launch_pdl(wkv_chunk_scan_kernel, dim3(a.H, 4), dim3(256), 0, st, a, a.chunk_ds, (const float *) a.chunk_pc, nchunks);
    if (e != cudaSuccess) return e;
    return launch_pdl(wkv_chunk_out_kernel, dim3(nchunks, a.H), dim3(256), 0, st, a, (const float *) a.chunk_ds, nchunks);
}

} // namespace rb
