// v5 / v6 WKV over long sequences, chunk-parallel form (head size 64, rows >= 64 of FP16 /
// quantised models -- the rows that also take the tensor-core MATMUL, i.e. the tolerance-mode
// sequence path; everything else keeps the exact scan of wkv.cu).
//
// Reference recurrence (ggml_compute_forward_rwkv_wkv6_f32, ggml-cpu.c:11945-11980), per head,
// key index i, value index j:
//     y_t[j]    = sum_i r_t[i] * (u[i] k_t[i] v_t[j] + S_{t-1}[i][j])
//     S_t[i][j] = S_{t-1}[i][j] w_t[i] + k_t[i] v_t[j]
// is sequential in t: the scan kernel spends 180 ns per token whatever the GPU.  Unrolling the
// recurrence over a chunk of C = 16 tokens, with W(a, b)[i] = prod_{a <= q < b} w_q[i]:
//     S_{t-1}  = S_in * W(0, t)  +  sum_{s<t} W(s+1, t) k_s v_s^T
//     y_t[j]   = sum_i r_t[i] W(0, t)[i] S_in[i][j]                               (inter)
//              + sum_{s<t} (sum_i r_t[i] k_s[i] W(s+1, t)[i]) v_s[j]               (intra)
//              + (sum_i r_t[i] u[i] k_t[i]) v_t[j]                                 (bonus)
//     S_out    = S_in * W(0, C) + sum_s W(s+1, C) k_s v_s^T
// Only products of decay factors in (0, 1] appear (no logarithms, no divisions), so nothing can
// overflow however strong the decay is, and an underflow is the correct limit.
// Three launches per layer, the first and last parallel over (head, chunk):
//   wkv_chunk_state_kernel   dS_c = sum_s W(s+1, C) k_s v_s^T and P_c = W(0, C)
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

__device__ __forceinline__ void cpa16(void * smem, const void * gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"((uint32_t) __cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cpa_commit_wait() {
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ float4 mul4(float4 a, float4 b) { return make_float4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
__device__ __forceinline__ float4 & f4(float * p) { return *reinterpret_cast<float4 *>(p); }
__device__ __forceinline__ const float4 & f4(const float * p) { return *reinterpret_cast<const float4 *>(p); }

// Thread layout of the loads: thread = (token t = tid / 16 of the chunk, channels c4 .. c4 + 3 of the head).
// decay factors of token t0 + t (1 past the end of the sequence: no decay, and k = v = r = 0 there)
__device__ __forceinline__ float4 load_decay4(const WkvArgs & a, int h, int t0) {
    const int t = threadIdx.x >> 4, c4 = (threadIdx.x & 15) * 4;
    if (t0 + t >= a.R) return make_float4(1.0f, 1.0f, 1.0f, 1.0f);
    if (a.w) return f4(a.w + (size_t) (t0 + t) * a.D + h * CS + c4);
    if (a.per_head_scalar) { const float v = a.decay[h]; return make_float4(v, v, v, v); }
    return f4(a.decay + h * CS + c4);
}
__device__ __forceinline__ float4 load_row4(const float * p, const WkvArgs & a, int h, int t0) {
    const int t = threadIdx.x >> 4, c4 = (threadIdx.x & 15) * 4;
    if (t0 + t >= a.R) return make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    return f4(p + (size_t) (t0 + t) * a.D + h * CS + c4);
}

// ---- 1: chunk-local state increment -------------------------------------------------------------
__global__ void __launch_bounds__(256) wkv_chunk_state_kernel(const __grid_constant__ WkvArgs a, float * ds, float * pc, int nchunks) {
    __shared__ __align__(16) float sw[CC][CS];
    __shared__ __align__(16) float kd[CC][CS];
    __shared__ __align__(16) float sv[CC][CS];
    pdl_trigger();
    const int tid = threadIdx.x, c = blockIdx.x, h = blockIdx.y, t0 = c * CC;
    pdl_wait();
    const int t = tid >> 4, c4 = (tid & 15) * 4;
    const float4 w4 = load_decay4(a, h, t0);
    f4(&sw[t][c4]) = w4;
    const float4 k4 = load_row4(a.k, a, h, t0);
    f4(&sv[t][c4]) = load_row4(a.v, a, h, t0);
    __syncthreads();
    {
        float4 sfx = make_float4(1.0f, 1.0f, 1.0f, 1.0f);       // W(t+1, C)
        for (int q = CC - 1; q > t; q--) sfx = mul4(sfx, f4(&sw[q][c4]));
        f4(&kd[t][c4]) = mul4(k4, sfx);
        if (t == 0) f4(pc + ((size_t) h * nchunks + c) * CS + c4) = mul4(sfx, w4);
    }
    __syncthreads();
    // dS[i][j] = sum_s kd[s][i] v[s][j]; thread = 4 x 4 tile
    const int i0 = (tid >> 4) * 4, j0 = (tid & 15) * 4;
    float acc[4][4] = {};
#pragma unroll
    for (int s = 0; s < CC; s++) {
        const float4 ka = f4(&kd[s][i0]);
        const float4 vb = f4(&sv[s][j0]);
        const float kk[4] = {ka.x, ka.y, ka.z, ka.w}, vv[4] = {vb.x, vb.y, vb.z, vb.w};
#pragma unroll
        for (int ii = 0; ii < 4; ii++)
#pragma unroll
            for (int jj = 0; jj < 4; jj++) acc[ii][jj] = fmaf(kk[ii], vv[jj], acc[ii][jj]);
    }
    float * out = ds + ((size_t) h * nchunks + c) * CS * CS;
#pragma unroll
    for (int ii = 0; ii < 4; ii++)
        f4(out + (size_t) (i0 + ii) * CS + j0) = make_float4(acc[ii][0], acc[ii][1], acc[ii][2], acc[ii][3]);
}

// ---- 2: state entering every chunk ----------------------------------------------------------------
// grid (H, 4): CTA = 16 key rows of a head, thread = one float4 of the state.  The loads of the
// next 8 chunks are in flight while the current 8 are chained.
__global__ void __launch_bounds__(256) wkv_chunk_scan_kernel(const __grid_constant__ WkvArgs a, float * ds, const float * pc /* written by the preceding kernel: NOT __restrict__ (invariant loads may be hoisted above griddepcontrol.wait) */, int nchunks) {
    pdl_trigger();
    const int tid = threadIdx.x, h = blockIdx.x;
    const int i = blockIdx.y * 16 + (tid >> 4), j0 = (tid & 15) * 4;
    const size_t so = (size_t) h * CS * CS + (size_t) i * CS + j0;
    float4 cur = f4(a.state_in + so);
    pdl_wait();
    float4 * p = reinterpret_cast<float4 *>(ds + (size_t) h * nchunks * CS * CS + (size_t) i * CS + j0);
    const float * pp = pc + (size_t) h * nchunks * CS + i;
    constexpr int U = 8;
    float4 d[U], dn[U]; float f[U], fn[U];
    auto fetch = [&](int c0, float4 (&dd)[U], float (&ff)[U]) {
#pragma unroll
        for (int q = 0; q < U; q++) {
            const int c = min(c0 + q, nchunks - 1);          // clamped: the surplus loads are ignored
            dd[q] = __ldcg(p + (size_t) c * (CS * CS / 4)); ff[q] = __ldcg(pp + (size_t) c * CS);
        }
    };
    fetch(0, d, f);
    for (int c0 = 0; c0 < nchunks; c0 += U) {
        if (c0 + U < nchunks) fetch(c0 + U, dn, fn);
#pragma unroll
        for (int q = 0; q < U; q++)
            if (c0 + q < nchunks) {
                p[(size_t) (c0 + q) * (CS * CS / 4)] = cur;
                cur = make_float4(fmaf(cur.x, f[q], d[q].x), fmaf(cur.y, f[q], d[q].y), fmaf(cur.z, f[q], d[q].z), fmaf(cur.w, f[q], d[q].w));
            }
#pragma unroll
        for (int q = 0; q < U; q++) { d[q] = dn[q]; f[q] = fn[q]; }
    }
    f4(a.state_out + so) = cur;
}

// ---- 3: outputs of a chunk ----------------------------------------------------------------------------
__global__ void __launch_bounds__(256) wkv_chunk_out_kernel(const __grid_constant__ WkvArgs a, const float * ds /* written by the preceding kernel: NOT __restrict__ (invariant loads may be hoisted above griddepcontrol.wait) */, int nchunks) {
    __shared__ __align__(16) float sst[CS * CS];          // S_start of the chunk; later the 4 partial y tiles, then y
    __shared__ __align__(16) float sw[CC][CS];
    __shared__ __align__(16) float sr[CC][CS];
    __shared__ __align__(16) float sk[CC][CS];
    __shared__ __align__(16) float sv[CC][CS];
    __shared__ __align__(16) float rq[CC][CS];            // r_t[i] W(0, t)[i]
    __shared__ __align__(16) float sa[CC][CC];            // sa[s][t] = A[t][s]
    __shared__ __align__(16) float su[CS];
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
    const float4 r4 = load_row4(a.r, a, h, t0);
    f4(&sw[t][c4]) = load_decay4(a, h, t0);
    f4(&sr[t][c4]) = r4;
    f4(&sk[t][c4]) = load_row4(a.k, a, h, t0);
    f4(&sv[t][c4]) = load_row4(a.v, a, h, t0);
    __syncthreads();
    // rq[t][i] = r_t[i] W(0, t)[i]
    {
        float4 pre = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
        for (int q = 0; q < t; q++) pre = mul4(pre, f4(&sw[q][c4]));
        f4(&rq[t][c4]) = mul4(r4, pre);
    }
    // A[.][s] for s = tid / 16 over this thread's 4 keys: g = k_s W(s+1, t) grows token by token
    {
        const int s = t;                                   // same thread layout: (s, 4 keys)
        const int smin = (tid >> 5) * 2;                   // smallest s of the warp
        float4 g = f4(&sk[s][c4]);
        float acc[CC];
#pragma unroll
        for (int q = 0; q < CC; q++) acc[q] = 0.0f;
        {
            const float4 uu = f4(&su[c4]);
            const float bonus = fmaf(r4.x * g.x, uu.x, fmaf(r4.y * g.y, uu.y, fmaf(r4.z * g.z, uu.z, r4.w * g.w * uu.w)));
#pragma unroll
            for (int q = 0; q < CC; q++) if (q == s) acc[q] = bonus;
        }
#pragma unroll
        for (int q = 1; q < CC; q++) {
            if (q > smin) {                                // warp-uniform
                const float4 rr = f4(&sr[q][c4]);
                const float dot = fmaf(rr.x, g.x, fmaf(rr.y, g.y, fmaf(rr.z, g.z, rr.w * g.w)));
                if (q > s) { acc[q] = dot; g = mul4(g, f4(&sw[q][c4])); }
            }
        }
        // sum over the 16 lanes that share s; lane l of the group ends up with the total of acc[l]
#pragma unroll
        for (int off = 8; off > 0; off >>= 1) {
            const bool up = (tid & off) != 0;
#pragma unroll
            for (int q = 0; q < off; q++) {
                const float send = up ? acc[q] : acc[q + off];
                const float keep = up ? acc[q + off] : acc[q];
                acc[q] = keep + __shfl_xor_sync(FULL, send, off);
            }
        }
        sa[s][tid & 15] = acc[0];
    }
    cpa_commit_wait();
    __syncthreads();
    // y[t][j]: thread = 4 tokens x 4 columns, K split over 4 thread groups (16 keys + 4 intra tokens each)
    const int g = tid >> 6, tq = ((tid >> 4) & 3) * 4, j0 = (tid & 15) * 4;
    float acc[4][4] = {};
#pragma unroll
    for (int ii = 0; ii < 4; ii++) {
        const int i = g * 16 + ii * 4;
        float ra[4][4], sb[4][4];
#pragma unroll
        for (int x = 0; x < 4; x++) { const float4 v = f4(&rq[tq + x][i]); ra[x][0] = v.x; ra[x][1] = v.y; ra[x][2] = v.z; ra[x][3] = v.w; }
#pragma unroll
        for (int z = 0; z < 4; z++) { const float4 v = f4(&sst[(i + z) * CS + j0]); sb[z][0] = v.x; sb[z][1] = v.y; sb[z][2] = v.z; sb[z][3] = v.w; }
#pragma unroll
        for (int z = 0; z < 4; z++)
#pragma unroll
            for (int x = 0; x < 4; x++)
#pragma unroll
                for (int y = 0; y < 4; y++) acc[x][y] = fmaf(ra[x][z], sb[z][y], acc[x][y]);
    }
#pragma unroll
    for (int ii = 0; ii < 4; ii++) {
        const int s = g * 4 + ii;
        const float4 aa = f4(&sa[s][tq]);
        const float4 vb = f4(&sv[s][j0]);
        const float av[4] = {aa.x, aa.y, aa.z, aa.w}, vv[4] = {vb.x, vb.y, vb.z, vb.w};
#pragma unroll
        for (int x = 0; x < 4; x++)
#pragma unroll
            for (int y = 0; y < 4; y++) acc[x][y] = fmaf(av[x], vv[y], acc[x][y]);
    }
    __syncthreads();                                       // every group is done reading S_start
#pragma unroll
    for (int x = 0; x < 4; x++)
        f4(&sst[(g * CC + tq + x) * CS + j0]) = make_float4(acc[x][0], acc[x][1], acc[x][2], acc[x][3]);
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
