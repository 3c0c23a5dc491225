// WKV state recurrences, fused with what surrounds them in the reference graph
// (group norm, ln_x affine, gate, v7 key/value preparation and bonus term), producing the
// activation rows of the att.output mat-vec.
//
//   v4      rwkv_att_wkv_v4           rwkv_graph.inc:119-161 (per channel, running-max form)
//   v5/v6   ggml_rwkv_wkv6            ggml.c:4627 -> ggml-cpu.c:11803-11980
//           + per-head ggml_norm, ln_x, gate   rwkv_graph.inc:276-290 (v5), :371-384 (v6)
//   v7      rwkv_wkv_v7_impl          rwkv_operators_wkv_v7.inc:37-121
//           + l2norm / k,v preparation / bonus  rwkv_graph.inc:428-477, rwkv_operators.inc:40-83
//
// Rows are either T consecutive tokens of one stream (the scan runs inside the kernel,
// state held in registers) or B independent streams (one step each).
#include "device_utils.cuh"
#include "launch.cuh"
#include "trace.cuh"

namespace rb {

// ----------------------------------------------------------------------------------------
// v4: one thread per channel.
// ----------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) wkv4_kernel(const __grid_constant__ WkvArgs a) {
    TRACE_DECL; TRACE(0);
    pdl_trigger();
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    const int D = a.D;                         // channels in the rows (a shard under tensor parallelism)
    const int SD = a.state_D ? a.state_D : D; // width of a state slot
    const bool valid = e < D;
    const int stream = a.batch ? blockIdx.y : 0;
    const size_t soff = (size_t) stream * a.state_stride + a.chan0;
    float aa = 0, bb = 0, pp = 0, u = 0, w = 0;
    if (valid) {
        aa = a.state_in[soff + e]; bb = a.state_in[soff + SD + e]; pp = a.state_in[soff + 2 * SD + e];
        u = a.u[a.chan0 + e]; w = a.decay[a.chan0 + e];
    }
    TRACE(1);
    pdl_wait();                                 // r, k, v come from the preceding mat-vec; the input state is read-only
    TRACE(2);
    const int steps = a.batch ? 1 : a.R;
    for (int s = 0; s < steps; s++) {
        const size_t row = a.batch ? stream : s;
        float out = 0.0f;
        if (valid) {
            const float k = a.k[row * D + e], v = a.v[row * D + e], r = a.r[row * D + e];
            float ww = __fadd_rn(u, k);
            float qq = fmaxf(pp, ww);
            float e1 = expf(__fsub_rn(pp, qq)), e2 = expf(__fsub_rn(ww, qq));
            const float num = __fadd_rn(__fmul_rn(e1, aa), __fmul_rn(e2, v));
            const float den = __fadd_rn(__fmul_rn(e1, bb), e2);
            ww = __fadd_rn(pp, w);
            qq = fmaxf(ww, k);
            e1 = expf(__fsub_rn(ww, qq)); e2 = expf(__fsub_rn(k, qq));
            aa = __fadd_rn(__fmul_rn(e1, aa), __fmul_rn(e2, v));
            bb = __fadd_rn(__fmul_rn(e1, bb), e2);
            pp = qq;
            out = __fmul_rn(r, __fdiv_rn(num, den));
        }
        act_store_block(a.y, row, e, valid, out);
    }
    if (valid) { a.state_out[soff + e] = aa; a.state_out[soff + SD + e] = bb; a.state_out[soff + 2 * SD + e] = pp; }
    TRACE(5);
    TRACE_FLUSH(200, gridDim.x * gridDim.y);
}

cudaError_t launch_wkv4(const WkvArgs & a, cudaStream_t st) {
    dim3 grid((a.D + 255) / 256, a.batch ? a.R : 1);
    return launch_pdl(wkv4_kernel, grid, dim3(256), 0, st, a);
}

// Sum over the S consecutive threads of a head (S = 8..64, groups aligned to S); every thread
// of the group receives the same double.  Butterfly inside the warp, then (S = 64) the two
// warp totals through shared memory.  All threads of the CTA must call it.
template <int S>
__device__ __forceinline__ double head_sum_d(double v, float * sy, int tid) {
    constexpr int W = S < 32 ? S : 32;
#pragma unroll
    for (int o = W / 2; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    if (S > 32) {
        double * red = reinterpret_cast<double *>(sy);
        __syncthreads();
        if ((tid & 31) == 0) red[tid >> 5] = v;
        __syncthreads();
        const int w0 = (tid >> 6) << 1;
        v = red[w0] + red[w0 + 1];
    }
    return v;
}

// Per-head normalisation of `y` over the S threads of a head (ggml norm: mean and variance
// accumulated in double, ggml-cpu.c:6906-6923).  sy: shared scratch of >= blockDim floats,
// 8-byte aligned; returns the normalised value for this thread.
template <int S>
__device__ __forceinline__ float head_norm(float y, float * sy, int tid, bool active, float eps) {
    const double s = head_sum_d<S>(active ? (double) y : 0.0, sy, tid);
    const float mean = (float) (s / S);
    const float vv = __fsub_rn(y, mean);
    const double s2 = head_sum_d<S>(active ? (double) __fmul_rn(vv, vv) : 0.0, sy, tid);
    const float var = (float) (s2 / S);
    const float scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(var, eps)));
    return __fmul_rn(vv, scale);
}

// ----------------------------------------------------------------------------------------
// v5.1 / v5.2 / v6.  CTA = HPB heads, thread (hl, j) owns column j (value index) of the
// head state S[i][j]; state[h*S*S + i*S + j] (i = key index).
// ----------------------------------------------------------------------------------------
// The output y[j] = sum_i r[i] * (u[i]*k[i]*v[j] + S[i][j]) is accumulated in a fixed order shared
// by the fused kernel below (one thread per column) and the sequence scan (8 lanes per column):
// 8 chains over consecutive key ranges of S/8, combined as ((0+1)+(2+3)) + ((4+5)+(6+7)).
constexpr int WKV_CHAINS = 8;
__device__ __forceinline__ float wkv_combine(const float (&yc)[WKV_CHAINS]) {
    return __fadd_rn(__fadd_rn(__fadd_rn(yc[0], yc[1]), __fadd_rn(yc[2], yc[3])), __fadd_rn(__fadd_rn(yc[4], yc[5]), __fadd_rn(yc[6], yc[7])));
}

template <int S>
__global__ void wkv5_kernel(const __grid_constant__ WkvArgs a) {
    extern __shared__ __align__(16) float sm[];
    constexpr int L = S / WKV_CHAINS > 0 ? S / WKV_CHAINS : 1;
    pdl_trigger();
    const int nthr = blockDim.x;
    float * sr = sm, * sk = sm + nthr, * sw = sm + 2 * nthr, * su = sm + 3 * nthr, * sy = sm + 4 * nthr;
    const int tid = threadIdx.x;
    const int hpb = nthr / S > 0 ? nthr / S : 1;
    const int h = blockIdx.x * hpb + tid / S, j = tid % S;
    const bool active = (tid < hpb * S) && h < a.H;
    const int e = h * S + j, D = a.D;
    const int hb = (tid / S) * S;
    const int stream = a.batch ? blockIdx.y : 0;
    const size_t soff = (size_t) stream * a.state_stride;

    float st[S];
    if (active) {
#pragma unroll
        for (int i = 0; i < S; i++) st[i] = a.state_in[soff + (size_t) h * S * S + (size_t) i * S + j];
        su[tid] = a.per_head_scalar ? a.u[h] : a.u[e];
        if (a.w == nullptr) sw[tid] = a.per_head_scalar ? a.decay[h] : a.decay[e];
    }
    pdl_wait();
    const int steps = a.batch ? 1 : a.R;
    for (int s = 0; s < steps; s++) {
        const size_t row = a.batch ? stream : s;
        __syncthreads();
        float vj = 0.0f;
        if (active) {
            sr[tid] = a.r[row * D + e]; sk[tid] = a.k[row * D + e]; vj = a.v[row * D + e];
            if (a.w) sw[tid] = a.w[row * D + e];
        }
        __syncthreads();
        float y = 0.0f;
        if (active) {
            float yc[WKV_CHAINS];
#pragma unroll
            for (int c = 0; c < WKV_CHAINS; c++) yc[c] = 0.0f;
#pragma unroll
            for (int i = 0; i < S; i++) {
                const float kv = __fmul_rn(vj, sk[hb + i]);
                const float tmp = __fmaf_rn(kv, su[hb + i], st[i]);
                yc[i / L] = __fmaf_rn(tmp, sr[hb + i], yc[i / L]);
                st[i] = __fmaf_rn(st[i], sw[hb + i], kv);
            }
            y = wkv_combine(yc);
        }
        float o = head_norm<S>(y, sy, tid, active, a.gn_eps);
        if (active) {
            o = __fadd_rn(__fmul_rn(o, a.lnx_w[e]), a.lnx_b[e]);
            if (a.g) o = __fmul_rn(o, a.g[row * D + e]);
        }
        act_store_block(a.y, row, e, active, o);
    }
    if (active) {
#pragma unroll
        for (int i = 0; i < S; i++) a.state_out[soff + (size_t) h * S * S + (size_t) i * S + j] = st[i];
    }
}

// ----------------------------------------------------------------------------------------
// v5 / v6 sequence scan (T tokens of one stream).  The recurrence is sequential in t but
// independent per state column, so a head is spread over S/8 CTAs of 8 columns, and each
// column over 8 lanes that own S/8 consecutive key rows: the per-step work of a lane is S/8
// multiply-adds on registers plus a 3-step butterfly.  r, k, w, v of the next 32 tokens are
// staged in shared memory with cp.async while the current chunk is processed.  The raw y goes
// to `yraw`; wkv5_norm_kernel applies the per-head norm, ln_x and the gate afterwards
// (parallel over tokens).  Same accumulation order as wkv5_kernel, so a sequence pass and
// token-by-token decode agree bit for bit.
// ----------------------------------------------------------------------------------------
constexpr int WKV_TCH = 32;       // tokens per staged chunk
#ifndef WKV_SEQ_LC
#define WKV_SEQ_LC 2     // measured on v6-1.6B T=1024: 1 -> 216 us, 2 -> ~125 us, 4 -> slower than 1 (the chain hand-off serialises)
#endif

__device__ __forceinline__ void cp_async16(void * smem, const void * gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" :: "r"((uint32_t) __cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }

// LC = lanes per accumulation chain (1 or 2): a column is spread over 8*LC lanes that own
// L = S/(8*LC) consecutive key rows each.  With LC = 2 the second lane of a chain continues the
// first lane's partial sum (one shuffle), so the chain is still summed key by key in order.
template <int S, int LC>
__global__ void __launch_bounds__(64 * LC) wkv5_seq_kernel(const __grid_constant__ WkvArgs a, float * yraw) {
    constexpr int P = WKV_CHAINS * LC;                 // lanes per column
    constexpr int L = S / P;                           // key rows per lane
    constexpr int NT = 8 * P;                          // threads: 8 columns per CTA
    constexpr int NV = S / 4;                          // float4 per token per array
    static_assert(L >= 1 && S % P == 0, "head size too small for this lane split");
    extern __shared__ __align__(16) float sm[];
    // per buffer: r, k, w [TCH][S], v [TCH][8]
    constexpr int BUF = WKV_TCH * (3 * S + 8);
    pdl_trigger();
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int h = blockIdx.x, jc = blockIdx.y * 8;      // this CTA's 8 columns of head h
    const int jl = warp * (32 / P) + lane / P, p = lane % P;
    const int j = jc + jl, D = a.D;
    const bool dyn_w = a.w != nullptr;

    float st[L], u[L], wst[L];
#pragma unroll
    for (int ii = 0; ii < L; ii++) {
        const int i = p * L + ii;
        st[ii] = a.state_in[(size_t) h * S * S + (size_t) i * S + j];
        u[ii] = a.per_head_scalar ? a.u[h] : a.u[h * S + i];
        wst[ii] = dyn_w ? 0.0f : (a.per_head_scalar ? a.decay[h] : a.decay[h * S + i]);
    }
    pdl_wait();

    auto stage = [&](int buf, int t0) {
        float * b = sm + buf * BUF;
        const int nt = min(WKV_TCH, a.R - t0);
        for (int x = tid; x < nt * NV; x += NT) {
            const int t = x / NV, c = x % NV;
            const size_t g = (size_t) (t0 + t) * D + h * S + c * 4;
            cp_async16(b + t * S + c * 4, a.r + g);
            cp_async16(b + WKV_TCH * S + t * S + c * 4, a.k + g);
            if (dyn_w) cp_async16(b + 2 * WKV_TCH * S + t * S + c * 4, a.w + g);
        }
        for (int x = tid; x < nt * 2; x += NT) {
            const int t = x >> 1, c = x & 1;
            cp_async16(b + 3 * WKV_TCH * S + t * 8 + c * 4, a.v + (size_t) (t0 + t) * D + h * S + jc + c * 4);
        }
        cp_async_commit();
    };

    const int nchunks = (a.R + WKV_TCH - 1) / WKV_TCH;
    stage(0, 0);
    for (int c = 0; c < nchunks; c++) {
        const int t0 = c * WKV_TCH, nt = min(WKV_TCH, a.R - t0);
        if (c + 1 < nchunks) { stage((c + 1) & 1, t0 + WKV_TCH); cp_async_wait<1>(); } else cp_async_wait<0>();
        __syncthreads();
        const float * b = sm + (c & 1) * BUF;
        const float * sr = b, * sk = b + WKV_TCH * S, * sw = b + 2 * WKV_TCH * S, * sv = b + 3 * WKV_TCH * S;
        // 4 tokens per trip: the state chain is the only dependency between tokens, so the loads,
        // the y chains and the butterflies of the 4 tokens overlap
        auto step = [&](int t, float & y) {
            const float vj = sv[t * 8 + jl];
            float kk[L], rr[L], ww[L];
            if constexpr (L % 4 == 0) {
#pragma unroll
                for (int q4 = 0; q4 < L / 4; q4++) {
                    const float4 k4 = *reinterpret_cast<const float4 *>(sk + t * S + p * L + q4 * 4);
                    const float4 r4 = *reinterpret_cast<const float4 *>(sr + t * S + p * L + q4 * 4);
                    kk[4 * q4] = k4.x; kk[4 * q4 + 1] = k4.y; kk[4 * q4 + 2] = k4.z; kk[4 * q4 + 3] = k4.w;
                    rr[4 * q4] = r4.x; rr[4 * q4 + 1] = r4.y; rr[4 * q4 + 2] = r4.z; rr[4 * q4 + 3] = r4.w;
                    if (dyn_w) {
                        const float4 w4 = *reinterpret_cast<const float4 *>(sw + t * S + p * L + q4 * 4);
                        ww[4 * q4] = w4.x; ww[4 * q4 + 1] = w4.y; ww[4 * q4 + 2] = w4.z; ww[4 * q4 + 3] = w4.w;
                    }
                }
            } else {
#pragma unroll
                for (int ii = 0; ii < L; ii++) { kk[ii] = sk[t * S + p * L + ii]; rr[ii] = sr[t * S + p * L + ii]; if (dyn_w) ww[ii] = sw[t * S + p * L + ii]; }
            }
            float tmp[L];
#pragma unroll
            for (int ii = 0; ii < L; ii++) {
                const float kv = __fmul_rn(vj, kk[ii]);
                tmp[ii] = __fmaf_rn(kv, u[ii], st[ii]);
                st[ii] = __fmaf_rn(st[ii], dyn_w ? ww[ii] : wst[ii], kv);
            }
            y = 0.0f;
#pragma unroll
            for (int ii = 0; ii < L; ii++) y = __fmaf_rn(tmp[ii], rr[ii], y);
            // lane q of a chain restarts from the partial of lane q-1: keys are still added in order
#pragma unroll
            for (int q = 1; q < LC; q++) {
                float y2 = __shfl_up_sync(FULL, y, 1);
#pragma unroll
                for (int ii = 0; ii < L; ii++) y2 = __fmaf_rn(tmp[ii], rr[ii], y2);
                y = (p % LC) == q ? y2 : y;
            }                                              // the last lane of each chain now holds its total
        };
        auto finish = [&](int t, float y) {
            y = __fadd_rn(y, __shfl_xor_sync(FULL, y, LC));
            y = __fadd_rn(y, __shfl_xor_sync(FULL, y, 2 * LC));
            y = __fadd_rn(y, __shfl_xor_sync(FULL, y, 4 * LC));
            if (p == LC - 1) yraw[(size_t) (t0 + t) * D + h * S + j] = y;
        };
        int t = 0;
        for (; t + 4 <= nt; t += 4) {
            float y0, y1, y2, y3;
            step(t, y0); step(t + 1, y1); step(t + 2, y2); step(t + 3, y3);
            finish(t, y0); finish(t + 1, y1); finish(t + 2, y2); finish(t + 3, y3);
        }
        for (; t < nt; t++) { float y; step(t, y); finish(t, y); }
        __syncthreads();                                  // everyone is done with this buffer before it is refilled
    }
#pragma unroll
    for (int ii = 0; ii < L; ii++) a.state_out[(size_t) h * S * S + (size_t) (p * L + ii) * S + j] = st[ii];
}

// per-head norm + ln_x + gate of raw scan outputs: CTA = 256 consecutive channels of one token
template <int S>
__global__ void __launch_bounds__(256) wkv5_norm_kernel(const __grid_constant__ WkvArgs a, const float * yraw /* written by the preceding kernel: NOT __restrict__ (invariant loads may be hoisted above griddepcontrol.wait) */) {
    __shared__ __align__(8) float sy[256];
    pdl_trigger();
    pdl_wait();
    const int tid = threadIdx.x, e = blockIdx.x * 256 + tid, D = a.D;
    const size_t row = blockIdx.y;
    const bool active = e < D;
    const float y = active ? yraw[row * D + e] : 0.0f;
    float o = head_norm<S>(y, sy, tid, active, a.gn_eps);
    if (active) {
        o = __fadd_rn(__fmul_rn(o, a.lnx_w[e]), a.lnx_b[e]);
        if (a.g) o = __fmul_rn(o, a.g[row * D + e]);
    }
    if (e - (tid & 31) < D) act_store_block(a.y, row, e, active, o);     // warps past the row own no q8 block
}

template <int S>
static cudaError_t launch_wkv5_t(const WkvArgs & a, cudaStream_t st) {
    if (S == 64 && wkv5_chunked_supported(a)) return launch_wkv5_chunked(a, st);
    if (!a.batch && a.R >= WKV_SEQ_MIN_ROWS && a.yraw != nullptr) {
        constexpr size_t smem = 2 * WKV_TCH * (3 * S + 8) * sizeof(float);
        constexpr int LC = S >= 32 ? WKV_SEQ_LC : (S >= 16 ? 2 : 1);   // 8*LC lanes per state column
        // > 48 KB of dynamic shared memory: opted in per device by wkv_init() (called from model_load)
        cudaError_t e = launch_pdl(wkv5_seq_kernel<S, LC>, dim3(a.H, S / 8), dim3(64 * LC), smem, st, a, a.yraw);
        if (e != cudaSuccess) return e;
        return launch_pdl(wkv5_norm_kernel<S>, dim3((a.D + 255) / 256, a.R), dim3(256), 0, st, a, (const float *) a.yraw);
    }
    const int hpb = S >= 32 ? 1 : 32 / S;
    const int threads = ((hpb * S + 31) / 32) * 32;
    dim3 grid((a.H + hpb - 1) / hpb, a.batch ? a.R : 1);
    return launch_pdl(wkv5_kernel<S>, grid, dim3(threads), 5 * threads * sizeof(float), st, a);
}

template <int S>
static cudaError_t wkv5_seq_attr() {
    constexpr size_t smem = 2 * WKV_TCH * (3 * S + 8) * sizeof(float);
    constexpr int LC = S >= 32 ? WKV_SEQ_LC : (S >= 16 ? 2 : 1);
    if (smem <= 48 * 1024) return cudaSuccess;
    return cudaFuncSetAttribute(wkv5_seq_kernel<S, LC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
}

// Function attributes are per device: called once per model load on the model's device (outside stream capture).
cudaError_t wkv_init() {
    cudaError_t e = wkv5_seq_attr<8>();
    if (e == cudaSuccess) e = wkv5_seq_attr<16>();
    if (e == cudaSuccess) e = wkv5_seq_attr<32>();
    if (e == cudaSuccess) e = wkv5_seq_attr<64>();
    return e;
}

cudaError_t launch_wkv5(const WkvArgs & a, cudaStream_t st) {
    switch (a.S) {
        case 8:  return launch_wkv5_t<8>(a, st);
        case 16: return launch_wkv5_t<16>(a, st);
        case 32: return launch_wkv5_t<32>(a, st);
        case 64: return launch_wkv5_t<64>(a, st);
        default: return cudaErrorInvalidValue;
    }
}

// ----------------------------------------------------------------------------------------
// v7.  Thread (hl, i) owns ROW i (value index) of the head state S[i][j]:
// state[h*S*S + i*S + j] (j = key index).
// ----------------------------------------------------------------------------------------
// CHUNK (sequence rows on the tensor path, head size 64): blockIdx.y is a chunk of W7C tokens; the state entering the chunk comes from
// the chunk scratch (written by wkv7_chunk_scan_kernel) and is not written back -- phase C of the chunk-parallel form below.
constexpr int W7C = 32;
template <int S, bool CHUNK = false>
__global__ void wkv7_kernel(const __grid_constant__ WkvArgs a) {
    extern __shared__ __align__(16) float sm[];
    pdl_trigger();
    const int nthr = blockDim.x;
    float * s_a = sm, * s_b = sm + nthr, * s_k = sm + 2 * nthr, * s_w = sm + 3 * nthr, * s_r = sm + 4 * nthr, * sy = sm + 5 * nthr;
    const int tid = threadIdx.x;
    const int hpb = nthr / S > 0 ? nthr / S : 1;
    const int h = blockIdx.x * hpb + tid / S, i = tid % S;
    const bool active = (tid < hpb * S) && h < a.H;
    const int e = h * S + i, D = a.D;
    const int hb = (tid / S) * S;
    const int stream = (a.batch && !CHUNK) ? blockIdx.y : 0;
    const size_t soff = (size_t) stream * a.state_stride;
    const int chunk = CHUNK ? (int) blockIdx.y : 0, nchunks = CHUNK ? (int) gridDim.y : 1;

    // The head states of the CTA are one contiguous run of hpb*S*S floats.  They move between global and
    // shared memory as consecutive 16-byte pieces (fully coalesced: with 256 streams this kernel is the
    // state traffic, 2 x 16 KB per head and step) and between shared memory and the owner's registers as
    // float4 rows; the row stride S + 4 keeps both access patterns free of bank conflicts.
    constexpr int RS = S + 4, Q = S / 4;
    float * s_st = sm + 6 * nthr;
    const int h0 = blockIdx.x * hpb, nh = min(hpb, a.H - h0);
    const float * gin = CHUNK ? a.chunk_ds + ((size_t) h0 * nchunks + chunk) * S * S : a.state_in + soff + (size_t) h0 * S * S;
    if (CHUNK) pdl_wait();                                // the chunk states come from the preceding kernel
    for (int f = tid; f < nh * S * Q; f += nthr)
        cp_async16(s_st + (f / Q) * RS + (f % Q) * 4, gin + (size_t) f * 4);
    cp_async_commit();
    float st[S];
    float kkc = 0, kac = 0, rkc = 0, lw = 0, lb = 0;
    if (active) { kkc = a.k_k[e]; kac = a.k_a[e]; rkc = a.r_k[e]; lw = a.lnx_w[e]; lb = a.lnx_b[e]; }
    cp_async_wait<0>();
    __syncthreads();
    if (active) {
        const float4 * p = reinterpret_cast<const float4 *>(s_st + (size_t) tid * RS);
#pragma unroll
        for (int j = 0; j < Q; j++) { const float4 f = p[j]; st[4 * j] = f.x; st[4 * j + 1] = f.y; st[4 * j + 2] = f.z; st[4 * j + 3] = f.w; }
    }
    if (!CHUNK) pdl_wait();
    const int s_begin = CHUNK ? chunk * W7C : 0;
    const int steps = CHUNK ? min(a.R, s_begin + W7C) : (a.batch ? 1 : a.R);
    for (int s = s_begin; s < steps; s++) {
        const size_t row = (a.batch && !CHUNK) ? stream : s;
        float k = 0, v = 0, r = 0, w = 0, av = 0, g = 0;
        if (active) {
            const size_t x = row * D + e;
            k = a.k[x]; v = a.v[x]; r = a.r[x]; w = a.w[x]; av = a.a[x]; g = a.g[x];
            if (a.layer == 0) a.v_first[x] = v;
            else v = __fadd_rn(v, __fmul_rn(__fsub_rn(a.v_first[x], v), a.vgate[x]));
        }
        // kk = l2norm_head(k * k_k)
        const float kkp = __fmul_rn(k, kkc);
        __syncthreads();
        sy[tid] = kkp;
        __syncthreads();
        float ss = 0.0f;
        if (active) {                                    // (the per-token vectors are read as float4: a quarter of the shared-memory instructions, same order of sums)
            const float4 * y4 = reinterpret_cast<const float4 *>(sy + hb);
#pragma unroll
            for (int j = 0; j < S / 4; j++) { const float4 f = y4[j]; ss = __fmaf_rn(f.x, f.x, ss); ss = __fmaf_rn(f.y, f.y, ss); ss = __fmaf_rn(f.z, f.z, ss); ss = __fmaf_rn(f.w, f.w, ss); }
        }
        const float kk = __fmul_rn(kkp, __fdiv_rn(1.0f, fmaxf(__fsqrt_rn(ss), 1e-12f)));
        const float ka = __fmul_rn(k, kac);
        k = __fadd_rn(k, __fsub_rn(__fmul_rn(av, ka), ka));
        __syncthreads();
        s_a[tid] = -kk; s_b[tid] = __fmul_rn(kk, av); s_k[tid] = k; s_w[tid] = w; s_r[tid] = r;
        sy[tid] = __fmul_rn(__fmul_rn(k, r), rkc);
        __syncthreads();
        float y = 0.0f, bonus = 0.0f;
        if (active) {
            float sa = 0.0f;
            const float4 * a4 = reinterpret_cast<const float4 *>(s_a + hb), * b4 = reinterpret_cast<const float4 *>(s_b + hb), * k4 = reinterpret_cast<const float4 *>(s_k + hb),
                         * w4 = reinterpret_cast<const float4 *>(s_w + hb), * r4 = reinterpret_cast<const float4 *>(s_r + hb), * y4 = reinterpret_cast<const float4 *>(sy + hb);
#pragma unroll
            for (int j = 0; j < S / 4; j++) {
                const float4 f = a4[j];
                sa = __fmaf_rn(f.x, st[4 * j], sa); sa = __fmaf_rn(f.y, st[4 * j + 1], sa); sa = __fmaf_rn(f.z, st[4 * j + 2], sa); sa = __fmaf_rn(f.w, st[4 * j + 3], sa);
            }
#pragma unroll
            for (int j = 0; j < S / 4; j++) {
                const float4 fk = k4[j], fb = b4[j], fw = w4[j], fr = r4[j];
                st[4 * j]     = __fmaf_rn(sa, fb.x, __fmaf_rn(st[4 * j],     fw.x, __fmul_rn(v, fk.x))); y = __fmaf_rn(st[4 * j],     fr.x, y);
                st[4 * j + 1] = __fmaf_rn(sa, fb.y, __fmaf_rn(st[4 * j + 1], fw.y, __fmul_rn(v, fk.y))); y = __fmaf_rn(st[4 * j + 1], fr.y, y);
                st[4 * j + 2] = __fmaf_rn(sa, fb.z, __fmaf_rn(st[4 * j + 2], fw.z, __fmul_rn(v, fk.z))); y = __fmaf_rn(st[4 * j + 2], fr.z, y);
                st[4 * j + 3] = __fmaf_rn(sa, fb.w, __fmaf_rn(st[4 * j + 3], fw.w, __fmul_rn(v, fk.w))); y = __fmaf_rn(st[4 * j + 3], fr.w, y);
            }
            double rk = 0.0;
#pragma unroll
            for (int j = 0; j < S / 4; j++) { const float4 f = y4[j]; rk += (double) f.x; rk += (double) f.y; rk += (double) f.z; rk += (double) f.w; }
            bonus = __fmul_rn(v, (float) rk);
        }
        float o = head_norm<S>(y, sy, tid, active, a.gn_eps);
        if (active) {
            o = __fadd_rn(__fmul_rn(o, lw), lb);
            o = __fadd_rn(o, bonus);
            o = __fmul_rn(o, g);
        }
        act_store_block(a.y, row, e, active, o);
    }
    if (CHUNK) return;                                    // the final state is written by the chunk scan
    __syncthreads();
    if (active) {
        float4 * p = reinterpret_cast<float4 *>(s_st + (size_t) tid * RS);
#pragma unroll
        for (int j = 0; j < Q; j++) p[j] = make_float4(st[4 * j], st[4 * j + 1], st[4 * j + 2], st[4 * j + 3]);
    }
    __syncthreads();
    float * gout = a.state_out + soff + (size_t) h0 * S * S;
    for (int f = tid; f < nh * S * Q; f += nthr)
        *reinterpret_cast<float4 *>(gout + (size_t) f * 4) = *reinterpret_cast<const float4 *>(s_st + (f / Q) * RS + (f % Q) * 4);
}

// ---- chunk-parallel WKV7 (head size 64; sequence rows on the tensor-core path) -----------------------------------------------
// The v7 recurrence is affine in the state: with G_t = diag(w_t) + a^_t b_t^T (a^ = -kk, b = kk * a; rwkv_operators_wkv_v7.inc:61-99)
//     S_t = S_{t-1} G_t + v_t k_t^T,        y_t = S_t r_t
// so over a chunk of W7C tokens   S_end = S_start P_c + dS_c   with   P_c = G_1 ... G_C (the identity pushed through the chunk) and
// dS_c = the chunk run from a zero state.  Both are the per-row update of wkv7_kernel (a state ROW evolves independently of the others).
//   A  wkv7_chunk_pd_kernel   grid (chunks, heads): P_c and dS_c of every chunk, all chunks in parallel
//   B  wkv7_chunk_scan_kernel grid (heads, 4): S_start[c+1] = S_start[c] P_c + dS_c, one 64 x 64 x 64 product per chunk (the only
//                             sequential part: T / 32 steps instead of T); S_start[c] replaces dS_c in the scratch; final state out
//   C  wkv7_kernel<64, true>  grid (heads, chunks): the exact recurrence from S_start[c] for the chunk's tokens, outputs fused as before
// FP32 sums in a different order than the token-by-token scan (S_start through products): tolerance mode, like the tensor-core
// mat-mul that feeds it; short sequences, FP32 files, batches and other head sizes keep the scan.
template <int S>
__global__ void __launch_bounds__(2 * S) wkv7_chunk_pd_kernel(const __grid_constant__ WkvArgs a, float * ds, float * pm, int nchunks) {
    __shared__ __align__(16) float s_a[S], s_b[S], s_k[S], s_w[S], sy[S];
    pdl_trigger();
    const int tid = threadIdx.x, i = tid % S, chunk = blockIdx.x, h = blockIdx.y;
    const bool is_p = tid >= S;                            // second half of the CTA pushes the identity through the chunk
    const int e = h * S + i, D = a.D;
    const float kkc = a.k_k[e], kac = a.k_a[e];
    float st[S];
#pragma unroll
    for (int j = 0; j < S; j++) st[j] = (is_p && j == i) ? 1.0f : 0.0f;
    pdl_wait();
    const int s_end = min(a.R, (chunk + 1) * W7C);
    for (int t = chunk * W7C; t < s_end; t++) {
        const size_t x = (size_t) t * D + e;
        float k = a.k[x], v = a.v[x];
        const float w = a.w[x], av = a.a[x];
        if (a.layer != 0) v = __fadd_rn(v, __fmul_rn(__fsub_rn(a.v_first[x], v), a.vgate[x]));
        const float kkp = __fmul_rn(k, kkc);
        __syncthreads();
        if (!is_p) sy[i] = kkp;
        __syncthreads();
        float ss = 0.0f;
        {
            const float4 * y4 = reinterpret_cast<const float4 *>(sy);
#pragma unroll
            for (int j = 0; j < S / 4; j++) { const float4 f = y4[j]; ss = __fmaf_rn(f.x, f.x, ss); ss = __fmaf_rn(f.y, f.y, ss); ss = __fmaf_rn(f.z, f.z, ss); ss = __fmaf_rn(f.w, f.w, ss); }
        }
        const float kk = __fmul_rn(kkp, __fdiv_rn(1.0f, fmaxf(__fsqrt_rn(ss), 1e-12f)));
        const float ka = __fmul_rn(k, kac);
        k = __fadd_rn(k, __fsub_rn(__fmul_rn(av, ka), ka));
        __syncthreads();
        if (!is_p) { s_a[i] = -kk; s_b[i] = __fmul_rn(kk, av); s_k[i] = k; s_w[i] = w; }
        __syncthreads();
        float sa = 0.0f;
        const float4 * a4 = reinterpret_cast<const float4 *>(s_a), * b4 = reinterpret_cast<const float4 *>(s_b), * k4 = reinterpret_cast<const float4 *>(s_k), * w4 = reinterpret_cast<const float4 *>(s_w);
#pragma unroll
        for (int j = 0; j < S / 4; j++) {
            const float4 f = a4[j];
            sa = __fmaf_rn(f.x, st[4 * j], sa); sa = __fmaf_rn(f.y, st[4 * j + 1], sa); sa = __fmaf_rn(f.z, st[4 * j + 2], sa); sa = __fmaf_rn(f.w, st[4 * j + 3], sa);
        }
        const float vv = is_p ? 0.0f : v;
#pragma unroll
        for (int j = 0; j < S / 4; j++) {
            const float4 fk = k4[j], fb = b4[j], fw = w4[j];
            st[4 * j]     = __fmaf_rn(sa, fb.x, __fmaf_rn(st[4 * j],     fw.x, __fmul_rn(vv, fk.x)));
            st[4 * j + 1] = __fmaf_rn(sa, fb.y, __fmaf_rn(st[4 * j + 1], fw.y, __fmul_rn(vv, fk.y)));
            st[4 * j + 2] = __fmaf_rn(sa, fb.z, __fmaf_rn(st[4 * j + 2], fw.z, __fmul_rn(vv, fk.z)));
            st[4 * j + 3] = __fmaf_rn(sa, fb.w, __fmaf_rn(st[4 * j + 3], fw.w, __fmul_rn(vv, fk.w)));
        }
    }
    float * out = (is_p ? pm : ds) + ((size_t) h * nchunks + chunk) * S * S + (size_t) i * S;
#pragma unroll
    for (int j = 0; j < S; j += 4) *reinterpret_cast<float4 *>(out + j) = make_float4(st[j], st[j + 1], st[j + 2], st[j + 3]);
}

// CTA = 16 rows of one head's state; thread (row il, 4 columns).  P_c tiles are double-buffered through shared memory (cp.async).
template <int S>
__global__ void __launch_bounds__(256) wkv7_chunk_scan_kernel(const __grid_constant__ WkvArgs a, float * ds, const float * pm, int nchunks) {
    static_assert(S == 64, "layout below is for 64 x 64 heads");
    __shared__ __align__(16) float s_P[2][S][S];
    __shared__ __align__(16) float s_S[16][S];
    pdl_trigger();
    const int tid = threadIdx.x, h = blockIdx.x, i0 = blockIdx.y * 16, il = tid / 16, j0 = (tid % 16) * 4;
    const float * gin = a.state_in + (size_t) h * S * S + (size_t) (i0 + il) * S + j0;
    *reinterpret_cast<float4 *>(&s_S[il][j0]) = *reinterpret_cast<const float4 *>(gin);      // input state: read-only, before the wait
    pdl_wait();
    auto fetch = [&](int c, int buf) {
        const float * src = pm + ((size_t) h * nchunks + c) * S * S;
        for (int f = tid; f < S * S / 4; f += 256) cp_async16(&s_P[buf][0][0] + f * 4, src + (size_t) f * 4);
        cp_async_commit();
    };
    fetch(0, 0);
    float * slot0 = ds + (size_t) h * nchunks * S * S + (size_t) (i0 + il) * S + j0;
    float4 d_next = *reinterpret_cast<const float4 *>(slot0);
    for (int c = 0; c < nchunks; c++) {
        if (c + 1 < nchunks) fetch(c + 1, (c + 1) & 1);
        float * slot = slot0 + (size_t) c * S * S;
        const float4 d = d_next;                                                            // dS_c (fetched one step ahead: off the critical path)
        if (c + 1 < nchunks) d_next = *reinterpret_cast<const float4 *>(slot + S * S);
        if (c + 1 < nchunks) cp_async_wait<1>(); else cp_async_wait<0>();
        __syncthreads();                                                                    // P_c landed, s_S holds S_start[c]
        const float4 cur = *reinterpret_cast<const float4 *>(&s_S[il][j0]);
        *reinterpret_cast<float4 *>(slot) = cur;                                            // S_start[c] takes dS_c's place
        float4 acc = d;
        const float (*P)[S] = s_P[c & 1];
#pragma unroll 16
        for (int m = 0; m < S; m++) {
            const float sv = s_S[il][m];
            const float4 pv = *reinterpret_cast<const float4 *>(&P[m][j0]);
            acc.x = __fmaf_rn(sv, pv.x, acc.x); acc.y = __fmaf_rn(sv, pv.y, acc.y); acc.z = __fmaf_rn(sv, pv.z, acc.z); acc.w = __fmaf_rn(sv, pv.w, acc.w);
        }
        __syncthreads();                                                                    // everyone has read s_S and this P buffer
        *reinterpret_cast<float4 *>(&s_S[il][j0]) = acc;
    }
    __syncthreads();
    *reinterpret_cast<float4 *>(a.state_out + (size_t) h * S * S + (size_t) (i0 + il) * S + j0) = *reinterpret_cast<const float4 *>(&s_S[il][j0]);
}

bool wkv7_chunked_supported(const WkvArgs & a) {
    static const bool off = [] { const char * e = getenv("RWKV_B200_NO_WKV_CHUNK"); return e && e[0] == '1'; }();
    return !off && !a.batch && a.S == 64 && a.D == a.H * 64 && a.chunk_ds != nullptr && a.chunk_pc != nullptr && a.R >= 2 * W7C;
}

static cudaError_t launch_wkv7_chunked(const WkvArgs & a, cudaStream_t st) {
    const int nchunks = (a.R + W7C - 1) / W7C;
    cudaError_t e = launch_pdl(wkv7_chunk_pd_kernel<64>, dim3(nchunks, a.H), dim3(128), 0, st, a, a.chunk_ds, a.chunk_pc, nchunks);
    if (e != cudaSuccess) return e;
    e = launch_pdl(wkv7_chunk_scan_kernel<64>, dim3(a.H, 4), dim3(256), 0, st, a, a.chunk_ds, (const float *) a.chunk_pc, nchunks);
    if (e != cudaSuccess) return e;
    constexpr int threads = 64;
    return launch_pdl(wkv7_kernel<64, true>, dim3(a.H, nchunks), dim3(threads), (6 * threads + 64 * (64 + 4)) * sizeof(float), st, a);
}

template <int S>
static cudaError_t launch_wkv7_t(const WkvArgs & a, cudaStream_t st) {
    const int hpb = S >= 32 ? 1 : 32 / S;
    const int threads = ((hpb * S + 31) / 32) * 32;
    dim3 grid((a.H + hpb - 1) / hpb, a.batch ? a.R : 1);
    return launch_pdl(wkv7_kernel<S>, grid, dim3(threads), (6 * threads + hpb * S * (S + 4)) * sizeof(float), st, a);
}

cudaError_t launch_wkv7(const WkvArgs & a, cudaStream_t st) {
    if (wkv7_chunked_supported(a)) return launch_wkv7_chunked(a, st);
    switch (a.S) {
        case 8:  return launch_wkv7_t<8>(a, st);
        case 16: return launch_wkv7_t<16>(a, st);
        case 32: return launch_wkv7_t<32>(a, st);
        case 64: return launch_wkv7_t<64>(a, st);
        default: return cudaErrorInvalidValue;
    }
}

// ---- tensor parallel: x += r * (Wv k) after the all-reduce of [partial Wv k | r shards] ----
__global__ void __launch_bounds__(256) tp_ffn_combine_kernel(float * x, const float * recv /* written by the preceding kernel: NOT __restrict__ (invariant loads may be hoisted above griddepcontrol.wait) */, int D) {
    TRACE_DECL; TRACE(0);
    pdl_trigger();
    TRACE(1);
    pdl_wait();
    TRACE(2);
    const int e = blockIdx.x * 256 + threadIdx.x;
    const size_t row = blockIdx.y;
    if (e < D) x[row * D + e] = __fadd_rn(x[row * D + e], __fmul_rn(recv[row * 2 * D + D + e], recv[row * 2 * D + e]));
    TRACE(5);
    TRACE_FLUSH(400, gridDim.x * gridDim.y);
}

cudaError_t launch_tp_ffn_combine(float * x, const float * recv, int R, int D, cudaStream_t st) {
    return launch_pdl(tp_ffn_combine_kernel, dim3((D + 255) / 256, R), dim3(256), 0, st, x, recv, D);
}

} // namespace rb
TRACE_EXPORT(rwkv_b200_debug_trace_wkv)
