// Device helpers shared by the kernels: warp/block reductions, the q8 activation
// quantiser, block-wide layer norm.  Arithmetic that the reference performs as separate
// ggml ops is written with the non-contracting intrinsics (__fmul_rn/__fadd_rn) so that
// nvcc cannot fuse what the reference rounds twice.
#pragma once
#include <cuda_fp16.h>
#include <cstdint>
#include "kernels.h"
#include "trace.cuh"

namespace rb {

constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return v;
}
__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(FULL, v, o));
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return v;
}

// Butterfly over the warp totals held by lanes < nw (zero elsewhere).  With nw <= 8 the xor-16 and
// xor-8 steps would only add zeros, so they are skipped: same value, two dependent shuffles less.
__device__ __forceinline__ double warp_totals_sum_d(double v, int nw) {
    if (nw > 8) return warp_sum_d(v);
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return __shfl_sync(FULL, v, 0);
}

// Block-wide double sum; `red` is >= 32 doubles of shared memory.  All threads get the result.
__device__ __forceinline__ double block_sum_d(double v, double * red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum_d(v);
    __syncthreads();                 // protect `red` from the previous use
    if (lane == 0) red[warp] = v;
    __syncthreads();
    double t = 0.0;
    for (int i = 0; i < nw; i++) t += red[i];      // same order in every thread -> identical result
    return t;
}

__device__ __forceinline__ float round_f16(float v) { return __half2float(__float2half_rn(v)); }

// Layer-norm statistics of NR rows held by a CTA of `nt` threads, ITEMS elements per thread
// (element it*nt + tid, 0.0f where that is past the row).  Mean and variance are accumulated in
// double like ggml_compute_forward_norm_f32 (ggml-cpu.c:6880-6930).  This is THE summation
// order of every kernel that normalises a residual row (ln_mix_kernel for sequences and
// batches, the mat-vec prologue for decode): per-thread sums in `it` order, xor-butterfly in the
// warp, butterfly over the warp totals -- so for a given (D, nt) they agree bit for bit.
// On return x holds x - mean and scale[k] = 1/sqrt(var + eps).  red: [2*NR][32] doubles shared.
template <int ITEMS, int NR>
__device__ __forceinline__ void ln_rows_stats(float (&x)[NR][ITEMS], int D, int nt, double (*red)[32], float (&scale)[NR], float eps TRACE_ARG) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = nt >> 5;
    double s[NR];
#pragma unroll
    for (int k = 0; k < NR; k++) {
        s[k] = 0.0;
#pragma unroll
        for (int it = 0; it < ITEMS; it++) s[k] += (double) x[k][it];
    }
    TRACE(6);                                                // (the per-thread sums consumed the row: its loads have returned)
#pragma unroll
    for (int k = 0; k < NR; k++) { s[k] = warp_sum_d(s[k]); if (lane == 0) red[k][warp] = s[k]; }
    __syncthreads();
    TRACE(7);
    float mean[NR];
#pragma unroll
    for (int k = 0; k < NR; k++) { const double t = warp_totals_sum_d(lane < nw ? red[k][lane] : 0.0, nw); mean[k] = (float) (t / D); }
    double s2[NR];
#pragma unroll
    for (int k = 0; k < NR; k++) {
        s2[k] = 0.0;
#pragma unroll
        for (int it = 0; it < ITEMS; it++) {
            const bool valid = it * nt + (int) threadIdx.x < D;
            const float v = valid ? __fsub_rn(x[k][it], mean[k]) : 0.0f;
            x[k][it] = v; s2[k] += (double) __fmul_rn(v, v);
        }
    }
    TRACE(8);
#pragma unroll
    for (int k = 0; k < NR; k++) { s2[k] = warp_sum_d(s2[k]); if (lane == 0) red[NR + k][warp] = s2[k]; }
    __syncthreads();
    TRACE(9);
#pragma unroll
    for (int k = 0; k < NR; k++) {
        const double t = warp_totals_sum_d(lane < nw ? red[NR + k][lane] : 0.0, nw);
        scale[k] = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn((float) (t / D), eps)));
    }
}

// The q8 block of 32 consecutive values held one per lane, exactly as act_store_block computes
// it; returns q, and d / s / isum (identical in every lane).
__device__ __forceinline__ int q8_block(float v, bool valid, float & d16, float & s16, int & isum) {
    const float amax = warp_max(valid ? fabsf(v) : 0.0f);
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = amax != 0.0f ? __fdiv_rn(127.0f, amax) : 0.0f;
    const int   q  = valid ? __float2int_rn(__fmul_rn(v, id)) : 0;
    isum = warp_sum_i(q);
    d16 = round_f16(d);
    s16 = round_f16(__fmul_rn(d, (float) isum));
    return q;
}

// One warp holds one block of 32 consecutive activations (lane = element).  Writes the
// FP32 value and -- when the row has quantised planes -- the q8 block exactly as ggml's
// x86 quantize_row_q8_0/q8_1 produce it (ggml-cpu-quants.c:783-850, 1083-1150):
//   d = amax/127 (stored as fp16), id = 127/amax, q = rint(v*id) (ties to even),
//   s = fp16(d * sum q) with the unrounded d.
// An Act may carry any subset of its forms: f (FP32), h (FP16 rows for the tensor-core MATMUL),
// q/d/s/isum (q8 blocks for the exact mat-vec); the producer writes the ones that are present.
__device__ __forceinline__ void act_store_block(const Act & a, size_t row, int e, bool valid, float v) {
    if (valid && a.f) a.f[row * a.K + e] = v;
    if (valid && a.h) a.h[row * a.K + e] = __half_as_ushort(__float2half_rn(v));
    if (a.q == nullptr) return;
    const float amax = warp_max(valid ? fabsf(v) : 0.0f);
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = amax != 0.0f ? __fdiv_rn(127.0f, amax) : 0.0f;
    const int   q  = valid ? __float2int_rn(__fmul_rn(v, id)) : 0;
    const int   sum = warp_sum_i(q);
    if (valid) a.q[row * a.K + e] = (int8_t) q;
    if ((threadIdx.x & 31) == 0 && e < a.K) {          // a warp past the end of the row owns no block (its index would alias the next row's)
        const size_t b = row * (size_t) (a.K >> 5) + (e >> 5);
        a.d[b] = round_f16(d);
        a.s[b] = round_f16(__fmul_rn(d, (float) sum));
        a.isum[b] = sum;
    }
}

// ggml_compute_forward_norm_f32 (ggml-cpu.c:6880-6930): mean and variance accumulated in
// double, y = (x-mean) * 1/sqrt(var+eps); then rwkv_layer_norm (rwkv_operators.inc:93):
// y*w + b as two roundings.  `buf` (shared, D floats) receives the result.
__device__ __forceinline__ void block_layer_norm(const float * x, int D, const float * __restrict__ w,
                                                 const float * __restrict__ b, float eps, float * buf, double * red) {
    double s = 0.0;
    for (int i = threadIdx.x; i < D; i += blockDim.x) { const float v = x[i]; buf[i] = v; s += (double) v; }
    s = block_sum_d(s, red);
    const float mean = (float) (s / D);
    double s2 = 0.0;
    for (int i = threadIdx.x; i < D; i += blockDim.x) { const float v = __fsub_rn(buf[i], mean); buf[i] = v; s2 += (double) __fmul_rn(v, v); }
    s2 = block_sum_d(s2, red);
    const float var = (float) (s2 / D);
    const float scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(var, eps)));
    for (int i = threadIdx.x; i < D; i += blockDim.x) {
        const float y = __fmul_rn(buf[i], scale);
        buf[i] = __fadd_rn(__fmul_rn(y, w[i]), b[i]);
    }
    __syncthreads();
}

__device__ __forceinline__ float sigmoidf_ref(float x) { return __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-x))); }
__device__ __forceinline__ float siluf_ref(float x)    { return __fdiv_rn(x, __fadd_rn(1.0f, expf(-x))); }

} // namespace rb
