// Layer norm + token shift + time-mix interpolation kernels.  Each produces the q8 / FP32
// activation rows (Act) the following mat-vec consumes, so no separate quantise launch.
//
// Reference arithmetic restated (paths relative to the reference tree):
//   rwkv_layer_norm            rwkv_operators.inc:93 over ggml norm (ggml-cpu.c:6880)
//   rwkv_carry_x (token shift) rwkv_graph.inc:56-82
//   v4/v5 mixes                rwkv_graph.inc:94-97,206-228   x*mu + (xp - xp*mu)
//   v6 mixes                   rwkv_graph.inc:301-336         x + (xp-x)*(maa + m)
//   v7 mixes                   rwkv_graph.inc:395-404         x + (xp-x)*mu
#include "device_utils.cuh"

namespace rb {

__global__ void __launch_bounds__(256) ln_mix_kernel(const __grid_constant__ LnMixArgs a) {
    extern __shared__ float smem[];
    __shared__ double red[32];
    const int r = blockIdx.x, D = a.D;
    float * s_x0 = smem, * s_xp = smem + D;
    block_layer_norm(a.x + (size_t) r * D, D, a.ln_w, a.ln_b, 1e-5f, s_x0, red);
    const size_t soff = a.batch ? (size_t) r * a.state_stride : 0;
    if (a.batch || r == 0) {
        for (int i = threadIdx.x; i < D; i += blockDim.x) s_xp[i] = a.state_in[soff + i];
        __syncthreads();
    } else {
        block_layer_norm(a.x + (size_t) (r - 1) * D, D, a.ln_w, a.ln_b, 1e-5f, s_xp, red);
    }
    if (a.batch || r == a.R - 1)
        for (int i = threadIdx.x; i < D; i += blockDim.x) a.state_out[soff + i] = s_x0[i];

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int e0 = warp * 32; e0 < D; e0 += nw * 32) {
        const int e = e0 + lane;
        const bool valid = e < D;
        const float x0 = valid ? s_x0[e] : 0.0f, xp = valid ? s_xp[e] : 0.0f;
        const float sx = __fsub_rn(xp, x0);
        if (valid && a.x0) a.x0[(size_t) r * D + e] = x0;
        if (valid && a.sx) a.sx[(size_t) r * D + e] = sx;
#pragma unroll
        for (int o = 0; o < MIX_MAX_OUT; o++) {
            if (o >= a.n_out) break;
            const float mu = valid ? a.mu[o][e] : 0.0f;
            float v;
            if (a.kind == MIX_LERP) v = __fadd_rn(__fmul_rn(x0, mu), __fsub_rn(xp, __fmul_rn(xp, mu)));
            else                    v = __fadd_rn(__fmul_rn(sx, mu), x0);
            act_store_block(a.out[o], (size_t) r, e, valid, v);
        }
    }
}

cudaError_t launch_ln_mix(const LnMixArgs & a, cudaStream_t st) {
    ln_mix_kernel<<<a.R, 256, 2 * a.D * sizeof(float), st>>>(a);
    return cudaGetLastError();
}

__global__ void __launch_bounds__(256) ln_act_kernel(const float * __restrict__ x, int ldx, int D, const float * __restrict__ w,
                                                     const float * __restrict__ b, Act out) {
    extern __shared__ float smem[];
    __shared__ double red[32];
    const int r = blockIdx.x;
    block_layer_norm(x + (size_t) r * ldx, D, w, b, 1e-5f, smem, red);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int e0 = warp * 32; e0 < D; e0 += nw * 32) {
        const int e = e0 + lane;
        act_store_block(out, (size_t) r, e, e < D, e < D ? smem[e] : 0.0f);
    }
}

cudaError_t launch_ln_act(const float * x, int ldx, int R, int D, const float * w, const float * b, Act out, cudaStream_t st) {
    ln_act_kernel<<<R, 256, D * sizeof(float), st>>>(x, ldx, D, w, b, out);
    return cudaGetLastError();
}

__global__ void __launch_bounds__(256) v6_mix2_kernel(const __grid_constant__ V6Mix2Args a) {
    const int r = blockIdx.x, D = a.D;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int e0 = warp * 32; e0 < D; e0 += nw * 32) {
        const int e = e0 + lane;
        const bool valid = e < D;
        const size_t i = (size_t) r * D + e;
        const float x0 = valid ? a.x0[i] : 0.0f, sx = valid ? a.sx[i] : 0.0f;
#pragma unroll
        for (int o = 0; o < 5; o++) {
            const float m = valid ? a.m[o][i] : 0.0f, maa = valid ? a.maa[o][e] : 0.0f;
            const float v = __fadd_rn(__fmul_rn(__fadd_rn(m, maa), sx), x0);
            act_store_block(a.out[o], (size_t) r, e, valid, v);
        }
    }
}

cudaError_t launch_v6_mix2(const V6Mix2Args & a, cudaStream_t st) {
    v6_mix2_kernel<<<a.R, 256, 0, st>>>(a);
    return cudaGetLastError();
}

__global__ void __launch_bounds__(256) quantize_act_kernel(Act a) {
    const int r = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int e0 = warp * 32; e0 < a.K; e0 += nw * 32) {
        const int e = e0 + lane;
        const bool valid = e < a.K;
        act_store_block(a, (size_t) r, e, valid, valid ? a.f[(size_t) r * a.K + e] : 0.0f);
    }
}

cudaError_t launch_quantize_act(Act a, int R, cudaStream_t st) {
    if (a.q == nullptr) return cudaSuccess;
    quantize_act_kernel<<<R, 256, 0, st>>>(a);
    return cudaGetLastError();
}

} // namespace rb
