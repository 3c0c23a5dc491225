// ggml block codecs on the device.
//
// File layout (ggml-common.h:161-208): block_q4_0 {f16 d; u8 qs[16]}, block_q4_1 {f16 d, m;
// u8 qs[16]}, block_q5_0 {f16 d; u8 qh[4]; u8 qs[16]}, block_q5_1 {f16 d, m; u8 qh[4];
// u8 qs[16]}, block_q8_0 {f16 d; i8 qs[32]} -- 18/20/22/24/34-byte records, i.e. no field
// is 16-byte aligned.  At upload they are split into planes (see DevMat in common.h) so
// that a warp reading 32 consecutive blocks issues one 512-byte fully coalesced request
// for the nibbles and one 128-/64-byte request each for qh and the scales.
//
// Q5 high bits: the file keeps bit e of qh for element e.  The mat-vec kernel handles the
// block as 8 words of 4 elements (word w = elements 4w..4w+3; words 0-3 are the low
// nibbles of qs, words 4-7 the high nibbles), so the plane stores qh bit-permuted:
//     plane bit (8*i + w)  =  file bit (4*w + i)        w = 0..7, i = 0..3
// and `(qh >> w) & 0x01010101` yields the four 5th bits of word w at byte LSBs.
#include "device_utils.cuh"

namespace rb {

__host__ __device__ __forceinline__ uint32_t permute_qh(uint32_t qh) {
    uint32_t out = 0;
#pragma unroll
    for (int w = 0; w < 8; w++)
#pragma unroll
        for (int i = 0; i < 4; i++) out |= ((qh >> (4 * w + i)) & 1u) << (8 * i + w);
    return out;
}

__global__ void repack_kernel(uint32_t type, const uint8_t * __restrict__ raw, size_t n_blocks,
                              uint8_t * __restrict__ qs, uint32_t * __restrict__ qh, __half * __restrict__ dm) {
    const size_t blk = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (blk >= n_blocks) return;
    const int bb = type == T_Q4_0 ? 18 : type == T_Q4_1 ? 20 : type == T_Q5_0 ? 22 : type == T_Q5_1 ? 24 : 34;
    const uint8_t * p = raw + blk * bb;
    const bool has_m = (type == T_Q4_1 || type == T_Q5_1);
    const bool has_h = (type == T_Q5_0 || type == T_Q5_1);
    uint16_t d16 = (uint16_t) p[0] | ((uint16_t) p[1] << 8);
    p += 2;
    if (has_m) {
        uint16_t m16 = (uint16_t) p[0] | ((uint16_t) p[1] << 8);
        p += 2;
        reinterpret_cast<uint16_t *>(dm)[2 * blk] = d16;
        reinterpret_cast<uint16_t *>(dm)[2 * blk + 1] = m16;
    } else {
        reinterpret_cast<uint16_t *>(dm)[blk] = d16;
    }
    if (has_h) {
        uint32_t h = (uint32_t) p[0] | ((uint32_t) p[1] << 8) | ((uint32_t) p[2] << 16) | ((uint32_t) p[3] << 24);
        p += 4;
        qh[blk] = permute_qh(h);
    }
    const int nq = type == T_Q8_0 ? 32 : 16;
    uint8_t * o = qs + blk * nq;
    for (int i = 0; i < nq; i++) o[i] = p[i];
}

cudaError_t launch_repack(uint32_t type, const uint8_t * raw, size_t n_blocks, void * qs, uint32_t * qh, void * dm, cudaStream_t st) {
    if (n_blocks == 0) return cudaSuccess;
    const int threads = 256;
    const unsigned grid = (unsigned) ((n_blocks + threads - 1) / threads);
    repack_kernel<<<grid, threads, 0, st>>>(type, raw, n_blocks, (uint8_t *) qs, qh, (__half *) dm);
    return cudaGetLastError();
}

// Integer value and scales of element j (0..31) of block `blk`, from the planes.
// y = q*d (+m) with one rounding (q*d is exact in FP32: <= 8 x 11 significant bits), which
// is what dequantize_row_* computes whether or not the host compiler contracts to an FMA.
__device__ __forceinline__ float plane_element(const DevMat & W, size_t blk, int j) {
    switch (W.type) {
        case T_Q8_0: {
            const int q = reinterpret_cast<const int8_t *>(W.qs)[blk * 32 + j];
            return __fmul_rn((float) q, __half2float(reinterpret_cast<const __half *>(W.dm)[blk]));
        }
        case T_Q4_0: case T_Q4_1: case T_Q5_0: case T_Q5_1: {
            const uint8_t byte = reinterpret_cast<const uint8_t *>(W.qs)[blk * 16 + (j & 15)];
            int q = (j < 16) ? (byte & 0x0F) : (byte >> 4);
            if (W.type == T_Q5_0 || W.type == T_Q5_1) {
                const int w = j >> 2, i = j & 3;
                q |= (int) ((W.qh[blk] >> (8 * i + w)) & 1u) << 4;
            }
            if (W.type == T_Q4_0) return __fmul_rn((float) (q - 8), __half2float(reinterpret_cast<const __half *>(W.dm)[blk]));
            if (W.type == T_Q5_0) return __fmul_rn((float) (q - 16), __half2float(reinterpret_cast<const __half *>(W.dm)[blk]));
            const __half2 dmv = reinterpret_cast<const __half2 *>(W.dm)[blk];
            return __fmaf_rn((float) q, __low2float(dmv), __high2float(dmv));
        }
        default: return 0.0f;
    }
}

__device__ __forceinline__ float mat_element(const DevMat & W, size_t row, int col) {
    if (W.type == T_F32) return reinterpret_cast<const float *>(W.qs)[row * W.K + col];
    if (W.type == T_F16) return __half2float(reinterpret_cast<const __half *>(W.qs)[row * W.K + col]);
    return plane_element(W, row * (size_t) (W.K >> 5) + (col >> 5), col & 31);
}

template <typename OutT>
__global__ void dequant_kernel(DevMat W, OutT * __restrict__ out) {
    const size_t n = (size_t) W.M * W.K;
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
        const float v = mat_element(W, i / W.K, (int) (i % W.K));
        if constexpr (sizeof(OutT) == 4) out[i] = v; else out[i] = __float2half_rn(v);
    }
}

cudaError_t launch_dequant(const DevMat & W, float * out, cudaStream_t st) {
    const size_t n = (size_t) W.M * W.K;
    if (n == 0) return cudaSuccess;
    const unsigned grid = (unsigned) std::min<size_t>((n + 255) / 256, 148 * 16);
    dequant_kernel<float><<<grid, 256, 0, st>>>(W, out);
    return cudaGetLastError();
}

cudaError_t launch_dequant_f16(const DevMat & W, void * out_half, cudaStream_t st) {
    const size_t n = (size_t) W.M * W.K;
    if (n == 0) return cudaSuccess;
    const unsigned grid = (unsigned) std::min<size_t>((n + 255) / 256, 148 * 16);
    dequant_kernel<__half><<<grid, 256, 0, st>>>(W, (__half *) out_half);
    return cudaGetLastError();
}

// ---- embedding row gather + ln0 (rwkv_graph.inc:655-658: ggml_get_rows then rwkv_layer_norm) ----
__global__ void embed_ln0_kernel(DevMat emb, const int * __restrict__ tokens, const float * __restrict__ w,
                                 const float * __restrict__ b, float * __restrict__ x) {
    extern __shared__ float smem[];
    __shared__ double red[32];
    __shared__ float  tmp_unused;
    (void) tmp_unused;
    const int t = blockIdx.x, D = emb.K;
    const int tok = tokens[t];
    float * raw = smem + D;          // second half: gathered row
    for (int i = threadIdx.x; i < D; i += blockDim.x) raw[i] = mat_element(emb, (size_t) tok, i);
    __syncthreads();
    block_layer_norm(raw, D, w, b, 1e-5f, smem, red);
    for (int i = threadIdx.x; i < D; i += blockDim.x) x[(size_t) t * D + i] = smem[i];
}

cudaError_t launch_embed_ln0(const DevMat & emb, const int * tokens, int T, const float * w, const float * b, float * x, cudaStream_t st) {
    embed_ln0_kernel<<<T, 256, 2 * emb.K * sizeof(float), st>>>(emb, tokens, w, b, x);
    return cudaGetLastError();
}

} // namespace rb
