// ggml block codecs on the device.
//
// File layout (ggml-common.h:161-208): block_q4_0 {f16 d; u8 qs[16]}, block_q4_1 {f16 d, m;
// u8 qs[16]}, block_q5_0 {f16 d; u8 qh[4]; u8 qs[16]}, block_q5_1 {f16 d, m; u8 qh[4];
// u8 qs[16]}, block_q8_0 {f16 d; i8 qs[32]} -- 18/20/22/24/34-byte records, i.e. no field
// is 16-byte aligned.  At upload they are regrouped into warp tiles (see DevMat in common.h):
// 4 rows x 32 blocks with nibbles, 5th bits and scales as separate 16-byte aligned planes, so
// that a warp fetches its next 4 x 1024 weights with ONE contiguous bulk copy and then reads
// lane-consecutive 16-byte / 4-byte / 2-byte items from shared memory without bank conflicts.
//
// Q5 high bits: the file keeps bit e of qh for element e.  The mat-vec kernel handles the
// block as 8 words of 4 elements (word w = elements 4w..4w+3; words 0-3 are the low
// nibbles of qs, words 4-7 the high nibbles), so the plane stores qh bit-permuted:
//     plane bit (8*i + w)  =  file bit (4*w + i)        w = 0..7, i = 0..3
// and `(qh >> w) & 0x01010101` yields the four 5th bits of word w at byte LSBs.
#include "device_utils.cuh"
#include "launch.cuh"

namespace rb {

__host__ __device__ __forceinline__ uint32_t permute_qh(uint32_t qh) {
    uint32_t out = 0;
    for (int w = 0; w < 8; w++)
        for (int i = 0; i < 4; i++) out |= ((qh >> (4 * w + i)) & 1u) << (8 * i + w);
    return out;
}

// Byte offsets of block (row, b) inside the tiled storage.
struct TilePos { size_t qs, qh, dm; };
__device__ __forceinline__ TilePos tile_pos(uint32_t type, uint32_t rg_stride, int nb, size_t row, int b) {
    const uint32_t QB = tile_qb(type), HB = tile_hb(type), DB = tile_db(type);
    const size_t rg = row / TILE_ROWS; const int r = (int) (row % TILE_ROWS);
    const int c = b / TILE_BLOCKS, jb = b % TILE_BLOCKS;
    const int cnt = min(TILE_BLOCKS, nb - c * TILE_BLOCKS);
    const size_t base = rg * rg_stride + (size_t) TILE_ROWS * TILE_BLOCKS * c * (QB + HB + DB);
    TilePos p;
    p.qs = base + (size_t) (r * cnt + jb) * QB;
    p.qh = base + (size_t) TILE_ROWS * cnt * QB + (size_t) (r * cnt + jb) * HB;
    p.dm = base + (size_t) TILE_ROWS * cnt * (QB + HB) + (size_t) (r * cnt + jb) * DB;
    return p;
}

// one thread per (padded row, block)
__global__ void repack_kernel(uint32_t type, const uint8_t * __restrict__ raw, uint32_t M, uint32_t nb, uint32_t rg_stride,
                              uint8_t * __restrict__ out) {
    const size_t idx = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    const size_t rows_padded = ((size_t) M + TILE_ROWS - 1) / TILE_ROWS * TILE_ROWS;
    if (idx >= rows_padded * nb) return;
    const size_t row = idx / nb; const int b = (int) (idx % nb);
    const int bb = type == T_Q4_0 ? 18 : type == T_Q4_1 ? 20 : type == T_Q5_0 ? 22 : type == T_Q5_1 ? 24 : 34;
    const bool has_m = (type == T_Q4_1 || type == T_Q5_1), has_h = (type == T_Q5_0 || type == T_Q5_1);
    const TilePos p = tile_pos(type, rg_stride, (int) nb, row, b);
    const int nq = type == T_Q8_0 ? 32 : 16;
    if (row >= M) {                                   // padding rows: zero blocks (d = 0)
        for (int i = 0; i < nq; i++) out[p.qs + i] = 0;
        if (has_h) *reinterpret_cast<uint32_t *>(out + p.qh) = 0;
        *reinterpret_cast<uint16_t *>(out + p.dm) = 0;
        if (has_m) *reinterpret_cast<uint16_t *>(out + p.dm + 2) = 0;
        return;
    }
    const uint8_t * s = raw + (row * nb + b) * bb;
    *reinterpret_cast<uint16_t *>(out + p.dm) = (uint16_t) s[0] | ((uint16_t) s[1] << 8);
    s += 2;
    if (has_m) { *reinterpret_cast<uint16_t *>(out + p.dm + 2) = (uint16_t) s[0] | ((uint16_t) s[1] << 8); s += 2; }
    if (has_h) {
        const uint32_t h = (uint32_t) s[0] | ((uint32_t) s[1] << 8) | ((uint32_t) s[2] << 16) | ((uint32_t) s[3] << 24);
        *reinterpret_cast<uint32_t *>(out + p.qh) = permute_qh(h);
        s += 4;
    }
    for (int i = 0; i < nq; i++) out[p.qs + i] = s[i];
}

cudaError_t launch_repack(uint32_t type, const uint8_t * raw, size_t M, size_t K, void * out, cudaStream_t st) {
    const size_t nb = K / 32, rows_padded = (M + TILE_ROWS - 1) / TILE_ROWS * TILE_ROWS;
    const size_t n = rows_padded * nb;
    if (n == 0) return cudaSuccess;
    const int threads = 256;
    repack_kernel<<<(unsigned) ((n + threads - 1) / threads), threads, 0, st>>>(type, raw, (uint32_t) M, (uint32_t) nb,
                                                                              (uint32_t) tiled_rg_stride(type, K), (uint8_t *) out);
    return cudaGetLastError();
}

// ---- tensor-core copy (DevMat::tc): one thread per (padded row, padded block) -----------------
__global__ void repack_tc_kernel(uint32_t type, const uint8_t * __restrict__ raw, uint32_t M, uint32_t nb, uint32_t kblocks,
                                 uint8_t * __restrict__ out) {
    const size_t idx = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    const size_t rows_padded = ((size_t) M + TC_ROWS - 1) / TC_ROWS * TC_ROWS, nbp = (size_t) kblocks * 2;
    if (idx >= rows_padded * nbp) return;
    const size_t row = idx / nbp; const int b = (int) (idx % nbp);
    const int bb = type == T_Q4_0 ? 18 : type == T_Q4_1 ? 20 : type == T_Q5_0 ? 22 : type == T_Q5_1 ? 24 : 34;
    const bool has_m = (type == T_Q4_1 || type == T_Q5_1), has_h = (type == T_Q5_0 || type == T_Q5_1);
    const uint32_t QB = tile_qb(type), HB = tile_hb(type), DB = tile_db(type);
    uint8_t * chunk = out + ((row / TC_ROWS) * kblocks + (size_t) (b >> 1)) * (size_t) (2 * TC_ROWS * bb);
    const int e = (b & 1) * TC_ROWS + (int) (row % TC_ROWS);
    uint8_t * pq = chunk + (size_t) e * QB;
    uint8_t * ph = chunk + (size_t) 2 * TC_ROWS * QB + (size_t) e * HB;
    uint8_t * pd = chunk + (size_t) 2 * TC_ROWS * (QB + HB) + (size_t) e * DB;
    const int nq = (int) QB;
    if (row >= M || b >= (int) nb) {                  // padding: zero blocks (d = 0 [, m = 0])
        for (int i = 0; i < nq; i++) pq[i] = 0;
        if (has_h) *reinterpret_cast<uint32_t *>(ph) = 0;
        *reinterpret_cast<uint16_t *>(pd) = 0;
        if (has_m) *reinterpret_cast<uint16_t *>(pd + 2) = 0;
        return;
    }
    const uint8_t * s = raw + (row * nb + b) * bb;
    *reinterpret_cast<uint16_t *>(pd) = (uint16_t) s[0] | ((uint16_t) s[1] << 8);
    s += 2;
    if (has_m) { *reinterpret_cast<uint16_t *>(pd + 2) = (uint16_t) s[0] | ((uint16_t) s[1] << 8); s += 2; }
    if (has_h) {
        const uint32_t h = (uint32_t) s[0] | ((uint32_t) s[1] << 8) | ((uint32_t) s[2] << 16) | ((uint32_t) s[3] << 24);
        *reinterpret_cast<uint32_t *>(ph) = permute_qh(h);
        s += 4;
    }
    for (int i = 0; i < nq; i++) pq[i] = s[i];
}

cudaError_t launch_repack_tc(uint32_t type, const uint8_t * raw, size_t M, size_t K, void * out, cudaStream_t st) {
    const size_t nb = K / 32, kblocks = (K + TC_KB - 1) / TC_KB, rows_padded = (M + TC_ROWS - 1) / TC_ROWS * TC_ROWS;
    const size_t n = rows_padded * kblocks * 2;
    if (n == 0) return cudaSuccess;
    const int threads = 256;
    repack_tc_kernel<<<(unsigned) ((n + threads - 1) / threads), threads, 0, st>>>(type, raw, (uint32_t) M, (uint32_t) nb, (uint32_t) kblocks, (uint8_t *) out);
    return cudaGetLastError();
}

// Value of weight (row, col) from the tiled storage.
// y = q*d (+m) with one rounding (q*d is exact in FP32: <= 8 x 11 significant bits), which
// is what dequantize_row_* computes whether or not the host compiler contracts to an FMA.
__device__ __forceinline__ float tiled_element(const DevMat & W, size_t row, int col) {
    const uint8_t * base = reinterpret_cast<const uint8_t *>(W.qs);
    const int j = col & 31;
    const TilePos p = tile_pos(W.type, W.rg_stride, (int) (W.K >> 5), row, col >> 5);
    const float d = __half2float(*reinterpret_cast<const __half *>(base + p.dm));
    if (W.type == T_Q8_0) return __fmul_rn((float) reinterpret_cast<const int8_t *>(base + p.qs)[j], d);
    const uint8_t byte = base[p.qs + (j & 15)];
    int q = (j < 16) ? (byte & 0x0F) : (byte >> 4);
    if (W.type == T_Q5_0 || W.type == T_Q5_1) {
        const int w = j >> 2, i = j & 3;
        q |= (int) ((*reinterpret_cast<const uint32_t *>(base + p.qh) >> (8 * i + w)) & 1u) << 4;
    }
    if (W.type == T_Q4_0) return __fmul_rn((float) (q - 8), d);
    if (W.type == T_Q5_0) return __fmul_rn((float) (q - 16), d);
    return __fmaf_rn((float) q, d, __half2float(*reinterpret_cast<const __half *>(base + p.dm + 2)));
}

__device__ __forceinline__ float mat_element(const DevMat & W, size_t row, int col) {
    if (W.type == T_F32) return reinterpret_cast<const float *>(W.qs)[row * W.K + col];
    if (W.type == T_F16) return __half2float(reinterpret_cast<const __half *>(W.qs)[row * W.K + col]);
    return tiled_element(W, row, col);
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
    const unsigned grid = (unsigned) std::min<size_t>((n + 255) / 256, (size_t) device_sm_count() * 16);
    dequant_kernel<float><<<grid, 256, 0, st>>>(W, out);
    return cudaGetLastError();
}

cudaError_t launch_dequant_f16(const DevMat & W, void * out_half, cudaStream_t st) {
    const size_t n = (size_t) W.M * W.K;
    if (n == 0) return cudaSuccess;
    const unsigned grid = (unsigned) std::min<size_t>((n + 255) / 256, (size_t) device_sm_count() * 16);
    dequant_kernel<__half><<<grid, 256, 0, st>>>(W, (__half *) out_half);
    return cudaGetLastError();
}

// ---- embedding row gather + ln0 (rwkv_graph.inc:655-658: ggml_get_rows then rwkv_layer_norm) ----
__global__ void embed_ln0_kernel(DevMat emb, const int * __restrict__ tokens, const float * __restrict__ w,
                                 const float * __restrict__ b, float * __restrict__ x, unsigned int * step) {
    extern __shared__ float smem[];
    __shared__ double red[32];
    pdl_trigger();
    pdl_wait();
    if (step && blockIdx.x == 0 && threadIdx.x == 0) *step = *step + 1;
    const int t = blockIdx.x, D = emb.K;
    const int tok = tokens[t];
    float * raw = smem + D;          // second half: gathered row
    for (int i = threadIdx.x; i < D; i += blockDim.x) raw[i] = mat_element(emb, (size_t) tok, i);
    __syncthreads();
    block_layer_norm(raw, D, w, b, 1e-5f, smem, red);
    for (int i = threadIdx.x; i < D; i += blockDim.x) x[(size_t) t * D + i] = smem[i];
}

// Register-resident variant for rows up to 256 * ITEMS wide: every load is issued up front and the
// statistics use the shared row reduction (device_utils.cuh ln_rows_stats).
template <int ITEMS>
__global__ void __launch_bounds__(256) embed_ln0_reg_kernel(DevMat emb, const int * __restrict__ tokens, const float * __restrict__ w,
                                                            const float * __restrict__ b, float * __restrict__ x, unsigned int * step) {
    __shared__ double red[2][32];
    TRACE_DECL_IF(false);
    pdl_trigger();
    if (step && blockIdx.x == 0 && threadIdx.x == 0) *step = *step + 1;     // plain stream-ordered launch: nothing of the previous pass is running
    const int t = blockIdx.x, D = emb.K;
    const int tok = tokens[t];
    float v[1][ITEMS], lw[ITEMS], lb[ITEMS];
#pragma unroll
    for (int it = 0; it < ITEMS; it++) {
        const int e = it * 256 + threadIdx.x;
        const bool valid = e < D;
        v[0][it] = valid ? mat_element(emb, (size_t) tok, e) : 0.0f;
        lw[it] = valid ? w[e] : 0.0f; lb[it] = valid ? b[e] : 0.0f;
    }
    float scale[1];
    ln_rows_stats<ITEMS, 1>(v, D, 256, red, scale, 1e-5f TRACE_PASS);
#pragma unroll
    for (int it = 0; it < ITEMS; it++) {
        const int e = it * 256 + threadIdx.x;
        if (e < D) x[(size_t) t * D + e] = __fadd_rn(__fmul_rn(__fmul_rn(v[0][it], scale[0]), lw[it]), lb[it]);
    }
}

cudaError_t launch_embed_ln0(const DevMat & emb, const int * tokens, int T, const float * w, const float * b, float * x, cudaStream_t st, unsigned int * step) {
    // Plain stream-ordered launch on purpose: the first kernel of a pass is a full fence, so nothing of
    // this pass (which may prefetch the input state before its pdl_wait) overlaps the previous pass.
    if (emb.K <= 1024)      embed_ln0_reg_kernel<4><<<T, 256, 0, st>>>(emb, tokens, w, b, x, step);
    else if (emb.K <= 2048) embed_ln0_reg_kernel<8><<<T, 256, 0, st>>>(emb, tokens, w, b, x, step);
    else if (emb.K <= 4096) embed_ln0_reg_kernel<16><<<T, 256, 0, st>>>(emb, tokens, w, b, x, step);
    else if (emb.K <= 8192) embed_ln0_reg_kernel<32><<<T, 256, 0, st>>>(emb, tokens, w, b, x, step);
    else embed_ln0_kernel<<<T, 256, 2 * emb.K * sizeof(float), st>>>(emb, tokens, w, b, x, step);
    return cudaGetLastError();
}

} // namespace rb
