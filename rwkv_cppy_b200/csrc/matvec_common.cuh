// Inner products and epilogues shared by the two mat-vec kernels (matvec.cu: register-staged,
// any shape / FP16 / FP32; matvec_q.cu: bulk-async staged, quantised).  See matvec.cu for the
// reference arithmetic these restate.
#pragma once
#include "device_utils.cuh"

namespace rb {

__device__ __forceinline__ int dp4a_(int a, int b, int c) { return __dp4a(a, b, c); }

// ---- per-type inner products -------------------------------------------------------------
template <int TYPE, int NCOL>
struct RowAcc {
    float acc[NCOL];
    float summ[NCOL];     // m_w * s_x sums (Q4_1 / Q5_1)
    __device__ __forceinline__ void clear() {
#pragma unroll
        for (int c = 0; c < NCOL; c++) { acc[c] = 0.0f; summ[c] = 0.0f; }
    }
};

template <int NCOL>
struct XBlock {             // the q8 activation block of NCOL columns held by one lane
    int   w[NCOL][8];
    float d[NCOL];
    float s[NCOL];          // q8_1: s ; q8_0 offset formats: (float) isum is not needed, keep int below
    int   isum[NCOL];
};

template <int TYPE, int NCOL>
__device__ __forceinline__ void load_xblock(XBlock<NCOL> & xb, const Act & x, int x_off, int col0, int T, int b) {
#pragma unroll
    for (int c = 0; c < NCOL; c++) {
        const int t = min(col0 + c, T - 1);
        const size_t e0 = (size_t) t * x.K + x_off + (size_t) b * 32;
        const uint4 lo = *reinterpret_cast<const uint4 *>(x.q + e0);
        const uint4 hi = *reinterpret_cast<const uint4 *>(x.q + e0 + 16);
        xb.w[c][0] = lo.x; xb.w[c][1] = lo.y; xb.w[c][2] = lo.z; xb.w[c][3] = lo.w;
        xb.w[c][4] = hi.x; xb.w[c][5] = hi.y; xb.w[c][6] = hi.z; xb.w[c][7] = hi.w;
        const size_t bi = (size_t) t * (x.K >> 5) + (x_off >> 5) + b;
        xb.d[c] = x.d[bi];
        if (TYPE == T_Q4_1 || TYPE == T_Q5_1) xb.s[c] = x.s[bi];
        if (TYPE == T_Q4_0 || TYPE == T_Q5_0) xb.isum[c] = x.isum[bi];
    }
}

// The 32 weights of one block as 8 words of 4 byte-sized integers each: word i (< 4) holds
// elements 4i..4i+3, word 4+i elements 16+4i..16+4i+3 (ggml block order: element j < 16 is the
// low nibble of byte j, element j+16 its high nibble).  Q4/Q5: unsigned 0..15 / 0..31 (the
// zero point is applied by the caller); Q8_0: the signed bytes as stored.
template <int TYPE>
__device__ __forceinline__ void unpack_block(int (&wv)[8], const uint4 q0, const uint4 q1, const uint32_t qh) {
    if (TYPE == T_Q8_0) {
        wv[0] = q0.x; wv[1] = q0.y; wv[2] = q0.z; wv[3] = q0.w; wv[4] = q1.x; wv[5] = q1.y; wv[6] = q1.z; wv[7] = q1.w;
    } else {
        const uint32_t v[4] = {q0.x, q0.y, q0.z, q0.w};
#pragma unroll
        for (int i = 0; i < 4; i++) {
            uint32_t lo = v[i] & 0x0F0F0F0Fu, hi = (v[i] >> 4) & 0x0F0F0F0Fu;
            if (TYPE == T_Q5_0 || TYPE == T_Q5_1) {
                lo |= (qh << (4 - i)) & 0x10101010u;            // plane bit (8*j + i)     -> bit 4 of byte j
                hi |= (qh >> i) & 0x10101010u;                  // plane bit (8*j + 4 + i) -> bit 4 of byte j
            }
            wv[i] = (int) lo; wv[4 + i] = (int) hi;
        }
    }
}

__device__ __forceinline__ int dp4a_us(uint32_t a, int b, int c) {          // unsigned bytes x signed bytes
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// one row x one block for all columns
template <int TYPE, int NCOL>
__device__ __forceinline__ void dot_block(RowAcc<TYPE, NCOL> & ra, const XBlock<NCOL> & xb,
                                          const uint4 q0, const uint4 q1, const uint32_t qh, const float dw, const float mw) {
    int wv[8];
    constexpr bool HI16 = (TYPE == T_Q4_0 || TYPE == T_Q4_1);
    // 4-bit formats: the high nibbles stay where they are (16 x their value, as unsigned bytes) and are summed in a chain of
    // their own; the chain total is an exact multiple of 16, so `>> 4` gives the same integer the shifted nibbles would --
    // four shifts fewer per block and two 4-deep dp4a chains instead of one 8-deep
    if (HI16) { wv[0] = q0.x; wv[1] = q0.y; wv[2] = q0.z; wv[3] = q0.w; }
    else unpack_block<TYPE>(wv, q0, q1, qh);
#pragma unroll
    for (int c = 0; c < NCOL; c++) {
        int dp = 0;
        if (HI16) {
            int hi = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                dp = dp4a_(wv[i] & 0x0F0F0F0F, xb.w[c][i], dp);
                hi = dp4a_us((uint32_t) wv[i] & 0xF0F0F0F0u, xb.w[c][4 + i], hi);
            }
            dp += hi >> 4;
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++) dp = dp4a_(wv[i], xb.w[c][i], dp);
        }
        if (TYPE == T_Q4_0) dp -= 8 * xb.isum[c];
        if (TYPE == T_Q5_0) dp -= 16 * xb.isum[c];
        ra.acc[c] = __fmaf_rn(__fmul_rn(dw, xb.d[c]), (float) dp, ra.acc[c]);
        if (TYPE == T_Q4_1 || TYPE == T_Q5_1) ra.summ[c] = __fmaf_rn(mw, xb.s[c], ra.summ[c]);
    }
}

// `EPI` known at compile time: the caller hoists the switch out of its inner loops.
template <int EPI, typename JobT>
__device__ __forceinline__ float apply_epilogue_c(const JobT & J, float y, int t, int m) {
    switch (EPI) {
        case EPI_STORE:        return y;
        case EPI_SIGMOID:      return sigmoidf_ref(y);
        case EPI_RELU_SQ:      { const float r = fmaxf(y, 0.0f); return __fmul_rn(r, r); }
        case EPI_SILU:         return siluf_ref(y);
        case EPI_TANH:         return tanhf(y);
        case EPI_ADD_TO:       return __fadd_rn(J.out[(size_t) t * J.ldo + m], y);
        case EPI_MULADD_TO:    return __fadd_rn(J.out[(size_t) t * J.ldo + m], __fmul_rn(J.aux[(size_t) t * J.ld_aux + m], y));
        case EPI_BIAS_SIGMOID: return sigmoidf_ref(__fadd_rn(y, J.vec[m]));
        case EPI_V6_DECAY:     return expf(-expf(__fadd_rn(y, J.vec[m])));
        case EPI_V7_DECAY:     return expf(__fmul_rn(sigmoidf_ref(__fadd_rn(y, J.vec[m])), -0.606531f));
        case EPI_ADD_AUX:      return __fadd_rn(J.aux[(size_t) t * J.ld_aux + m], y);
        default:               return y;
    }
}

// The epilogue's global operands, fetched ahead of the dot products (right after the dependency
// wait) so that the tail of a short mat-vec is not another L2 round trip.
struct EpiOperands { float o = 0.0f, a = 0.0f, v = 0.0f; };
template <typename JobT>
__device__ __forceinline__ EpiOperands load_epilogue_operands(const JobT & J, int t, int m) {
    EpiOperands p;
    const int e = J.epi;
    if (e == EPI_ADD_TO || e == EPI_MULADD_TO) p.o = J.out[(size_t) t * J.ldo + m];
    if (e == EPI_MULADD_TO || e == EPI_ADD_AUX) p.a = J.aux[(size_t) t * J.ld_aux + m];
    if (e == EPI_BIAS_SIGMOID || e == EPI_V6_DECAY || e == EPI_V7_DECAY) p.v = J.vec[m];
    return p;
}
template <typename JobT>
__device__ __forceinline__ float apply_epilogue_pre(const JobT & J, float y, const EpiOperands & p) {
    switch (J.epi) {
        case EPI_SIGMOID:      return sigmoidf_ref(y);
        case EPI_RELU_SQ:      { const float r = fmaxf(y, 0.0f); return __fmul_rn(r, r); }
        case EPI_SILU:         return siluf_ref(y);
        case EPI_TANH:         return tanhf(y);
        case EPI_ADD_TO:       return __fadd_rn(p.o, y);
        case EPI_MULADD_TO:    return __fadd_rn(p.o, __fmul_rn(p.a, y));
        case EPI_BIAS_SIGMOID: return sigmoidf_ref(__fadd_rn(y, p.v));
        case EPI_V6_DECAY:     return expf(-expf(__fadd_rn(y, p.v)));
        case EPI_V7_DECAY:     return expf(__fmul_rn(sigmoidf_ref(__fadd_rn(y, p.v)), -0.606531f));
        case EPI_ADD_AUX:      return __fadd_rn(p.a, y);
        default:               return y;
    }
}

template <typename JobT>
__device__ __forceinline__ float apply_epilogue(const JobT & J, float y, int t, int m) {
    switch (J.epi) {
        case EPI_STORE:        return apply_epilogue_c<EPI_STORE>(J, y, t, m);
        case EPI_SIGMOID:      return apply_epilogue_c<EPI_SIGMOID>(J, y, t, m);
        case EPI_RELU_SQ:      return apply_epilogue_c<EPI_RELU_SQ>(J, y, t, m);
        case EPI_SILU:         return apply_epilogue_c<EPI_SILU>(J, y, t, m);
        case EPI_TANH:         return apply_epilogue_c<EPI_TANH>(J, y, t, m);
        case EPI_ADD_TO:       return apply_epilogue_c<EPI_ADD_TO>(J, y, t, m);
        case EPI_MULADD_TO:    return apply_epilogue_c<EPI_MULADD_TO>(J, y, t, m);
        case EPI_BIAS_SIGMOID: return apply_epilogue_c<EPI_BIAS_SIGMOID>(J, y, t, m);
        case EPI_V6_DECAY:     return apply_epilogue_c<EPI_V6_DECAY>(J, y, t, m);
        case EPI_V7_DECAY:     return apply_epilogue_c<EPI_V7_DECAY>(J, y, t, m);
        case EPI_ADD_AUX:      return apply_epilogue_c<EPI_ADD_AUX>(J, y, t, m);
        default:               return y;
    }
}

} // namespace rb
