// rwkv_quantize_model_file: FP32/FP16 ggmf -> quantised ggmf, host only.
// Which tensors are converted follows rwkv_quantize.inc:1-13,137-140 (2-D FP32/FP16 tensors
// except emb / head and the v7 low-rank + r_k matrices); the block encoders restate
// quantize_row_q{4_0,4_1,5_0,5_1,8_0}_ref (ggml-quants.c:31-217), which is what
// ggml_quantize_chunk runs without an importance matrix, so output files are
// byte-identical to the reference's (tests/golden/quant_sha256.json).
// Compile this file with -ffp-contract=off: the encoders round every product and sum.
#include "common.h"

#include <cmath>
#include <cstring>
#include <memory>

namespace rb {

static inline float h2f(uint16_t h) {
    const uint32_t sign = (uint32_t) (h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1F, man = h & 0x3FF, bits;
    if (exp == 0) {
        if (man == 0) bits = sign;
        else { int e = -1; do { man <<= 1; e++; } while (!(man & 0x400)); bits = sign | ((uint32_t) (127 - 15 - e) << 23) | ((man & 0x3FF) << 13); }
    } else if (exp == 31) bits = sign | 0x7F800000u | (man << 13);
    else bits = sign | ((exp + 112) << 23) | (man << 13);
    float f; memcpy(&f, &bits, 4); return f;
}

// round-to-nearest-even, as F16C / ggml's software path
static inline uint16_t f2h(float f) {
    uint32_t x; memcpy(&x, &f, 4);
    const uint16_t sign = (uint16_t) ((x >> 16) & 0x8000u);
    x &= 0x7FFFFFFFu;
    if (x >= 0x7F800000u) return (uint16_t) (sign | 0x7C00u | (x > 0x7F800000u ? (0x200u | ((x >> 13) & 0x3FFu)) : 0u));
    if (x >= 0x477FF000u) return (uint16_t) (sign | 0x7C00u);             // >= 65520 rounds to inf
    if (x < 0x38800000u) {                                               // below the smallest normal half
        if (x < 0x33000000u) return sign;                                // < 2^-25
        const int e = (int) (x >> 23);
        const uint32_t m = (x & 0x7FFFFFu) | 0x800000u;
        const int shift = 126 - e;                                       // 14..24
        uint32_t r = m >> shift;
        const uint32_t rem = m & ((1u << shift) - 1u), half = 1u << (shift - 1);
        if (rem > half || (rem == half && (r & 1u))) r++;
        return (uint16_t) (sign | r);
    }
    uint32_t r = (((x >> 23) - 112u) << 10) | ((x & 0x7FFFFFu) >> 13);
    const uint32_t rem = x & 0x1FFFu;
    if (rem > 0x1000u || (rem == 0x1000u && (r & 1u))) r++;
    return (uint16_t) (sign | r);
}

static inline void put16(uint8_t * p, uint16_t v) { memcpy(p, &v, 2); }
static inline int imin(int a, int b) { return a < b ? a : b; }

static void encode_block(uint32_t type, const float * x, uint8_t * o) {
    switch (type) {
        case T_Q4_0: case T_Q5_0: {
            float amax = 0.0f, mx = 0.0f;
            for (int j = 0; j < 32; j++) { const float v = x[j]; if (amax < fabsf(v)) { amax = fabsf(v); mx = v; } }
            const bool q5 = type == T_Q5_0;
            const float d = mx / (q5 ? -16.0f : -8.0f);
            const float id = d ? 1.0f / d : 0.0f;
            put16(o, f2h(d));
            uint8_t * qs = o + (q5 ? 6 : 2);
            uint32_t qh = 0;
            for (int j = 0; j < 16; j++) {
                const float x0 = x[j] * id, x1 = x[16 + j] * id;
                if (q5) {
                    const uint8_t xi0 = (uint8_t) imin(31, (int8_t) (x0 + 16.5f)), xi1 = (uint8_t) imin(31, (int8_t) (x1 + 16.5f));
                    qs[j] = (uint8_t) ((xi0 & 0x0F) | ((xi1 & 0x0F) << 4));
                    qh |= ((xi0 & 0x10u) >> 4) << (j + 0);
                    qh |= ((xi1 & 0x10u) >> 4) << (j + 16);
                } else {
                    const uint8_t xi0 = (uint8_t) imin(15, (int8_t) (x0 + 8.5f)), xi1 = (uint8_t) imin(15, (int8_t) (x1 + 8.5f));
                    qs[j] = (uint8_t) (xi0 | (xi1 << 4));
                }
            }
            if (q5) memcpy(o + 2, &qh, 4);
            break;
        }
        case T_Q4_1: case T_Q5_1: {
            float mn = INFINITY, mxv = -INFINITY;
            // FLT_MAX start values in the reference; identical result for finite inputs
            mn = 3.402823466e+38f; mxv = -3.402823466e+38f;
            for (int j = 0; j < 32; j++) { const float v = x[j]; if (v < mn) mn = v; if (v > mxv) mxv = v; }
            const bool q5 = type == T_Q5_1;
            const float d = (mxv - mn) / (q5 ? 31.0f : 15.0f);
            const float id = d ? 1.0f / d : 0.0f;
            put16(o, f2h(d)); put16(o + 2, f2h(mn));
            uint8_t * qs = o + (q5 ? 8 : 4);
            uint32_t qh = 0;
            for (int j = 0; j < 16; j++) {
                const float x0 = (x[j] - mn) * id, x1 = (x[16 + j] - mn) * id;
                if (q5) {
                    const uint8_t xi0 = (uint8_t) (x0 + 0.5f), xi1 = (uint8_t) (x1 + 0.5f);
                    qs[j] = (uint8_t) ((xi0 & 0x0F) | ((xi1 & 0x0F) << 4));
                    qh |= ((xi0 & 0x10u) >> 4) << (j + 0);
                    qh |= ((xi1 & 0x10u) >> 4) << (j + 16);
                } else {
                    const uint8_t xi0 = (uint8_t) imin(15, (int8_t) (x0 + 0.5f)), xi1 = (uint8_t) imin(15, (int8_t) (x1 + 0.5f));
                    qs[j] = (uint8_t) (xi0 | (xi1 << 4));
                }
            }
            if (q5) memcpy(o + 4, &qh, 4);
            break;
        }
        case T_Q8_0: {
            float amax = 0.0f;
            for (int j = 0; j < 32; j++) { const float v = fabsf(x[j]); if (v > amax) amax = v; }
            const float d = amax / 127.0f;
            const float id = d ? 1.0f / d : 0.0f;
            put16(o, f2h(d));
            for (int j = 0; j < 32; j++) o[2 + j] = (uint8_t) (int8_t) roundf(x[j] * id);
            break;
        }
        default: break;
    }
}

static bool needs_quant(const std::string & name) {
    static const char * const skip[] = {"att.v1", "att.v2", "att.g1", "att.g2", "att.a1", "att.a2", "att.w1", "att.w2", "att.r_k"};
    if (name == "emb.weight" || name == "head.weight") return false;
    for (const char * s : skip) if (name.find(s) != std::string::npos) return false;
    return true;
}

} // namespace rb

using namespace rb;

extern "C" bool rwkv_quantize_model_file(const char * in_path, const char * out_path, const char * format_name) {
    g_last_error = RWKV_ERROR_NONE;
    RB_CHECK(in_path && out_path && format_name, RWKV_ERROR_ARGS, false, "NULL argument");
    const int out_type = type_from_name(format_name);
    RB_CHECK(out_type >= 0 && type_quantized((uint32_t) out_type), RWKV_ERROR_ARGS | RWKV_ERROR_DATA_TYPE, false, "Unsupported output data type (%s)", format_name);
    if (g_print_errors) fprintf(stderr, "Loading model from '%s'\n", in_path);
    GgmfFile f;
    if (!ggmf_open(in_path, f)) { g_last_error |= RWKV_ERROR_FILE; return false; }
    RB_CHECK(f.data_type == T_F32 || f.data_type == T_F16, RWKV_ERROR_FILE, false, "Unsupported input data type (%s); needs to be FP32 or FP16", type_name(f.data_type));
    FILE * out = fopen(out_path, "wb");
    RB_CHECK(out != nullptr, RWKV_ERROR_FILE | RWKV_ERROR_FILE_OPEN, false, "Failed to open %s for writing", out_path);
    struct Closer { FILE * f; ~Closer() { if (f) fclose(f); } } closer{out};
    const uint32_t hdr[6] = {RWKV_FILE_MAGIC, RWKV_FILE_VERSION, f.n_vocab, f.n_embed, f.n_layer, (uint32_t) out_type};
    RB_CHECK(fwrite(hdr, 4, 6, out) == 6, RWKV_ERROR_FILE | RWKV_ERROR_FILE_WRITE, false, "Failed to write file header");
    size_t orig_total = 0, new_total = 0;
    std::vector<float> f32;
    std::vector<uint8_t> q;
    const size_t obb = type_block_bytes((uint32_t) out_type);
    for (const FileTensor & t : f.tensors) {
        uint32_t type = t.type;
        const uint8_t * data = t.data;
        size_t nbytes = t.nbytes;
        if ((t.type == T_F32 || t.type == T_F16) && t.n_dims == 2 && needs_quant(t.name)) {
            const size_t n = t.n_elements();
            RB_CHECK(t.ne[0] % 32 == 0, RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_SHAPE, false, "Tensor %s: row length %llu is not a multiple of 32",
                     t.name.c_str(), (unsigned long long) t.ne[0]);
            f32.resize(n);
            if (t.type == T_F32) memcpy(f32.data(), t.data, n * 4);
            else { const uint8_t * p = t.data; for (size_t i = 0; i < n; i++) { uint16_t h; memcpy(&h, p + 2 * i, 2); f32[i] = h2f(h); } }
            q.resize(n / 32 * obb);
            for (size_t b = 0; b < n / 32; b++) encode_block((uint32_t) out_type, f32.data() + b * 32, q.data() + b * obb);
            type = (uint32_t) out_type; data = q.data(); nbytes = q.size();
        }
        const uint32_t th[3] = {t.n_dims, (uint32_t) t.name.size(), type};
        bool ok = fwrite(th, 4, 3, out) == 3;
        for (uint32_t d = 0; d < t.n_dims && ok; d++) { const uint32_t ne = (uint32_t) t.ne[d]; ok = fwrite(&ne, 4, 1, out) == 1; }
        ok = ok && fwrite(t.name.data(), 1, t.name.size(), out) == t.name.size();
        ok = ok && fwrite(data, 1, nbytes, out) == nbytes;
        RB_CHECK(ok, RWKV_ERROR_FILE | RWKV_ERROR_FILE_WRITE, false, "Failed to write tensor %s", t.name.c_str());
        orig_total += t.nbytes; new_total += nbytes;
    }
    if (g_print_errors) {
        fprintf(stderr, "original size     = %8.2f MB\n", orig_total / 1024.0 / 1024.0);
        fprintf(stderr, "quantized size    = %8.2f MB\n", new_total / 1024.0 / 1024.0);
        fprintf(stderr, "compression ratio = %8.2f\n", orig_total / (double) new_total);
    }
    return true;
}
