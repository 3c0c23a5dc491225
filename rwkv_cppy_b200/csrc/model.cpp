// Model loading: ggmf file -> HBM.  Replaces rwkv_model_loading.inc:128-419 (parameter
// binding per architecture, architecture detection :319-340, head geometry :403-409).
// Quantised 2-D tensors are uploaded raw and repacked on the device into the plane layout
// the mat-vec kernels read (codec.cu); 1-D / 3-D tensors become FP32 vectors.
#include "engine.h"

#include <cstring>
#include <cmath>
#include <algorithm>
#include <thread>

namespace rb {

Model::~Model() {
    int prev = -1;
    cudaGetDevice(&prev);                                 // leave the calling thread's current device as it was
    cudaSetDevice(device);
    tp_comm_destroy(tp);
    for (void * p : allocations) cudaFree(p);
    if (prev >= 0 && prev != device) cudaSetDevice(prev);
}

static float half_bits_to_float(uint16_t h) {
    const uint32_t sign = (uint32_t) (h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1F, man = h & 0x3FF, bits;
    if (exp == 0) {
        if (man == 0) bits = sign;
        else { int e = -1; do { man <<= 1; e++; } while (!(man & 0x400)); bits = sign | ((uint32_t) (127 - 15 - e) << 23) | ((man & 0x3FF) << 13); }
    } else if (exp == 31) bits = sign | 0x7F800000u | (man << 13);
    else bits = sign | ((exp + 112) << 23) | (man << 13);
    float f; memcpy(&f, &bits, 4); return f;
}

namespace {

// Host -> device staging for the weight upload (replaces the reference's malloc / fread / ggml_backend_tensor_set / free per tensor,
// rwkv_file_format.inc:300-313): the file is an mmap (page cache), a few threads copy slices of it into a ring of PINNED buffers, and
// each filled buffer goes to the GPU with an asynchronous DMA on the load stream while the threads fill the next one.
struct Stager {
    static constexpr size_t BUF = (size_t) 32 << 20;
    static constexpr int NB = 4, NT = 4;
    uint8_t * buf[NB] = {};
    cudaEvent_t ev[NB] = {};
    int next = 0;
    bool ok = true;
    bool init() {
        for (int i = 0; i < NB && ok; i++) ok = cudaMallocHost((void **) &buf[i], BUF) == cudaSuccess && cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming) == cudaSuccess;
        return ok;
    }
    ~Stager() { for (int i = 0; i < NB; i++) { if (buf[i]) cudaFreeHost(buf[i]); if (ev[i]) cudaEventDestroy(ev[i]); } }
    static void fill(uint8_t * dst, const uint8_t * src, size_t n) {
        if (n < ((size_t) 4 << 20)) { memcpy(dst, src, n); return; }
        std::thread th[NT];
        const size_t per = ((n + NT - 1) / NT + 4095) & ~(size_t) 4095;
        for (int t = 0; t < NT; t++) {
            const size_t o = std::min(n, per * t), l = std::min(per, n - o);
            th[t] = std::thread([=] { if (l) memcpy(dst + o, src + o, l); });
        }
        for (int t = 0; t < NT; t++) th[t].join();
    }
    // [src, src + n) of the mapping -> device memory `dst`, ordered on `st`
    bool copy(void * dst, const uint8_t * src, size_t n, cudaStream_t st) {
        for (size_t off = 0; off < n && ok; off += BUF) {
            const size_t len = std::min(BUF, n - off);
            const int b = next++ % NB;
            ok = cudaEventSynchronize(ev[b]) == cudaSuccess;          // the DMA that last used this buffer has finished
            if (!ok) break;
            fill(buf[b], src + off, len);
            ok = cudaMemcpyAsync((uint8_t *) dst + off, buf[b], len, cudaMemcpyHostToDevice, st) == cudaSuccess && cudaEventRecord(ev[b], st) == cudaSuccess;
        }
        return ok;
    }
};

struct Loader {
    const GgmfFile & f;
    Model & m;
    uint8_t * arena = nullptr; size_t arena_size = 0, arena_used = 0;
    uint8_t * scratch = nullptr; size_t scratch_size = 0;
    cudaStream_t st = nullptr;
    bool ok = true;
    Stager stager;

    Loader(const GgmfFile & f_, Model & m_) : f(f_), m(m_) {}

    void * bump(size_t bytes) {
        const size_t a = (arena_used + 255) & ~(size_t) 255;
        if (a + bytes > arena_size) { ok = false; global_fail(RWKV_ERROR_MODEL | RWKV_ERROR_ALLOC, __FILE__, __LINE__, "weight arena exhausted"); return nullptr; }
        arena_used = a + bytes;
        return arena + a;
    }

    const FileTensor * need(const std::string & name) {
        const FileTensor * t = f.find(name);
        if (!t) { ok = false; global_fail(RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_PARAM_MISSING, __FILE__, __LINE__, "Parameter %s not found", name.c_str()); }
        return t;
    }

    // FP32 device vector of exactly `n` elements (n == 0: any length)
    const float * vec(const std::string & name, size_t n) {
        const FileTensor * t = need(name);
        if (!t) return nullptr;
        if (n && t->n_elements() != n) {
            ok = false; global_fail(RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_SHAPE, __FILE__, __LINE__, "Parameter %s has %llu elements, expected %llu",
                                    name.c_str(), (unsigned long long) t->n_elements(), (unsigned long long) n);
            return nullptr;
        }
        if (t->type != T_F32 && t->type != T_F16) {
            ok = false; global_fail(RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_DATA_TYPE, __FILE__, __LINE__, "Parameter %s must be FP32 or FP16", name.c_str());
            return nullptr;
        }
        float * d = (float *) bump(t->n_elements() * 4);
        if (!d) return nullptr;
        if (t->type == T_F32) {
            if (cudaMemcpy(d, t->data, t->nbytes, cudaMemcpyHostToDevice) != cudaSuccess) ok = false;
        } else {
            std::vector<float> tmp(t->n_elements());
            const uint16_t * h = (const uint16_t *) t->data;
            for (size_t i = 0; i < tmp.size(); i++) { uint16_t b; memcpy(&b, h + i, 2); tmp[i] = half_bits_to_float(b); }
            if (cudaMemcpy(d, tmp.data(), tmp.size() * 4, cudaMemcpyHostToDevice) != cudaSuccess) ok = false;
        }
        m.weight_bytes += t->nbytes;
        return d;
    }

    // Matrix view: M rows of K weights starting `elem_off` elements into the tensor.
    bool mat_from(const FileTensor * t, size_t M, size_t K, size_t elem_off, DevMat & out) {
        const size_t blk = type_block(t->type), bb = type_block_bytes(t->type);
        if (K % blk != 0 || elem_off % blk != 0) {
            ok = false; global_fail(RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_SHAPE, __FILE__, __LINE__, "Parameter %s: row length %llu not a multiple of %llu",
                                    t->name.c_str(), (unsigned long long) K, (unsigned long long) blk);
            return false;
        }
        if ((elem_off + M * K) / blk * bb > t->nbytes) {      // the view must stay inside the tensor's bytes in the mapping
            ok = false; global_fail(RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_SHAPE, __FILE__, __LINE__, "Parameter %s: matrix view runs past the tensor data", t->name.c_str());
            return false;
        }
        return upload(t->type, t->data + elem_off / blk * bb, M, K, out);
    }

    // Column slice [k0, k0 + Kl) of every row (tensor parallel row-split matrices): gathered on the host.
    bool mat_cols(const FileTensor * t, size_t M, size_t K, size_t k0, size_t Kl, DevMat & out) {
        const size_t blk = type_block(t->type), bb = type_block_bytes(t->type);
        if (K % blk != 0 || k0 % blk != 0 || Kl % blk != 0 || k0 + Kl > K) {
            ok = false; global_fail(RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_SHAPE, __FILE__, __LINE__, "Parameter %s: cannot slice columns [%llu, +%llu)",
                                    t->name.c_str(), (unsigned long long) k0, (unsigned long long) Kl);
            return false;
        }
        const size_t row_b = K / blk * bb, off_b = k0 / blk * bb, len_b = Kl / blk * bb;
        std::vector<uint8_t> tmp(M * len_b);
        for (size_t r = 0; r < M; r++) memcpy(tmp.data() + r * len_b, t->data + r * row_b + off_b, len_b);
        return upload(t->type, tmp.data(), M, Kl, out);
    }

    bool upload(uint32_t type, const uint8_t * src, size_t M, size_t K, DevMat & out) {
        out.type = type; out.M = (uint32_t) M; out.K = (uint32_t) K;
        const size_t blk = type_block(type), bb = type_block_bytes(type);
        const size_t n_blocks = M * K / blk, nbytes = n_blocks * bb;
        out.bytes = nbytes;
        m.weight_bytes += nbytes;
        if (!type_quantized(type)) {
            void * d = bump(nbytes);
            if (!d) return false;
            if (!stager.copy(d, src, nbytes, st)) { ok = false; return false; }      // stream-ordered; model_load synchronises at the end
            out.qs = d;
            return true;
        }
        void * tiled = bump(tiled_bytes(type, M, K));
        if (!tiled) return false;
        if (nbytes > scratch_size) { ok = false; return false; }
        if (!stager.copy(scratch, src, nbytes, st)) { ok = false; return false; }
        if (launch_repack(type, scratch, M, K, tiled, st) != cudaSuccess) { ok = false; return false; }
        if (cudaStreamSynchronize(st) != cudaSuccess) { ok = false; return false; }
        out.qs = tiled; out.rg_stride = (uint32_t) tiled_rg_stride(type, K);
        if (matmul_tc_enabled() && K >= TC_KB) {                 // second copy in the tensor-core chunk layout (scratch still holds the raw rows)
            void * tc = bump(tc_bytes(type, M, K));
            if (!tc) return false;
            if (launch_repack_tc(type, scratch, M, K, tc, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) { ok = false; return false; }
            out.tc = tc; out.tc_kblocks = (uint32_t) ((K + TC_KB - 1) / TC_KB);
        }
        return true;
    }

    // Tensor-parallel views of a 2-D tensor (K, M): this rank's rows / this rank's columns.
    bool mat_row_shard(const std::string & name, size_t expect_K, size_t expect_M, DevMat & out) {
        const FileTensor * t = checked(name, expect_K, expect_M);
        if (!t) return false;
        const size_t K = t->ne[0], M = t->ne[1] * t->ne[2], n = (size_t) m.tp.world;
        const size_t Ml = (M + n - 1) / n, r0 = std::min(M, Ml * m.tp.rank), cnt = std::min(Ml, M - r0);
        return mat_from(t, cnt, K, r0 * K, out);
    }
    bool mat_col_shard(const std::string & name, size_t expect_K, size_t expect_M, DevMat & out) {
        const FileTensor * t = checked(name, expect_K, expect_M);
        if (!t) return false;
        const size_t K = t->ne[0], M = t->ne[1] * t->ne[2], n = (size_t) m.tp.world;
        if (K % (n * 32) != 0) { ok = false; global_fail(RWKV_ERROR_MODEL | RWKV_ERROR_UNSUPPORTED, __FILE__, __LINE__, "Parameter %s: %llu columns do not split over %d ranks in blocks of 32", name.c_str(), (unsigned long long) K, m.tp.world); return false; }
        return mat_cols(t, M, K, K / n * m.tp.rank, K / n, out);
    }
    const FileTensor * checked(const std::string & name, size_t expect_K, size_t expect_M) {
        const FileTensor * t = need(name);
        if (!t) return nullptr;
        const size_t K = t->ne[0], M = t->ne[1] * t->ne[2];
        if ((expect_K && K != expect_K) || (expect_M && M != expect_M)) {
            ok = false; global_fail(RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_SHAPE, __FILE__, __LINE__, "Parameter %s has shape (%llu, %llu), expected (%llu, %llu)",
                                    name.c_str(), (unsigned long long) K, (unsigned long long) M, (unsigned long long) expect_K, (unsigned long long) expect_M);
            return nullptr;
        }
        return t;
    }

    // 2-D ggml tensor (K, M); expect_K / expect_M == 0 means "any".
    bool mat(const std::string & name, size_t expect_K, size_t expect_M, DevMat & out) {
        const FileTensor * t = need(name);
        if (!t) return false;
        const size_t K = t->ne[0], M = t->ne[1] * t->ne[2];
        if ((expect_K && K != expect_K) || (expect_M && M != expect_M)) {
            ok = false; global_fail(RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_SHAPE, __FILE__, __LINE__, "Parameter %s has shape (%llu, %llu), expected (%llu, %llu)",
                                    name.c_str(), (unsigned long long) K, (unsigned long long) M, (unsigned long long) expect_K, (unsigned long long) expect_M);
            return false;
        }
        return mat_from(t, M, K, 0, out);
    }
};

} // namespace

Model * model_load(const char * path, const TpInfo & tp) {
    GgmfFile f;
    if (!ggmf_open(path, f)) { g_last_error |= RWKV_ERROR_FILE; return nullptr; }

    int dev = 0;
    RB_CUDA(cudaGetDevice(&dev), RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, nullptr);
    std::unique_ptr<Model> mp(new Model());
    Model & m = *mp;
    m.device = dev; m.tp = tp;
    m.n_vocab = f.n_vocab; m.n_embed = f.n_embed; m.n_layer = f.n_layer; m.data_type = f.data_type;
    RB_CHECK(m.n_layer > 0 && m.n_embed > 0 && m.n_vocab > 0, RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_SHAPE, nullptr, "Empty model");

    // architecture detection by key presence (rwkv_model_loading.inc:319-340)
    if (f.find("blocks.0.att.ln_x.weight")) { m.arch_major = 5; m.arch_minor = f.find("blocks.0.att.gate.weight") ? 2 : 1; }
    if (f.find("blocks.0.att.time_maa_x")) { m.arch_major = 6; m.arch_minor = 0; }
    if (f.find("blocks.0.att.r_k"))        { m.arch_major = 7; m.arch_minor = 0; }

    // head geometry (rwkv_model_loading.inc:403-409)
    if (m.arch_major == 7) {
        const FileTensor * t = f.find("blocks.0.att.r_k");
        m.head_count = (int) t->ne[1];
    } else if (m.arch_major >= 5) {
        const FileTensor * t = f.find("blocks.0.att.time_decay");
        RB_CHECK(t != nullptr, RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_PARAM_MISSING, nullptr, "Parameter blocks.0.att.time_decay not found");
        m.head_count = (int) t->ne[2];
    }
    if (m.arch_major >= 5) {
        RB_CHECK(m.head_count > 0 && m.n_embed % m.head_count == 0, RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_SHAPE, nullptr, "Bad head count");
        m.head_size = m.n_embed / m.head_count;
        RB_CHECK(m.head_size == 8 || m.head_size == 16 || m.head_size == 32 || m.head_size == 64,
                 RWKV_ERROR_MODEL | RWKV_ERROR_UNSUPPORTED, nullptr, "Head size %d is not supported (8, 16, 32, 64)", m.head_size);
    }

    const bool TP = tp.world > 1;
    if (TP) {
        RB_CHECK(m.arch_major == 4, RWKV_ERROR_MODEL | RWKV_ERROR_UNSUPPORTED, nullptr, "Tensor parallelism is implemented for RWKV v4 models");
        RB_CHECK(m.n_embed % (32 * tp.world) == 0, RWKV_ERROR_MODEL | RWKV_ERROR_UNSUPPORTED, nullptr, "n_embed %u does not split over %d ranks in blocks of 32", m.n_embed, tp.world);
    }
    Loader L(f, m);
    size_t arena = 0, scratch = 0;
    for (const FileTensor & t : f.tensors) {
        const size_t as_vec = t.n_elements() * 4;
        arena += std::max(t.nbytes, t.n_dims == 2 ? t.nbytes : as_vec) + 4 * 256;
        if (type_quantized(t.type)) arena += TILE_ROWS * (t.ne[0] / 32) * type_block_bytes(t.type) + 64;   // row padding of the tiled layout
        if (type_quantized(t.type) && t.n_dims == 2 && matmul_tc_enabled()) arena += tc_bytes(t.type, t.ne[1] * t.ne[2], t.ne[0]) + 256;
        if (type_quantized(t.type)) scratch = std::max(scratch, t.nbytes);
    }
    RB_CUDA(cudaMalloc((void **) &L.arena, arena), RWKV_ERROR_MODEL | RWKV_ERROR_ALLOC, nullptr);
    m.allocations.push_back(L.arena);
    L.arena_size = arena;
    if (scratch) {
        RB_CUDA(cudaMalloc((void **) &L.scratch, scratch), RWKV_ERROR_MODEL | RWKV_ERROR_ALLOC, nullptr);
        L.scratch_size = scratch;
    }
    RB_CUDA(cudaStreamCreate(&L.st), RWKV_ERROR_MODEL | RWKV_ERROR_ALLOC, nullptr);
    RB_CHECK(L.stager.init(), RWKV_ERROR_MODEL | RWKV_ERROR_ALLOC, nullptr, "Failed to allocate the pinned staging buffers");
    RB_CUDA(matvec_async_init(), RWKV_ERROR_MODEL | RWKV_ERROR_UNSUPPORTED, nullptr);
    RB_CUDA(wkv_init(), RWKV_ERROR_MODEL | RWKV_ERROR_UNSUPPORTED, nullptr);
    RB_CUDA(elementwise_init(), RWKV_ERROR_MODEL | RWKV_ERROR_UNSUPPORTED, nullptr);
    if (matmul_tc_enabled()) RB_CUDA(matmul_tc_init(), RWKV_ERROR_MODEL | RWKV_ERROR_UNSUPPORTED, nullptr);

    const size_t D = m.n_embed, V = m.n_vocab;
    L.mat("emb.weight", D, V, m.emb);
    m.ln0_w = L.vec("blocks.0.ln0.weight", D); m.ln0_b = L.vec("blocks.0.ln0.bias", D);
    m.layers.resize(m.n_layer);
    m.ffn_dim = (int) D;
    for (uint32_t i = 0; i < m.n_layer && L.ok; i++) {
        LayerW & l = m.layers[i];
        const std::string p = "blocks." + std::to_string(i) + ".";
        l.ln1_w = L.vec(p + "ln1.weight", D); l.ln1_b = L.vec(p + "ln1.bias", D);
        l.ln2_w = L.vec(p + "ln2.weight", D); l.ln2_b = L.vec(p + "ln2.bias", D);
        if (TP) {     // column-parallel k, v, r (output channels), row-parallel output (input channels)
            L.mat_row_shard(p + "att.key.weight", D, D, l.att_key);
            L.mat_row_shard(p + "att.value.weight", D, D, l.att_value);
            L.mat_row_shard(p + "att.receptance.weight", D, D, l.att_receptance);
            L.mat_col_shard(p + "att.output.weight", D, D, l.att_output);
        } else {
            L.mat(p + "att.key.weight", D, D, l.att_key);
            L.mat(p + "att.value.weight", D, D, l.att_value);
            L.mat(p + "att.receptance.weight", D, D, l.att_receptance);
            L.mat(p + "att.output.weight", D, D, l.att_output);
        }
        switch (m.arch_major) {
            case 4:
                l.att_mu[0] = L.vec(p + "att.time_mix_k", D); l.att_mu[1] = L.vec(p + "att.time_mix_v", D); l.att_mu[2] = L.vec(p + "att.time_mix_r", D);
                l.time_first = L.vec(p + "att.time_first", D); l.time_decay = L.vec(p + "att.time_decay", D);
                break;
            case 5: {
                l.att_mu[0] = L.vec(p + "att.time_mix_k", D); l.att_mu[1] = L.vec(p + "att.time_mix_v", D); l.att_mu[2] = L.vec(p + "att.time_mix_r", D);
                const size_t n = m.arch_minor >= 2 ? D : (size_t) m.head_count;
                l.time_first = L.vec(p + (m.arch_minor >= 2 ? "att.time_faaaa" : "att.time_first"), n);
                l.time_decay = L.vec(p + "att.time_decay", n);
                l.lnx_w = L.vec(p + "att.ln_x.weight", D); l.lnx_b = L.vec(p + "att.ln_x.bias", D);
                if (m.arch_minor >= 2) { l.att_mu[3] = L.vec(p + "att.time_mix_g", D); L.mat(p + "att.gate.weight", D, D, l.att_gate); }
                break;
            }
            case 6: {
                static const char * const names[6] = {"x", "w", "k", "v", "r", "g"};
                for (int n = 0; n < 6; n++) l.att_mu[n] = L.vec(p + "att.time_maa_" + names[n], D);
                L.mat(p + "att.time_maa_w1", D, 0, l.maa_w1);
                const FileTensor * w2 = L.need(p + "att.time_maa_w2");
                if (w2) {
                    const size_t E = w2->ne[0];
                    if (w2->ne[1] != D || w2->ne[2] != 5 || l.maa_w1.M != 5 * E) {
                        L.ok = false; global_fail(RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_SHAPE, __FILE__, __LINE__, "Bad time_maa_w1/w2 shapes");
                    } else {
                        for (int n = 0; n < 5; n++) L.mat_from(w2, D, E, (size_t) n * D * E, l.maa_w2[n]);
                        m.max_lora = std::max(m.max_lora, (int) (5 * E));
                    }
                }
                l.time_first = L.vec(p + "att.time_faaaa", D); l.time_decay = L.vec(p + "att.time_decay", D);
                L.mat(p + "att.time_decay_w1", D, 0, l.decay_w1);
                L.mat(p + "att.time_decay_w2", l.decay_w1.M, D, l.decay_w2);
                m.max_lora = std::max(m.max_lora, (int) l.decay_w1.M);
                L.mat(p + "att.gate.weight", D, D, l.att_gate);
                l.lnx_w = L.vec(p + "att.ln_x.weight", D); l.lnx_b = L.vec(p + "att.ln_x.bias", D);
                break;
            }
            case 7: {
                const float * mu = L.vec(p + "att.x_rwkvag", 6 * D);
                for (int n = 0; n < 6 && mu; n++) l.att_mu[n] = mu + n * D;
                l.w0 = L.vec(p + "att.w0", D); L.mat(p + "att.w1", D, 0, l.w1); L.mat(p + "att.w2", l.w1.M, D, l.w2);
                l.a0 = L.vec(p + "att.a0", D); L.mat(p + "att.a1", D, 0, l.a1); L.mat(p + "att.a2", l.a1.M, D, l.a2);
                L.mat(p + "att.g1", D, 0, l.g1); L.mat(p + "att.g2", l.g1.M, D, l.g2);
                if (i != 0) { l.v0 = L.vec(p + "att.v0", D); L.mat(p + "att.v1", D, 0, l.v1); L.mat(p + "att.v2", l.v1.M, D, l.v2); }
                l.r_k = L.vec(p + "att.r_k", D); l.k_k = L.vec(p + "att.k_k", D); l.k_a = L.vec(p + "att.k_a", D);
                l.lnx_w = L.vec(p + "att.ln_x.weight", D); l.lnx_b = L.vec(p + "att.ln_x.bias", D);
                m.max_lora = std::max({m.max_lora, (int) l.w1.M, (int) l.a1.M, (int) l.g1.M, (int) l.v1.M});
                break;
            }
        }
        if (m.arch_major == 7) {
            l.ffn_mu[0] = L.vec(p + "ffn.x_k", D);
        } else if (m.arch_major == 6) {
            l.ffn_mu[0] = L.vec(p + "ffn.time_maa_k", D); l.ffn_mu[1] = L.vec(p + "ffn.time_maa_r", D);
        } else {
            l.ffn_mu[0] = L.vec(p + "ffn.time_mix_k", D); l.ffn_mu[1] = L.vec(p + "ffn.time_mix_r", D);
        }
        if (TP) {
            L.mat_row_shard(p + "ffn.key.weight", D, 0, l.ffn_key);
            L.mat_col_shard(p + "ffn.value.weight", 0, D, l.ffn_value);
            L.mat_row_shard(p + "ffn.receptance.weight", D, D, l.ffn_receptance);
            if (L.ok && l.ffn_value.K != l.ffn_key.M) { L.ok = false; global_fail(RWKV_ERROR_MODEL | RWKV_ERROR_UNSUPPORTED, __FILE__, __LINE__, "ffn width does not split evenly over %d ranks", tp.world); }
        } else {
            L.mat(p + "ffn.key.weight", D, 0, l.ffn_key);
            L.mat(p + "ffn.value.weight", l.ffn_key.M, D, l.ffn_value);
            if (m.arch_major != 7) L.mat(p + "ffn.receptance.weight", D, D, l.ffn_receptance);
        }
        m.ffn_dim = std::max(m.ffn_dim, (int) l.ffn_key.M);
    }
    if (L.ok) {
        m.ln_out_w = L.vec("ln_out.weight", D); m.ln_out_b = L.vec("ln_out.bias", D);
        if (TP) L.mat_row_shard("head.weight", D, V, m.head); else L.mat("head.weight", D, V, m.head);
    }
    cudaStreamSynchronize(L.st);
    cudaStreamDestroy(L.st);
    if (L.scratch) cudaFree(L.scratch);
    if (!L.ok) { g_last_error |= RWKV_ERROR_MODEL_PARAMS; return nullptr; }
    if (cudaDeviceSynchronize() != cudaSuccess) RB_FAIL(RWKV_ERROR_MODEL | RWKV_ERROR_ALLOC, nullptr, "weight upload failed: %s", cudaGetErrorString(cudaGetLastError()));
    return mp.release();
}

} // namespace rb
