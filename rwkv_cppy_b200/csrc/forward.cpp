// The fixed per-layer launch sequence that replaces the reference's ggml graphs
// (rwkv_build_serial_graph rwkv_graph.inc:611-720, rwkv_build_sequential_graph :744-870,
// block builders :84-543).  One function serves single-token decode (R = 1), sequence mode
// (R = T tokens of one stream) and batched decode (R independent streams): the kernels take
// the row count and the "rows are streams" flag.
#include "engine.h"
#include "launch.cuh"
#include <cstdlib>
#include <initializer_list>
#include <unordered_map>

namespace rb {

Workspace::~Workspace() { for (void * p : allocations) cudaFree(p); }

namespace {
template <typename T>
bool dalloc(Workspace & ws, T *& p, size_t n) {
    void * q = nullptr;
    if (cudaMalloc(&q, std::max<size_t>(n, 1) * sizeof(T)) != cudaSuccess) return false;
    ws.allocations.push_back(q);
    p = (T *) q;
    return true;
}
bool act_alloc(Workspace & ws, Act & a, int rows, int K) {
    a.K = K;
    if (!dalloc(ws, a.f, (size_t) rows * K)) return false;
    if (matmul_tc_enabled() && rows >= matmul_tc_min_rows(true) && !dalloc(ws, a.h, (size_t) rows * K)) return false;
    if (K % 32 == 0) {
        const size_t nb = (size_t) rows * (K / 32);
        if (!dalloc(ws, a.q, (size_t) rows * K) || !dalloc(ws, a.d, nb) || !dalloc(ws, a.s, nb) || !dalloc(ws, a.isum, nb)) return false;
    }
    return true;
}
} // namespace

bool workspace_reserve(const Model & m, Workspace & ws, int rows, int logit_rows) {
    if (rows <= ws.rows && logit_rows <= ws.logit_rows) return true;
    rows = std::max(rows, ws.rows); logit_rows = std::max(logit_rows, ws.logit_rows);
    for (void * p : ws.allocations) cudaFree(p);
    ws.allocations.clear();
    ws.rows = 0; ws.logit_rows = 0;
    const size_t D = m.n_embed, n = (size_t) rows * D;
    bool ok = dalloc(ws, ws.x, n) && dalloc(ws, ws.x0, n) && dalloc(ws, ws.sx, n);
    for (int i = 0; i < 6 && ok; i++) ok = act_alloc(ws, ws.mix[i], rows, (int) D);
    ok = ok && act_alloc(ws, ws.y, rows, (int) D) && act_alloc(ws, ws.ffn_k, rows, m.ffn_dim);
    for (int i = 0; i < 4 && ok; i++) ok = act_alloc(ws, ws.lora[i], rows, std::max(m.max_lora, 32));
    ok = ok && act_alloc(ws, ws.head_in, logit_rows, (int) D);
    ok = ok && dalloc(ws, ws.r, n) && dalloc(ws, ws.k, n) && dalloc(ws, ws.v, n) && dalloc(ws, ws.g, n) && dalloc(ws, ws.w, n)
            && dalloc(ws, ws.a, n) && dalloc(ws, ws.vgate, n) && dalloc(ws, ws.ffn_r, n) && dalloc(ws, ws.v_first, n);
    if (m.arch_major == 6) ok = ok && dalloc(ws, ws.m5, 5 * n);
    if (m.arch_major >= 5 && rows >= WKV_SEQ_MIN_ROWS) ok = ok && dalloc(ws, ws.yraw, n);
    ws.matmul_scratch = MatmulScratch();
    if (matmul_tc_enabled() && rows >= matmul_tc_min_rows(true)) {
        ok = ok && dalloc(ws, ws.matmul_scratch.buf, MATMUL_SCRATCH_FLOATS) && dalloc(ws, ws.matmul_scratch.flags, MATMUL_SCRATCH_FLAGS);
        ok = ok && cudaMemset(ws.matmul_scratch.flags, 0, MATMUL_SCRATCH_FLAGS * sizeof(unsigned int)) == cudaSuccess;
        if (ok) ws.matmul_scratch.floats = MATMUL_SCRATCH_FLOATS;
    }
    if ((m.arch_major == 5 || m.arch_major == 6) && m.head_size == 64 && matmul_tc_enabled() && rows >= matmul_tc_min_rows()) {
        const size_t chunks = ((size_t) rows + WKV_CHUNK_TOKENS - 1) / WKV_CHUNK_TOKENS;
        ok = ok && dalloc(ws, ws.wkv_ds, chunks * D * 64) && dalloc(ws, ws.wkv_pc, chunks * D);
    }
    if (m.arch_major == 7 && m.head_size == 64 && matmul_tc_enabled() && rows >= matmul_tc_min_rows()) {      // chunk-parallel WKV7: dS_c / S_start[c] and P_c per head and chunk
        const size_t chunks = ((size_t) rows + WKV7_CHUNK_TOKENS - 1) / WKV7_CHUNK_TOKENS;
        ok = ok && dalloc(ws, ws.wkv_ds, chunks * D * 64) && dalloc(ws, ws.wkv_pc, chunks * D * 64);
    }
    if (m.tp.world > 1) {
        ok = ok && dalloc(ws, ws.tp_send, 2 * n) && dalloc(ws, ws.tp_recv, 2 * n) && dalloc(ws, ws.tp_att, n) && dalloc(ws, ws.x_alt, n);
        ok = ok && cudaMemset(ws.tp_send, 0, 2 * n * sizeof(float)) == cudaSuccess;     // the r-shard half stays zero outside this rank's slice
    }
    const size_t v_pad = m.tp.world > 1 ? (size_t) m.tp.world * ((m.n_vocab + m.tp.world - 1) / m.tp.world) : m.n_vocab;
    ok = ok && dalloc(ws, ws.logits, (size_t) logit_rows * v_pad) && dalloc(ws, ws.tokens, (size_t) rows);
    if (m.tp.world > 1) ok = ok && dalloc(ws, ws.logits_part, (size_t) logit_rows * (v_pad / m.tp.world));
    if (!ok) return false;
    ws.rows = rows; ws.logit_rows = logit_rows;
    return true;
}

namespace {

struct Pass {
    const Model & m; Workspace & ws; const ForwardArgs & fa; cudaStream_t st;
    cudaError_t err = cudaSuccess;
    int R; size_t D;
    size_t sstride;

    TpFused tpf;                  // fused tensor-parallel collectives: single-row decode over connected peer memory
    bool tp_fused = false;
    float * xres = nullptr;       // the residual row(s); under fused tensor parallelism it alternates between ws.x and ws.x_alt

    Pass(const Model & m_, Workspace & ws_, const ForwardArgs & fa_, cudaStream_t st_)
        : m(m_), ws(ws_), fa(fa_), st(st_), R(fa_.R), D(m_.n_embed), sstride(m_.state_len()), xres(ws_.x) {
        // opt-in: measured slower than the one-shot collective kernels on the 14B shape at 2 / 4 / 8 GPUs (profiles/r02/tp_scaling_midround.md)
        const char * e = getenv("RWKV_B200_TP_FUSE");
        const bool on = e && e[0] == '1';
        if (m.tp.world > 1 && R == 1 && on && D <= TP_LL_MAX_D) { tpf = tp_fused_desc(m.tp); tp_fused = tpf.ll[0] != nullptr; }
    }

    // RWKV_B200_SYNC_DEBUG=1 (never under graph capture: use sequence calls): synchronise after every launch and name the first that fails
    int n_launch = 0;
    void chk(cudaError_t e) {
        static const bool dbg = [] { const char * v = getenv("RWKV_B200_SYNC_DEBUG"); return v && v[0] == '1'; }();
        n_launch++;
        if (dbg && err == cudaSuccess && e == cudaSuccess) {
            cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
            cudaStreamIsCapturing(st, &cs);
            if (cs == cudaStreamCaptureStatusNone) {
                e = cudaStreamSynchronize(st);
                if (e != cudaSuccess) fprintf(stderr, "[rwkv_b200] launch group %d (site %d, R = %d) failed: %s\n", n_launch, pdl_site(), R, cudaGetErrorString(e));
            }
        }
        if (err == cudaSuccess && e != cudaSuccess) err = e;
    }

    static MatvecJob job(const DevMat & W, const Act & x, float * out, int ldo, int epi) {
        MatvecJob j; j.W = W; j.x = x; j.out = out; j.ldo = ldo; j.epi = epi; return j;
    }
    // Acts last written by a tensor-core MATMUL: their FP16 copy is current (value true) but their
    // q8 blocks are stale.  Every other producer writes FP32 + q8.
    std::unordered_map<const float *, bool> tc_written;

    // launches the jobs grouped by weight type, `rows` rows each; wide row counts go to the
    // tensor-core MATMUL (FP16 operands), everything else to the exact mat-vec kernels
    void matvec(std::vector<MatvecJob> jobs, int rows, const MatvecPrologue * pro = nullptr, int tp_site = -1) {
        std::vector<bool> done(jobs.size(), false);
        // ---- tensor-core path ----
        for (size_t i = 0; i < jobs.size(); i++) {
            if (done[i] || !tc_job(jobs[i], rows)) continue;
            MatmulJob g[MATMUL_MAX_JOBS]; int n = 0;
            for (size_t k = i; k < jobs.size() && n < MATMUL_MAX_JOBS; k++) {
                if (done[k] || jobs[k].W.type != jobs[i].W.type || !tc_job(jobs[k], rows)) continue;
                const MatvecJob & j = jobs[k];
                auto it = tc_written.find(j.x.f);
                if (it == tc_written.end() || !it->second) {            // FP16 copy of the input rows
                    chk(launch_to_half(j.x.f, j.x.h, (size_t) rows * j.x.K, st));
                    tc_written[j.x.f] = true;                           // (q8 state of x is irrelevant from here on within this call)
                    to_half_marks.push_back(j.x.f);
                }
                MatmulJob & o = g[n++];
                o.W = j.W; o.x16 = j.x.h + j.x_off; o.ldx = j.x.K;
                o.out = j.out; o.ldo = j.ldo; o.epi = j.epi; o.aux = j.aux; o.ld_aux = j.ld_aux; o.vec = j.vec;
                if (j.out_act.f) {
                    if (!o.out) { o.out = j.out_act.f; o.ldo = j.out_act.K; }
                    o.out_h = j.out_act.h; o.ld_h = j.out_act.K;
                }
                done[k] = true;
            }
            chk(launch_matmul_tc(g, n, rows, st, &ws.matmul_scratch));
            // inputs converted for this launch keep a valid q8 form: only the marks of MATMUL outputs persist
            for (const float * p : to_half_marks) tc_written.erase(p);
            to_half_marks.clear();
            for (size_t k = i; k < jobs.size(); k++)
                if (done[k] && jobs[k].out_act.f && tc_job(jobs[k], rows) && jobs[k].W.type == jobs[i].W.type) tc_written[jobs[k].out_act.f] = jobs[k].out_act.h != nullptr;
        }
        // ---- exact path ----
        for (size_t i = 0; i < jobs.size(); i++) {
            if (done[i]) continue;
            MatvecBatch b; b.T = rows;
            if (pro) b.pro = *pro;
            for (const DevMat * W : pf_next) {                 // weights of the next dependent launch: prefetched into the L2 by this one
                if (b.n_pf >= 3 || !type_quantized(jobs[i].W.type)) break;
                const size_t bytes = type_quantized(W->type) ? tiled_bytes(W->type, W->M, W->K) : (size_t) W->M * W->K * type_block_bytes(W->type);
                b.pf_ptr[b.n_pf] = W->qs; b.pf_bytes[b.n_pf] = (unsigned) std::min<size_t>(bytes & ~(size_t) 15, MATVEC_PF_MAX_BYTES); b.n_pf++;
            }
            pf_next.clear();
            if (tp_site >= 0) { b.tp = tpf; b.tp.site_push = tp_site; }
            // jobs of one launch share the K split, unless every split job of the launch can be split inside its CTAs (kernels.h
            // MatvecBatch::ks_warps): then each job keeps the split of its own shape
            auto flexible = [&](const MatvecJob & j) { return !pro && (matvec_ksplit(j.W.M, j.W.K) == 1 || matvec_split_in_cta(j)); };
            const bool quant = type_quantized(jobs[i].W.type);
            const int ks_i = quant ? matvec_ksplit(jobs[i].W.M, jobs[i].W.K) : 1;
            bool grp_flex = quant && flexible(jobs[i]), grp_uniform = true;
            for (size_t k = i; k < jobs.size() && b.n_jobs < MATVEC_MAX_JOBS; k++) {
                if (done[k] || jobs[k].W.type != jobs[i].W.type) continue;
                if (quant) {
                    const int ks_k = matvec_ksplit(jobs[k].W.M, jobs[k].W.K);
                    if (grp_flex && flexible(jobs[k])) { if (ks_k != ks_i) grp_uniform = false; }
                    else if (grp_uniform && ks_k == ks_i) grp_flex = grp_flex && flexible(jobs[k]);
                    else continue;
                }
                {
                    const MatvecJob & j = jobs[k];
                    if (type_quantized(j.W.type) && tc_written.count(j.x.f)) {   // produced by a MATMUL: q8 blocks are stale
                        Act a = j.x; a.h = nullptr; chk(launch_quantize_act(a, rows, st)); tc_written.erase(j.x.f);
                    }
                    if (j.out_act.f) tc_written.erase(j.out_act.f);
                    b.job[b.n_jobs] = j; b.job[b.n_jobs].out_act.h = nullptr; b.n_jobs++; done[k] = true;
                }
            }
            chk(launch_matvec(b, st));
        }
    }
    std::vector<const float *> to_half_marks;
    // set right before a matvec() call: the matrices of the launch that depends on it (single-row decode only; see MatvecBatch::pf_ptr)
    std::vector<const DevMat *> pf_next;
    void prefetch_next(std::initializer_list<const DevMat *> mats) {
        // opt-in (RWKV_B200_L2_PREFETCH=1): on the 14B shape it measured SLOWER (2.93 ms against 2.78 ms per token): the prefetch of up to
        // 48 MB competes with the 50 - 80 MB stream of the running launch for the 126 MB L2 and for HBM, and costs more than the hand-off it hides
        static const bool on = [] { const char * e = getenv("RWKV_B200_L2_PREFETCH"); return e && e[0] == '1'; }();
        pf_next.clear();
        if (R == 1 && on) pf_next.assign(mats.begin(), mats.end());
    }
    bool tc_job(const MatvecJob & j, int rows) const {
        return j.x.h != nullptr && (j.x_off % 8) == 0 && (j.x.K % 8) == 0 && matmul_tc_supported(j.W, rows, fa.batch != 0);
    }
    // The form a producer should emit `a` in: when every consumer is a tensor-core MATMUL the rows are
    // written once as FP16 (no FP32 rows, no q8 blocks, no conversion launch); otherwise FP32 + q8.
    Act produced(const Act & a, std::initializer_list<const DevMat *> consumers, int rows = -1) {
        if (rows < 0) rows = R;
        bool all_tc = a.h != nullptr && (a.K % 8) == 0 && consumers.size() > 0;
        for (const DevMat * W : consumers) all_tc = all_tc && matmul_tc_supported(*W, rows, fa.batch != 0);
        Act o = a;
        if (all_tc) { o.f = nullptr; o.q = nullptr; o.d = nullptr; o.s = nullptr; o.isum = nullptr; tc_written[a.f] = true; }
        else        { o.h = nullptr; tc_written.erase(a.f); }
        return o;
    }

    // Decode of narrow models is a chain of dependent launches: fold the layer norm + token-shift mix
    // (and the v4 recurrence) into the prologue of the mat-vec that consumes them.  Possible when
    // one row is evaluated and every consumer is a quantised matrix of D columns.
    static bool fusion_enabled() {
        static const bool on = [] { const char * e = getenv("RWKV_B200_NO_FUSE"); return !(e && e[0] == '1'); }();
        return on;
    }
    bool can_fuse(std::initializer_list<const DevMat *> mats) const {
        // Rows wider than LN_SMALL_D: every CTA of a fused launch would rebuild a row whose operands (x, norm vectors, shift state,
        // mixes: 5 x 4 D bytes from L2) rival its share of the weights -- measured on the 14B shape (D = 5120): 7.18 ms per token
        // fused against 3.44 ms with the stand-alone layer-norm launch.  Fusion pays where the launch chain dominates, not bytes.
        if (!fusion_enabled() || R != 1 || D > (size_t) LN_SMALL_D || D % 32 != 0 || m.tp.world > 1) return false;
        for (const DevMat * W : mats) if (!type_quantized(W->type) || W->K != D || matvec_ksplit(W->M, W->K) != 1) return false;   // (prologue launches do not split K)
        return true;
    }
    MatvecPrologue ln_prologue(uint32_t layer, bool att, int kind) const {
        const LayerW & l = m.layers[layer];
        MatvecPrologue P;
        P.kind = PRO_LNMIX; P.D = (int) D; P.mix_kind = kind;
        P.x = ws.x; P.ln_w = att ? l.ln1_w : l.ln2_w; P.ln_b = att ? l.ln1_b : l.ln2_b;
        const size_t slot = att ? D : 0;
        P.state_in = sin(layer, slot); P.state_out = sout(layer, slot);
        return P;
    }

    const float * sin(uint32_t layer, size_t off) const { return fa.state_in + (size_t) layer * m.layer_state_len() + off; }
    float * sout(uint32_t layer, size_t off) const { return fa.state_out + (size_t) layer * m.layer_state_len() + off; }

    TpWait tp_wait(int site, bool gated) const {
        TpWait w; w.site = site; w.gated = gated ? 1 : 0; w.n = (int) D; w.n_slice = (int) (D / m.tp.world);
        w.x_out = xres == ws.x ? ws.x_alt : ws.x;             // never in place: several CTAs gather the same base row
        return w;
    }
    void ln_mix(uint32_t layer, bool att, int kind, int n_out, const float * const * mu, bool want_sx, const Act * outs = nullptr, const TpWait * wait = nullptr) {
        const LayerW & l = m.layers[layer];
        LnMixArgs a;
        if (wait) { a.tp = tpf; a.wait = *wait; }
        a.x = xres;
        if (wait) xres = wait->x_out;                         // from here on the gathered row is the residual
        a.ln_w = att ? l.ln1_w : l.ln2_w; a.ln_b = att ? l.ln1_b : l.ln2_b;
        const size_t slot = att ? D : 0;                       // [ffn_xx, att_xx, ...] (rwkv_graph.inc:545-606)
        a.state_in = sin(layer, slot); a.state_out = sout(layer, slot);
        a.state_stride = sstride; a.batch = fa.batch; a.R = R; a.D = (int) D; a.kind = kind; a.n_out = n_out;
        for (int i = 0; i < n_out; i++) {
            a.mu[i] = mu[i];
            if (outs) a.out[i] = outs[i];
            else { a.out[i] = ws.mix[i]; a.out[i].h = nullptr; tc_written.erase(ws.mix[i].f); }
        }
        if (want_sx) { a.x0 = ws.x0; a.sx = ws.sx; }
        chk(launch_ln_mix(a, st));
    }

    WkvArgs wkv_base(uint32_t layer) {
        WkvArgs a;
        a.R = R; a.D = (int) D; a.H = m.head_count; a.S = m.head_size; a.batch = fa.batch; a.state_stride = sstride;
        a.state_in = sin(layer, 2 * D); a.state_out = sout(layer, 2 * D);
        a.r = ws.r; a.k = ws.k; a.v = ws.v; a.y = produced(ws.y, {&m.layers[layer].att_output}); a.layer = (int) layer; a.yraw = ws.yraw;
        // rows whose projections run on the tensor cores (FP16 operands) are in tolerance mode anyway:
        // their v5 / v6 recurrence takes the chunk-parallel form; exact-path rows keep the serial order
        if (!fa.batch && ws.wkv_ds && matmul_tc_supported(m.layers[layer].att_output, R)) { a.chunk_ds = ws.wkv_ds; a.chunk_pc = ws.wkv_pc; }
        return a;
    }

    // ---- tensor parallel (v4): column-parallel k / v / r and ffn key / receptance, row-parallel
    // att output and ffn value, one all-reduce per block -------------------------------
    void tp_all_reduce(const float * send, float * recv, size_t n) { chk(tp_allreduce(m.tp, send, recv, n, st, /* low_latency = */ R == 1)); }
    static bool tp_nccl_only() {
        static const bool v = [] { const char * e = getenv("RWKV_B200_TP_NCCL_ONLY"); return e && e[0] == '1'; }();
        return v;
    }

    // Fused form (single-row decode, peer memory connected): no collective launches.  The row-split mat-vecs push their partial
    // sums into every peer's low-latency inbox from their epilogue (site 2*layer: att, 2*layer + 1: ffn), and the layer norm that
    // follows gathers them in rank order on top of the residual row it already holds (kernels.h TpFused / TpWait).
    void att_v4_tp_fused(uint32_t layer) {
        const LayerW & l = m.layers[layer];
        const int Dl = (int) l.att_key.M;
        const TpWait w = tp_wait(2 * (int) layer - 1, true);
        ln_mix(layer, true, MIX_LERP, 3, l.att_mu, false, nullptr, layer > 0 ? &w : nullptr);
        pdl_site() = 1;
        matvec({job(l.att_key, ws.mix[0], ws.k, Dl, EPI_STORE), job(l.att_value, ws.mix[1], ws.v, Dl, EPI_STORE),
                job(l.att_receptance, ws.mix[2], ws.r, Dl, EPI_SIGMOID)}, R);
        WkvArgs a = wkv_base(layer);
        a.D = Dl; a.state_D = (int) D; a.chan0 = m.tp.rank * Dl;
        a.y = ws.y; a.y.K = Dl; a.y.h = nullptr; tc_written.erase(ws.y.f);
        a.u = l.time_first; a.decay = l.time_decay;
        pdl_site() = 2;
        chk(launch_wkv4(a, st));
        Act yl = ws.y; yl.K = Dl;
        pdl_site() = 4;
        matvec({job(l.att_output, yl, nullptr, 0, EPI_TP_PUSH)}, R, nullptr, 2 * (int) layer);
    }
    void ffn_v4_tp_fused(uint32_t layer) {
        const LayerW & l = m.layers[layer];
        const int Fl = (int) l.ffn_key.M;
        const TpWait w = tp_wait(2 * (int) layer, false);
        pdl_site() = 8;
        ln_mix(layer, false, MIX_LERP, 2, l.ffn_mu, false, nullptr, &w);
        MatvecJob k = job(l.ffn_key, ws.mix[0], nullptr, 0, EPI_RELU_SQ); k.out_act = ws.ffn_k; k.out_act.K = Fl;
        MatvecJob r = job(l.ffn_receptance, ws.mix[1], nullptr, 0, EPI_TP_PUSH_SIGMOID); r.tp_off = (int) D;     // this rank's slice of r rides in its slot behind the partial sums
        matvec({k, r}, R, nullptr, 2 * (int) layer + 1);
        Act kin = ws.ffn_k; kin.K = Fl;
        pdl_site() = 16;
        matvec({job(l.ffn_value, kin, nullptr, 0, EPI_TP_PUSH)}, R, nullptr, 2 * (int) layer + 1);
    }

    void att_v4_tp(uint32_t layer) {
        if (tp_fused) { att_v4_tp_fused(layer); return; }
        const LayerW & l = m.layers[layer];
        const int Dl = (int) l.att_key.M;
        ln_mix(layer, true, MIX_LERP, 3, l.att_mu, false);
        matvec({job(l.att_key, ws.mix[0], ws.k, Dl, EPI_STORE), job(l.att_value, ws.mix[1], ws.v, Dl, EPI_STORE),
              job(l.att_receptance, ws.mix[2], ws.r, Dl, EPI_SIGMOID)}, R);
        WkvArgs a = wkv_base(layer);
        a.D = Dl; a.state_D = (int) D; a.chan0 = m.tp.rank * Dl;
        a.y = ws.y; a.y.K = Dl; a.y.h = nullptr; tc_written.erase(ws.y.f);
        a.u = l.time_first; a.decay = l.time_decay;
        chk(launch_wkv4(a, st));
        Act yl = ws.y; yl.K = Dl;
        MatvecJob o = job(l.att_output, yl, ws.tp_att, (int) D, m.tp.rank == 0 ? EPI_ADD_AUX : EPI_STORE);   // rank 0 carries the residual
        o.aux = ws.x; o.ld_aux = (int) D;
        matvec({o}, R);
        tp_all_reduce(ws.tp_att, ws.x, (size_t) R * D);
    }

    void ffn_v4_tp(uint32_t layer) {
        if (tp_fused) { ffn_v4_tp_fused(layer); return; }
        const LayerW & l = m.layers[layer];
        const int Fl = (int) l.ffn_key.M, Dl = (int) l.ffn_receptance.M;
        ln_mix(layer, false, MIX_LERP, 2, l.ffn_mu, false);
        MatvecJob k = job(l.ffn_key, ws.mix[0], nullptr, 0, EPI_RELU_SQ); k.out_act = ws.ffn_k; k.out_act.K = Fl;
        // this rank's slice of r goes straight into the second half of the send buffer (zero elsewhere)
        MatvecJob r = job(l.ffn_receptance, ws.mix[1], ws.tp_send + D + (size_t) m.tp.rank * Dl, 2 * (int) D, EPI_SIGMOID);
        matvec({k, r}, R);
        Act kin = ws.ffn_k; kin.K = Fl;
        matvec({job(l.ffn_value, kin, ws.tp_send, 2 * (int) D, EPI_STORE)}, R);
        if (R == 1 && tp_ll_allreduce_usable(m.tp, 2 * D) && !tp_nccl_only()) {
            // one row: the low-latency all-reduce finishes the channel mix itself (x += sum(r shards) * sum(partial Wv k)): no combine launch
            chk(tp_ll_allreduce(m.tp, ws.tp_send, nullptr, 2 * D, ws.x, st));
        } else {
            tp_all_reduce(ws.tp_send, ws.tp_recv, (size_t) R * 2 * D);
            chk(launch_tp_ffn_combine(ws.x, ws.tp_recv, R, (int) D, st));
        }
    }

    void ffn(uint32_t layer) {
        const LayerW & l = m.layers[layer];
        pdl_site() = 8;
        if (m.tp.world > 1) { ffn_v4_tp(layer); return; }
        if (m.arch_major == 7) {
            MatvecJob k = job(l.ffn_key, ws.mix[0], nullptr, 0, EPI_RELU_SQ); k.out_act = ws.ffn_k; k.out_act.K = (int) l.ffn_key.M;
            if (can_fuse({&l.ffn_key})) {
                const MatvecPrologue P = ln_prologue(layer, false, MIX_SX);
                k.pro_mu = l.ffn_mu[0];
                matvec({k}, R, &P);
            } else {
                const Act outs[1] = {produced(ws.mix[0], {&l.ffn_key})};
                ln_mix(layer, false, MIX_SX, 1, l.ffn_mu, false, outs);
                matvec({k}, R);
            }
            Act kin = ws.ffn_k; kin.K = (int) l.ffn_key.M;
            pdl_site() = 16;
            matvec({job(l.ffn_value, kin, ws.x, (int) D, EPI_ADD_TO)}, R);
            return;
        }
        const int kind = m.arch_major == 6 ? MIX_SX : MIX_LERP;
        MatvecJob k = job(l.ffn_key, ws.mix[0], nullptr, 0, EPI_RELU_SQ); k.out_act = ws.ffn_k; k.out_act.K = (int) l.ffn_key.M;
        MatvecJob r = job(l.ffn_receptance, ws.mix[1], ws.ffn_r, (int) D, EPI_SIGMOID);
        prefetch_next({&l.ffn_value});
        if (can_fuse({&l.ffn_key, &l.ffn_receptance})) {
            const MatvecPrologue P = ln_prologue(layer, false, kind);
            k.pro_mu = l.ffn_mu[0]; r.pro_mu = l.ffn_mu[1];
            matvec({k, r}, R, &P);
        } else {
            const Act outs[2] = {produced(ws.mix[0], {&l.ffn_key}), produced(ws.mix[1], {&l.ffn_receptance})};
            ln_mix(layer, false, kind, 2, l.ffn_mu, false, outs);
            matvec({k, r}, R);
        }
        Act kin = ws.ffn_k; kin.K = (int) l.ffn_key.M;
        MatvecJob v = job(l.ffn_value, kin, ws.x, (int) D, EPI_MULADD_TO); v.aux = ws.ffn_r; v.ld_aux = (int) D;
        pdl_site() = 16;
        if (layer + 1 < m.n_layer) {
            const LayerW & nx = m.layers[layer + 1];
            if (m.arch_major == 6) prefetch_next({&nx.maa_w1}); else prefetch_next({&nx.att_key, &nx.att_value, &nx.att_receptance});
        } else if (fa.logits) prefetch_next({&m.head});
        matvec({v}, R);
    }

    void att_v4(uint32_t layer) {
        const LayerW & l = m.layers[layer];
        std::vector<MatvecJob> kvr = {job(l.att_key, ws.mix[0], ws.k, (int) D, EPI_STORE), job(l.att_value, ws.mix[1], ws.v, (int) D, EPI_STORE),
                                    job(l.att_receptance, ws.mix[2], ws.r, (int) D, EPI_SIGMOID)};
        prefetch_next({&l.att_output});
        if (can_fuse({&l.att_key, &l.att_value, &l.att_receptance})) {
            const MatvecPrologue P = ln_prologue(layer, true, MIX_LERP);
            for (int i = 0; i < 3; i++) kvr[i].pro_mu = l.att_mu[i];
            matvec(kvr, R, &P);
        } else {
            const Act outs[3] = {produced(ws.mix[0], {&l.att_key}), produced(ws.mix[1], {&l.att_value}), produced(ws.mix[2], {&l.att_receptance})};
            ln_mix(layer, true, MIX_LERP, 3, l.att_mu, false, outs);
            matvec(kvr, R);
        }
        WkvArgs a = wkv_base(layer);
        a.u = l.time_first; a.decay = l.time_decay;
        MatvecJob o = job(l.att_output, ws.y, ws.x, (int) D, EPI_ADD_TO);
        prefetch_next({&l.ffn_key, &l.ffn_receptance});
        if (can_fuse({&l.att_output})) {
            MatvecPrologue P;
            P.kind = PRO_WKV4; P.D = (int) D;
            P.k = ws.k; P.v = ws.v; P.r = ws.r; P.u = a.u; P.decay = a.decay;
            P.state_in = a.state_in; P.state_out = a.state_out;
            matvec({o}, R, &P);
        } else {
            chk(launch_wkv4(a, st));
            matvec({o}, R);
        }
    }

    void att_v5(uint32_t layer) {
        const LayerW & l = m.layers[layer];
        const bool gate = m.arch_minor >= 2;
        std::vector<MatvecJob> jobs = {job(l.att_key, ws.mix[0], ws.k, (int) D, EPI_STORE), job(l.att_value, ws.mix[1], ws.v, (int) D, EPI_STORE),
                                     job(l.att_receptance, ws.mix[2], ws.r, (int) D, EPI_STORE)};
        if (gate) jobs.push_back(job(l.att_gate, ws.mix[3], ws.g, (int) D, EPI_SILU));
        if (gate ? can_fuse({&l.att_key, &l.att_value, &l.att_receptance, &l.att_gate}) : can_fuse({&l.att_key, &l.att_value, &l.att_receptance})) {
            const MatvecPrologue P = ln_prologue(layer, true, MIX_LERP);
            for (size_t i = 0; i < jobs.size(); i++) jobs[i].pro_mu = l.att_mu[i];
            matvec(jobs, R, &P);
        } else {
            const Act outs[4] = {produced(ws.mix[0], {&l.att_key}), produced(ws.mix[1], {&l.att_value}), produced(ws.mix[2], {&l.att_receptance}),
                                 gate ? produced(ws.mix[3], {&l.att_gate}) : Act()};
            ln_mix(layer, true, MIX_LERP, gate ? 4 : 3, l.att_mu, false, outs);
            matvec(jobs, R);
        }
        WkvArgs a = wkv_base(layer);
        a.u = l.time_first; a.decay = l.time_decay; a.per_head_scalar = gate ? 0 : 1;
        a.g = gate ? ws.g : nullptr; a.lnx_w = l.lnx_w; a.lnx_b = l.lnx_b; a.gn_eps = 1e-5f;
        chk(launch_wkv5(a, st));
        matvec({job(l.att_output, ws.y, ws.x, (int) D, EPI_ADD_TO)}, R);
    }

    void att_v6(uint32_t layer) {
        const LayerW & l = m.layers[layer];
        const Act xxx[1] = {produced(ws.mix[0], {&l.maa_w1})};
        ln_mix(layer, true, MIX_SX, 1, l.att_mu, true, xxx);                 // xxx = x + sx*maa_x
        const int E5 = (int) l.maa_w1.M, E = E5 / 5;
        MatvecJob j1 = job(l.maa_w1, ws.mix[0], nullptr, 0, EPI_TANH);
        j1.out_act = ws.lora[0]; j1.out_act.K = E5;
        if (E5 % 32 != 0) { j1.out = ws.lora[0].f; j1.ldo = E5; j1.out_act = Act(); }
        matvec({j1}, R);
        Act t5 = ws.lora[0]; t5.K = E5; if (E5 % 32 != 0) t5.q = nullptr;
        V6Mix2Args mx; mx.x0 = ws.x0; mx.sx = ws.sx; mx.R = R; mx.D = (int) D;
        const DevMat * cons5[5] = {&l.decay_w1, &l.att_key, &l.att_value, &l.att_receptance, &l.att_gate};
        bool fold = E == 32 && D % 2 == 0 && ws.lora[0].f != nullptr;          // all five mixes feed tensor-core MATMULs: fold the rank-32 stage in
        for (int n = 0; n < 5; n++) {
            mx.m[n] = ws.m5 + (size_t) n * R * D; mx.maa[n] = l.att_mu[1 + n]; mx.out[n] = produced(ws.mix[1 + n], {cons5[n]});
            fold = fold && mx.out[n].f == nullptr && mx.out[n].h != nullptr && l.maa_w2[n].type == T_F32 && l.maa_w2[n].K == 32;
        }
        if (fold) {
            V6MixLoraArgs ml; ml.x0 = ws.x0; ml.sx = ws.sx; ml.lora = ws.lora[0].f; ml.ld_lora = E5; ml.R = R; ml.D = (int) D;
            for (int n = 0; n < 5; n++) { ml.w2[n] = (const float *) l.maa_w2[n].qs; ml.maa[n] = l.att_mu[1 + n]; ml.out_h[n] = mx.out[n].h; }
            chk(launch_v6_mix_lora(ml, st));
        } else {
            std::vector<MatvecJob> jobs;
            for (int n = 0; n < 5; n++) { MatvecJob j = job(l.maa_w2[n], t5, ws.m5 + (size_t) n * R * D, (int) D, EPI_STORE); j.x_off = n * E; jobs.push_back(j); }
            matvec(jobs, R);
            chk(launch_v6_mix2(mx, st));                                      // mix[1..5] = xw, xk, xv, xr, xg
        }
        const int DW = (int) l.decay_w1.M;
        MatvecJob jd = job(l.decay_w1, ws.mix[1], nullptr, 0, EPI_TANH); jd.out_act = ws.lora[1]; jd.out_act.K = DW;
        if (DW % 32 != 0) { jd.out = ws.lora[1].f; jd.ldo = DW; jd.out_act = Act(); }
        matvec({job(l.att_key, ws.mix[2], ws.k, (int) D, EPI_STORE), job(l.att_value, ws.mix[3], ws.v, (int) D, EPI_STORE),
              job(l.att_receptance, ws.mix[4], ws.r, (int) D, EPI_STORE), job(l.att_gate, ws.mix[5], ws.g, (int) D, EPI_SILU), jd}, R);
        Act td = ws.lora[1]; td.K = DW; if (DW % 32 != 0) td.q = nullptr;
        MatvecJob jw = job(l.decay_w2, td, ws.w, (int) D, EPI_V6_DECAY); jw.vec = l.time_decay;
        matvec({jw}, R);
        WkvArgs a = wkv_base(layer);
        a.u = l.time_first; a.w = ws.w; a.g = ws.g; a.lnx_w = l.lnx_w; a.lnx_b = l.lnx_b; a.gn_eps = 64e-5f;
        chk(launch_wkv5(a, st));
        matvec({job(l.att_output, ws.y, ws.x, (int) D, EPI_ADD_TO)}, R);
    }

    void att_v7(uint32_t layer) {
        const LayerW & l = m.layers[layer];
        const Act outs[6] = {produced(ws.mix[0], {&l.att_receptance}), produced(ws.mix[1], {&l.w1}), produced(ws.mix[2], {&l.att_key}),
                             layer != 0 ? produced(ws.mix[3], {&l.att_value, &l.v1}) : produced(ws.mix[3], {&l.att_value}),
                             produced(ws.mix[4], {&l.a1}), produced(ws.mix[5], {&l.g1})};
        ln_mix(layer, true, MIX_SX, 6, l.att_mu, false, outs);               // mix = xr, xw, xk, xv, xa, xg
        // low-rank pairs (never quantised).  Wide row counts chain the two stages on the tensor cores: the first
        // stage then also leaves an FP16 copy of its rows for the second (no q8 form: the second stage is FP16 / FP32).
        auto lora_tc = [&](const DevMat & W1, const DevMat & W2) {
            return ws.lora[0].h != nullptr && W1.M % 8 == 0 && matmul_tc_supported(W1, R, fa.batch != 0) && matmul_tc_supported(W2, R, fa.batch != 0);
        };
        const DevMat * second[4] = {&l.w2, &l.a2, &l.g2, &l.v2};
        auto lora_in = [&](int slot, const DevMat & W, int epi) {
            MatvecJob j = job(W, ws.mix[slot == 0 ? 1 : slot == 1 ? 4 : slot == 2 ? 5 : 3], nullptr, 0, epi);
            if (lora_tc(W, *second[slot])) { j.out_act.f = ws.lora[slot].f; j.out_act.h = ws.lora[slot].h; j.out_act.K = (int) W.M; }
            else { j.out = ws.lora[slot].f; j.ldo = (int) W.M; tc_written.erase(ws.lora[slot].f); }
            return j;
        };
        std::vector<MatvecJob> jobs = {job(l.att_receptance, ws.mix[0], ws.r, (int) D, EPI_STORE), job(l.att_key, ws.mix[2], ws.k, (int) D, EPI_STORE),
                                     job(l.att_value, ws.mix[3], ws.v, (int) D, EPI_STORE),
                                     lora_in(0, l.w1, EPI_TANH), lora_in(1, l.a1, EPI_STORE), lora_in(2, l.g1, EPI_SIGMOID)};
        if (layer != 0) jobs.push_back(lora_in(3, l.v1, EPI_STORE));
        matvec(jobs, R);
        auto lora_out = [&](int slot, const DevMat & W, float * out, int epi, const float * vec) {
            Act x; x.f = ws.lora[slot].f; x.K = (int) W.K;
            if (tc_written.count(ws.lora[slot].f)) x.h = ws.lora[slot].h;     // first stage ran on the tensor cores and left FP16 rows
            MatvecJob j = job(W, x, out, (int) D, epi); j.vec = vec;
            return j;
        };
        jobs = {lora_out(0, l.w2, ws.w, EPI_V7_DECAY, l.w0), lora_out(1, l.a2, ws.a, EPI_BIAS_SIGMOID, l.a0), lora_out(2, l.g2, ws.g, EPI_STORE, nullptr)};
        if (layer != 0) jobs.push_back(lora_out(3, l.v2, ws.vgate, EPI_BIAS_SIGMOID, l.v0));
        matvec(jobs, R);
        WkvArgs a = wkv_base(layer);
        a.w = ws.w; a.a = ws.a; a.g = ws.g; a.vgate = ws.vgate; a.v_first = ws.v_first;
        a.k_k = l.k_k; a.k_a = l.k_a; a.r_k = l.r_k; a.lnx_w = l.lnx_w; a.lnx_b = l.lnx_b; a.gn_eps = 64e-5f;
        chk(launch_wkv7(a, st));
        matvec({job(l.att_output, ws.y, ws.x, (int) D, EPI_ADD_TO)}, R);
    }

    void run() {
        chk(launch_embed_ln0(m.emb, ws.tokens, R, m.ln0_w, m.ln0_b, ws.x, st,
                             tp_fused ? reinterpret_cast<unsigned int *>(m.tp.p2p.base) + TP_STEP_WORD : nullptr));
        for (uint32_t i = 0; i < m.n_layer; i++) {
            pdl_site() = 1;
            switch (m.arch_major) {
                case 4: if (m.tp.world > 1) att_v4_tp(i); else att_v4(i); break;
                case 5: att_v5(i); break;
                case 6: att_v6(i); break;
                default: att_v7(i); break;
            }
            ffn(i);
        }
        pdl_site() = 0;
        if (tp_fused && fa.logits) {                          // the last ffn's partial sums are still in the inbox: materialise the final residual row
            const TpWait w = tp_wait(2 * (int) m.n_layer - 1, true);
            chk(launch_tp_gather_row(tpf, w, xres, st));
            xres = w.x_out;
        }
        if (fa.logits) {
            const int rows = fa.batch ? R : 1;
            const float * x = fa.batch ? xres : xres + (size_t) (R - 1) * D;
            pdl_site() = 32;
            chk(launch_ln_act(x, (int) D, rows, (int) D, m.ln_out_w, m.ln_out_b, produced(ws.head_in, {&m.head}, rows), st));
            if (m.tp.world > 1) {      // vocabulary rows are sharded: gather the slices (equal-sized, zero-padded at the end)
                const int Vl = (int) ((m.n_vocab + m.tp.world - 1) / m.tp.world);
                matvec({job(m.head, ws.head_in, ws.logits_part, Vl, EPI_STORE)}, rows);
                if (rows == 1) chk(tp_allgather(m.tp, ws.logits_part, ws.logits, (size_t) Vl, st));
                else err = err == cudaSuccess ? cudaErrorNotSupported : err;     // batched logits under TP: not needed by any caller yet
            } else {
                pdl_site() = 64;
                matvec({job(m.head, ws.head_in, ws.logits, (int) m.n_vocab, EPI_STORE)}, rows);
            }
        }
    }
};

} // namespace

cudaError_t forward(const Model & m, Workspace & ws, const ForwardArgs & fa, cudaStream_t st) {
    Pass p(m, ws, fa, st);
    p.run();
    return p.err;
}

} // namespace rb
