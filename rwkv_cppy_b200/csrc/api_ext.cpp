// Additive entry points (include/rwkv_b200.h): batched state-resident decode, device-timed
// runs for bench.py, and the codec / mat-vec hooks the parity tests drive.
#include "context.h"
#include "../../include/rwkv_b200.h"
#include "matrix_handle.h"

#include <cstring>
#include <new>

using namespace rb;
using namespace rbapi;

struct rwkv_b200_batch {
    rwkv_context * ctx = nullptr;          // private clone: own stream, shares the weights
    size_t n = 0;
    Workspace ws;
    float * d_state[2] = {nullptr, nullptr};     // [n][state_len]
    int cur = 0;
    int * h_tokens = nullptr;                    // pinned [n]
    cudaGraphExec_t graph[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};   // [state parity][logits]
    size_t graph_nodes[2] = {0, 0};
    // on-device sampling (allocated on first use)
    float * d_probs = nullptr, * d_u = nullptr, * h_u = nullptr;
    uint32_t * d_picked = nullptr, * h_picked = nullptr;
};

namespace {

size_t g_flush_bytes = 0;      // rwkv_b200_set_l2_flush
void * g_flush_buf = nullptr;

bool upload_matrix(int type, const void * raw, size_t M, size_t K, cudaStream_t st, std::vector<void *> & allocs, DevMat & out) {
    RB_CHECK(type_supported((uint32_t) type), RWKV_ERROR_ARGS | RWKV_ERROR_DATA_TYPE, false, "Unsupported type id %d", type);
    const size_t blk = type_block(type), bb = type_block_bytes(type);
    RB_CHECK(K % blk == 0 && M > 0 && K > 0, RWKV_ERROR_ARGS | RWKV_ERROR_SHAPE, false, "Bad matrix shape");
    const size_t n_blocks = M * K / blk, nbytes = n_blocks * bb;
    out.type = type; out.M = (uint32_t) M; out.K = (uint32_t) K; out.bytes = nbytes;
    auto alloc = [&](size_t bytes) -> void * { void * p = nullptr; if (cudaMalloc(&p, bytes) != cudaSuccess) return nullptr; allocs.push_back(p); return p; };
    if (!type_quantized(type)) {
        void * d = alloc(nbytes);
        RB_CHECK(d != nullptr, RWKV_ERROR_ALLOC, false, "cudaMalloc failed");
        RB_CUDA(cudaMemcpyAsync(d, raw, nbytes, cudaMemcpyHostToDevice, st), RWKV_ERROR_ALLOC, false);
        RB_CUDA(cudaStreamSynchronize(st), RWKV_ERROR_ALLOC, false);
        out.qs = d;
        return true;
    }
    void * scratch = nullptr;
    RB_CUDA(cudaMalloc(&scratch, nbytes), RWKV_ERROR_ALLOC, false);
    void * tiled = alloc(tiled_bytes(type, M, K));
    bool ok = tiled != nullptr;
    ok = ok && cudaMemcpyAsync(scratch, raw, nbytes, cudaMemcpyHostToDevice, st) == cudaSuccess;
    ok = ok && launch_repack(type, (const uint8_t *) scratch, M, K, tiled, st) == cudaSuccess;
    void * tc = nullptr;
    if (ok && matmul_tc_enabled() && K >= TC_KB) {
        tc = alloc(tc_bytes(type, M, K));
        ok = tc != nullptr && launch_repack_tc(type, (const uint8_t *) scratch, M, K, tc, st) == cudaSuccess;
    }
    ok = ok && cudaStreamSynchronize(st) == cudaSuccess;
    cudaFree(scratch);
    RB_CHECK(ok, RWKV_ERROR_ALLOC, false, "matrix upload failed: %s", cudaGetErrorString(cudaGetLastError()));
    out.qs = tiled; out.rg_stride = (uint32_t) tiled_rg_stride(type, K);
    if (tc) { out.tc = tc; out.tc_kblocks = (uint32_t) ((K + TC_KB - 1) / TC_KB); }
    return true;
}

bool act_dev_alloc(std::vector<void *> & allocs, Act & a, int rows, int K) {
    auto alloc = [&](size_t bytes) -> void * { void * p = nullptr; if (cudaMalloc(&p, std::max<size_t>(bytes, 16)) != cudaSuccess) return nullptr; allocs.push_back(p); return p; };
    a.K = K;
    a.f = (float *) alloc((size_t) rows * K * 4);
    if (!a.f) return false;
    if (matmul_tc_enabled()) { a.h = (uint16_t *) alloc((size_t) rows * K * 2); if (!a.h) return false; }
    if (K % 32 == 0) {
        const size_t nb = (size_t) rows * (K / 32);
        a.q = (int8_t *) alloc((size_t) rows * K); a.d = (float *) alloc(nb * 4); a.s = (float *) alloc(nb * 4); a.isum = (int *) alloc(nb * 4);
        if (!a.q || !a.d || !a.s || !a.isum) return false;
    }
    return true;
}

bool batch_graph(rwkv_b200_batch * b, int parity, int logits) {
    if (b->graph[parity][logits]) return true;
    ForwardArgs fa;
    fa.R = (int) b->n; fa.batch = 1; fa.state_in = b->d_state[parity]; fa.state_out = b->d_state[parity ^ 1]; fa.logits = logits != 0;
    // count nodes while capturing (a second, throw-away capture is cheaper than keeping the cudaGraph_t around)
    rwkv_context * ctx = b->ctx;
    cudaGraph_t g = nullptr;
    CTX_CUDA(ctx, cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal), false);
    const cudaError_t fe = forward(*ctx->model, b->ws, fa, ctx->stream);
    const cudaError_t ce = cudaStreamEndCapture(ctx->stream, &g);
    CTX_CHECK(ctx, fe == cudaSuccess && ce == cudaSuccess && g, RWKV_ERROR_GRAPH, false, "Failed to capture the batch graph: %s / %s",
              cudaGetErrorString(fe), cudaGetErrorString(ce));
    size_t nodes = 0;
    cudaGraphGetNodes(g, nullptr, &nodes);
    b->graph_nodes[logits] = nodes;
    const cudaError_t ie = cudaGraphInstantiate(&b->graph[parity][logits], g, 0);
    cudaGraphDestroy(g);
    CTX_CHECK(ctx, ie == cudaSuccess, RWKV_ERROR_GRAPH, false, "Failed to instantiate the batch graph: %s", cudaGetErrorString(ie));
    return true;
}

} // namespace

extern "C" {

int rwkv_b200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

bool rwkv_b200_set_device(int device) {
    int n = rwkv_b200_device_count();
    RB_CHECK(device >= 0 && device < n, RWKV_ERROR_ARGS, false, "Device %d out of range (%d devices)", device, n);
    t_device = device;
    return true;
}

void rwkv_b200_set_l2_flush(size_t bytes) {
    if (g_flush_buf) { cudaFree(g_flush_buf); g_flush_buf = nullptr; }
    g_flush_bytes = bytes;
}

// ---- batch ------------------------------------------------------------------------------------
struct rwkv_b200_batch * rwkv_b200_batch_create(struct rwkv_context * ctx, size_t n_streams) {
    RB_CHECK(ctx != nullptr && n_streams > 0, RWKV_ERROR_ARGS, nullptr, "Bad arguments");
    rwkv_context * c = rwkv_clone_context(ctx, ctx->n_threads);
    if (!c) return nullptr;
    rwkv_b200_batch * b = new (std::nothrow) rwkv_b200_batch();
    if (!b) { rwkv_free(c); return nullptr; }
    b->ctx = c; b->n = n_streams;
    const Model & m = *c->model;
    const size_t bytes = m.state_len() * sizeof(float) * n_streams;
    bool ok = workspace_reserve(m, b->ws, (int) n_streams, (int) n_streams);
    for (int i = 0; i < 2 && ok; i++) ok = cudaMalloc((void **) &b->d_state[i], bytes) == cudaSuccess;
    ok = ok && cudaMallocHost((void **) &b->h_tokens, n_streams * sizeof(int)) == cudaSuccess;
    if (!ok) { rwkv_b200_batch_free(b); RB_FAIL(RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, nullptr, "Failed to allocate batch buffers"); }
    for (size_t s = 0; s < n_streams; s++) rwkv_b200_batch_set_state(b, s, nullptr);
    cudaStreamSynchronize(c->stream);
    return b;
}

void rwkv_b200_batch_free(struct rwkv_b200_batch * b) {
    if (!b) return;
    if (b->ctx) { cudaSetDevice(b->ctx->model->device); cudaStreamSynchronize(b->ctx->stream); }
    for (int p = 0; p < 2; p++) for (int l = 0; l < 2; l++) if (b->graph[p][l]) cudaGraphExecDestroy(b->graph[p][l]);
    for (int i = 0; i < 2; i++) if (b->d_state[i]) cudaFree(b->d_state[i]);
    if (b->h_tokens) cudaFreeHost(b->h_tokens);
    if (b->d_probs) cudaFree(b->d_probs);
    if (b->d_u) cudaFree(b->d_u);
    if (b->d_picked) cudaFree(b->d_picked);
    if (b->h_u) cudaFreeHost(b->h_u);
    if (b->h_picked) cudaFreeHost(b->h_picked);
    b->ws.~Workspace(); new (&b->ws) Workspace();
    rwkv_free(b->ctx);
    delete b;
}

size_t rwkv_b200_batch_size(const struct rwkv_b200_batch * b) { return b ? b->n : 0; }

bool rwkv_b200_batch_set_state(struct rwkv_b200_batch * b, size_t stream, const float * state) {
    if (!b) return false;
    rwkv_context * ctx = b->ctx;
    CTX_CHECK(ctx, stream < b->n, RWKV_ERROR_ARGS, false, "Stream %zu out of range", stream);
    CTX_CUDA(ctx, cudaSetDevice(ctx->model->device), false);
    const size_t len = ctx->model->state_len();
    float * dst = b->d_state[b->cur] + stream * len;
    if (state) CTX_CUDA(ctx, cudaMemcpyAsync(dst, state, len * sizeof(float), cudaMemcpyHostToDevice, ctx->stream), false);
    else       CTX_CUDA(ctx, cudaMemcpyAsync(dst, ctx->d_init_state, len * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream), false);
    CTX_CUDA(ctx, cudaStreamSynchronize(ctx->stream), false);
    return true;
}

bool rwkv_b200_batch_get_state(struct rwkv_b200_batch * b, size_t stream, float * state) {
    if (!b) return false;
    rwkv_context * ctx = b->ctx;
    CTX_CHECK(ctx, stream < b->n && state, RWKV_ERROR_ARGS, false, "Bad arguments");
    CTX_CUDA(ctx, cudaSetDevice(ctx->model->device), false);
    const size_t len = ctx->model->state_len();
    CTX_CUDA(ctx, cudaMemcpyAsync(state, b->d_state[b->cur] + stream * len, len * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream), false);
    CTX_CUDA(ctx, cudaStreamSynchronize(ctx->stream), false);
    return true;
}

bool rwkv_b200_batch_get_state_prefix(struct rwkv_b200_batch * b, size_t n, float * out) {
    if (!b) return false;
    rwkv_context * ctx = b->ctx;
    const size_t len = ctx->model->state_len();
    CTX_CHECK(ctx, n <= len && out, RWKV_ERROR_ARGS, false, "Bad arguments");
    CTX_CUDA(ctx, cudaSetDevice(ctx->model->device), false);
    CTX_CUDA(ctx, cudaMemcpy2DAsync(out, n * sizeof(float), b->d_state[b->cur], len * sizeof(float), n * sizeof(float), b->n,
                                    cudaMemcpyDeviceToHost, ctx->stream), false);
    CTX_CUDA(ctx, cudaStreamSynchronize(ctx->stream), false);
    return true;
}

bool rwkv_b200_batch_eval(struct rwkv_b200_batch * b, const uint32_t * tokens, float * logits_out) {
    if (!b) return false;
    rwkv_context * ctx = b->ctx;
    const Model & m = *ctx->model;
    CTX_CHECK(ctx, tokens != nullptr, RWKV_ERROR_ARGS, false, "tokens is NULL");
    for (size_t i = 0; i < b->n; i++) CTX_CHECK(ctx, tokens[i] < m.n_vocab, RWKV_ERROR_ARGS, false, "Token of stream %zu (%u) is out of range", i, tokens[i]);
    CTX_CUDA(ctx, cudaSetDevice(m.device), false);
    const int lg = logits_out ? 1 : 0;
    if (!batch_graph(b, b->cur, lg)) return false;
    for (size_t i = 0; i < b->n; i++) b->h_tokens[i] = (int) tokens[i];
    CTX_CUDA(ctx, cudaMemcpyAsync(b->ws.tokens, b->h_tokens, b->n * sizeof(int), cudaMemcpyHostToDevice, ctx->stream), false);
    CTX_CUDA(ctx, cudaGraphLaunch(b->graph[b->cur][lg], ctx->stream), false);
    b->cur ^= 1;
    if (logits_out) CTX_CUDA(ctx, cudaMemcpyAsync(logits_out, b->ws.logits, b->n * (size_t) m.n_vocab * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream), false);
    CTX_CUDA(ctx, cudaStreamSynchronize(ctx->stream), false);
    CTX_CHECK(ctx, !tp_error_pending(m.tp), RWKV_ERROR_GRAPH, false, "Tensor-parallel collective timed out waiting for a peer rank; results of this call are invalid");
    return true;
}

// ---- on-device sampling ------------------------------------------------------------------------
static bool batch_sample_buffers(rwkv_b200_batch * b) {
    rwkv_context * ctx = b->ctx;
    if (b->d_probs) return true;
    const size_t V = ctx->model->n_vocab;
    CTX_CUDA(ctx, cudaMalloc((void **) &b->d_probs, b->n * V * sizeof(float)), false);
    CTX_CUDA(ctx, cudaMalloc((void **) &b->d_u, b->n * sizeof(float)), false);
    CTX_CUDA(ctx, cudaMalloc((void **) &b->d_picked, b->n * sizeof(uint32_t)), false);
    CTX_CUDA(ctx, cudaMallocHost((void **) &b->h_u, b->n * sizeof(float)), false);
    CTX_CUDA(ctx, cudaMallocHost((void **) &b->h_picked, b->n * sizeof(uint32_t)), false);
    return true;
}

bool rwkv_b200_batch_eval_sample(struct rwkv_b200_batch * b, const uint32_t * tokens, float temperature, float top_p, const float * uniforms, uint32_t * tokens_out) {
    if (!b) return false;
    rwkv_context * ctx = b->ctx;
    const Model & m = *ctx->model;
    CTX_CHECK(ctx, tokens && uniforms && tokens_out, RWKV_ERROR_ARGS, false, "NULL argument");
    CTX_CHECK(ctx, temperature >= 0.0f && top_p >= 0.0f && top_p <= 1.0f, RWKV_ERROR_ARGS, false, "temperature must be >= 0 and top_p in [0, 1]");
    for (size_t i = 0; i < b->n; i++) {
        CTX_CHECK(ctx, tokens[i] < m.n_vocab, RWKV_ERROR_ARGS, false, "Token of stream %zu (%u) is out of range", i, tokens[i]);
        CTX_CHECK(ctx, uniforms[i] >= 0.0f && uniforms[i] < 1.0f, RWKV_ERROR_ARGS, false, "uniforms[%zu] must be in [0, 1)", i);
    }
    CTX_CUDA(ctx, cudaSetDevice(m.device), false);
    if (!batch_graph(b, b->cur, 1) || !batch_sample_buffers(b)) return false;
    for (size_t i = 0; i < b->n; i++) { b->h_tokens[i] = (int) tokens[i]; b->h_u[i] = uniforms[i]; }
    CTX_CUDA(ctx, cudaMemcpyAsync(b->ws.tokens, b->h_tokens, b->n * sizeof(int), cudaMemcpyHostToDevice, ctx->stream), false);
    CTX_CUDA(ctx, cudaMemcpyAsync(b->d_u, b->h_u, b->n * sizeof(float), cudaMemcpyHostToDevice, ctx->stream), false);
    CTX_CUDA(ctx, cudaGraphLaunch(b->graph[b->cur][1], ctx->stream), false);
    b->cur ^= 1;
    CTX_CUDA(ctx, launch_sample(b->ws.logits, b->d_probs, (int) b->n, (int) m.n_vocab, temperature, top_p, b->d_u, b->d_picked, ctx->stream), false);
    CTX_CUDA(ctx, cudaMemcpyAsync(b->h_picked, b->d_picked, b->n * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream), false);
    CTX_CUDA(ctx, cudaStreamSynchronize(ctx->stream), false);
    CTX_CHECK(ctx, !tp_error_pending(m.tp), RWKV_ERROR_GRAPH, false, "Tensor-parallel collective timed out waiting for a peer rank; results of this call are invalid");
    memcpy(tokens_out, b->h_picked, b->n * sizeof(uint32_t));
    return true;
}

bool rwkv_b200_sample_logits(const float * logits, size_t n_rows, size_t n_vocab, float temperature, float top_p, const float * uniforms, uint32_t * tokens_out) {
    g_last_error = RWKV_ERROR_NONE;
    RB_CHECK(logits && uniforms && tokens_out && n_rows > 0 && n_vocab > 0, RWKV_ERROR_ARGS, false, "Bad arguments");
    RB_CHECK(temperature >= 0.0f && top_p >= 0.0f && top_p <= 1.0f, RWKV_ERROR_ARGS, false, "temperature must be >= 0 and top_p in [0, 1]");
    const int dev = pick_device();
    RB_CHECK(dev >= 0, RWKV_ERROR_CTX | RWKV_ERROR_UNSUPPORTED, false, "No CUDA device");
    RB_CUDA(cudaSetDevice(dev), RWKV_ERROR_CTX, false);
    float * d_l = nullptr, * d_p = nullptr, * d_u = nullptr; uint32_t * d_o = nullptr;
    const size_t nb = n_rows * n_vocab * sizeof(float);
    bool ok = cudaMalloc((void **) &d_l, nb) == cudaSuccess && cudaMalloc((void **) &d_p, nb) == cudaSuccess
              && cudaMalloc((void **) &d_u, n_rows * sizeof(float)) == cudaSuccess && cudaMalloc((void **) &d_o, n_rows * sizeof(uint32_t)) == cudaSuccess;
    ok = ok && cudaMemcpy(d_l, logits, nb, cudaMemcpyHostToDevice) == cudaSuccess && cudaMemcpy(d_u, uniforms, n_rows * sizeof(float), cudaMemcpyHostToDevice) == cudaSuccess;
    ok = ok && launch_sample(d_l, d_p, (int) n_rows, (int) n_vocab, temperature, top_p, d_u, d_o, nullptr) == cudaSuccess;
    ok = ok && cudaMemcpy(tokens_out, d_o, n_rows * sizeof(uint32_t), cudaMemcpyDeviceToHost) == cudaSuccess;
    cudaFree(d_l); cudaFree(d_p); cudaFree(d_u); cudaFree(d_o);
    RB_CHECK(ok, RWKV_ERROR_GRAPH, false, "sampling failed: %s", cudaGetErrorString(cudaGetLastError()));
    return true;
}

bool rwkv_b200_batch_run_timed(struct rwkv_b200_batch * b, const uint32_t * tokens, size_t n_steps, int with_logits, double * ms) {
    if (!b) return false;
    rwkv_context * ctx = b->ctx;
    const Model & m = *ctx->model;
    CTX_CHECK(ctx, tokens != nullptr && n_steps > 0 && ms, RWKV_ERROR_ARGS, false, "Bad arguments");
    for (size_t i = 0; i < b->n * n_steps; i++) CTX_CHECK(ctx, tokens[i] < m.n_vocab, RWKV_ERROR_ARGS, false, "Token %zu out of range", i);
    CTX_CUDA(ctx, cudaSetDevice(m.device), false);
    const int lg = with_logits ? 1 : 0;
    if (!batch_graph(b, 0, lg) || !batch_graph(b, 1, lg)) return false;
    int * d_all = nullptr;
    CTX_CUDA(ctx, cudaMalloc((void **) &d_all, b->n * n_steps * sizeof(int)), false);
    std::vector<int> h(b->n * n_steps);
    for (size_t i = 0; i < h.size(); i++) h[i] = (int) tokens[i];
    cudaMemcpyAsync(d_all, h.data(), h.size() * sizeof(int), cudaMemcpyHostToDevice, ctx->stream);
    cudaStreamSynchronize(ctx->stream);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaError_t err = cudaSuccess;
    float t = 0.0f;
    if (g_flush_bytes == 0) {
        cudaEventRecord(e0, ctx->stream);
        for (size_t s = 0; s < n_steps && err == cudaSuccess; s++) {
            err = cudaMemcpyAsync(b->ws.tokens, d_all + s * b->n, b->n * sizeof(int), cudaMemcpyDeviceToDevice, ctx->stream);
            if (err == cudaSuccess) err = cudaGraphLaunch(b->graph[b->cur][lg], ctx->stream);
            b->cur ^= 1;
        }
        cudaEventRecord(e1, ctx->stream);
        if (err == cudaSuccess) err = cudaStreamSynchronize(ctx->stream);
        cudaEventElapsedTime(&t, e0, e1);
    } else {
        // L2 flushed before every step; only the steps are timed (one event pair per step)
        if (!g_flush_buf) err = cudaMalloc(&g_flush_buf, g_flush_bytes);
        for (size_t s = 0; s < n_steps && err == cudaSuccess; s++) {
            cudaMemsetAsync(g_flush_buf, (int) (s & 0xFF), g_flush_bytes, ctx->stream);
            err = cudaMemcpyAsync(b->ws.tokens, d_all + s * b->n, b->n * sizeof(int), cudaMemcpyDeviceToDevice, ctx->stream);
            cudaEventRecord(e0, ctx->stream);
            if (err == cudaSuccess) err = cudaGraphLaunch(b->graph[b->cur][lg], ctx->stream);
            cudaEventRecord(e1, ctx->stream);
            b->cur ^= 1;
            if (err == cudaSuccess) err = cudaStreamSynchronize(ctx->stream);
            float dt = 0.0f;
            cudaEventElapsedTime(&dt, e0, e1);
            t += dt;
        }
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(d_all);
    CTX_CHECK(ctx, err == cudaSuccess, RWKV_ERROR_GRAPH, false, "timed run failed: %s", cudaGetErrorString(err));
    *ms = (double) t;
    return true;
}

size_t rwkv_b200_batch_launches_per_step(const struct rwkv_b200_batch * b, int with_logits) {
    if (!b) return 0;
    rwkv_b200_batch * bb = const_cast<rwkv_b200_batch *>(b);
    if (!batch_graph(bb, bb->cur, with_logits ? 1 : 0)) return 0;
    return b->graph_nodes[with_logits ? 1 : 0];
}

bool rwkv_b200_sequence_run_timed(struct rwkv_context * ctx, const uint32_t * tokens, size_t n, int with_logits, int iters, double * ms) {
    if (!ctx) return false;
    const Model & m = *ctx->model;
    CTX_CHECK(ctx, tokens && n > 0 && n <= SEQ_MAX_ROWS && iters > 0 && ms, RWKV_ERROR_ARGS, false, "Bad arguments (n must be <= %d)", SEQ_MAX_ROWS);
    for (size_t i = 0; i < n; i++) CTX_CHECK(ctx, tokens[i] < m.n_vocab, RWKV_ERROR_ARGS, false, "Token %zu out of range", i);
    CTX_CUDA(ctx, cudaSetDevice(m.device), false);
    CTX_CHECK(ctx, workspace_reserve(m, ctx->wsn, (int) n, 1), RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, false, "Failed to allocate the sequence workspace");
    std::vector<int> h(n);
    for (size_t i = 0; i < n; i++) h[i] = (int) tokens[i];
    CTX_CUDA(ctx, cudaMemcpyAsync(ctx->wsn.tokens, h.data(), n * sizeof(int), cudaMemcpyHostToDevice, ctx->stream), false);
    CTX_CUDA(ctx, cudaStreamSynchronize(ctx->stream), false);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaError_t err = cudaSuccess;
    cudaEventRecord(e0, ctx->stream);
    for (int it = 0; it < iters && err == cudaSuccess; it++) {
        err = cudaMemcpyAsync(ctx->d_state[0], ctx->d_init_state, m.state_len() * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream);
        ForwardArgs fa; fa.R = (int) n; fa.batch = 0; fa.state_in = ctx->d_state[0]; fa.state_out = ctx->d_state[1]; fa.logits = with_logits != 0;
        if (err == cudaSuccess) err = forward(m, ctx->wsn, fa, ctx->stream);
    }
    cudaEventRecord(e1, ctx->stream);
    if (err == cudaSuccess) err = cudaStreamSynchronize(ctx->stream);
    float t = 0.0f;
    cudaEventElapsedTime(&t, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    CTX_CHECK(ctx, err == cudaSuccess, RWKV_ERROR_GRAPH, false, "timed run failed: %s", cudaGetErrorString(err));
    *ms = (double) t / iters;
    return true;
}

// ---- tensor parallel ---------------------------------------------------------------------------
bool rwkv_b200_tp_unique_id(void * out_128_bytes) {
    g_last_error = RWKV_ERROR_NONE;
    RB_CHECK(out_128_bytes != nullptr, RWKV_ERROR_ARGS, false, "out is NULL");
    RB_CHECK(tp_available(), RWKV_ERROR_CTX | RWKV_ERROR_UNSUPPORTED, false, "NCCL (libnccl.so.2) not found");
    RB_CHECK(tp_unique_id(out_128_bytes), RWKV_ERROR_CTX, false, "ncclGetUniqueId failed");
    return true;
}

struct rwkv_context * rwkv_b200_tp_init_from_file(const char * model_file_path, int rank, int world_size, const void * unique_id_128_bytes) {
    g_last_error = RWKV_ERROR_NONE;
    RB_CHECK(model_file_path && unique_id_128_bytes, RWKV_ERROR_ARGS, nullptr, "NULL argument");
    RB_CHECK(world_size >= 2 && rank >= 0 && rank < world_size, RWKV_ERROR_ARGS, nullptr, "Bad rank %d / world size %d", rank, world_size);
    RB_CHECK(tp_available(), RWKV_ERROR_CTX | RWKV_ERROR_UNSUPPORTED, nullptr, "NCCL (libnccl.so.2) not found");
    const int dev = pick_device();
    RB_CHECK(dev != -2, RWKV_ERROR_ARGS | RWKV_ERROR_CTX, nullptr, "LOCAL_RANK / rwkv_b200_set_device names a CUDA device that does not exist");
    RB_CHECK(dev >= 0, RWKV_ERROR_CTX | RWKV_ERROR_UNSUPPORTED, nullptr, "No CUDA device: this engine has no CPU path");
    RB_CUDA(cudaSetDevice(dev), RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, nullptr);
    TpInfo tp; tp.rank = rank; tp.world = world_size;
    RB_CHECK(tp_comm_init(tp, unique_id_128_bytes), RWKV_ERROR_CTX, nullptr, "ncclCommInitRank failed");
    Model * model = model_load(model_file_path, tp);
    if (!model) { tp_comm_destroy(tp); return nullptr; }
    rwkv_context * ctx = new (std::nothrow) rwkv_context();
    if (!ctx) { model_release(model); RB_FAIL(RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, nullptr, "Failed to allocate context"); }
    ctx->model = model; ctx->print_errors = g_print_errors;
    if (!ctx_setup(ctx)) { ctx_destroy(ctx); return nullptr; }
    return ctx;
}

bool rwkv_b200_tp_ipc_handle(struct rwkv_context * ctx, void * out_64_bytes) {
    RB_CHECK(ctx && out_64_bytes, RWKV_ERROR_ARGS, false, "NULL argument");
    CTX_CHECK(ctx, ctx->model->tp.world > 1, RWKV_ERROR_ARGS, false, "Not a tensor-parallel context");
    CTX_CUDA(ctx, cudaSetDevice(ctx->model->device), false);
    CTX_CHECK(ctx, tp_p2p_alloc(ctx->model->tp, out_64_bytes), RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, false, "Failed to allocate / export the peer-memory region");
    return true;
}

bool rwkv_b200_tp_connect(struct rwkv_context * ctx, const void * handles) {
    RB_CHECK(ctx && handles, RWKV_ERROR_ARGS, false, "NULL argument");
    CTX_CHECK(ctx, ctx->model->tp.world > 1, RWKV_ERROR_ARGS, false, "Not a tensor-parallel context");
    CTX_CUDA(ctx, cudaSetDevice(ctx->model->device), false);
    CTX_CHECK(ctx, tp_p2p_connect(ctx->model->tp, handles), RWKV_ERROR_CTX, false, "Failed to map the peers' memory (no P2P / NVLink access?)");
    for (int i = 0; i < 2; i++) if (ctx->graph[i]) { cudaGraphExecDestroy(ctx->graph[i]); ctx->graph[i] = nullptr; }   // captured with NCCL nodes
    return true;
}

// ---- codec / mat-vec hooks ----------------------------------------------------------------------
bool rwkv_b200_dequantize(int type, const void * raw, size_t n_rows, size_t n_cols, float * out) {
    g_last_error = RWKV_ERROR_NONE;
    const int dev = pick_device();
    RB_CHECK(dev >= 0, RWKV_ERROR_CTX | RWKV_ERROR_UNSUPPORTED, false, "No CUDA device");
    RB_CUDA(cudaSetDevice(dev), RWKV_ERROR_CTX, false);
    std::vector<void *> allocs;
    DevMat W;
    bool ok = upload_matrix(type, raw, n_rows, n_cols, nullptr, allocs, W);
    float * d = nullptr;
    ok = ok && cudaMalloc((void **) &d, n_rows * n_cols * sizeof(float)) == cudaSuccess;
    ok = ok && launch_dequant(W, d, nullptr) == cudaSuccess;
    ok = ok && cudaMemcpy(out, d, n_rows * n_cols * sizeof(float), cudaMemcpyDeviceToHost) == cudaSuccess;
    if (d) cudaFree(d);
    for (void * p : allocs) cudaFree(p);
    RB_CHECK(ok, RWKV_ERROR_GRAPH, false, "dequantize failed: %s", cudaGetErrorString(cudaGetLastError()));
    return true;
}

struct rwkv_b200_matrix * rwkv_b200_matrix_create(int type, const void * raw, size_t n_rows, size_t n_cols, size_t copies) {
    g_last_error = RWKV_ERROR_NONE;
    const int dev = pick_device();
    RB_CHECK(dev >= 0, RWKV_ERROR_CTX | RWKV_ERROR_UNSUPPORTED, nullptr, "No CUDA device");
    RB_CUDA(cudaSetDevice(dev), RWKV_ERROR_CTX, nullptr);
    RB_CHECK(copies > 0, RWKV_ERROR_ARGS, nullptr, "copies must be > 0");
    rwkv_b200_matrix * m = new (std::nothrow) rwkv_b200_matrix();
    RB_CHECK(m != nullptr, RWKV_ERROR_ALLOC, nullptr, "alloc failed");
    m->device = dev;
    bool ok = cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking) == cudaSuccess && matvec_async_init() == cudaSuccess
              && (!matmul_tc_enabled() || matmul_tc_init() == cudaSuccess);
    m->copies.resize(copies);
    for (size_t c = 0; c < copies && ok; c++) ok = upload_matrix(type, raw, n_rows, n_cols, m->stream, m->allocations, m->copies[c]);
    if (!ok) { rwkv_b200_matrix_free(m); return nullptr; }
    return m;
}

void rwkv_b200_matrix_free(struct rwkv_b200_matrix * m) {
    if (!m) return;
    cudaSetDevice(m->device);
    if (m->stream) { cudaStreamSynchronize(m->stream); cudaStreamDestroy(m->stream); }
    for (void * p : m->allocations) cudaFree(p);
    if (m->y) cudaFree(m->y);
    delete m;
}

size_t rwkv_b200_matrix_bytes(const struct rwkv_b200_matrix * m) { return m && !m->copies.empty() ? m->copies[0].bytes : 0; }

int rwkv_b200_matvec_k_split(size_t n_rows, size_t n_cols) { return matvec_ksplit((uint32_t) n_rows, (uint32_t) n_cols); }

static bool matrix_prepare(rwkv_b200_matrix * m, size_t n_vec) {
    const DevMat & W = m->copies[0];
    if (!m->scratch.buf && matmul_tc_enabled()) {
        void * p = nullptr, * f = nullptr;
        RB_CUDA(cudaMalloc(&p, MATMUL_SCRATCH_FLOATS * sizeof(float)), RWKV_ERROR_ALLOC, false);
        m->allocations.push_back(p);
        RB_CUDA(cudaMalloc(&f, MATMUL_SCRATCH_FLAGS * sizeof(unsigned int)), RWKV_ERROR_ALLOC, false);
        m->allocations.push_back(f);
        RB_CUDA(cudaMemset(f, 0, MATMUL_SCRATCH_FLAGS * sizeof(unsigned int)), RWKV_ERROR_ALLOC, false);
        m->scratch.buf = (float *) p; m->scratch.flags = (unsigned int *) f; m->scratch.floats = MATMUL_SCRATCH_FLOATS;
    }
    if ((int) n_vec > m->x_rows) {
        RB_CHECK(act_dev_alloc(m->allocations, m->x, (int) n_vec, (int) W.K), RWKV_ERROR_ALLOC, false, "alloc failed");
        m->x_rows = (int) n_vec;
    }
    if (n_vec * W.M > m->y_cap) {
        if (m->y) cudaFree(m->y);
        m->y = nullptr; m->y_cap = 0;
        RB_CUDA(cudaMalloc((void **) &m->y, n_vec * W.M * sizeof(float)), RWKV_ERROR_ALLOC, false);
        m->y_cap = n_vec * W.M;
    }
    return true;
}

bool rwkv_b200_matrix_mul(struct rwkv_b200_matrix * m, const float * x, size_t n_vec, float * y) {
    RB_CHECK(m && x && y && n_vec > 0, RWKV_ERROR_ARGS, false, "Bad arguments");
    RB_CUDA(cudaSetDevice(m->device), RWKV_ERROR_CTX, false);
    const DevMat & W = m->copies[m->next % m->copies.size()];
    m->next++;
    if (!matrix_prepare(m, n_vec)) return false;
    RB_CUDA(cudaMemcpyAsync(m->x.f, x, n_vec * W.K * sizeof(float), cudaMemcpyHostToDevice, m->stream), RWKV_ERROR_GRAPH, false);
    RB_CUDA(launch_quantize_act(m->x, (int) n_vec, m->stream), RWKV_ERROR_GRAPH, false);
    MatvecBatch b; b.T = (int) n_vec; b.n_jobs = 1;
    b.job[0].W = W; b.job[0].x = m->x; b.job[0].out = m->y; b.job[0].ldo = (int) W.M; b.job[0].epi = EPI_STORE;
    RB_CUDA(launch_matvec(b, m->stream), RWKV_ERROR_GRAPH, false);
    RB_CUDA(cudaMemcpyAsync(y, m->y, n_vec * W.M * sizeof(float), cudaMemcpyDeviceToHost, m->stream), RWKV_ERROR_GRAPH, false);
    RB_CUDA(cudaStreamSynchronize(m->stream), RWKV_ERROR_GRAPH, false);
    return true;
}

bool rwkv_b200_matrix_mul_tc(struct rwkv_b200_matrix * m, const float * x, size_t n_vec, float * y) {
    RB_CHECK(m && x && y && n_vec > 0, RWKV_ERROR_ARGS, false, "Bad arguments");
    RB_CUDA(cudaSetDevice(m->device), RWKV_ERROR_CTX, false);
    const DevMat & W = m->copies[m->next % m->copies.size()];
    m->next++;
    RB_CHECK(matmul_tc_enabled() && (W.type == T_F16 || W.tc != nullptr) && W.K >= 64 && W.K % 8 == 0, RWKV_ERROR_ARGS | RWKV_ERROR_UNSUPPORTED, false,
             "Matrix not eligible for the tensor-core path");
    if (!matrix_prepare(m, n_vec)) return false;
    RB_CUDA(cudaMemcpyAsync(m->x.f, x, n_vec * W.K * sizeof(float), cudaMemcpyHostToDevice, m->stream), RWKV_ERROR_GRAPH, false);
    RB_CUDA(launch_to_half(m->x.f, m->x.h, n_vec * W.K, m->stream), RWKV_ERROR_GRAPH, false);
    MatmulJob g; g.W = W; g.x16 = m->x.h; g.ldx = (int) W.K; g.out = m->y; g.ldo = (int) W.M; g.epi = EPI_STORE;
    RB_CUDA(launch_matmul_tc(&g, 1, (int) n_vec, m->stream, &m->scratch), RWKV_ERROR_GRAPH, false);
    RB_CUDA(cudaMemcpyAsync(y, m->y, n_vec * W.M * sizeof(float), cudaMemcpyDeviceToHost, m->stream), RWKV_ERROR_GRAPH, false);
    RB_CUDA(cudaStreamSynchronize(m->stream), RWKV_ERROR_GRAPH, false);
    return true;
}

bool rwkv_b200_matrix_bench_tc(struct rwkv_b200_matrix * m, size_t n_vec, int iters, double * ms_per_launch) {
    RB_CHECK(m && n_vec > 0 && iters > 0 && ms_per_launch, RWKV_ERROR_ARGS, false, "Bad arguments");
    RB_CUDA(cudaSetDevice(m->device), RWKV_ERROR_CTX, false);
    const DevMat & W0 = m->copies[0];
    RB_CHECK(matmul_tc_enabled() && (W0.type == T_F16 || W0.tc != nullptr) && W0.K >= 64 && W0.K % 8 == 0, RWKV_ERROR_ARGS | RWKV_ERROR_UNSUPPORTED, false,
             "Matrix not eligible for the tensor-core path");
    if (!matrix_prepare(m, n_vec)) return false;
    RB_CUDA(cudaMemsetAsync(m->x.h, 0, n_vec * W0.K * 2, m->stream), RWKV_ERROR_GRAPH, false);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaError_t err = cudaSuccess;
    auto one = [&](size_t c) {
        MatmulJob g; g.W = m->copies[c % m->copies.size()]; g.x16 = m->x.h; g.ldx = (int) W0.K; g.out = m->y; g.ldo = (int) W0.M; g.epi = EPI_STORE;
        return launch_matmul_tc(&g, 1, (int) n_vec, m->stream, &m->scratch);
    };
    for (size_t c = 0; c < m->copies.size() && err == cudaSuccess; c++) err = one(c);
    cudaEventRecord(e0, m->stream);
    for (int i = 0; i < iters && err == cudaSuccess; i++) err = one((size_t) i);
    cudaEventRecord(e1, m->stream);
    if (err == cudaSuccess) err = cudaStreamSynchronize(m->stream);
    float t = 0.0f;
    cudaEventElapsedTime(&t, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    RB_CHECK(err == cudaSuccess, RWKV_ERROR_GRAPH, false, "bench failed: %s", cudaGetErrorString(err));
    *ms_per_launch = (double) t / iters;
    return true;
}

bool rwkv_b200_matrix_bench(struct rwkv_b200_matrix * m, size_t n_vec, int iters, double * ms_per_launch) {
    RB_CHECK(m && n_vec > 0 && iters > 0 && ms_per_launch, RWKV_ERROR_ARGS, false, "Bad arguments");
    RB_CUDA(cudaSetDevice(m->device), RWKV_ERROR_CTX, false);
    if (!matrix_prepare(m, n_vec)) return false;
    const DevMat & W0 = m->copies[0];
    RB_CUDA(cudaMemsetAsync(m->x.f, 0, n_vec * W0.K * sizeof(float), m->stream), RWKV_ERROR_GRAPH, false);
    RB_CUDA(launch_quantize_act(m->x, (int) n_vec, m->stream), RWKV_ERROR_GRAPH, false);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaError_t err = cudaSuccess;
    auto one = [&](size_t c) {
        MatvecBatch b; b.T = (int) n_vec; b.n_jobs = 1;
        b.job[0].W = m->copies[c % m->copies.size()]; b.job[0].x = m->x; b.job[0].out = m->y; b.job[0].ldo = (int) W0.M; b.job[0].epi = EPI_STORE;
        return launch_matvec(b, m->stream);
    };
    for (size_t c = 0; c < m->copies.size() && err == cudaSuccess; c++) err = one(c);     // warm-up, touches every copy
    cudaEventRecord(e0, m->stream);
    for (int i = 0; i < iters && err == cudaSuccess; i++) err = one((size_t) i);
    cudaEventRecord(e1, m->stream);
    if (err == cudaSuccess) err = cudaStreamSynchronize(m->stream);
    float t = 0.0f;
    cudaEventElapsedTime(&t, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    RB_CHECK(err == cudaSuccess, RWKV_ERROR_GRAPH, false, "bench failed: %s", cudaGetErrorString(err));
    *ms_per_launch = (double) t / iters;
    return true;
}

} // extern "C"
