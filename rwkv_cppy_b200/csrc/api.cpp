// The rwkv.h C ABI over the engine.  Mirrors the behaviour of rwkv.cpp:71-258 and
// rwkv_eval.inc:2-241 (argument checks, error words, NULL conventions); the compute is the
// CUDA launch sequence in forward.cpp, captured as a CUDA graph for single-token decode.
#include "context.h"

#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>

using namespace rb;

namespace rbapi {

void ctx_fail(rwkv_context * ctx, int flags, const char * file, int line, const char * fmt, ...) {
    ctx->last_error |= flags;
    if (!ctx->print_errors) return;
    va_list ap; va_start(ap, fmt); vfprintf(stderr, fmt, ap); va_end(ap);
    fprintf(stderr, "\n%s:%d: error 0x%x\n", file, line, flags);
}
thread_local int t_device = -1;

// -1: no CUDA device; -2: LOCAL_RANK names a device that does not exist (several ranks silently sharing GPU 0 would
// "work" and, under tensor parallelism, deadlock on each other).
int pick_device() {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) { cudaGetLastError(); return -1; }
    if (t_device >= 0) return t_device < n ? t_device : -2;
    const char * lr = getenv("LOCAL_RANK");
    const int d = lr ? atoi(lr) : 0;
    if (d < 0 || d >= n) return -2;
    return d;
}

void host_init_state(const Model & m, float * s) {
    // rwkv_eval.inc:224-241: zeros; v4 pp slots (5th vector of every layer) = -1e30
    const size_t n = m.state_len();
    memset(s, 0, n * sizeof(float));
    if (m.arch_major >= 5) return;
    const size_t D = m.n_embed;
    for (uint32_t l = 0; l < m.n_layer; l++)
        for (size_t i = 0; i < D; i++) s[(size_t) l * 5 * D + 4 * D + i] = -1e30f;
}

bool ctx_setup(rwkv_context * ctx) {
    const Model & m = *ctx->model;
    RB_CUDA(cudaSetDevice(m.device), RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, false);
    RB_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking), RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, false);
    const size_t n = m.state_len();
    for (int i = 0; i < 2; i++) {
        RB_CUDA(cudaMalloc((void **) &ctx->d_state[i], n * sizeof(float)), RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, false);
        // tensor parallel: a rank only writes the aa / bb / pp slices of its own channels; the rest of state_out must not be uninitialised memory
        RB_CUDA(cudaMemset(ctx->d_state[i], 0, n * sizeof(float)), RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, false);
    }
    RB_CUDA(cudaMalloc((void **) &ctx->d_init_state, n * sizeof(float)), RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, false);
    std::vector<float> init(n);
    host_init_state(m, init.data());
    RB_CUDA(cudaMemcpy(ctx->d_init_state, init.data(), n * sizeof(float), cudaMemcpyHostToDevice), RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, false);
    RB_CHECK(workspace_reserve(m, ctx->ws1, 1, 1), RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, false, "Failed to allocate the decode workspace");
    ctx->h_tokens_cap = 64;
    RB_CUDA(cudaMallocHost((void **) &ctx->h_tokens, ctx->h_tokens_cap * sizeof(int)), RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, false);
    return true;
}

void model_release(Model * m) {
    if (m && m->refcount.fetch_sub(1) == 1) delete m;
}

void ctx_destroy(rwkv_context * ctx) {
    if (!ctx) return;
    if (ctx->model) cudaSetDevice(ctx->model->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    for (int i = 0; i < 2; i++) if (ctx->graph[i]) cudaGraphExecDestroy(ctx->graph[i]);
    for (int i = 0; i < 2; i++) if (ctx->d_state[i]) cudaFree(ctx->d_state[i]);
    if (ctx->d_init_state) cudaFree(ctx->d_init_state);
    if (ctx->h_tokens) cudaFreeHost(ctx->h_tokens);
    ctx->ws1.~Workspace(); new (&ctx->ws1) Workspace();
    ctx->wsn.~Workspace(); new (&ctx->wsn) Workspace();
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    model_release(ctx->model);
    delete ctx;
}

bool capture_graph(rwkv_context * ctx, cudaStream_t st, const Model & m, Workspace & ws, const ForwardArgs & fa, cudaGraphExec_t * out) {
    cudaGraph_t g = nullptr;
    CTX_CUDA(ctx, cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal), false);
    const cudaError_t fe = forward(m, ws, fa, st);
    const cudaError_t ce = cudaStreamEndCapture(st, &g);
    CTX_CHECK(ctx, fe == cudaSuccess && ce == cudaSuccess && g, RWKV_ERROR_GRAPH, false, "Failed to capture the launch graph: %s / %s",
              cudaGetErrorString(fe), cudaGetErrorString(ce));
    const cudaError_t ie = cudaGraphInstantiate(out, g, 0);
    cudaGraphDestroy(g);
    CTX_CHECK(ctx, ie == cudaSuccess, RWKV_ERROR_GRAPH, false, "Failed to instantiate the launch graph: %s", cudaGetErrorString(ie));
    return true;
}

bool load_state(rwkv_context * ctx, const float * state_in, float * dst) {
    const size_t bytes = ctx->model->state_len() * sizeof(float);
    if (state_in) CTX_CUDA(ctx, cudaMemcpyAsync(dst, state_in, bytes, cudaMemcpyHostToDevice, ctx->stream), false);
    else          CTX_CUDA(ctx, cudaMemcpyAsync(dst, ctx->d_init_state, bytes, cudaMemcpyDeviceToDevice, ctx->stream), false);
    return true;
}

bool ensure_tokens(rwkv_context * ctx, size_t n) {
    if (n <= ctx->h_tokens_cap) return true;
    cudaFreeHost(ctx->h_tokens); ctx->h_tokens = nullptr; ctx->h_tokens_cap = 0;
    CTX_CUDA(ctx, cudaMallocHost((void **) &ctx->h_tokens, n * sizeof(int)), false);
    ctx->h_tokens_cap = n;
    return true;
}

// Sequence pass over tokens[0..n) tiled by `tile` rows; state in d_state[cur] on entry, result index returned in cur.
bool run_sequence(rwkv_context * ctx, const uint32_t * tokens, size_t n, size_t tile, int & cur, bool want_logits) {
    const Model & m = *ctx->model;
    tile = std::min<size_t>(tile, SEQ_MAX_ROWS);
    CTX_CHECK(ctx, workspace_reserve(m, ctx->wsn, (int) std::min(tile, n), 1), RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, false, "Failed to allocate the sequence workspace");
    if (!ensure_tokens(ctx, std::min(tile, n))) return false;
    for (size_t pos = 0; pos < n; pos += tile) {
        const size_t cnt = std::min(tile, n - pos);
        // the pinned token buffer is reused: wait for the previous tile's copy
        if (pos) CTX_CUDA(ctx, cudaStreamSynchronize(ctx->stream), false);
        for (size_t i = 0; i < cnt; i++) ctx->h_tokens[i] = (int) tokens[pos + i];
        CTX_CUDA(ctx, cudaMemcpyAsync(ctx->wsn.tokens, ctx->h_tokens, cnt * sizeof(int), cudaMemcpyHostToDevice, ctx->stream), false);
        ForwardArgs fa;
        fa.R = (int) cnt; fa.batch = 0; fa.state_in = ctx->d_state[cur]; fa.state_out = ctx->d_state[cur ^ 1];
        fa.logits = want_logits && (pos + cnt == n);
        CTX_CUDA(ctx, forward(m, ctx->wsn, fa, ctx->stream), false);
        cur ^= 1;
    }
    return true;
}

bool store_outputs(rwkv_context * ctx, const float * d_state, const float * d_logits, float * state_out, float * logits_out) {
    const Model & m = *ctx->model;
    if (state_out)  CTX_CUDA(ctx, cudaMemcpyAsync(state_out, d_state, m.state_len() * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream), false);
    if (logits_out) CTX_CUDA(ctx, cudaMemcpyAsync(logits_out, d_logits, (size_t) m.n_vocab * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream), false);
    CTX_CUDA(ctx, cudaStreamSynchronize(ctx->stream), false);
    // a collective kernel that gave up waiting for a peer (dead or > 2 s late) has summed garbage: fail the call instead of returning it
    CTX_CHECK(ctx, !tp_error_pending(m.tp), RWKV_ERROR_GRAPH, false, "Tensor-parallel collective timed out waiting for a peer rank; results of this call are invalid");
    return true;
}

} // namespace rbapi
using namespace rbapi;

// ============================================================================================
// rwkv.h
// ============================================================================================
extern "C" {

void rwkv_set_print_errors(struct rwkv_context * ctx, const bool print_errors) {
    bool * p = ctx ? &ctx->print_errors : &g_print_errors;
    *p = print_errors;
}

bool rwkv_get_print_errors(const struct rwkv_context * ctx) { return ctx ? ctx->print_errors : g_print_errors; }

enum rwkv_error_flags rwkv_get_last_error(struct rwkv_context * ctx) {
    int * p = ctx ? &ctx->last_error : &g_last_error;
    const int v = *p;
    *p = RWKV_ERROR_NONE;
    return (enum rwkv_error_flags) v;
}

struct rwkv_context * rwkv_init_from_file(const char * model_file_path, const uint32_t n_threads, const uint32_t n_gpu_layers) {
    (void) n_gpu_layers;                 // the whole model always lives on the GPU
    g_last_error = RWKV_ERROR_NONE;
    RB_CHECK(model_file_path != nullptr, RWKV_ERROR_ARGS, nullptr, "model_file_path is NULL");
    const int dev = pick_device();
    RB_CHECK(dev != -2, RWKV_ERROR_ARGS | RWKV_ERROR_CTX, nullptr, "LOCAL_RANK / rwkv_b200_set_device names a CUDA device that does not exist");
    RB_CHECK(dev >= 0, RWKV_ERROR_CTX | RWKV_ERROR_UNSUPPORTED, nullptr, "No CUDA device: this engine has no CPU path");
    RB_CUDA(cudaSetDevice(dev), RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, nullptr);
    Model * model = model_load(model_file_path);
    if (!model) return nullptr;
    rwkv_context * ctx = new (std::nothrow) rwkv_context();
    if (!ctx) { model_release(model); RB_FAIL(RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, nullptr, "Failed to allocate context"); }
    ctx->model = model; ctx->n_threads = n_threads; ctx->print_errors = g_print_errors;
    if (!ctx_setup(ctx)) { ctx_destroy(ctx); return nullptr; }
    return ctx;
}

struct rwkv_context * rwkv_clone_context(struct rwkv_context * ctx, const uint32_t n_threads) {
    RB_CHECK(ctx != nullptr, RWKV_ERROR_ARGS, nullptr, "ctx is NULL");
    rwkv_context * c = new (std::nothrow) rwkv_context();
    RB_CHECK(c != nullptr, RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, nullptr, "Failed to allocate context");
    ctx->model->refcount.fetch_add(1);
    c->model = ctx->model; c->n_threads = n_threads; c->print_errors = ctx->print_errors;
    if (!ctx_setup(c)) { ctx_destroy(c); return nullptr; }
    return c;
}

bool rwkv_eval(struct rwkv_context * ctx, const uint32_t token, const float * state_in, float * state_out, float * logits_out) {
    if (!ctx) return false;
    const Model & m = *ctx->model;
    ctx->last_error = RWKV_ERROR_NONE;
    CTX_CHECK(ctx, token < m.n_vocab, RWKV_ERROR_ARGS, false, "Token (%u) is out of range (0 .. %u)", token, m.n_vocab - 1);
    CTX_CUDA(ctx, cudaSetDevice(m.device), false);
    const int gi = logits_out ? 1 : 0;
    if (!ctx->graph[gi]) {
        ForwardArgs fa; fa.R = 1; fa.batch = 0; fa.state_in = ctx->d_state[0]; fa.state_out = ctx->d_state[1]; fa.logits = gi == 1;
        if (!capture_graph(ctx, ctx->stream, m, ctx->ws1, fa, &ctx->graph[gi])) return false;
    }
    if (!load_state(ctx, state_in, ctx->d_state[0])) return false;
    ctx->h_tokens[0] = (int) token;
    CTX_CUDA(ctx, cudaMemcpyAsync(ctx->ws1.tokens, ctx->h_tokens, sizeof(int), cudaMemcpyHostToDevice, ctx->stream), false);
    CTX_CUDA(ctx, cudaGraphLaunch(ctx->graph[gi], ctx->stream), false);
    return store_outputs(ctx, ctx->d_state[1], ctx->ws1.logits, state_out, logits_out);
}

bool rwkv_eval_sequence(struct rwkv_context * ctx, const uint32_t * tokens, const size_t sequence_len, const float * state_in, float * state_out, float * logits_out) {
    if (!ctx) return false;
    const Model & m = *ctx->model;
    ctx->last_error = RWKV_ERROR_NONE;
    CTX_CHECK(ctx, sequence_len > 0, RWKV_ERROR_ARGS, false, "Sequence length is 0");
    if (!tokens) return true;            // "prepare only" (rwkv.h:135): nothing to build ahead of time here
    for (size_t i = 0; i < sequence_len; i++)
        CTX_CHECK(ctx, tokens[i] < m.n_vocab, RWKV_ERROR_ARGS, false, "Token at index %zu (%u) is out of range (0 .. %u)", i, tokens[i], m.n_vocab - 1);
    if (sequence_len == 1) return rwkv_eval(ctx, tokens[0], state_in, state_out, logits_out);
    CTX_CUDA(ctx, cudaSetDevice(m.device), false);
    int cur = 0;
    if (!load_state(ctx, state_in, ctx->d_state[cur])) return false;
    if (!run_sequence(ctx, tokens, sequence_len, SEQ_MAX_ROWS, cur, logits_out != nullptr)) return false;
    return store_outputs(ctx, ctx->d_state[cur], ctx->wsn.logits, state_out, logits_out);
}

bool rwkv_eval_sequence_in_chunks(struct rwkv_context * ctx, const uint32_t * tokens, const size_t sequence_len, const size_t chunk_size,
                                  const float * state_in, float * state_out, float * logits_out) {
    if (!ctx) return false;
    const Model & m = *ctx->model;
    ctx->last_error = RWKV_ERROR_NONE;
    CTX_CHECK(ctx, sequence_len > 0, RWKV_ERROR_ARGS, false, "Sequence length is 0");
    CTX_CHECK(ctx, chunk_size > 0, RWKV_ERROR_ARGS, false, "Chunk size is 0");
    CTX_CHECK(ctx, tokens != nullptr, RWKV_ERROR_ARGS, false, "tokens is NULL");
    for (size_t i = 0; i < sequence_len; i++)
        CTX_CHECK(ctx, tokens[i] < m.n_vocab, RWKV_ERROR_ARGS, false, "Token at index %zu (%u) is out of range (0 .. %u)", i, tokens[i], m.n_vocab - 1);
    if (sequence_len == 1) return rwkv_eval(ctx, tokens[0], state_in, state_out, logits_out);
    CTX_CUDA(ctx, cudaSetDevice(m.device), false);
    int cur = 0;
    if (!load_state(ctx, state_in, ctx->d_state[cur])) return false;
    // The caller's chunk size only exists to stay under ggml's node limit (rwkv.h:121-155); there is no such limit here, and tiling by it
    // would make FP16 / quantised results depend on it (rows >= 64 take the tensor-core path).  The engine's own tiling is used whatever
    // the chunk size, so the call equals rwkv_eval_sequence bit for bit.
    if (!run_sequence(ctx, tokens, sequence_len, SEQ_MAX_ROWS, cur, logits_out != nullptr)) return false;
    return store_outputs(ctx, ctx->d_state[cur], ctx->wsn.logits, state_out, logits_out);
}

size_t rwkv_get_n_vocab(const struct rwkv_context * ctx) { return ctx->model->n_vocab; }
size_t rwkv_get_n_embed(const struct rwkv_context * ctx) { return ctx->model->n_embed; }
size_t rwkv_get_n_layer(const struct rwkv_context * ctx) { return ctx->model->n_layer; }
size_t rwkv_get_state_len(const struct rwkv_context * ctx) { return ctx->model->state_len(); }
size_t rwkv_get_logits_len(const struct rwkv_context * ctx) { return ctx->model->n_vocab; }
uint32_t rwkv_get_state_buffer_element_count(const struct rwkv_context * ctx) { return (uint32_t) rwkv_get_state_len(ctx); }
uint32_t rwkv_get_logits_buffer_element_count(const struct rwkv_context * ctx) { return (uint32_t) rwkv_get_logits_len(ctx); }

void rwkv_init_state(const struct rwkv_context * ctx, float * state) { host_init_state(*ctx->model, state); }

void rwkv_free(struct rwkv_context * ctx) { ctx_destroy(ctx); }

bool rwkv_gpu_offload_layers(struct rwkv_context * ctx, const uint32_t n_layers) { (void) n_layers; return ctx != nullptr; }

const char * rwkv_get_system_info_string(void) {
    static std::string s;
    static std::once_flag once;
    std::call_once(once, [] {
        int n = 0, sm = 0, nsm = 0; size_t free_b = 0, total_b = 0;
        if (cudaGetDeviceCount(&n) != cudaSuccess) n = 0;
        if (n > 0) {
            cudaDeviceProp p;
            int dev = 0; cudaGetDevice(&dev);
            if (cudaGetDeviceProperties(&p, dev) == cudaSuccess) { sm = p.major * 10 + p.minor; nsm = p.multiProcessorCount; }
            cudaMemGetInfo(&free_b, &total_b);
        }
        s = "CUDA=" + std::to_string(n > 0 ? 1 : 0) + " GPUS=" + std::to_string(n) + " SM=" + std::to_string(sm) + " SMS=" + std::to_string(nsm) +
            " HBM_GB=" + std::to_string(total_b >> 30) + " CPU_FALLBACK=0";
    });
    return s.c_str();
}

} // extern "C"
