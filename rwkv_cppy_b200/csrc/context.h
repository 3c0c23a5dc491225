// The opaque context behind `struct rwkv_context *` (shared by api.cpp and api_ext.cpp).
#pragma once
#include "engine.h"

constexpr int SEQ_MAX_ROWS = 2048;      // rows per launch sequence in sequence mode (longer inputs are tiled; state stays in HBM)

struct rwkv_context {
    rb::Model * model = nullptr;
    int     last_error = RWKV_ERROR_NONE;
    bool    print_errors = true;
    uint32_t n_threads = 0;
    cudaStream_t stream = nullptr;
    rb::Workspace ws1;                       // single-row workspace the decode graphs are captured over
    rb::Workspace wsn;                       // sequence-mode workspace (grows)
    float * d_state[2] = {nullptr, nullptr};
    float * d_init_state = nullptr;
    int *   h_tokens = nullptr; size_t h_tokens_cap = 0;     // pinned
    cudaGraphExec_t graph[2] = {nullptr, nullptr};            // [with logits]
};


namespace rbapi {
extern thread_local int t_device;
int  pick_device();
void ctx_fail(rwkv_context * ctx, int flags, const char * file, int line, const char * fmt, ...);
void host_init_state(const rb::Model & m, float * s);
void model_release(rb::Model * m);
bool ctx_setup(rwkv_context * ctx);
void ctx_destroy(rwkv_context * ctx);
bool capture_graph(rwkv_context * ctx, cudaStream_t st, const rb::Model & m, rb::Workspace & ws, const rb::ForwardArgs & fa, cudaGraphExec_t * out);
}
#define CTX_CHECK(ctx, cond, flags, ret, ...) do { if (!(cond)) { rbapi::ctx_fail((ctx), (flags), __FILE__, __LINE__, __VA_ARGS__); return ret; } } while (0)
#define CTX_CUDA(ctx, expr, ret) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { \
    rbapi::ctx_fail((ctx), RWKV_ERROR_GRAPH, __FILE__, __LINE__, "CUDA error %s at %s", cudaGetErrorString(e_), #expr); return ret; } } while (0)
