// Model (weights in HBM, shared by clones) and Context (stream, workspace, launch graphs).
#pragma once
#include "kernels.h"
#include "tp.h"
#include <memory>

namespace rb {

struct LayerW {
    const float * ln1_w = nullptr, * ln1_b = nullptr, * ln2_w = nullptr, * ln2_b = nullptr;
    // time-mix interpolation vectors.  v4/v5: [k, v, r, g]; v6: [x, w, k, v, r, g]; v7: x_rwkvag rows [r, w, k, v, a, g]
    const float * att_mu[6] = {};
    const float * ffn_mu[2] = {};            // v4-v6: [k, r]; v7: [x_k]
    const float * time_first = nullptr;      // v4 [D]; v5.1 [H]; v5.2/v6 time_faaaa [D]
    const float * time_decay = nullptr;      // v4 [D]; v5.1 [H]; v5.2/v6 [D]
    const float * lnx_w = nullptr, * lnx_b = nullptr;
    DevMat att_key, att_value, att_receptance, att_output, att_gate;
    DevMat ffn_key, ffn_value, ffn_receptance;
    // v6
    DevMat maa_w1, maa_w2[5], decay_w1, decay_w2;
    // v7
    DevMat w1, w2, a1, a2, v1, v2, g1, g2;
    const float * w0 = nullptr, * a0 = nullptr, * v0 = nullptr, * k_k = nullptr, * k_a = nullptr, * r_k = nullptr;
};

struct Model {
    std::atomic<int> refcount{1};
    int device = 0;
    TpInfo tp;                               // tensor-parallel rank / world / communicator (world == 1: none)
    uint32_t n_vocab = 0, n_embed = 0, n_layer = 0, data_type = 0;
    int arch_major = 4, arch_minor = 0;
    int head_count = 0, head_size = 0;
    int ffn_dim = 0;                         // widest mat-vec input in the model
    int max_lora = 0;                        // widest low-rank intermediate (v6/v7)
    DevMat emb, head;
    const float * ln0_w = nullptr, * ln0_b = nullptr, * ln_out_w = nullptr, * ln_out_b = nullptr;
    std::vector<LayerW> layers;
    std::vector<void *> allocations;         // cudaMalloc'ed blocks owned by the model
    size_t weight_bytes = 0;                 // algorithmic bytes of all tensors (as in the file)
    size_t state_len() const {
        return arch_major >= 5 ? (size_t) n_embed * (2 + head_size) * n_layer : (size_t) n_embed * 5 * n_layer;
    }
    size_t layer_state_len() const { return state_len() / n_layer; }
    ~Model();
};

// Returns nullptr with the thread-local error word set on failure.
Model * model_load(const char * path, const TpInfo & tp = TpInfo());

// Activation workspace for up to `rows` rows.
struct Workspace {
    int rows = 0;
    std::vector<void *> allocations;
    float * x = nullptr, * x0 = nullptr, * sx = nullptr;
    Act   mix[6];                 // mat-vec inputs of width D
    Act   y;                      // WKV output rows (input of att.output)
    Act   ffn_k;                  // relu(k)^2 rows, width F
    Act   lora[4];                // low-rank intermediates (v6: [5E tanh, decay tanh]; v7: w,a,g,v)
    Act   head_in;                // ln_out rows
    float * r = nullptr, * k = nullptr, * v = nullptr, * g = nullptr, * w = nullptr, * a = nullptr, * vgate = nullptr;
    float * ffn_r = nullptr, * m5 = nullptr /* [5][rows][D] */, * v_first = nullptr;
    float * yraw = nullptr;       // [rows][D] raw WKV scan outputs (sequence kernels)
    MatmulScratch matmul_scratch;     // K-split partial tiles of the tensor-core MATMUL
    float * wkv_ds = nullptr, * wkv_pc = nullptr;   // chunk-parallel WKV5/6 scratch (head size 64, rows on the tensor-core path)
    float * tp_send = nullptr, * tp_recv = nullptr;   // tensor parallel: [rows][2*D] all-reduce buffers of the ffn ([partial Wv k | r shards])
    float * tp_att = nullptr;                         // [rows][D] partial att output
    float * x_alt = nullptr;                          // fused tensor parallelism: second residual row (a gather reads one and writes the other)
    float * logits = nullptr;     // [logit_rows][V]
    float * logits_part = nullptr; // tensor parallel: this rank's vocabulary slice
    int   logit_rows = 0;
    int * tokens = nullptr;
    ~Workspace();
};

struct rwkv_context_impl;

// What one forward pass covers.
struct ForwardArgs {
    int   R = 1;                  // rows
    int   batch = 0;              // 0: R consecutive tokens of one stream, 1: R independent streams
    const float * state_in = nullptr; float * state_out = nullptr;   // device; stream stride = state_len
    bool  logits = true;          // sequence: last row only; batch: every row
};

// Enqueue the whole pass on `st`.  Tokens must already be in ws.tokens.
cudaError_t forward(const Model & m, Workspace & ws, const ForwardArgs & fa, cudaStream_t st);
bool workspace_reserve(const Model & m, Workspace & ws, int rows, int logit_rows);

} // namespace rb
