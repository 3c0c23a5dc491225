// Shared host-side declarations of the B200 engine: error plumbing, ggmf types,
// device tensor descriptors.  Nothing here is visible across the C ABI.
#pragma once

#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>
#include <unordered_map>
#include <atomic>
#include <mutex>

#include <cuda_runtime.h>

#include "../../include/rwkv.h"

// ---------------------------------------------------------------------------------------
// Error words.  Same convention as the reference (rwkv_error_handling.inc:1-54): init /
// quantize record into a thread-local global word, eval paths into the context's word;
// messages go to stderr only when printing is enabled.
// ---------------------------------------------------------------------------------------
namespace rb {

extern thread_local int  g_last_error;
extern thread_local bool g_print_errors;

void global_fail(int flags, const char * file, int line, const char * fmt, ...);

#define RB_FAIL(flags, ret, ...) do { rb::global_fail((flags), __FILE__, __LINE__, __VA_ARGS__); return ret; } while (0)
#define RB_CHECK(cond, flags, ret, ...) do { if (!(cond)) RB_FAIL(flags, ret, __VA_ARGS__); } while (0)
#define RB_CUDA(expr, flags, ret) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) \
    RB_FAIL(flags, ret, "CUDA error %s at %s", cudaGetErrorString(e_), #expr); } while (0)

// ---------------------------------------------------------------------------------------
// ggmf container (rwkv_file_format.inc:5-26 type ids, :102-197 headers).
// ---------------------------------------------------------------------------------------
enum FileType : uint32_t {
    T_F32 = 0, T_F16 = 1, T_Q4_0 = 2, T_Q4_1 = 3,
    T_Q4_1_O = 4, T_Q4_2 = 5, T_Q4_3 = 6,     // removed formats: rejected like the reference does
    T_Q5_0 = 7, T_Q5_1 = 8, T_Q8_0 = 9, T_COUNT = 10
};

inline bool type_supported(uint32_t t) { return t == T_F32 || t == T_F16 || t == T_Q4_0 || t == T_Q4_1 || t == T_Q5_0 || t == T_Q5_1 || t == T_Q8_0; }
inline bool type_quantized(uint32_t t) { return t == T_Q4_0 || t == T_Q4_1 || t == T_Q5_0 || t == T_Q5_1 || t == T_Q8_0; }
// bytes per block of `type_block(t)` elements (ggml-common.h:161-208)
inline size_t type_block(uint32_t t) { return type_quantized(t) ? 32 : 1; }
inline size_t type_block_bytes(uint32_t t) {
    switch (t) { case T_F32: return 4; case T_F16: return 2; case T_Q4_0: return 18; case T_Q4_1: return 20;
                 case T_Q5_0: return 22; case T_Q5_1: return 24; case T_Q8_0: return 34; default: return 0; }
}
const char * type_name(uint32_t t);
int type_from_name(const char * s);   // -1 when unknown

struct FileTensor {
    std::string name;
    uint32_t    type = 0;
    uint32_t    n_dims = 0;
    uint64_t    ne[3] = {1, 1, 1};        // ggml order: ne[0] fastest
    const uint8_t * data = nullptr;       // into the mapping
    size_t      nbytes = 0;
    uint64_t    n_elements() const { return ne[0] * ne[1] * ne[2]; }
};

struct GgmfFile {
    uint32_t version = 0, n_vocab = 0, n_embed = 0, n_layer = 0, data_type = 0;
    std::vector<FileTensor> tensors;
    std::unordered_map<std::string, size_t> index;
    // mapping
    void * map_base = nullptr; size_t map_size = 0; int fd = -1;
    ~GgmfFile();
    const FileTensor * find(const std::string & name) const {
        auto it = index.find(name); return it == index.end() ? nullptr : &tensors[it->second];
    }
};
// Returns false (with global error flags set) on any format violation.
bool ggmf_open(const char * path, GgmfFile & out);

// ---------------------------------------------------------------------------------------
// Device-side matrix: a ggml (K, M) matrix = M rows of K weights.
//   F16 / F32: `qs` is the dense row-major matrix.
//   quantised: `qs` is the WARP-TILE layout the mat-vec kernel streams with one bulk copy per
//     tile.  A tile = 4 consecutive rows (a row group) x one chunk of up to 32 blocks
//     (1024 weights) and stores three planes back to back:
//        nibbles  [4][cnt] x 16 B (Q8_0: 32 B)   bytes exactly as in the ggml block
//        qh       [4][cnt] x u32  (Q5 only)       5th bits, bit-permuted (see codec.cu)
//        scales   [4][cnt] x half (_0) | half2 d,m (_1)
//     cnt = 32 except for the last chunk of a row (nb % 32).  Tiles of a row group are
//     consecutive (chunk 0, 1, ...); row groups are `rg_stride` bytes apart (4*nb*block
//     bytes rounded up to 16).  Rows M..4*ceil(M/4)-1 are zero blocks.  Same bytes per
//     weight as the file format.
// ---------------------------------------------------------------------------------------
constexpr int TILE_ROWS = 4, TILE_BLOCKS = 32;

struct DevMat {
    uint32_t type = T_F32;
    uint32_t M = 0, K = 0;
    const void * qs = nullptr;
    uint32_t rg_stride = 0;                // bytes between row groups (quantised only)
    size_t bytes = 0;                      // algorithmic bytes of the matrix (file bytes)
    // Second copy for the tensor-core MATMUL (quantised only; null when not built): chunks of
    // 128 rows x 64 weights (2 blocks), chunk (mt, kb) at (mt * tc_kblocks + kb) * 256 * block
    // bytes.  Inside a chunk, entry e = blk * 128 + row and three planes like the warp tile:
    // nibbles [256] x 16 B (Q8_0 32 B) | qh [256] x u32 (Q5, permuted) | scales [256] x half or half2.
    // Rows past M and blocks past K/32 are zero blocks.
    const void * tc = nullptr;
    uint32_t tc_kblocks = 0;               // ceil(K / 64)
};
constexpr int TC_ROWS = 128, TC_KB = 64;
inline size_t tc_chunk_bytes(uint32_t t) { return (size_t) 2 * TC_ROWS * type_block_bytes(t); }
inline size_t tc_bytes(uint32_t t, size_t M, size_t K) { return ((M + TC_ROWS - 1) / TC_ROWS) * ((K + TC_KB - 1) / TC_KB) * tc_chunk_bytes(t); }

// SM count of the current device (cudaDevAttrMultiProcessorCount, cached per device): grids and
// split rules are sized against it instead of a literal 148.
inline int device_sm_count() {
    static std::atomic<int> cache[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    int n = cache[dev].load(std::memory_order_relaxed);
    if (n > 0) return n;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cache[dev].store(n, std::memory_order_relaxed);
    return n;
}

#if defined(__CUDACC__)
#define RB_HD __host__ __device__
#else
#define RB_HD
#endif
RB_HD inline uint32_t tile_qb(uint32_t t) { return t == T_Q8_0 ? 32u : 16u; }
RB_HD inline uint32_t tile_hb(uint32_t t) { return (t == T_Q5_0 || t == T_Q5_1) ? 4u : 0u; }
RB_HD inline uint32_t tile_db(uint32_t t) { return (t == T_Q4_1 || t == T_Q5_1) ? 4u : 2u; }
inline size_t tiled_rg_stride(uint32_t t, size_t K) { return ((size_t) TILE_ROWS * (K / 32) * type_block_bytes(t) + 15) & ~(size_t) 15; }
inline size_t tiled_bytes(uint32_t t, size_t M, size_t K) { return ((M + TILE_ROWS - 1) / TILE_ROWS) * tiled_rg_stride(t, K) + 16; }

} // namespace rb
