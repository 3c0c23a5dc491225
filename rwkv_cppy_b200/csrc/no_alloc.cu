// rwkv_b200_matrix_mul_no_alloc: the quantised mat-vec for callers that cannot allocate (code under
// CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// The mat-vec kernel reads its activations as q8 blocks (kernels.h Act: int8 values + per-block d, s,
// isum) that a quantiser launch writes first: SCRATCH memory the existing entry point
// (rwkv_b200_matrix_mul, api_ext.cpp) cudaMallocs per matrix handle.  Here that scratch is a file-scope
// __device__ array of a fixed size, resident on every GPU that loads the library until the process
// exits:
//     NA_MAX_ELEMS = 262 144 activation values per call (n_vec * n_cols), 352 KiB in total
//     [ q  int8  NA_MAX_ELEMS | d f32 NA_MAX_ELEMS/32 | s f32 NA_MAX_ELEMS/32 | isum i32 NA_MAX_ELEMS/32 ]
// Host code cannot use the array's name as an address (it is a host-side stand-in): the device address
// is fetched with cudaGetSymbolAddress once per device, with that device current, and handed to the
// kernels.  Every call shares the array and it keeps the previous call's contents; nothing in it has to
// start at zero because the quantiser launch rewrites every byte the mat-vec launch of the same call
// reads.  Calls that overlap on the GPU (two host threads, two streams) race on it: serialise them.
#include "matrix_handle.h"
#include "device_utils.cuh"
#include "../../include/rwkv_b200.h"

#include <mutex>

using namespace rb;

namespace {

constexpr size_t NA_MAX_ELEMS = RWKV_B200_NO_ALLOC_MAX_ELEMS;
constexpr size_t NA_BLOCKS = NA_MAX_ELEMS / 32;
constexpr size_t NA_BYTES = NA_MAX_ELEMS + 3 * NA_BLOCKS * 4;
static_assert(NA_MAX_ELEMS % 256 == 0, "plane alignment");

static __device__ __align__(256) unsigned char g_no_alloc_scratch[NA_BYTES];

// The activation quantiser writing straight into the static array (device code may name it directly; only
// host code needs cudaGetSymbolAddress).  Same q8 blocks as quantize_act_kernel / act_store_block
// (device_utils.cuh q8_block: ggml's x86 quantize_row_q8_0 / q8_1), but x is only read.
__global__ void __launch_bounds__(256) quantize_act_no_alloc_kernel(const float * __restrict__ x, int K) {
    int8_t * q = reinterpret_cast<int8_t *>(g_no_alloc_scratch);
    float * d = reinterpret_cast<float *>(g_no_alloc_scratch + NA_MAX_ELEMS);
    float * s = d + NA_BLOCKS;
    int * isum = reinterpret_cast<int *>(s + NA_BLOCKS);
    const int r = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int e0 = warp * 32; e0 < K; e0 += nw * 32) {        // K % 32 == 0: every warp iteration is one whole block
        const int e = e0 + lane;
        float d16, s16; int sum;
        const int qv = q8_block(x[(size_t) r * K + e], true, d16, s16, sum);
        q[(size_t) r * K + e] = (int8_t) qv;
        if (lane == 0) {
            const size_t b = (size_t) r * (K >> 5) + (e0 >> 5);
            d[b] = d16; s[b] = s16; isum[b] = sum;
        }
    }
}

constexpr int NA_MAX_DEVICES = 64;
std::mutex g_na_mutex;
unsigned char * g_na_ptr[NA_MAX_DEVICES] = {};      // device address of the array, per device

// Device address of the scratch array on `device` (which must be current); nullptr + error word on failure.
unsigned char * scratch_address(int device) {
    if (device < 0 || device >= NA_MAX_DEVICES) return nullptr;
    std::lock_guard<std::mutex> lock(g_na_mutex);
    if (g_na_ptr[device]) return g_na_ptr[device];
    void * p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_no_alloc_scratch);
    if (e != cudaSuccess || p == nullptr) {
        global_fail(RWKV_ERROR_CTX | RWKV_ERROR_ALLOC, __FILE__, __LINE__, "cudaGetSymbolAddress(scratch) failed on device %d: %s", device, cudaGetErrorString(e));
        return nullptr;
    }
    g_na_ptr[device] = (unsigned char *) p;
    return g_na_ptr[device];
}

bool is_device_pointer(const void * p, int device) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return (a.type == cudaMemoryTypeDevice && a.device == device) || a.type == cudaMemoryTypeManaged;
}

} // namespace

extern "C" {

size_t rwkv_b200_no_alloc_max_elements(void) { return NA_MAX_ELEMS; }

bool rwkv_b200_matrix_mul_no_alloc(struct rwkv_b200_matrix * m, const float * x_dev, size_t n_vec, float * y_dev) {
    g_last_error = RWKV_ERROR_NONE;
    RB_CHECK(m && x_dev && y_dev && n_vec > 0, RWKV_ERROR_ARGS, false, "Bad arguments");
    const DevMat & W = m->copies[0];
    RB_CHECK(type_quantized(W.type), RWKV_ERROR_ARGS | RWKV_ERROR_DATA_TYPE, false,
             "rwkv_b200_matrix_mul_no_alloc covers the quantised mat-vec (the kernel that reads q8 scratch); matrix type is %s", type_name(W.type));
    RB_CHECK(n_vec <= NA_MAX_ELEMS / W.K && n_vec * (size_t) W.K <= NA_MAX_ELEMS, RWKV_ERROR_ARGS | RWKV_ERROR_SHAPE, false,
             "n_vec * n_cols = %zu x %u exceeds the static scratch of %zu activation values", n_vec, W.K, NA_MAX_ELEMS);
    RB_CUDA(cudaSetDevice(m->device), RWKV_ERROR_CTX, false);
    RB_CHECK(is_device_pointer(x_dev, m->device) && is_device_pointer(y_dev, m->device), RWKV_ERROR_ARGS, false,
             "x and y must be device memory of GPU %d (this entry point stages nothing)", m->device);
    unsigned char * base = scratch_address(m->device);
    if (!base) return false;
    quantize_act_no_alloc_kernel<<<(unsigned) n_vec, 256, 0, m->stream>>>(x_dev, (int) W.K);
    RB_CUDA(cudaGetLastError(), RWKV_ERROR_GRAPH, false);
    Act x;                                                  // the same planes, as the mat-vec launcher wants them: through the fetched address
    x.K = (int) W.K;
    x.f = const_cast<float *>(x_dev);                       // (not read by the quantised mat-vec)
    x.q = reinterpret_cast<int8_t *>(base);
    x.d = reinterpret_cast<float *>(base + NA_MAX_ELEMS);
    x.s = x.d + NA_BLOCKS;
    x.isum = reinterpret_cast<int *>(x.s + NA_BLOCKS);
    MatvecBatch b; b.T = (int) n_vec; b.n_jobs = 1;
    b.job[0].W = W; b.job[0].x = x; b.job[0].out = y_dev; b.job[0].ldo = (int) W.M; b.job[0].epi = EPI_STORE;
    RB_CUDA(launch_matvec(b, m->stream), RWKV_ERROR_GRAPH, false);
    RB_CUDA(cudaStreamSynchronize(m->stream), RWKV_ERROR_GRAPH, false);
    return true;
}

} // extern "C"
