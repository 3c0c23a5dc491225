// Optional device-side timeline (RWKV_B200_TRACE builds only; tools/decode_trace.py): the first and the last CTA of a launch
// stamp %globaltimer at a few points and append one record of 12 words.  Every translation unit that includes this header keeps a
// buffer of its own (no relocatable device code) and exports a reader with TRACE_EXPORT; the records carry absolute nanoseconds, so
// the host merges them by time.  Record: [0..9] stamps, [10] kind, [11] grid size | (linear CTA index << 24).
#pragma once
#ifdef RWKV_B200_TRACE
#include <cuda_runtime.h>
namespace rb {
constexpr unsigned TRACE_RECORDS = 8192;
static __device__ unsigned long long g_trace[TRACE_RECORDS * 12];
static __device__ unsigned int g_trace_n;
__device__ __forceinline__ unsigned long long gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ bool trace_first() { return blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0; }
__device__ __forceinline__ bool trace_last()  { return blockIdx.x == gridDim.x - 1 && blockIdx.y == gridDim.y - 1 && blockIdx.z == gridDim.z - 1; }
static inline int trace_read(unsigned long long * out, int max_records) {
    unsigned int n = 0;
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(&n, g_trace_n, sizeof(n));
    const int have = (int) (n < TRACE_RECORDS ? n : TRACE_RECORDS);
    const int k = have < max_records ? have : max_records;
    cudaMemcpyFromSymbol(out, g_trace, (size_t) k * 12 * sizeof(unsigned long long));
    const unsigned int zero = 0;
    cudaMemcpyToSymbol(g_trace_n, &zero, sizeof(zero));
    return k;
}
} // namespace rb
#define TRACE_DECL_IF(cond) unsigned long long tr_[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0}; const bool tr_on_ = threadIdx.x == 0 && (cond)
#define TRACE_DECL TRACE_DECL_IF(rb::trace_first() || rb::trace_last())
#define TRACE(i) do { if (tr_on_) tr_[i] = rb::gtimer(); } while (0)
#define TRACE_FLUSH(kind, grid) do { if (tr_on_) { \
    const unsigned int n_ = atomicAdd(&rb::g_trace_n, 1u) & (rb::TRACE_RECORDS - 1); \
    for (int i_ = 0; i_ < 10; i_++) rb::g_trace[n_ * 12 + i_] = tr_[i_]; rb::g_trace[n_ * 12 + 10] = (unsigned long long) (kind); \
    rb::g_trace[n_ * 12 + 11] = (unsigned long long) (grid) | ((unsigned long long) ((blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) << 24); } } while (0)
#define TRACE_ARG , unsigned long long * tr_, bool tr_on_
#define TRACE_PASS , tr_, tr_on_
#define TRACE_EXPORT(name) extern "C" __attribute__((visibility("default"))) int name(unsigned long long * out, int max_records) { return rb::trace_read(out, max_records); }
#else
#define TRACE_DECL
#define TRACE_DECL_IF(cond)
#define TRACE(i)
#define TRACE_FLUSH(kind, grid)
#define TRACE_ARG
#define TRACE_PASS
#define TRACE_EXPORT(name)
#endif
