// Programmatic dependent launch (PDL) plumbing.
//
// Every kernel of the forward pass is launched with the programmatic-stream-serialization
// attribute and follows one protocol:
//   pdl_trigger()  first thing: the NEXT kernel in the stream may start its prologue now;
//   ... prologue that touches only constants (weights, norm vectors) and the read-only
//       input state: bulk-copy / prefetch them ...
//   pdl_wait()     before the first access to anything a preceding kernel writes, and before
//                  the first global write.
// RULE: a pointer to anything a preceding kernel writes must not be `const T * __restrict__` (nor read with __ldg): nvcc
// marks such loads invariant (LDG.E.CONSTANT) and may hoist them above griddepcontrol.wait -- seen in ln_act_reg_kernel<8>,
// which read the residual row before the last ffn mat-vec had added to it (caught by tests/test_gpu_fast_paths.py).
// So a mat-vec streams its first weight tiles while the producer of its activations is
// still running, and the tail of one kernel overlaps the fill of the next.  pdl_wait()
// returns when the preceding kernel has completed and flushed; since every kernel waits
// before it finishes, completion is transitive along the stream.  Captured into the CUDA
// graphs as programmatic edges.  RWKV_B200_NO_PDL=1 falls back to plain stream order.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdlib>
#include <utility>

namespace rb {

__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait()    { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// Debug aid: forward.cpp tags the launches of a pass with a site bit (1 att input, 2 wkv, 4 att output, 8 ffn key / receptance,
// 16 ffn value, 32 ln_out, 64 head); RWKV_B200_PDL_OFF_SITES=<mask> launches those sites in plain stream order.
inline int & pdl_site() { static thread_local int s = 0; return s; }
inline bool pdl_enabled() {
    static const bool on = [] { const char * e = getenv("RWKV_B200_NO_PDL"); return !(e && e[0] == '1'); }();
    static const int off_sites = [] { const char * e = getenv("RWKV_B200_PDL_OFF_SITES"); return e ? atoi(e) : 0; }();
    return on && !(off_sites & pdl_site());
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args &&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

// Same, with a thread-block cluster of `cluster_x` CTAs along x (0 or 1: no cluster attribute).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                      int cluster_x, Args &&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[2];
    unsigned n = 0;
    if (pdl_enabled()) {
        at[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[n].val.programmaticStreamSerializationAllowed = 1;
        n++;
    }
    if (cluster_x > 1) {
        at[n].id = cudaLaunchAttributeClusterDimension;
        at[n].val.clusterDim.x = (unsigned) cluster_x; at[n].val.clusterDim.y = 1; at[n].val.clusterDim.z = 1;
        n++;
    }
    cfg.attrs = at; cfg.numAttrs = n;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

} // namespace rb
