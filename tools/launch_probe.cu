// Measures kernel-to-kernel latency of dependent launches inside a CUDA graph on this GPU:
// plain stream order vs programmatic dependent launch.  Build: nvcc -arch=sm_100a -O3.
#include <cstdio>
#include <cuda_runtime.h>

__global__ void empty_kernel(int * p) { if (p && threadIdx.x == 9999) *p = 1; }
__global__ void pdl_kernel(int * p) {
    cudaGridDependencySynchronize();
    if (p && threadIdx.x == 9999) *p = 1;
    cudaTriggerProgrammaticLaunchCompletion();
}
__global__ void chain_kernel(const float * in, float * out, int n) {   // one dependent global round trip + block reduce
    __shared__ float s[256];
    float v = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) v += in[i];
    s[threadIdx.x] = v; __syncthreads();
    for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o]; __syncthreads(); }
    for (int i = threadIdx.x; i < n; i += blockDim.x) out[i] = in[i] * 0.5f + s[0] * 1e-9f;
}

static float run_graph(cudaGraphExec_t g, cudaStream_t st, int iters) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 5; i++) cudaGraphLaunch(g, st);
    cudaEventRecord(a, st);
    for (int i = 0; i < iters; i++) cudaGraphLaunch(g, st);
    cudaEventRecord(b, st); cudaStreamSynchronize(st);
    float ms; cudaEventElapsedTime(&ms, a, b); return ms / iters;
}

int main() {
    cudaStream_t st; cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
    const int N = 100;
    float * x; cudaMalloc(&x, 2 * 768 * 4); cudaMemset(x, 0, 2 * 768 * 4);
    for (int mode = 0; mode < 4; mode++) {
        for (int grid : {1, 148, 592}) {
            cudaGraph_t g; cudaGraphExec_t ge;
            cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
            for (int i = 0; i < N; i++) {
                if (mode == 0) empty_kernel<<<grid, 256, 0, st>>>(nullptr);
                else if (mode == 1 || mode == 3) {
                    cudaLaunchConfig_t cfg = {}; cfg.gridDim = grid; cfg.blockDim = 256; cfg.stream = st;
                    cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
                    cfg.attrs = at; cfg.numAttrs = 1;
                    if (mode == 1) cudaLaunchKernelEx(&cfg, pdl_kernel, (int *) nullptr);
                    else cudaLaunchKernelEx(&cfg, empty_kernel, (int *) nullptr);
                } else chain_kernel<<<grid, 256, 0, st>>>(x + (i & 1) * 768, x + ((i + 1) & 1) * 768, 768);
            }
            cudaError_t e = cudaStreamEndCapture(st, &g);
            if (e != cudaSuccess) { printf("capture failed mode %d: %s\n", mode, cudaGetErrorString(e)); continue; }
            e = cudaGraphInstantiate(&ge, g, 0);
            if (e != cudaSuccess) { printf("instantiate failed mode %d: %s\n", mode, cudaGetErrorString(e)); continue; }
            float ms = run_graph(ge, st, 50);
            const char * names[4] = {"empty, stream order", "empty, PDL (sync+trigger)", "1-roundtrip reduce kernel, stream order", "empty, PDL attr only"};
            printf("%-42s grid=%4d : %.2f us per kernel\n", names[mode], grid, ms * 1000.f / N);
            cudaGraphExecDestroy(ge); cudaGraphDestroy(g);
        }
    }
    return 0;
}
