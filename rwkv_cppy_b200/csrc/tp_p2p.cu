// One-shot all-reduce / all-gather over NVLink peer memory for the small vectors of
// tensor-parallel decode ([D] or [2D] floats twice per layer, the logits shards once per token).
// NCCL needs ~20 us for these latency-bound collectives; here a rank pushes its vector straight
// into an inbox in every peer's HBM (P2P stores), publishes a per-CTA epoch flag with release
// semantics and polls its own flags, so the cost is one NVLink round trip.  Each CTA owns a
// fixed slice of the vector, which removes any grid-wide synchronisation: CTA c of rank a only
// waits for CTA c of the other ranks.  The sum runs over the slots in rank order on every rank,
// so all ranks hold bit-identical results (the replicated layer norms stay in lockstep).
// Two inbox generations alternate by epoch parity: a peer can be at most one call ahead.
#include "tp.h"
#include "launch.cuh"
#include "trace.cuh"
#include "tp_ll.cuh"
#include <cstdio>
#include <cstring>

namespace rb {

struct P2pArgs {
    uint8_t * peer[TP_MAX_WORLD];
    int rank, world;
    const float * send; float * recv;
    unsigned n;            // floats contributed by this rank
    int gather;            // 0: recv[i] = sum_r slot_r[i];  1: recv[r * n + i] = slot_r[i]
    unsigned int * err;    // host-mapped error word
};

__device__ __forceinline__ uint32_t * region_flags(uint8_t * base) { return reinterpret_cast<uint32_t *>(base); }
__device__ __forceinline__ uint32_t * region_epoch(uint8_t * base) { return reinterpret_cast<uint32_t *>(base) + TP_MAX_WORLD * TP_CTAS; }
__device__ __forceinline__ uint32_t * region_error(uint8_t * base) { return reinterpret_cast<uint32_t *>(base) + TP_MAX_WORLD * TP_CTAS + TP_CTAS; }
__device__ __forceinline__ float * region_inbox(uint8_t * base, int world, int parity, int src) {
    return reinterpret_cast<float *>(base + 4096) + ((size_t) parity * world + src) * TP_NMAX;
}

__global__ void __launch_bounds__(256) tp_p2p_kernel(const __grid_constant__ P2pArgs a) {
    TRACE_DECL; TRACE(0);
    pdl_trigger();
    const int cta = blockIdx.x, tid = threadIdx.x;
    uint8_t * mine = a.peer[a.rank];
    // this CTA's slice, in float4 units
    const unsigned n4 = (a.n + 3) / 4, per = (n4 + TP_CTAS - 1) / TP_CTAS;
    const unsigned lo = min(n4, per * cta), hi = min(n4, lo + per);
    TRACE(1);
    pdl_wait();                                              // `send` comes from the preceding kernel; the previous collective
                                                             // (a few launches back, possibly still running under PDL) owns the epoch until it exits
    TRACE(2);
    const uint32_t e = region_epoch(mine)[cta] + 1;          // this CTA's call counter (only this CTA writes it)
    const int par = e & 1;
    for (unsigned i = lo + tid; i < hi; i += 256) {
        float4 v;
        if (4 * i + 3 < a.n) v = reinterpret_cast<const float4 *>(a.send)[i];
        else { v.x = 4 * i < a.n ? a.send[4 * i] : 0.0f; v.y = 4 * i + 1 < a.n ? a.send[4 * i + 1] : 0.0f; v.z = 4 * i + 2 < a.n ? a.send[4 * i + 2] : 0.0f; v.w = 0.0f; }
        for (int p = 0; p < a.world; p++) reinterpret_cast<float4 *>(region_inbox(a.peer[p], a.world, par, a.rank))[i] = v;
    }
    __syncthreads();
    TRACE(3);
    if (tid < a.world) {
        __threadfence_system();
        // publish: flags[src = me][cta] on peer `tid`
        uint32_t * f = region_flags(a.peer[tid]) + a.rank * TP_CTAS + cta;
        asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(f), "r"(e) : "memory");
        // wait for source rank `tid` to have delivered epoch e into my inbox
        const uint32_t * w = region_flags(mine) + tid * TP_CTAS + cta;
        uint32_t seen = 0;
        const long long t0 = clock64();
        for (;;) {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(w) : "memory");
            if ((int) (seen - e) >= 0) break;
            if (clock64() - t0 > 4000000000LL) { *region_error(mine) = 1; if (a.err) *a.err = 1u; break; }      // ~2 s: a peer is gone; never hang the GPU
        }
    }
    __syncthreads();
    TRACE(4);
    for (unsigned i = lo + tid; i < hi; i += 256) {
        if (a.gather) {
            for (int r = 0; r < a.world; r++) {
                const float4 v = reinterpret_cast<const float4 *>(region_inbox(mine, a.world, par, r))[i];
                float * o = a.recv + (size_t) r * a.n + 4 * i;
                if (4 * i + 3 < a.n && (((size_t) r * a.n) & 3) == 0) *reinterpret_cast<float4 *>(o) = v;
                else { if (4 * i < a.n) o[0] = v.x; if (4 * i + 1 < a.n) o[1] = v.y; if (4 * i + 2 < a.n) o[2] = v.z; if (4 * i + 3 < a.n) o[3] = v.w; }
            }
        } else {
            float4 s = reinterpret_cast<const float4 *>(region_inbox(mine, a.world, par, 0))[i];
            for (int r = 1; r < a.world; r++) {
                const float4 v = reinterpret_cast<const float4 *>(region_inbox(mine, a.world, par, r))[i];
                s.x = __fadd_rn(s.x, v.x); s.y = __fadd_rn(s.y, v.y); s.z = __fadd_rn(s.z, v.z); s.w = __fadd_rn(s.w, v.w);
            }
            if (4 * i + 3 < a.n) reinterpret_cast<float4 *>(a.recv)[i] = s;
            else { if (4 * i < a.n) a.recv[4 * i] = s.x; if (4 * i + 1 < a.n) a.recv[4 * i + 1] = s.y; if (4 * i + 2 < a.n) a.recv[4 * i + 2] = s.z; }
        }
    }
    if (tid == 0) region_epoch(mine)[cta] = e;
    TRACE(5);
    TRACE_FLUSH(300 + a.gather, gridDim.x);
}

// Low-latency all-reduce: every entry travels as one 8-byte {value, call counter} store, so the receiver that sees the counter of
// this call also holds the value -- no __threadfence_system(), no flag word, no second round trip (the flag protocol above measured
// 9.5 us per call on 2 GPUs, 7.8 us of it between a rank's own stores and the peers' flags: tools/decode_trace.py).  One thread per
// element: it stores its value into slot `rank` of every peer (a warp writes 256 contiguous bytes per peer), then polls the world
// slots of its own inbox and adds them in rank order.  Counters are per CTA (a CTA always owns the same elements, whatever the
// vector length of the call), generations alternate with the counter's parity: a peer can be at most one call ahead per element.
// Only single-row vectors take this kernel (tp.h tp_allreduce): then the plain form (entries [0, D), thread e) and the channel-mix form
// (entries e and D + e, thread e) agree on which CTA's counter validates every entry.
struct LlArArgs {
    uint8_t * peer[TP_MAX_WORLD];      // regions of all ranks
    int rank, world;
    const float * send; float * recv; float * combine_x;
    unsigned n, half;                  // elements of the vector; half = n / 2 when combine_x is set
    size_t ll2_off;                    // byte offset of the area inside a region
    unsigned int * err;
};

__global__ void __launch_bounds__(256) tp_ll_allreduce_kernel(const __grid_constant__ LlArArgs a) {
    TRACE_DECL; TRACE(0);
    pdl_trigger();
    const unsigned e = blockIdx.x * 256 + threadIdx.x;
    const unsigned n_thr = a.combine_x ? a.half : a.n;             // a thread owns element e (and e + half under combine)
    uint32_t * counter = reinterpret_cast<uint32_t *>(a.peer[a.rank]) + TP_LL2_COUNTER_WORD + blockIdx.x;
    TRACE(1);
    pdl_wait();                                                      // `send` comes from the preceding kernel; the previous call of this kernel has completed
    TRACE(2);
    const uint32_t call = *reinterpret_cast<volatile uint32_t *>(counter) + 1;
    const size_t gen = (size_t) (call & 1) * a.world * TP_LL_SLOT;
    if (e < n_thr) {
        const float v0 = a.send[e], v1 = a.combine_x ? a.send[a.half + e] : 0.0f;
        for (int p = 0; p < a.world; p++) {
            uint2 * slot = reinterpret_cast<uint2 *>(a.peer[p] + a.ll2_off) + gen + (size_t) a.rank * TP_LL_SLOT;
            asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" :: "l"(slot + e), "r"(__float_as_uint(v0)), "r"(call) : "memory");
            if (a.combine_x) asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" :: "l"(slot + a.half + e), "r"(__float_as_uint(v1)), "r"(call) : "memory");
        }
        TRACE(3);
        const uint2 * inbox = reinterpret_cast<const uint2 *>(a.peer[a.rank] + a.ll2_off) + gen;
        uint2 w0[TP_MAX_WORLD], w1[TP_MAX_WORLD];
#pragma unroll
        for (int r = 0; r < TP_MAX_WORLD; r++) if (r < a.world) {
            w0[r] = ll_load(inbox + (size_t) r * TP_LL_SLOT + e);
            if (a.combine_x) w1[r] = ll_load(inbox + (size_t) r * TP_LL_SLOT + a.half + e);
        }
        float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
        for (int r = 0; r < TP_MAX_WORLD; r++) if (r < a.world) {
            const float p0 = ll_wait(inbox + (size_t) r * TP_LL_SLOT + e, w0[r], call, a.err);
            s0 = r == 0 ? p0 : __fadd_rn(s0, p0);
            if (a.combine_x) {
                const float p1 = ll_wait(inbox + (size_t) r * TP_LL_SLOT + a.half + e, w1[r], call, a.err);
                s1 = r == 0 ? p1 : __fadd_rn(s1, p1);
            }
        }
        TRACE(4);
        if (a.combine_x) a.combine_x[e] = __fadd_rn(a.combine_x[e], __fmul_rn(s1, s0));
        else a.recv[e] = s0;
    }
    __syncthreads();                                                 // every thread of the CTA has read the counter
    if (threadIdx.x == 0) *counter = call;
    TRACE(5);
    TRACE_FLUSH(310 + (a.combine_x ? 1 : 0), gridDim.x);
}

bool tp_ll_allreduce_usable(const TpInfo & tp, size_t n) {
    static const bool off = [] { const char * e = getenv("RWKV_B200_TP_NO_LL"); return e && e[0] == '1'; }();     // back to the flag protocol (timing comparisons)
    return !off && tp.p2p.connected && n > 0 && n <= TP_LL_SLOT && tp.world <= TP_MAX_WORLD;
}

cudaError_t tp_ll_allreduce(const TpInfo & tp, const float * send, float * recv, size_t n, float * combine_x, cudaStream_t st) {
    if (!tp_ll_allreduce_usable(tp, n) || (combine_x && (n & 1))) return cudaErrorInvalidValue;
    LlArArgs a;
    for (int r = 0; r < TP_MAX_WORLD; r++) a.peer[r] = r < tp.world ? tp.p2p.peer[r] : nullptr;
    a.rank = tp.rank; a.world = tp.world; a.send = send; a.recv = recv; a.combine_x = combine_x;
    a.n = (unsigned) n; a.half = (unsigned) (n / 2); a.ll2_off = tp_region_ll2_offset(tp.world); a.err = tp.p2p.d_err;
    const unsigned n_thr = combine_x ? a.half : a.n;
    return launch_pdl(tp_ll_allreduce_kernel, dim3((n_thr + 255) / 256), dim3(256), 0, st, a);
}

// Fused tensor parallelism: the residual row of a site materialised outside a mat-vec prologue (the last
// ffn site of a pass, in front of ln_out): same arithmetic as the prologue gather (tp_ll.cuh).
__global__ void __launch_bounds__(256) tp_gather_row_kernel(const __grid_constant__ TpFused tp, const __grid_constant__ TpWait w, const float * base) {
    pdl_trigger();
    pdl_wait();
    const int e = blockIdx.x * 256 + threadIdx.x;
    if (e >= w.n) return;
    w.x_out[e] = tp_gather_element(tp, w, tp_inbox(tp, w.site), tp_flag(tp, w.site), base[e], e);
}

cudaError_t launch_tp_gather_row(const TpFused & tp, const TpWait & w, const float * base, cudaStream_t st) {
    if (!tp.ll[0] || w.site < 0 || !w.x_out || w.n <= 0) return cudaErrorInvalidValue;
    return launch_pdl(tp_gather_row_kernel, dim3((w.n + 255) / 256), dim3(256), 0, st, tp, w, base);
}

TpFused tp_fused_desc(const TpInfo & tp) {
    TpFused f;
    if (!tp.p2p.connected || tp.world < 2 || tp.world > TP_LL_WORLD) return f;
    for (int r = 0; r < tp.world; r++) f.ll[r] = tp.p2p.peer[r] + tp_region_ll_offset(tp.world);
    f.rank = tp.rank; f.world = tp.world;
    f.step = reinterpret_cast<const unsigned int *>(tp.p2p.base) + TP_STEP_WORD;
    f.err = tp.p2p.d_err;
    f.slot_entries = (unsigned) TP_LL_SLOT;
    return f;
}

bool tp_error_pending(const TpInfo & tp) {
    if (!tp.p2p.h_err) return false;
    volatile unsigned int * w = tp.p2p.h_err;
    const unsigned int v = *w;
    if (v) *w = 0u;
    return v != 0;
}

bool tp_p2p_alloc(TpInfo & tp, void * handle_out_64) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    if (tp.world < 2 || tp.world > TP_MAX_WORLD) return false;
    if (!tp.p2p.base) {
        if (cudaMalloc(&tp.p2p.base, tp_region_bytes(tp.world)) != cudaSuccess) { tp.p2p.base = nullptr; return false; }
        if (cudaMemset(tp.p2p.base, 0, tp_region_bytes(tp.world)) != cudaSuccess) return false;
        if (cudaDeviceSynchronize() != cudaSuccess) return false;
        if (cudaHostAlloc((void **) &tp.p2p.h_err, sizeof(unsigned int), cudaHostAllocMapped) != cudaSuccess) { tp.p2p.h_err = nullptr; return false; }
        *tp.p2p.h_err = 0u;
        if (cudaHostGetDevicePointer((void **) &tp.p2p.d_err, tp.p2p.h_err, 0) != cudaSuccess) return false;
    }
    cudaIpcMemHandle_t h;
    if (cudaIpcGetMemHandle(&h, tp.p2p.base) != cudaSuccess) return false;
    memcpy(handle_out_64, &h, 64);
    return true;
}

bool tp_p2p_connect(TpInfo & tp, const void * handles) {
    if (!tp.p2p.base) return false;
    for (int r = 0; r < tp.world; r++) {
        if (r == tp.rank) { tp.p2p.peer[r] = (uint8_t *) tp.p2p.base; continue; }
        cudaIpcMemHandle_t h;
        memcpy(&h, (const uint8_t *) handles + (size_t) r * 64, 64);
        void * p = nullptr;
        const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) { fprintf(stderr, "cudaIpcOpenMemHandle(rank %d) failed: %s\n", r, cudaGetErrorString(e)); return false; }
        tp.p2p.peer[r] = (uint8_t *) p;
    }
    tp.p2p.connected = true;
    return true;
}

void tp_p2p_destroy(TpInfo & tp) {
    for (int r = 0; r < tp.world; r++)
        if (r != tp.rank && tp.p2p.peer[r]) cudaIpcCloseMemHandle(tp.p2p.peer[r]);
    if (tp.p2p.base) cudaFree(tp.p2p.base);
    if (tp.p2p.h_err) cudaFreeHost(tp.p2p.h_err);
    tp.p2p = TpPeer();
}

cudaError_t tp_p2p_collective(const TpInfo & tp, const float * send, float * recv, size_t n, bool gather, cudaStream_t st) {
    if (!tp.p2p.connected || n > TP_NMAX) return cudaErrorInvalidValue;
    P2pArgs a;
    for (int r = 0; r < TP_MAX_WORLD; r++) a.peer[r] = r < tp.world ? tp.p2p.peer[r] : nullptr;
    a.rank = tp.rank; a.world = tp.world; a.send = send; a.recv = recv; a.n = (unsigned) n; a.gather = gather ? 1 : 0; a.err = tp.p2p.d_err;
    return launch_pdl(tp_p2p_kernel, dim3(TP_CTAS), dim3(256), 0, st, a);
}

} // namespace rb
TRACE_EXPORT(rwkv_b200_debug_trace_tp)
