// Device side of the FUSED tensor-parallel all-reduce (kernels.h TpFused / TpWait; region layout in tp.h):
// {value, flag} entries in the peers' low-latency inboxes.  Shared by the mat-vec kernel (producer epilogue,
// consumer prologue) and the stand-alone row gather before ln_out.
#pragma once
#include "device_utils.cuh"

namespace rb {

// One {value, flag} entry of the local inbox.  An 8-byte store is a single NVLink transaction, so a reader
// that sees the flag of this (pass, site) also sees the value: no fence, no separate flag word.
__device__ __forceinline__ uint2 ll_load(const uint2 * p) {
    uint2 v;
    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
    return v;
}
// Spins until the flag shows up.  A peer that does not deliver within ~2 s raises the host-visible error
// word instead of hanging the GPU (the API then fails the call: api.cpp store_outputs).
__device__ __forceinline__ float ll_wait(const uint2 * p, uint2 v, unsigned want, unsigned int * err) {
    if (v.y != want) {
        const long long t0 = clock64();
        do {
            v = ll_load(p);
            if (clock64() - t0 > 4000000000LL) { if (err) *err = 1u; break; }
        } while (v.y != want);
    }
    return __uint_as_float(v.x);
}
__device__ __forceinline__ const uint2 * tp_inbox(const TpFused & tp, int site) {
    return reinterpret_cast<const uint2 *>(tp.ll[tp.rank]) + (size_t) (site & 1) * tp.world * tp.slot_entries;
}
__device__ __forceinline__ unsigned tp_flag(const TpFused & tp, int site) {
    const unsigned step = *reinterpret_cast<const volatile unsigned int *>(tp.step);
    return (step << 8) | (unsigned) (site & 0xFF);
}
// x_new[e] of the site a launch waits for (see TpWait): every CTA rebuilds the replicated residual row from
// the partial sums the ranks pushed, in rank order, so all ranks (and all CTAs) obtain the same bits.
__device__ __forceinline__ float tp_gather_element(const TpFused & tp, const TpWait & w, const uint2 * inbox, unsigned want, float base, int e) {
    uint2 v[TP_LL_WORLD];
#pragma unroll
    for (int r = 0; r < TP_LL_WORLD; r++) if (r < tp.world) v[r] = ll_load(inbox + (size_t) r * tp.slot_entries + e);
    uint2 rv = make_uint2(0u, 0u);
    const uint2 * rp = nullptr;
    if (w.gated) { const int owner = e / w.n_slice; rp = inbox + (size_t) owner * tp.slot_entries + w.n + (e - owner * w.n_slice); rv = ll_load(rp); }
    float acc = w.gated ? 0.0f : base;
#pragma unroll
    for (int r = 0; r < TP_LL_WORLD; r++) if (r < tp.world) {
        const float p = ll_wait(inbox + (size_t) r * tp.slot_entries + e, v[r], want, tp.err);
        acc = (w.gated && r == 0) ? p : __fadd_rn(acc, p);
    }
    if (w.gated) acc = __fadd_rn(base, __fmul_rn(ll_wait(rp, rv, want, tp.err), acc));
    return acc;
}
// producer: {value, flag} into slot `rank` of peer `peer` at entry `entry` of the generation of `site`
__device__ __forceinline__ void tp_push(const TpFused & tp, int site, unsigned flag, int peer, size_t entry, float val) {
    uint2 * dst = reinterpret_cast<uint2 *>(tp.ll[peer]) + ((size_t) (site & 1) * tp.world + tp.rank) * tp.slot_entries + entry;
    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" :: "l"(dst), "r"(__float_as_uint(val)), "r"(flag) : "memory");
}

} // namespace rb
