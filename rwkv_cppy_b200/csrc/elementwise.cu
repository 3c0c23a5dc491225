// Layer norm + token shift + time-mix interpolation kernels.  Each produces the q8 / FP32
// activation rows (Act) the following mat-vec consumes, so no separate quantise launch.
//
// Reference arithmetic restated (paths relative to the reference tree):
//   rwkv_layer_norm            rwkv_operators.inc:93 over ggml norm (ggml-cpu.c:6880)
//   rwkv_carry_x (token shift) rwkv_graph.inc:56-82
//   v4/v5 mixes                rwkv_graph.inc:94-97,206-228   x*mu + (xp - xp*mu)
//   v6 mixes                   rwkv_graph.inc:301-336         x + (xp-x)*(maa + m)
//   v7 mixes                   rwkv_graph.inc:395-404         x + (xp-x)*mu
#include "device_utils.cuh"
#include "launch.cuh"
#include "tp_ll.cuh"
#include "trace.cuh"

namespace rb {

// N activation blocks at once (one per output), shuffle chains interleaved for ILP.
template <int N>
__device__ __forceinline__ void act_store_blocks(const Act * outs, size_t row, int e, bool valid, const float (&v)[N]) {
    // rows whose consumers all run on the tensor cores carry no q8 planes: the two shuffle chains and the divisions of every block
    // would be computed for nothing (256-stream v7 step: the layer norm was issue-bound on exactly this code)
    bool need_q = false;
#pragma unroll
    for (int o = 0; o < N; o++) need_q = need_q || outs[o].q != nullptr;                   // uniform
    if (!need_q) {
#pragma unroll
        for (int o = 0; o < N; o++) {
            const Act & a = outs[o];
            if (valid && a.f) a.f[row * a.K + e] = v[o];
            if (valid && a.h) a.h[row * a.K + e] = __half_as_ushort(__float2half_rn(v[o]));
        }
        return;
    }
    float amax[N]; int q[N], sum[N];
#pragma unroll
    for (int o = 0; o < N; o++) amax[o] = valid ? fabsf(v[o]) : 0.0f;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1)
#pragma unroll
        for (int o = 0; o < N; o++) amax[o] = fmaxf(amax[o], __shfl_xor_sync(FULL, amax[o], off));
#pragma unroll
    for (int o = 0; o < N; o++) {
        const float id = amax[o] != 0.0f ? __fdiv_rn(127.0f, amax[o]) : 0.0f;
        q[o] = valid ? __float2int_rn(__fmul_rn(v[o], id)) : 0;
        sum[o] = q[o];
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1)
#pragma unroll
        for (int o = 0; o < N; o++) sum[o] += __shfl_xor_sync(FULL, sum[o], off);
#pragma unroll
    for (int o = 0; o < N; o++) {
        const Act & a = outs[o];
        if (valid && a.f) a.f[row * a.K + e] = v[o];
        if (valid && a.h) a.h[row * a.K + e] = __half_as_ushort(__float2half_rn(v[o]));
        if (a.q == nullptr) continue;
        if (valid) a.q[row * a.K + e] = (int8_t) q[o];
        if ((threadIdx.x & 31) == 0 && e < a.K) {      // warps past the end of the row own no block
            const float d = __fdiv_rn(amax[o], 127.0f);
            const size_t b = row * (size_t) (a.K >> 5) + (e >> 5);
            a.d[b] = round_f16(d);
            a.s[b] = round_f16(__fmul_rn(d, (float) sum[o]));
            a.isum[b] = sum[o];
        }
    }
}

// One CTA per row, thread t owns elements t, t + blockDim, ... (ITEMS of them), so every
// global load of the kernel is issued up front (one round trip) and each warp quantises
// aligned 32-element blocks.  LN(x[r]) and -- in sequence mode -- LN(x[r-1]) (the token-shift
// source, recomputed instead of waiting for the neighbouring CTA) run through the same
// reductions; both use identical arithmetic, so the recomputed row is bit-identical.
// TPG (tensor-parallel decode, one row): the row is first gathered from the peers' partial sums (kernels.h TpWait); that variant
// fetches its constants after the gather (it waits for the peers anyway) to stay within the register budget.
template <int ITEMS, int NOUT, int NT, bool TPG = false>
__global__ void __launch_bounds__(NT) ln_mix_kernel(const __grid_constant__ LnMixArgs a) {
    __shared__ double red[4][32];
    const int r = blockIdx.x, D = a.D, nt = blockDim.x;
    const bool from_state = a.batch || r == 0;
    const size_t soff = a.batch ? (size_t) r * a.state_stride : 0;

    TRACE_DECL_IF(true); TRACE(0);
    pdl_trigger();
    float x[2][ITEMS], xs[ITEMS], w[ITEMS], bia[ITEMS], mu[NOUT][ITEMS];
    // SPLIT (few wide rows, i.e. decode of a wide model): gridDim.y = ITEMS CTAs share a row; each computes the row statistics
    // (same reduction, same bits) but mixes / quantises only item blockIdx.y, so the serial q8 shuffle chains of a row run in parallel.
    const int my_it = gridDim.y > 1 ? (int) blockIdx.y : -1;
    // ... and gridDim.z = n_out CTAs per item, one output each (the kernel is then the NOUT = 1 instantiation and o0 selects the
    // output): the q8 blocks of an item cost ~200 warp instructions per output and warp, issue-bound on the one SM that runs the CTA
    // (tools/decode_trace.py: 1.7 us for two outputs, 3.1 us for three on the 14B shape).  Plane z = 0 stores the shift state.
    const int o0 = (int) blockIdx.z;
    auto load_constants = [&]() {
#pragma unroll
        for (int it = 0; it < ITEMS; it++) {
            if (my_it >= 0 && it != my_it) continue;
            const int e = it * nt + threadIdx.x;
            const bool valid = e < D;
            xs[it] = (valid && from_state) ? a.state_in[soff + e] : 0.0f;
            w[it] = valid ? a.ln_w[e] : 0.0f; bia[it] = valid ? a.ln_b[e] : 0.0f;
#pragma unroll
            for (int o = 0; o < NOUT; o++) mu[o][it] = valid ? a.mu[o0 + o][e] : 0.0f;
        }
    };
    if constexpr (!TPG) load_constants();      // constants and the (read-only) input state first, then wait for the producer of x
    TRACE(1);
    pdl_wait();
    TRACE(2);
#pragma unroll
    for (int it = 0; it < ITEMS; it++) {
        const int e = it * nt + threadIdx.x;
        const bool valid = e < D;
        x[0][it] = valid ? a.x[(size_t) r * D + e] : 0.0f;
        x[1][it] = (valid && !from_state) ? a.x[(size_t) (r - 1) * D + e] : 0.0f;
    }
    if constexpr (TPG) {                                     // x = base row (+) the ranks' partial sums, still in flight
        const uint2 * inbox = tp_inbox(a.tp, a.wait.site);
        const unsigned want = tp_flag(a.tp, a.wait.site);
#pragma unroll
        for (int it = 0; it < ITEMS; it++) {
            const int e = it * nt + threadIdx.x;
            if (e < D) {
                x[0][it] = tp_gather_element(a.tp, a.wait, inbox, want, x[0][it], e);
                if (my_it < 0 || it == my_it) a.wait.x_out[e] = x[0][it];
            }
        }
        load_constants();
    }
    const int nrow = from_state ? 1 : 2;                     // rows to normalise (uniform per CTA)
    float scale[2];
    if (a.dbg_skip_stats) { scale[0] = scale[1] = 1.0f; }    // timing experiments only (RWKV_B200_DBG_LN_FAST): wrong results
    else if (from_state) {                                   // decode / batch rows: one row to normalise, half the reduction chains (same sums for that row)
        float s1[1];
        ln_rows_stats<ITEMS, 1>(reinterpret_cast<float (&)[1][ITEMS]>(x[0]), D, nt, red, s1, 1e-5f TRACE_PASS);
        scale[0] = s1[0]; scale[1] = 1.0f;
    } else ln_rows_stats<ITEMS, 2>(x, D, nt, red, scale, 1e-5f TRACE_PASS);
    TRACE(3);
    // ---- affine, shift state, mixes ----
#pragma unroll
    for (int it = 0; it < ITEMS; it++) {
        const int e = it * nt + threadIdx.x;
        const bool valid = e < D;
        const int p = it;
        if (my_it >= 0 && it != my_it) continue;
        const float x0 = __fadd_rn(__fmul_rn(__fmul_rn(x[0][it], scale[0]), w[p]), bia[p]);
        float xp = xs[p];                                    // state value ...
        if (nrow == 2) xp = __fadd_rn(__fmul_rn(__fmul_rn(x[1][it], scale[1]), w[p]), bia[p]);   // ... or LN(x[r-1])
        const float sx = __fsub_rn(xp, x0);
        if (valid && o0 == 0 && (a.batch || r == a.R - 1)) a.state_out[soff + e] = x0;
        if (valid && o0 == 0 && a.x0) a.x0[(size_t) r * D + e] = x0;
        if (valid && o0 == 0 && a.sx) a.sx[(size_t) r * D + e] = sx;
        float v[NOUT];
#pragma unroll
        for (int o = 0; o < NOUT; o++) {
            if (a.kind == MIX_LERP) v[o] = __fadd_rn(__fmul_rn(x0, mu[o][p]), __fsub_rn(xp, __fmul_rn(xp, mu[o][p])));
            else                    v[o] = __fadd_rn(__fmul_rn(sx, mu[o][p]), x0);
        }
        TRACE(4);
        act_store_blocks<NOUT>(a.out + o0, (size_t) r, e, valid, v);
    }
    TRACE(5);
    TRACE_FLUSH(100 + NOUT * gridDim.z, gridDim.x * gridDim.y * gridDim.z);
}

// Sequence rows whose consumers are all tensor-core MATMULs (FP16 outputs only: the tolerance-mode path).
// CTA = RPC (4 or 8) consecutive rows, one warp per row with the row in registers (two-pass statistics in FP32, no
// block-wide synchronisation), plus one more warp that normalises the row before the tile (or fetches the
// shift state for the first row of the sequence).  The normalised rows meet in shared memory, so every
// row is normalised once instead of twice; outputs leave as 8-byte FP16 quads.  N4 = D / 128.
template <int N4, int RPC>
__global__ void __launch_bounds__((RPC + 1) * 32, 2) ln_mix_seq_kernel(const __grid_constant__ LnMixArgs a) {
    extern __shared__ __align__(16) float sm_rows[];          // [RPC + 1][D]
    pdl_trigger();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, D = a.D;
    const int r0 = blockIdx.x * RPC;
    const int r = warp < RPC ? r0 + warp : r0 - 1;            // last warp: the row before the tile
    float * mine = sm_rows + (size_t) warp * D;
    pdl_wait();
    float4 x[N4];
    if (r >= 0 && r < a.R) {
#pragma unroll
        for (int j = 0; j < N4; j++) x[j] = *reinterpret_cast<const float4 *>(a.x + (size_t) r * D + (j * 32 + lane) * 4);
        float s = 0.0f;
#pragma unroll
        for (int j = 0; j < N4; j++) s += (x[j].x + x[j].y) + (x[j].z + x[j].w);
        s = warp_sum(s);
        const float mean = s / (float) D;
        float s2 = 0.0f;
#pragma unroll
        for (int j = 0; j < N4; j++) {
            x[j].x -= mean; x[j].y -= mean; x[j].z -= mean; x[j].w -= mean;
            s2 += (x[j].x * x[j].x + x[j].y * x[j].y) + (x[j].z * x[j].z + x[j].w * x[j].w);
        }
        s2 = warp_sum(s2);
        const float scale = 1.0f / sqrtf(s2 / (float) D + 1e-5f);
#pragma unroll
        for (int j = 0; j < N4; j++) {
            const int e = (j * 32 + lane) * 4;
            const float4 w = *reinterpret_cast<const float4 *>(a.ln_w + e), b = *reinterpret_cast<const float4 *>(a.ln_b + e);
            x[j].x = __fadd_rn(__fmul_rn(__fmul_rn(x[j].x, scale), w.x), b.x); x[j].y = __fadd_rn(__fmul_rn(__fmul_rn(x[j].y, scale), w.y), b.y);
            x[j].z = __fadd_rn(__fmul_rn(__fmul_rn(x[j].z, scale), w.z), b.z); x[j].w = __fadd_rn(__fmul_rn(__fmul_rn(x[j].w, scale), w.w), b.w);
            *reinterpret_cast<float4 *>(mine + e) = x[j];
        }
    } else if (r < 0) {                                       // first tile: the token-shift source is the input state
#pragma unroll
        for (int j = 0; j < N4; j++) {
            const int e = (j * 32 + lane) * 4;
            *reinterpret_cast<float4 *>(mine + e) = *reinterpret_cast<const float4 *>(a.state_in + e);
        }
    }
    __syncthreads();
    if (warp == RPC || r >= a.R) return;
    const float * prev = sm_rows + (size_t) (warp == 0 ? RPC : warp - 1) * D;
#pragma unroll
    for (int j = 0; j < N4; j++) {
        const int e = (j * 32 + lane) * 4;
        const float4 x0 = x[j], xp = *reinterpret_cast<const float4 *>(prev + e);
        const float4 sx = make_float4(__fsub_rn(xp.x, x0.x), __fsub_rn(xp.y, x0.y), __fsub_rn(xp.z, x0.z), __fsub_rn(xp.w, x0.w));
        if (r == a.R - 1) *reinterpret_cast<float4 *>(a.state_out + e) = x0;
        if (a.x0) *reinterpret_cast<float4 *>(a.x0 + (size_t) r * D + e) = x0;
        if (a.sx) *reinterpret_cast<float4 *>(a.sx + (size_t) r * D + e) = sx;
#pragma unroll 1
        for (int o = 0; o < a.n_out; o++) {
            const float4 mu = *reinterpret_cast<const float4 *>(a.mu[o] + e);
            float4 v;
            if (a.kind == MIX_LERP) {
                v.x = __fadd_rn(__fmul_rn(x0.x, mu.x), __fsub_rn(xp.x, __fmul_rn(xp.x, mu.x))); v.y = __fadd_rn(__fmul_rn(x0.y, mu.y), __fsub_rn(xp.y, __fmul_rn(xp.y, mu.y)));
                v.z = __fadd_rn(__fmul_rn(x0.z, mu.z), __fsub_rn(xp.z, __fmul_rn(xp.z, mu.z))); v.w = __fadd_rn(__fmul_rn(x0.w, mu.w), __fsub_rn(xp.w, __fmul_rn(xp.w, mu.w)));
            } else {
                v.x = __fadd_rn(__fmul_rn(sx.x, mu.x), x0.x); v.y = __fadd_rn(__fmul_rn(sx.y, mu.y), x0.y);
                v.z = __fadd_rn(__fmul_rn(sx.z, mu.z), x0.z); v.w = __fadd_rn(__fmul_rn(sx.w, mu.w), x0.w);
            }
            const __half2 h0 = __floats2half2_rn(v.x, v.y), h1 = __floats2half2_rn(v.z, v.w);
            uint2 pk; pk.x = *reinterpret_cast<const uint32_t *>(&h0); pk.y = *reinterpret_cast<const uint32_t *>(&h1);
            *reinterpret_cast<uint2 *>(a.out[o].h + (size_t) r * a.out[o].K + e) = pk;
        }
    }
}

template <int N4>
static cudaError_t launch_ln_mix_seq_n(const LnMixArgs & a, cudaStream_t st) {
    // 4 rows per CTA fill the SMs better for sequences up to a few thousand rows (25 % halo work instead of 12.5 %)
    constexpr int RPC = 4;
    const size_t smem = (size_t) (RPC + 1) * a.D * sizeof(float);
    // rows wider than 2400 floats need > 48 KB: opted in per device by elementwise_init() (called from model_load)
    return launch_pdl(ln_mix_seq_kernel<N4, RPC>, dim3((a.R + RPC - 1) / RPC), dim3((RPC + 1) * 32), smem, st, a);
}

template <int N4>
static cudaError_t ln_mix_seq_attr() {
    constexpr int RPC = 4;
    return cudaFuncSetAttribute(ln_mix_seq_kernel<N4, RPC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (RPC + 1) * N4 * 128 * (int) sizeof(float));
}

// Function attributes are per device: called once per model load on the model's device (outside stream capture).
cudaError_t elementwise_init() {
    cudaError_t e = ln_mix_seq_attr<2>();
    if (e == cudaSuccess) e = ln_mix_seq_attr<4>();
    if (e == cudaSuccess) e = ln_mix_seq_attr<6>();
    if (e == cudaSuccess) e = ln_mix_seq_attr<8>();
    if (e == cudaSuccess) e = ln_mix_seq_attr<12>();
    if (e == cudaSuccess) e = ln_mix_seq_attr<16>();
    if (e == cudaSuccess) e = ln_mix_seq_attr<20>();
    return e;
}

static bool ln_mix_seq_ok(const LnMixArgs & a) {
    static const bool off = [] { const char * e = getenv("RWKV_B200_NO_LN_SEQ"); return e && e[0] == '1'; }();
    if (off || a.batch || a.R < 16 || a.D % 128 != 0 || a.D > 2560 || a.n_out < 1) return false;   // (a warp per row starves the GPU when the rows are only a few hundred streams)
    for (int o = 0; o < a.n_out; o++)
        if (a.out[o].h == nullptr || a.out[o].f != nullptr || a.out[o].q != nullptr || a.out[o].K % 4 != 0) return false;
    return true;
}

template <int ITEMS, int NT>
static cudaError_t launch_ln_mix_i(const LnMixArgs & a, int threads, cudaStream_t st) {
    // a handful of wide rows: split every row over ITEMS CTAs (see the kernel).  A TP gather must then not be in place: every CTA
    // reads the whole base row while its siblings write their items of the gathered row.
    const int items_used = (a.D + threads - 1) / threads;
    const bool split = a.D > LN_SMALL_D && a.R <= 8 && (a.wait.site < 0 || a.wait.x_out != a.x);
    dim3 grid(a.R, split ? items_used : 1);
    if (split && a.wait.site < 0 && a.n_out > 1) {          // one CTA per (row, item, output): see the kernel
        grid.z = a.n_out;
        return launch_pdl(ln_mix_kernel<ITEMS, 1, NT>, grid, dim3(threads), 0, st, a);
    }
    if (a.wait.site >= 0) {                                  // tensor-parallel decode (v4: 3 att mixes, 2 ffn mixes)
        if (a.n_out == 3) return launch_pdl(ln_mix_kernel<ITEMS, 3, NT, true>, grid, dim3(threads), 0, st, a);
        if (a.n_out == 2) return launch_pdl(ln_mix_kernel<ITEMS, 2, NT, true>, grid, dim3(threads), 0, st, a);
        return cudaErrorInvalidValue;
    }
    switch (a.n_out) {
        case 1: return launch_pdl(ln_mix_kernel<ITEMS, 1, NT>, grid, dim3(threads), 0, st, a);
        case 2: return launch_pdl(ln_mix_kernel<ITEMS, 2, NT>, grid, dim3(threads), 0, st, a);
        case 3: return launch_pdl(ln_mix_kernel<ITEMS, 3, NT>, grid, dim3(threads), 0, st, a);
        case 4: return launch_pdl(ln_mix_kernel<ITEMS, 4, NT>, grid, dim3(threads), 0, st, a);
        case 6: return launch_pdl(ln_mix_kernel<ITEMS, 6, NT>, grid, dim3(threads), 0, st, a);
        default: return cudaErrorInvalidValue;
    }
}

cudaError_t launch_ln_mix(const LnMixArgs & a_in, cudaStream_t st) {
    static const int dbg_fast = [] { const char * e = getenv("RWKV_B200_DBG_LN_FAST"); return e ? atoi(e) : 0; }();
    LnMixArgs a = a_in;
    a.dbg_skip_stats = dbg_fast;
    if (a.wait.site >= 0 && (a.R != 1 || !a.tp.ll[0] || !a.wait.x_out)) return cudaErrorInvalidValue;
    if (a.D > LN_PRO_MAX_D) return cudaErrorInvalidValue;
    if (ln_mix_seq_ok(a)) {
        switch (a.D / 128) {
            case 2:  return launch_ln_mix_seq_n<2>(a, st);
            case 4:  return launch_ln_mix_seq_n<4>(a, st);
            case 6:  return launch_ln_mix_seq_n<6>(a, st);
            case 8:  return launch_ln_mix_seq_n<8>(a, st);
            case 12: return launch_ln_mix_seq_n<12>(a, st);
            case 16: return launch_ln_mix_seq_n<16>(a, st);
            case 20: return launch_ln_mix_seq_n<20>(a, st);
            default: break;                                   // other widths: the general kernel below
        }
    }
    if (a.D <= LN_SMALL_D) {
        // Rows up to 2048 wide: 256 threads, element it*256 + tid.  The decode mat-vecs rebuild these
        // rows in their prologue (matvec_q.cu) with the same thread shape, hence the same sums.
        const int items = (a.D + 255) / 256;
        if (items <= 1) return launch_ln_mix_i<1, 256>(a, 256, st);
        if (items <= 2) return launch_ln_mix_i<2, 256>(a, 256, st);
        if (items <= 4) return launch_ln_mix_i<4, 256>(a, 256, st);
        return launch_ln_mix_i<8, 256>(a, 256, st);
    }
    // Wider rows (no fused prologue exists for them): up to 1024 threads with every operand in registers before the dependency
    // wait.  (A 256-thread shape with operands fetched item by item was tried for these widths: the fetches serialise behind the
    // warp shuffles of the q8 blocks, 88 us per launch on the 14B shape.)
    const int items = a.D <= 4096 ? 4 : 8;
    const int threads = (((a.D + items - 1) / items) + 31) / 32 * 32;
    if (items == 4) return launch_ln_mix_i<4, 1024>(a, threads, st);
    if (threads <= 768) return launch_ln_mix_i<8, 768>(a, threads, st);      // 85 registers per thread instead of 64: no spills at D = 5120 (640 threads)
    return launch_ln_mix_i<8, 1024>(a, threads, st);
}

__global__ void __launch_bounds__(256) ln_act_kernel(const float * x /* written by the preceding kernel: NOT __restrict__ (invariant loads may be hoisted above griddepcontrol.wait) */, int ldx, int D, const float * __restrict__ w,
                                                     const float * __restrict__ b, Act out) {
    extern __shared__ float smem[];
    __shared__ double red[32];
    const int r = blockIdx.x;
    pdl_trigger();
    pdl_wait();
    block_layer_norm(x + (size_t) r * ldx, D, w, b, 1e-5f, smem, red);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int e0 = warp * 32; e0 < D; e0 += nw * 32) {
        const int e = e0 + lane;
        act_store_block(out, (size_t) r, e, e < D, e < D ? smem[e] : 0.0f);
    }
}

// Register-resident variant (rows up to 256 * ITEMS wide), same structure as ln_mix_kernel.
template <int ITEMS>
__global__ void __launch_bounds__(256) ln_act_reg_kernel(const float * x /* written by the preceding kernel: NOT __restrict__ (invariant loads may be hoisted above griddepcontrol.wait) */, int ldx, int D, const float * __restrict__ w,
                                                         const float * __restrict__ b, Act out) {
    __shared__ double red[2][32];
    const int r = blockIdx.x;
    TRACE_DECL_IF(false);
    pdl_trigger();
    float v[1][ITEMS], lw[ITEMS], lb[ITEMS];
#pragma unroll
    for (int it = 0; it < ITEMS; it++) {
        const int e = it * 256 + threadIdx.x;
        lw[it] = e < D ? w[e] : 0.0f; lb[it] = e < D ? b[e] : 0.0f;
    }
    pdl_wait();
#pragma unroll
    for (int it = 0; it < ITEMS; it++) {
        const int e = it * 256 + threadIdx.x;
        v[0][it] = e < D ? x[(size_t) r * ldx + e] : 0.0f;
    }
    float scale[1];
    ln_rows_stats<ITEMS, 1>(v, D, 256, red, scale, 1e-5f TRACE_PASS);
#pragma unroll
    for (int it = 0; it < ITEMS; it++) {
        const int e = it * 256 + threadIdx.x;
        const float y = __fadd_rn(__fmul_rn(__fmul_rn(v[0][it], scale[0]), lw[it]), lb[it]);
        if (e - (int) (threadIdx.x & 31) < D) act_store_block(out, (size_t) r, e, e < D, e < D ? y : 0.0f);
    }
}

cudaError_t launch_ln_act(const float * x, int ldx, int R, int D, const float * w, const float * b, Act out, cudaStream_t st) {
    if (D <= 1024) return launch_pdl(ln_act_reg_kernel<4>, dim3(R), dim3(256), 0, st, x, ldx, D, w, b, out);
    if (D <= 2048) return launch_pdl(ln_act_reg_kernel<8>, dim3(R), dim3(256), 0, st, x, ldx, D, w, b, out);
    if (D <= 4096) return launch_pdl(ln_act_reg_kernel<16>, dim3(R), dim3(256), 0, st, x, ldx, D, w, b, out);
    if (D <= 8192) return launch_pdl(ln_act_reg_kernel<32>, dim3(R), dim3(256), 0, st, x, ldx, D, w, b, out);
    return launch_pdl(ln_act_kernel, dim3(R), dim3(256), D * sizeof(float), st, x, ldx, D, w, b, out);
}

__global__ void __launch_bounds__(256) v6_mix2_kernel(const __grid_constant__ V6Mix2Args a) {
    pdl_trigger();
    pdl_wait();
    const int r = blockIdx.x, D = a.D;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int e0 = warp * 32; e0 < D; e0 += nw * 32) {
        const int e = e0 + lane;
        const bool valid = e < D;
        const size_t i = (size_t) r * D + e;
        const float x0 = valid ? a.x0[i] : 0.0f, sx = valid ? a.sx[i] : 0.0f;
#pragma unroll
        for (int o = 0; o < 5; o++) {
            const float m = valid ? a.m[o][i] : 0.0f, maa = valid ? a.maa[o][e] : 0.0f;
            const float v = __fadd_rn(__fmul_rn(__fadd_rn(m, maa), sx), x0);
            act_store_block(a.out[o], (size_t) r, e, valid, v);
        }
    }
}

cudaError_t launch_v6_mix2(const V6Mix2Args & a, cudaStream_t st) {
    return launch_pdl(v6_mix2_kernel, dim3(a.R), dim3(256), 0, st, a);
}

// CTA = 64 channels x 32 tokens; thread = channels (c0 + lane, c0 + lane + 32) x 4 tokens.  The low-rank
// rows of the 32 tokens sit in shared memory (broadcast reads); the W2 rows of the current mix are staged
// through shared memory (coalesced loads, transposed so that lanes read consecutive channels) into registers.
__global__ void __launch_bounds__(256) v6_mix_lora_kernel(const __grid_constant__ V6MixLoraArgs a) {
    __shared__ __align__(16) float sl[32][160];
    __shared__ float sw[32][65];                             // sw[j][channel]
    pdl_trigger();
    const int tid = threadIdx.x, lane = tid & 31, tg = tid >> 5;
    const int c0 = blockIdx.x * 64, t0 = blockIdx.y * 32, D = a.D;
    const int ea = c0 + lane, eb = c0 + lane + 32;
    const bool a_ok = ea < D, b_ok = eb < D;
    // W2 tile of one mix: 64 rows x 32 floats, two float4 per thread, fetched into registers one mix ahead
    // (the L2 latency hides behind the arithmetic of the current mix) and transposed into sw when its turn comes
    float4 wreg[2]; float ma, mb;
    auto fetch_w = [&](int n) {
#pragma unroll
        for (int q = 0; q < 2; q++) {
            const int idx = q * 256 + tid, ch = idx >> 3, j4 = (idx & 7) * 4;
            wreg[q] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            if (c0 + ch < D) wreg[q] = *reinterpret_cast<const float4 *>(a.w2[n] + (size_t) (c0 + ch) * 32 + j4);
        }
        ma = a_ok ? a.maa[n][ea] : 0.0f; mb = b_ok ? a.maa[n][eb] : 0.0f;
    };
    auto store_w = [&]() {
#pragma unroll
        for (int q = 0; q < 2; q++) {
            const int idx = q * 256 + tid, ch = idx >> 3, j4 = (idx & 7) * 4;
            sw[j4][ch] = wreg[q].x; sw[j4 + 1][ch] = wreg[q].y; sw[j4 + 2][ch] = wreg[q].z; sw[j4 + 3][ch] = wreg[q].w;
        }
    };
    fetch_w(0);                                              // constants: before the wait
    pdl_wait();
#pragma unroll
    for (int q = 0; q < 5; q++) {
        const int idx = q * 256 + tid, t = idx / 40, c = idx % 40;
        float4 v = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (t0 + t < a.R) v = *reinterpret_cast<const float4 *>(a.lora + (size_t) (t0 + t) * a.ld_lora + c * 4);
        *reinterpret_cast<float4 *>(&sl[t][c * 4]) = v;
    }
    float x0a[4], x0b[4], sxa[4], sxb[4];
#pragma unroll
    for (int tt = 0; tt < 4; tt++) {
        const int t = t0 + tg * 4 + tt;
        const bool ok = t < a.R;
        x0a[tt] = (ok && a_ok) ? a.x0[(size_t) t * D + ea] : 0.0f; sxa[tt] = (ok && a_ok) ? a.sx[(size_t) t * D + ea] : 0.0f;
        x0b[tt] = (ok && b_ok) ? a.x0[(size_t) t * D + eb] : 0.0f; sxb[tt] = (ok && b_ok) ? a.sx[(size_t) t * D + eb] : 0.0f;
    }
#pragma unroll 1
    for (int n = 0; n < 5; n++) {
        if (n > 0) __syncthreads();                          // everyone holds the rows of the previous mix: sw may be refilled
        store_w();
        const float ma_n = ma, mb_n = mb;
        __syncthreads();                                     // sw (and, first time, sl) staged
        float wa[32], wb[32];
#pragma unroll
        for (int j = 0; j < 32; j++) { wa[j] = sw[j][lane]; wb[j] = sw[j][lane + 32]; }
        if (n + 1 < 5) fetch_w(n + 1);
#pragma unroll
        for (int tt = 0; tt < 4; tt++) {
            const int t = t0 + tg * 4 + tt;
            const float * lp = &sl[tg * 4 + tt][n * 32];
            float m0 = 0.0f, m1 = 0.0f;
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const float4 l4 = *reinterpret_cast<const float4 *>(lp + 4 * j);
                m0 = fmaf(l4.x, wa[4 * j], m0); m0 = fmaf(l4.y, wa[4 * j + 1], m0); m0 = fmaf(l4.z, wa[4 * j + 2], m0); m0 = fmaf(l4.w, wa[4 * j + 3], m0);
                m1 = fmaf(l4.x, wb[4 * j], m1); m1 = fmaf(l4.y, wb[4 * j + 1], m1); m1 = fmaf(l4.z, wb[4 * j + 2], m1); m1 = fmaf(l4.w, wb[4 * j + 3], m1);
            }
            if (t < a.R) {
                __half * o = reinterpret_cast<__half *>(a.out_h[n]) + (size_t) t * D;
                if (a_ok) o[ea] = __float2half_rn(__fadd_rn(__fmul_rn(__fadd_rn(m0, ma_n), sxa[tt]), x0a[tt]));
                if (b_ok) o[eb] = __float2half_rn(__fadd_rn(__fmul_rn(__fadd_rn(m1, mb_n), sxb[tt]), x0b[tt]));
            }
        }
    }
}

cudaError_t launch_v6_mix_lora(const V6MixLoraArgs & a, cudaStream_t st) {
    if (a.ld_lora != 160) return cudaErrorInvalidValue;
    return launch_pdl(v6_mix_lora_kernel, dim3((a.D + 63) / 64, (a.R + 31) / 32), dim3(256), 0, st, a);
}

__global__ void __launch_bounds__(256) quantize_act_kernel(Act a) {
    pdl_trigger();
    pdl_wait();
    const int r = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int e0 = warp * 32; e0 < a.K; e0 += nw * 32) {
        const int e = e0 + lane;
        const bool valid = e < a.K;
        act_store_block(a, (size_t) r, e, valid, valid ? a.f[(size_t) r * a.K + e] : 0.0f);
    }
}

cudaError_t launch_quantize_act(Act a, int R, cudaStream_t st) {
    if (a.q == nullptr) return cudaSuccess;
    return launch_pdl(quantize_act_kernel, dim3(R), dim3(256), 0, st, a);
}

} // namespace rb
TRACE_EXPORT(rwkv_b200_debug_trace_ew)
