// On-device temperature / top-p sampling: restates python/sampling.py:10-52 of the reference so that a decode
// step of B streams returns B token ids (4 B each) instead of B x n_vocab logits (67 MB per step at 256 streams of V = 65536).
//
//   probs = softmax(logits)                                             sampling.py:5-16
//   temperature == 0 -> argmax                                          :38-39
//   top_p < 1: cutoff = sorted_desc[first k with cumsum_k > top_p]; probs[probs < cutoff] = 0      :41-45
//   temperature != 1: probs = probs ** (1 / temperature)                :47-48
//   token = np.random.choice(len(probs), p = probs / sum(probs))        :50-52
// np.random.choice draws ONE uniform u and returns searchsorted(cumsum(p) / sum(p), u, side = 'right'); the device cannot share
// numpy's generator, so the caller passes u per row (any generator it likes) and everything else happens here.  (logit_bias: not
// supported on the device path.)
//
// One CTA per row.  The cutoff is found without sorting: it is the LARGEST probability value v with sum(p[p >= v]) > top_p, located
// by bisection on the 31 value bits of the positive floats (each step one deterministic block-wide double sum); the final pick is a
// block-wide exclusive scan over per-thread segment sums in index order (double, like numpy's cumsum of p) and a walk inside one segment.
#include "device_utils.cuh"
#include "launch.cuh"

namespace rb {

constexpr int SMP_NT = 1024;

__device__ __forceinline__ double smp_block_sum(double v, double * red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = warp_sum_d(v);
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    double t = 0.0;
    for (int i = 0; i < SMP_NT / 32; i++) t += red[i];
    return t;
}

// logits [rows][V] (device), probs: scratch [rows][V], u [rows], out [rows]
__global__ void __launch_bounds__(SMP_NT) sample_kernel(const float * logits, float * probs, int V, float temperature, float top_p,
                                                        const float * __restrict__ u, uint32_t * __restrict__ out) {
    __shared__ double red[SMP_NT / 32];
    __shared__ float fred[SMP_NT / 32];
    __shared__ int ired[SMP_NT / 32];
    __shared__ double seg[SMP_NT];
    pdl_trigger();
    pdl_wait();
    const int row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float * x = logits + (size_t) row * V;
    float * p = probs + (size_t) row * V;
    // ---- softmax (x - max, exp, / sum: sampling.py:5-8) ----
    float mx = -INFINITY; int arg = 0;
    for (int i = tid; i < V; i += SMP_NT) { const float v = x[i]; if (v > mx) { mx = v; arg = i; } }
    for (int o = 16; o > 0; o >>= 1) {
        const float om = __shfl_xor_sync(FULL, mx, o); const int oa = __shfl_xor_sync(FULL, arg, o);
        if (om > mx || (om == mx && oa < arg)) { mx = om; arg = oa; }
    }
    if (lane == 0) { fred[warp] = mx; ired[warp] = arg; }
    __syncthreads();
    mx = fred[0]; arg = ired[0];
    for (int w = 1; w < SMP_NT / 32; w++) if (fred[w] > mx || (fred[w] == mx && ired[w] < arg)) { mx = fred[w]; arg = ired[w]; }
    if (temperature == 0.0f) { if (tid == 0) out[row] = (uint32_t) arg; return; }      // np.argmax: first maximum
    double s = 0.0;
    for (int i = tid; i < V; i += SMP_NT) { const float e = expf(__fsub_rn(x[i], mx)); p[i] = e; s += (double) e; }
    const float total = (float) smp_block_sum(s, red);
    for (int i = tid; i < V; i += SMP_NT) p[i] = __fdiv_rn(p[i], total);
    __syncthreads();
    if (top_p == 0.0f) top_p = 1.0f;                                                    // sampling.py:25-26
    // ---- top-p cutoff ----
    if (top_p < 1.0f) {
        uint32_t lo = 0u;                                                               // invariant: S(lo) > top_p (S(0) = 1 > top_p unless the sum rounds below it)
        for (int bit = 30; bit >= 0; bit--) {
            const uint32_t cand = lo | (1u << bit);
            double m = 0.0;
            for (int i = tid; i < V; i += SMP_NT) { const float v = p[i]; if (__float_as_uint(v) >= cand) m += (double) v; }
            m = smp_block_sum(m, red);
            if ((float) m > top_p) lo = cand;                                           // np.cumsum of float32 probs is float32
        }
        // lo = largest bit pattern with mass above top_p: it is a value of the set (the sum only changes at values of the set)
        const float cutoff = __uint_as_float(lo);
        for (int i = tid; i < V; i += SMP_NT) if (p[i] < cutoff) p[i] = 0.0f;
        __syncthreads();
    }
    if (temperature != 1.0f) {
        const float inv = __fdiv_rn(1.0f, temperature);
        for (int i = tid; i < V; i += SMP_NT) p[i] = powf(p[i], inv);
        __syncthreads();
    }
    // ---- np.random.choice(p = probs / sum): first index whose cumulative probability exceeds u ----
    const int per = (V + SMP_NT - 1) / SMP_NT, i0 = tid * per, i1 = min(V, i0 + per);
    double mine = 0.0;
    for (int i = i0; i < i1; i++) mine += (double) p[i];
    seg[tid] = mine;
    __syncthreads();
    if (tid == 0) {
        double tot = 0.0;
        for (int t = 0; t < SMP_NT; t++) tot += seg[t];
        const double target = (double) u[row] * tot;                                    // cdf / cdf[-1] > u  <=>  cum > u * total
        double cum = 0.0; int t = 0;
        while (t < SMP_NT - 1 && cum + seg[t] <= target) { cum += seg[t]; t++; }
        int pick = min(V, (t + 1) * per) - 1;
        for (int i = t * per; i < min(V, (t + 1) * per); i++) { cum += (double) p[i]; if (cum > target) { pick = i; break; } }
        while (pick > 0 && p[pick] == 0.0f) pick--;                                     // rounding at the very end of the row: never return a zero-probability id
        out[row] = (uint32_t) pick;
    }
}

cudaError_t launch_sample(const float * logits, float * probs, int rows, int V, float temperature, float top_p, const float * u, uint32_t * out, cudaStream_t st) {
    if (rows <= 0 || V <= 0) return cudaErrorInvalidValue;
    return launch_pdl(sample_kernel, dim3(rows), dim3(SMP_NT), 0, st, logits, probs, V, temperature, top_p, u, out);
}

} // namespace rb
