// kernels.cu — the non-GEMM kernels of the hot path: input projection (gather),
// local monotonic attention, softmax + greedy selection.  All are HBM/L2-bound row kernels:
// one warp per row, coalesced 128-bit accesses, no shared memory needed.
#include "engine.h"
#include "epilogue.cuh"
#include <math.h>
#include <algorithm>

namespace cor {

namespace {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// ---------------------------------------------------------------- K1: input projection
// seq2seq.py:237-244 — x_t = X_t . E with X_t given by its non-zeros.  One block per (t,row).
__global__ void embed_kernel(const int32_t* __restrict__ idx, const float* __restrict__ val,
                             const int32_t* __restrict__ row_T, int n_rows, int t_max, int k_alt,
                             const float* __restrict__ E, int W,
                             float* X, float* X_lo, float* Xr, float* Xr_lo) {
    const int t = blockIdx.x, r = blockIdx.y;
    const int T = row_T[r];
    if (t >= T) return;
    const int32_t* ip = idx + ((size_t)r * t_max + t) * k_alt;
    const float* vp = val ? val + ((size_t)r * t_max + t) * k_alt : nullptr;
    const size_t o_fw = ((size_t)t * n_rows + r) * W;
    const size_t o_bw = ((size_t)(T - 1 - t) * n_rows + r) * W;
    for (int u = threadIdx.x; u < W; u += blockDim.x) {
        float acc = 0.f;
        for (int k = 0; k < k_alt; ++k) {
            int v = ip[k];
            if (v >= 0) acc = fmaf(vp ? vp[k] : 1.0f, E[(size_t)v * W + u], acc);
        }
        store_split(X, X_lo, o_fw + u, acc);
        if (Xr) store_split(Xr, Xr_lo, o_bw + u, acc);
    }
}

__global__ void live_rows_kernel(const int32_t* __restrict__ idx, const float* __restrict__ val,
                                 int n_rows, int t_max, int k_alt, uint8_t* live) {
    // not np.any(encoder_input_data[j])  (seq2seq.py:1255): any non-zero entry in the row
    const int r = blockIdx.x;
    const size_t n = (size_t)t_max * k_alt;
    int any = 0;
    for (size_t i = threadIdx.x; i < n; i += blockDim.x) {
        int v = idx[(size_t)r * n + i];
        if (v >= 0 && (!val || val[(size_t)r * n + i] != 0.0f)) any = 1;
    }
    any = __syncthreads_or(any);
    if (threadIdx.x == 0) live[r] = (uint8_t)any;
}

// ---------------------------------------------------------------- K5c-e: attention
// attention.py:526-575 restricted to the window: positions outside |centre-j|<=w have e*0 = 0
// exactly, so evaluating only the window gives the same sums.  One warp per hypothesis row.
// HBM-bound (per row: <=13 positions x (W + D) x 4 B, each read once): the loads of all window
// positions of a 128-column chunk are issued together, the energies stay in registers (static
// indices only) and the normalised weights go through 16 floats of shared memory per warp.
constexpr int MAXWIN = 13;       // 2*window_width+1 with window_width <= 6
__global__ void __launch_bounds__(128, 5) attention_kernel(AttnArgs a) {
    __shared__ float s_w[4][16];
    const int R = a.r_dev ? min(*a.r_dev, a.R) : a.R;
    const int wid = threadIdx.x >> 5;
    const int row = blockIdx.x * (blockDim.x >> 5) + wid;
    if (row >= R) return;
    const int lane = threadIdx.x & 31;
    const int line = a.row_line ? a.row_line[row] : row;
    const int T = a.line_T[line];
    const float* ap = a.a_prev + (size_t)row * a.t_max;
    // centre = sum_j a_prev[j]*j + 1  (fp32, attention.py:551-553)
    float part = 0.f;
    for (int j = lane; j < T; j += 32) part = fmaf(ap[j], (float)j, part);
    const float centre = warp_sum(part) + 1.0f;
    // window: all j with |centre - j| <= w (inclusive) or < w — a contiguous range [j0, j0+nj)
    const int j_lo = max(0, (int)ceilf(centre - (float)a.window) - 1);
    const int j_hi = min(T - 1, (int)floorf(centre + (float)a.window) + 1);
    int j0 = 0, nj = 0;
    for (int j = j_lo; j <= j_hi && nj < MAXWIN; ++j) {
        const float dist = fabsf(centre - (float)j);
        const bool keep = a.inclusive ? (dist <= (float)a.window) : (dist < (float)a.window);
        if (keep) { if (nj == 0) j0 = j; ++nj; }
    }
    const float* wq = a.wq + (size_t)row * a.W;
    const size_t ustride = (size_t)a.W;                          // line-major: [L][t_max][W]
    const float* ub = a.u + ((size_t)line * a.t_max + j0) * a.W;
    float s[MAXWIN];
#pragma unroll
    for (int q = 0; q < MAXWIN; ++q) s[q] = 0.f;
    for (int k = lane * 4; k < a.W; k += 128) {
        const float4 w4 = *reinterpret_cast<const float4*>(wq + k);
        const float4 v4 = *reinterpret_cast<const float4*>(a.v_a + k);
        // loads of four positions in flight per lane; small enough for 6 blocks (24 warps) per SM
#pragma unroll
        for (int q0 = 0; q0 < MAXWIN; q0 += 4) {
            float4 u4[4];
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if (q0 + q < MAXWIN && q0 + q < nj) u4[q] = *reinterpret_cast<const float4*>(ub + (q0 + q) * ustride + k);
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if (q0 + q < MAXWIN && q0 + q < nj) {
                    float t = s[q0 + q];
                    t = fmaf(tanhf_fast(w4.x + u4[q].x), v4.x, t);
                    t = fmaf(tanhf_fast(w4.y + u4[q].y), v4.y, t);
                    t = fmaf(tanhf_fast(w4.z + u4[q].z), v4.z, t);
                    t = fmaf(tanhf_fast(w4.w + u4[q].w), v4.w, t);
                    s[q0 + q] = t;
                }
        }
    }
    float esum = 0.f;
#pragma unroll
    for (int q = 0; q < MAXWIN; ++q)
        if (q < nj) { s[q] = expf(warp_sum(s[q]) + a.b_v); esum += s[q]; }
#pragma unroll
    for (int q = 0; q < MAXWIN; ++q)
        if (q < nj && lane == 0) s_w[wid][q] = s[q] / esum;
    __syncwarp();
    // alignment row: zero everywhere, a_j = e_j / sum in the window
    float* ao = a.a_out + (size_t)row * a.t_max;
    float* al = a.align_log ? a.align_log + (a.step_dev ? (size_t)(*a.step_dev + a.step_off) * (size_t)a.align_stride : 0) + (size_t)(a.align_by_line ? line : row) * a.t_max : nullptr;
    for (int j = lane; j < a.t_max; j += 32) {
        const int q = j - j0;
        const float v = (q >= 0 && q < nj) ? s_w[wid][q] : 0.f;
        ao[j] = v;
        if (al) al[j] = v;
    }
    // context = sum_j a_j * enc_j
    const size_t estride = (size_t)a.D;
    const size_t eb = ((size_t)line * a.t_max + j0) * a.D;
    for (int k = lane * 4; k < a.D; k += 128) {
        float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int q0 = 0; q0 < MAXWIN; q0 += 4) {
            float4 x4[4];
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if (q0 + q < MAXWIN && q0 + q < nj) x4[q] = load_split4(a.enc, a.enc_lo, eb + (q0 + q) * estride + k);
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if (q0 + q < MAXWIN && q0 + q < nj) {
                    const float w = s_w[wid][q0 + q];
                    c.x = fmaf(w, x4[q].x, c.x); c.y = fmaf(w, x4[q].y, c.y);
                    c.z = fmaf(w, x4[q].z, c.z); c.w = fmaf(w, x4[q].w, c.w);
                }
        }
        store_split4(a.ctx, a.ctx_lo, (size_t)row * a.D + k, c);
    }
}

// Variant that shares an SM with the persistent lower-cell GEMM (decoder_core, overlap mode): no shared-memory slab
// (the GEMM holds ~197 KB), at most 128 registers for 128 threads (the GEMM leaves 16 K), and therefore ALL window
// positions of a 128-column chunk in flight per lane (13 x 16 B) -- with one resident block per SM that is what keeps
// enough bytes in flight (4 warps x 6.5 KB) to stream the windows while the tensor pipe works.
__global__ void __launch_bounds__(128, 4) attention_side_kernel(AttnArgs a) {
    __shared__ float s_w[4][16];
    const int R = a.r_dev ? min(*a.r_dev, a.R) : a.R;
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int row = blockIdx.x * 4 + wid; row < R; row += gridDim.x * 4) {
        const int line = a.row_line ? a.row_line[row] : row;
        const int T = a.line_T[line];
        const float* ap = a.a_prev + (size_t)row * a.t_max;
        float part = 0.f;
        for (int j = lane; j < T; j += 32) part = fmaf(ap[j], (float)j, part);
        const float centre = warp_sum(part) + 1.0f;
        const int j_lo = max(0, (int)ceilf(centre - (float)a.window) - 1);
        const int j_hi = min(T - 1, (int)floorf(centre + (float)a.window) + 1);
        int j0 = 0, nj = 0;
        for (int j = j_lo; j <= j_hi && nj < MAXWIN; ++j) {
            const float dist = fabsf(centre - (float)j);
            const bool keep = a.inclusive ? (dist <= (float)a.window) : (dist < (float)a.window);
            if (keep) { if (nj == 0) j0 = j; ++nj; }
        }
        const float* wq = a.wq + (size_t)row * a.W;
        const float* ub = a.u + ((size_t)line * a.t_max + j0) * a.W;
        float s[MAXWIN];
#pragma unroll
        for (int q = 0; q < MAXWIN; ++q) s[q] = 0.f;
        for (int k = lane * 4; k < a.W; k += 128) {
            const float4 w4 = *reinterpret_cast<const float4*>(wq + k);
            const float4 v4 = __ldg(reinterpret_cast<const float4*>(a.v_a + k));
            constexpr int HALF = 7;
#pragma unroll
            for (int q0 = 0; q0 < MAXWIN; q0 += HALF) {
                float4 u4[HALF];
#pragma unroll
                for (int q = 0; q < HALF; ++q)
                    if (q0 + q < MAXWIN && q0 + q < nj) u4[q] = __ldcs(reinterpret_cast<const float4*>(ub + (size_t)(q0 + q) * a.W + k));
#pragma unroll
                for (int q = 0; q < HALF; ++q)
                    if (q0 + q < MAXWIN && q0 + q < nj) {
                        float t = s[q0 + q];
                        t = fmaf(tanhf_fast(w4.x + u4[q].x), v4.x, t);
                        t = fmaf(tanhf_fast(w4.y + u4[q].y), v4.y, t);
                        t = fmaf(tanhf_fast(w4.z + u4[q].z), v4.z, t);
                        t = fmaf(tanhf_fast(w4.w + u4[q].w), v4.w, t);
                        s[q0 + q] = t;
                    }
            }
        }
        float esum = 0.f;
#pragma unroll
        for (int q = 0; q < MAXWIN; ++q)
            if (q < nj) { s[q] = expf(warp_sum(s[q]) + a.b_v); esum += s[q]; }
        __syncwarp();
#pragma unroll
        for (int q = 0; q < MAXWIN; ++q)
            if (q < nj && lane == 0) s_w[wid][q] = s[q] / esum;
        __syncwarp();
        float* ao = a.a_out + (size_t)row * a.t_max;
        float* al = a.align_log ? a.align_log + (a.step_dev ? (size_t)(*a.step_dev + a.step_off) * (size_t)a.align_stride : 0) + (size_t)(a.align_by_line ? line : row) * a.t_max : nullptr;
        for (int j = lane; j < a.t_max; j += 32) {
            const int q = j - j0;
            const float v = (q >= 0 && q < nj) ? s_w[wid][q] : 0.f;
            ao[j] = v;
            if (al) al[j] = v;
        }
        const size_t eb = ((size_t)line * a.t_max + j0) * a.D;
        for (int k = lane * 4; k < a.D; k += 128) {
            float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
            constexpr int HALF = 7;
            if (a.enc_lo) {
#pragma unroll
                for (int q0 = 0; q0 < MAXWIN; q0 += HALF) {
                    uint2 h[HALF], l[HALF];
#pragma unroll
                    for (int q = 0; q < HALF; ++q)
                        if (q0 + q < MAXWIN && q0 + q < nj) {
                            h[q] = __ldcs(reinterpret_cast<const uint2*>(reinterpret_cast<const __half*>(a.enc) + eb + (size_t)(q0 + q) * a.D + k));
                            l[q] = __ldcs(reinterpret_cast<const uint2*>(reinterpret_cast<const __half*>(a.enc_lo) + eb + (size_t)(q0 + q) * a.D + k));
                        }
#pragma unroll
                    for (int q = 0; q < HALF; ++q)
                        if (q0 + q < MAXWIN && q0 + q < nj) {
                            const float w = s_w[wid][q0 + q];
                            const float2 h0 = __half22float2(*reinterpret_cast<const __half2*>(&h[q].x)), h1 = __half22float2(*reinterpret_cast<const __half2*>(&h[q].y));
                            const float2 l0 = __half22float2(*reinterpret_cast<const __half2*>(&l[q].x)), l1 = __half22float2(*reinterpret_cast<const __half2*>(&l[q].y));
                            c.x = fmaf(w, h0.x + l0.x, c.x); c.y = fmaf(w, h0.y + l0.y, c.y);
                            c.z = fmaf(w, h1.x + l1.x, c.z); c.w = fmaf(w, h1.y + l1.y, c.w);
                        }
                }
            } else {
#pragma unroll
                for (int q0 = 0; q0 < MAXWIN; q0 += HALF) {
                    float4 x4[HALF];
#pragma unroll
                    for (int q = 0; q < HALF; ++q)
                        if (q0 + q < MAXWIN && q0 + q < nj) x4[q] = __ldcs(reinterpret_cast<const float4*>(a.enc + eb + (size_t)(q0 + q) * a.D + k));
#pragma unroll
                    for (int q = 0; q < HALF; ++q)
                        if (q0 + q < MAXWIN && q0 + q < nj) {
                            const float w = s_w[wid][q0 + q];
                            c.x = fmaf(w, x4[q].x, c.x); c.y = fmaf(w, x4[q].y, c.y);
                            c.z = fmaf(w, x4[q].z, c.z); c.w = fmaf(w, x4[q].w, c.w);
                        }
                }
            }
            store_split4(a.ctx, a.ctx_lo, (size_t)row * a.D + k, c);
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------- K5g softmax (+K6 greedy)
// softmax(h.E^T) (seq2seq.py:379); greedy bookkeeping of decode_batch_greedy (:1250-1260).
// One warp per row.  The row stride of logits / p is V rounded up to 128; the pad stays zero.
// NVL > 0: Vp == 32 * NVL and the row lives in registers (one pass over the logits, one expf per element --
// the same values the generic path computes twice, so the bits are identical); NVL == 0: generic path.
// best and second-best probability over v >= 1 of one lane's elements, merged across the warp: the runner-up
// is only used to LOG near-ties (north star: "excluding exact-tie argmaxes, which are logged")
__device__ __forceinline__ void top2_merge(float& best, int& best_v, float& second) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int ov = __shfl_xor_sync(0xffffffffu, best_v, o);
        const float os = __shfl_xor_sync(0xffffffffu, second, o);
        const bool take = ob > best || (ob == best && ov < best_v);
        const float loser = take ? best : ob;
        second = fmaxf(fmaxf(second, os), loser);
        if (take) { best = ob; best_v = ov; }
    }
}
template <bool GREEDY, int NVL>
__global__ void __launch_bounds__(128) softmax_kernel(GreedyArgs g, const int* r_dev) {
    const int R = r_dev ? min(*r_dev, g.R) : g.R;
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= R) return;
    const int lane = threadIdx.x & 31;
    const int ld = (g.V + 127) & ~127;
    const float* z = g.logits + (size_t)row * ld;
    float best = -INFINITY, second = -INFINITY;
    int best_v = 0x7fffffff;
    if constexpr (NVL > 0) {
        float e[NVL];
        float m = -INFINITY;
#pragma unroll
        for (int k = 0; k < NVL; ++k) { const int v = lane + 32 * k; e[k] = v < g.V ? z[v] : -INFINITY; }
#pragma unroll
        for (int k = 0; k < NVL; ++k) m = fmaxf(m, e[k]);
        m = warp_max(m);
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < NVL; ++k)
            if (lane + 32 * k < g.V) { e[k] = expf(e[k] - m); s += e[k]; } else e[k] = 0.f;
        s = warp_sum(s);
#pragma unroll
        for (int k = 0; k < NVL; ++k) {
            const int v = lane + 32 * k;
            const float pv = v < g.V ? e[k] / s : 0.f;
            if (GREEDY && v < g.V && v >= 1) {
                if (pv > best) { second = best; best = pv; best_v = v; }   // first maximum among v>=1
                else second = fmaxf(second, pv);
            }
            store_split(g.p, g.p_lo, (size_t)row * ld + v, pv);
        }
    } else {
        float m = -INFINITY;
        for (int v = lane; v < g.V; v += 32) m = fmaxf(m, z[v]);
        m = warp_max(m);
        float s = 0.f;
        for (int v = lane; v < g.V; v += 32) s += expf(z[v] - m);
        s = warp_sum(s);
        for (int v = lane; v < ld; v += 32) {
            float pv = 0.f;
            if (v < g.V) {
                pv = expf(z[v] - m) / s;
                if (GREEDY && v >= 1) {
                    if (pv > best) { second = best; best = pv; best_v = v; }   // first maximum among v>=1
                    else second = fmaxf(second, pv);
                }
            }
            store_split(g.p, g.p_lo, (size_t)row * ld + v, pv);
        }
    }
    if (!GREEDY) return;
    top2_merge(best, best_v, second);
    if (lane == 0) {
        const int line = g.row_line ? g.row_line[row] : row;
        const int step = g.step_dev ? *g.step_dev + g.step : g.step;
        const int s_row = 2 * g.row_T[line];
        const bool alive = !g.done[line] && g.live[line] && step < s_row;
        // loop graphs: a line is counted ONCE, at the step it finishes (newline, last step, or never started) -- the
        // loop ends when the count reaches the number of rows; one atomic per line and call, not per step
        if (g.done_count && (alive ? (best_v == g.eos || step + 1 >= s_row) : step == 0)) atomicAdd(g.done_count, 1);
        if (alive) {
            const int n = g.out_len[line];
            g.out_chars[(size_t)line * g.s_max + n] = best_v;
            g.out_probs[(size_t)line * g.s_max + n] = best;
            g.score_sum[line] += (double)(-logf(best));
            g.out_len[line] = n + 1;
            if (best_v == g.eos) g.done[line] = 1;
            if (g.out_ties && best - second <= kTieMargin * best) g.out_ties[line] += 1;
        }
    }
}

__global__ void split_kernel(const float* __restrict__ x, float* hi, float* lo, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) store_split(hi, lo, i, x[i]);
}

__global__ void join_kernel(const float* __restrict__ hi, const float* __restrict__ lo, float* out, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) out[i] = load_split(hi, lo, i);
}

// out[t][l][:] (fp32, time-major: the C-ABI layout of cor_encode) = hi/lo[l][t][:] (line-major, internal)
__global__ void join_lt_kernel(const float* __restrict__ hi, const float* __restrict__ lo, float* out, int L, int T, int D) {
    const size_t n = (size_t)L * T * D;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int dd = (int)(i % D); const size_t lt = i / D; const int t = (int)(lt % T), l = (int)(lt / T);
        out[((size_t)t * L + l) * D + dd] = load_split(hi, lo, i);
    }
}

// hi/lo[l][t][:] (line-major, internal) = in[t][l][:] (fp32, time-major: the C-ABI layout of cor_encode)
__global__ void split_tl_kernel(const float* __restrict__ in, float* hi, float* lo, int L, int T, int D) {
    const size_t n = (size_t)L * T * D;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int dd = (int)(i % D); const size_t lt = i / D; const int t = (int)(lt % T), l = (int)(lt / T);
        store_split(hi, lo, i, in[((size_t)t * L + l) * D + dd]);
    }
}

__global__ void finish_scores_kernel(double* score_sum, const int32_t* len, int R) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < R && len[r] > 0) score_sum[r] /= (double)len[r];
}

}  // namespace

int launch_embed(cudaStream_t st, const int32_t* idx, const float* val, const int32_t* row_T,
                 int n_rows, int t_max, int k_alt, const float* E, int W, int V,
                 float* X, float* X_lo, float* Xr, float* Xr_lo) {
    (void)V;
    dim3 grid(t_max, n_rows);
    embed_kernel<<<grid, 128, 0, st>>>(idx, val, row_T, n_rows, t_max, k_alt, E, W, X, X_lo, Xr, Xr_lo);
    ++ctx().launches;
    COR_CUDA(cudaGetLastError());
    return 0;
}

int launch_live_rows(cudaStream_t st, const int32_t* idx, const float* val, int n_rows, int t_max, int k_alt, uint8_t* live) {
    live_rows_kernel<<<n_rows, 128, 0, st>>>(idx, val, n_rows, t_max, k_alt, live);
    ++ctx().launches;
    COR_CUDA(cudaGetLastError());
    return 0;
}

int launch_attention_reg(cudaStream_t st, const AttnArgs& a) {
    COR_CHECK(a.window >= 0 && a.window <= 6, "window_width above 6 is not supported");
    COR_CHECK(a.W % 4 == 0 && a.D % 4 == 0, "W and D must be multiples of 4");
    if (a.R <= 0) return 0;
    attention_kernel<<<(a.R + 3) / 4, 128, 0, st>>>(a);
    ++ctx().launches;
    COR_CUDA(cudaGetLastError());
    return 0;
}

int launch_attention_side(cudaStream_t st, const AttnArgs& a) {
    COR_CHECK(a.window >= 0 && a.window <= 6, "window_width above 6 is not supported");
    COR_CHECK(a.W % 4 == 0 && a.D % 4 == 0, "W and D must be multiples of 4");
    if (a.R <= 0) return 0;
    // one block per SM fits beside the GEMM: a persistent grid of a few blocks per SM lets the block scheduler refill
    // the slot the moment a block ends, whatever the row count
    const int blocks = std::min((a.R + 3) / 4, 148 * 4);
    attention_side_kernel<<<blocks, 128, 0, st>>>(a);
    ++ctx().launches;
    COR_CUDA(cudaGetLastError());
    return 0;
}

int launch_softmax_greedy(cudaStream_t st, const GreedyArgs& a) {
    if (a.R <= 0) return 0;
    const int ld = (a.V + 127) & ~127;
    if (ld == 512) softmax_kernel<true, 16><<<(a.R + 3) / 4, 128, 0, st>>>(a, a.r_dev);
    else if (ld == 128) softmax_kernel<true, 4><<<(a.R + 3) / 4, 128, 0, st>>>(a, a.r_dev);
    else softmax_kernel<true, 0><<<(a.R + 3) / 4, 128, 0, st>>>(a, a.r_dev);
    ++ctx().launches;
    COR_CUDA(cudaGetLastError());
    return 0;
}

int launch_softmax(cudaStream_t st, const float* logits, float* p, float* p_lo, int R, const int* r_dev, int V) {
    if (R <= 0) return 0;
    GreedyArgs g{};
    g.logits = logits; g.p = p; g.p_lo = p_lo; g.R = R; g.V = V;
    const int ld = (V + 127) & ~127;
    if (ld == 512) softmax_kernel<false, 16><<<(R + 3) / 4, 128, 0, st>>>(g, r_dev);
    else if (ld == 128) softmax_kernel<false, 4><<<(R + 3) / 4, 128, 0, st>>>(g, r_dev);
    else softmax_kernel<false, 0><<<(R + 3) / 4, 128, 0, st>>>(g, r_dev);
    ++ctx().launches;
    COR_CUDA(cudaGetLastError());
    return 0;
}

int launch_split(cudaStream_t st, const float* x, float* hi, float* lo, size_t n) {
    if (!n) return 0;
    int blocks = (int)((n + 255) / 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    split_kernel<<<blocks, 256, 0, st>>>(x, hi, lo, n);
    ++ctx().launches;
    COR_CUDA(cudaGetLastError());
    return 0;
}

int launch_join(cudaStream_t st, const float* hi, const float* lo, float* out, size_t n) {
    if (!n) return 0;
    int blocks = (int)((n + 255) / 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    join_kernel<<<blocks, 256, 0, st>>>(hi, lo, out, n);
    ++ctx().launches;
    COR_CUDA(cudaGetLastError());
    return 0;
}

int launch_join_lt(cudaStream_t st, const float* hi, const float* lo, float* out, int L, int T, int D) {
    const size_t n = (size_t)L * T * D;
    if (!n) return 0;
    int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 16);
    join_lt_kernel<<<blocks, 256, 0, st>>>(hi, lo, out, L, T, D);
    ++ctx().launches;
    COR_CUDA(cudaGetLastError());
    return 0;
}

int launch_split_tl(cudaStream_t st, const float* in, float* hi, float* lo, int L, int T, int D) {
    const size_t n = (size_t)L * T * D;
    if (!n) return 0;
    int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 16);
    split_tl_kernel<<<blocks, 256, 0, st>>>(in, hi, lo, L, T, D);
    ++ctx().launches;
    COR_CUDA(cudaGetLastError());
    return 0;
}

int launch_finish_scores(cudaStream_t st, double* score_sum, const int32_t* len, int R) {
    finish_scores_kernel<<<(R + 255) / 256, 256, 0, st>>>(score_sum, len, R);
    ++ctx().launches;
    COR_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace cor
