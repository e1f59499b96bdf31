// train.cu — row T (config 5): one training step of the encoder-attention-decoder model on one GPU.
//
//   forward  = the training graph of ocrd_cor_asv_ann/lib/seq2seq.py:316-390 (teacher forcing over the whole target
//              sequence, decoder residual adds, time-constant dropout masks handed in by the caller),
//   loss     = Keras categorical cross-entropy on probabilities with temporal weights (:494-497, :1111-1112),
//   backward = hand-derived BPTT through the tied projections, the LSTM stacks, the bidirectional bottom layer and the
//              local monotonic attention cell (attention.py:526-575; the window is a K.equal/K.cast mask: no gradient),
//   update   = embedding regulariser (:530-553) + Adam(clipnorm=5) as stand-alone Keras 2.3.1 does it (global norm).
// Data parallel use: every rank runs cor_train_forward_backward on its share (gradients of the SUM of weighted
// cross-entropies, not yet divided by the number of target steps), the flat gradient buffer and the loss triple are
// all-reduced by the caller (NCCL through torch.distributed), and every rank runs cor_train_apply with the GLOBAL
// number of target steps: identical weights everywhere, equal to a single-GPU step on the union of the batches.
//
// Everything is fp32.  The dense contractions are plain library GEMMs (cuBLAS SGEMM, TF32 off): at 64 lines per GPU a
// timestep is a 64-row GEMM, far below any tensor-core tile that would pay, and every weight changes every step, which
// rules out the load-time fp16 (hi, lo) split the inference kernels live on.  All other work -- LSTM cells forward and
// backward, attention forward and backward over the window only, softmax + cross-entropy, embedding gather / scatter,
// column sums, global norm, Adam -- is hand-written below.
#include "model.h"
#include <cublas_v2.h>
#include <math.h>
#include <string.h>
#include <stdlib.h>
#include <algorithm>
#include <map>
#include <string>
#include <vector>

namespace cor {

namespace {

constexpr int MAXWIN = 13;

#define COR_BLAS(expr)                                                                               \
    do {                                                                                             \
        cublasStatus_t _s = (expr);                                                                  \
        if (_s != CUBLAS_STATUS_SUCCESS) {                                                           \
            cor::set_error(std::string(#expr) + ": cuBLAS status " + std::to_string((int)_s) + " at " + __FILE__ + ":" + std::to_string(__LINE__)); \
            return 1;                                                                                \
        }                                                                                            \
    } while (0)

struct PSlot { size_t off; size_t n; int rows, cols; };      // one Keras weight inside the flat parameter vector

struct LstmP { size_t K, R, b; int in; };                    // offsets of kernel (in,4W), recurrent (W,4W), bias (4W)

// activations of one LSTM scan over T steps of B rows (time-major), kept for the backward pass
struct LstmAct {
    float* G = nullptr;      // [T][B][4W] pre-activations, then gates (i,f,g,o), then dZ (backward) -- in place
    float* C = nullptr;      // [T+1][B][W] cell state, slot 0 = initial (scan order: slot s+1 = after scan step s)
    float* H = nullptr;      // [T+1][B][ldh] hidden state by scan step, slot 0 = initial; ldh = W
};

}  // namespace

}  // namespace cor

using namespace cor;

struct cor_trainer {
    cor_config cfg{};
    int device = 0;
    int W = 0, d = 0, V = 0, D = 0;
    cublasHandle_t blas = nullptr, blas2 = nullptr, cur_blas = nullptr;
    cudaStream_t side = nullptr; cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    float* d_x2 = nullptr;                                               // input gradient of the backward direction (added after the join)
    std::map<std::string, std::vector<PSlot>> slots;      // Keras layer -> weights
    std::vector<std::string> layer_order;
    size_t n_params = 0;
    float *P = nullptr, *Gd = nullptr, *M1 = nullptr, *M2 = nullptr;     // params, grads, Adam moments
    double* d_scal = nullptr;                                            // [8] device scalars: loss sum, weight count, |g|^2, ...
    long long adam_t = 0;
    std::vector<void*> allocs;                                           // activation store of the current geometry
    int cap_B = 0, cap_Ts = 0, cap_Td = 0, cap_k = 0;
    // parameter offsets
    size_t E = 0, U = 0;
    std::vector<LstmP> enc_fw, enc_bw, dec;
    size_t Wa = 0, va = 0, bUW = 0, bv = 0;
    // activation store
    float *X0 = nullptr, *D0 = nullptr;                    // embedded source [Ts][B][W], embedded decoder input [Td][B][W]
    std::vector<LstmAct> a_fw, a_bw, a_dec;
    std::vector<float*> enc_in, dec_in;                    // input sequence of layer n (after residual / dropout of the layer below)
    std::vector<float*> enc_out_raw;                       // [Ts][B][D_n] layer output before dropout (concat for the bottom layer)
    float *enc_out = nullptr, *u = nullptr;                // [Ts][B][D], [Ts][B][W]
    float *wq = nullptr, *al = nullptr, *ctx = nullptr, *ctxm = nullptr, *xm = nullptr;   // top cell: [Td][B][W], [Td+1][B][Ts], [Td][B][D]
    int32_t* win = nullptr;                                // [Td][B][2] window start, count
    float *top_out = nullptr, *prob = nullptr;             // [Td][B][W], [Td][B][V]
    // backward scratch
    float *d_top = nullptr, *d_encout = nullptr, *d_u = nullptr, *d_wq = nullptr, *d_ctx = nullptr, *dv_part = nullptr, *dbv_part = nullptr;
    float *d_seq_a = nullptr, *d_seq_b = nullptr;          // [T][B][2W] gradient of a layer's output / input sequence
    float *dh = nullptr, *dc = nullptr, *dh_tmp = nullptr; // [B][W] running recurrent gradients
    float *dh_r = nullptr, *dc_r = nullptr;                // ... of a backward-direction scan
    float *dh_s = nullptr, *dc_s = nullptr;                // ... of a decoder layer running beside an encoder layer
    float *e_seq_a = nullptr, *e_seq_b = nullptr;          // [Ts][B][2W] encoder-side input gradients (the decoder's ping-pong pair is busy beside them)
    std::vector<float*> dh_init, dc_init;                  // gradient wrt the decoder layers' initial states [B][W]
    float* ones = nullptr; size_t ones_n = 0;
    // the batch, staged into buffers of the activation store (a recorded step reads fixed addresses)
    int32_t *s_src = nullptr, *s_din = nullptr, *s_dout = nullptr;
    float *s_val = nullptr, *s_menc = nullptr, *s_mdec = nullptr, *s_mcell = nullptr;
    // one executable graph per batch geometry (a step is ~2 500 launches at T = 81: replayed, not re-issued)
    struct StepGraph { std::vector<int> key; cudaGraph_t graph; cudaGraphExec_t exec; long long nodes; };
    long long launches = 0;                               // graph nodes (kernels, copies, memsets) run since the last reset
    std::vector<StepGraph> graphs;
    cudaStream_t cap_stream = nullptr;
};

namespace cor {
namespace {

// ---------------------------------------------------------------------------------------------- small kernels
__device__ __forceinline__ float sigm(float x) { return 1.0f / (1.0f + expf(-x)); }

// rows of E gathered (and weighted) into X[t][b][:]; idx [B][T][k] (src) -- idx < 0: nothing
__global__ void tr_embed_kernel(const int32_t* idx, const float* val, int B, int T, int k, const float* E, int W, float* X) {
    const int t = blockIdx.x, b = blockIdx.y;
    const int32_t* ip = idx + ((size_t)b * T + t) * k;
    const float* vp = val ? val + ((size_t)b * T + t) * k : nullptr;
    for (int u = threadIdx.x; u < W; u += blockDim.x) {
        float acc = 0.f;
        for (int a = 0; a < k; ++a) { const int v = ip[a]; if (v >= 0) acc = fmaf(vp ? vp[a] : 1.0f, E[(size_t)v * W + u], acc); }
        X[((size_t)t * B + b) * W + u] = acc;
    }
}
// dE[idx] += val * dX[t][b][:]
__global__ void tr_embed_bwd_kernel(const int32_t* idx, const float* val, int B, int T, int k, const float* dX, int W, float* dE) {
    const int t = blockIdx.x, b = blockIdx.y;
    const int32_t* ip = idx + ((size_t)b * T + t) * k;
    const float* vp = val ? val + ((size_t)b * T + t) * k : nullptr;
    for (int a = 0; a < k; ++a) {
        const int v = ip[a];
        if (v < 0) continue;
        const float w = vp ? vp[a] : 1.0f;
        for (int u = threadIdx.x; u < W; u += blockDim.x) atomicAdd(dE + (size_t)v * W + u, w * dX[((size_t)t * B + b) * W + u]);
    }
}

// out[r][:] = (a[r][:] + (resid ? resid[r][:] : 0)) * (mask ? mask[(r % mask_rows)][:] : 1); lda/ldr/ldo leading dimensions
__global__ void tr_combine_kernel(const float* a, int lda, const float* resid, int ldr, const float* mask, int mask_rows, int B,
                                  float* out, int ldo, size_t rows, int cols) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < rows * cols; i += (size_t)gridDim.x * blockDim.x) {
        const size_t r = i / cols; const int c = (int)(i % cols);
        float v = a[r * lda + c];
        if (resid) v += resid[r * ldr + c];
        if (mask) v *= mask[(mask_rows > 1 ? (r % B) : 0) * (size_t)cols + c];
        out[r * ldo + c] = v;
    }
}
// acc[r][:] += g[r][:] * mask
__global__ void tr_accum_kernel(float* acc, int lda, const float* g, int ldg, const float* mask, int mask_rows, int B, size_t rows, int cols) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < rows * cols; i += (size_t)gridDim.x * blockDim.x) {
        const size_t r = i / cols; const int c = (int)(i % cols);
        float v = g[r * ldg + c];
        if (mask) v *= mask[(mask_rows > 1 ? (r % B) : 0) * (size_t)cols + c];
        acc[r * lda + c] += v;
    }
}

// LSTM cell forward of one timestep: Z [B][4W] (gate blocks i|f|g|o, bias already in) -> gates in place, c, h
__global__ void tr_cell_fwd_kernel(float* Z, const float* c_prev, float* c_out, float* h_out, int ldh, float* h_copy, int ldc2, int B, int W) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * W) return;
    const int b = i / W, u = i % W;
    float* z = Z + (size_t)b * 4 * W;
    const float gi = sigm(z[u]), gf = sigm(z[W + u]), gg = tanhf(z[2 * W + u]), go = sigm(z[3 * W + u]);
    const float c = gf * c_prev[i] + gi * gg;
    const float h = go * tanhf(c);
    z[u] = gi; z[W + u] = gf; z[2 * W + u] = gg; z[3 * W + u] = go;
    c_out[i] = c;
    h_out[(size_t)b * ldh + u] = h;
    if (h_copy) h_copy[(size_t)b * ldc2 + u] = h;
}
// LSTM cell backward of one timestep: gates G (in place -> dZ), dh (total), dc (in: from the future, out: to the past)
__global__ void tr_cell_bwd_kernel(float* G, const float* c_prev, const float* c_cur, const float* dh, const float* dh2, int ld2, float* dc, int B, int W) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * W) return;
    const int b = i / W, u = i % W;
    float* g = G + (size_t)b * 4 * W;
    const float gi = g[u], gf = g[W + u], gg = g[2 * W + u], go = g[3 * W + u];
    const float tc = tanhf(c_cur[i]);
    float d_h = dh[i];
    if (dh2) d_h += dh2[(size_t)b * ld2 + u];
    const float d_c = dc[i] + d_h * go * (1.0f - tc * tc);
    g[u] = d_c * gg * gi * (1.0f - gi);
    g[W + u] = d_c * c_prev[i] * gf * (1.0f - gf);
    g[2 * W + u] = d_c * gi * (1.0f - gg * gg);
    g[3 * W + u] = d_h * tc * go * (1.0f - go);
    dc[i] = d_c * gf;
}

// out[c] (+)= sum_r x[r][c]   (deterministic: one block per 32 columns, fixed summation order)
__global__ void tr_colsum_kernel(const float* x, int ld, size_t rows, int cols, float* out, int accumulate) {
    __shared__ float sm[8][33];
    const int c = blockIdx.x * 32 + threadIdx.x;
    float acc = 0.f;
    if (c < cols) for (size_t r = threadIdx.y; r < rows; r += 8) acc += x[r * ld + c];
    sm[threadIdx.y][threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.y == 0 && c < cols) {
        float s = 0.f;
        for (int k = 0; k < 8; ++k) s += sm[k][threadIdx.x];
        out[c] = accumulate ? out[c] + s : s;
    }
}
__global__ void tr_bias_rows_kernel(float* Z, const float* bias, size_t rows, int cols) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < rows * cols; i += (size_t)gridDim.x * blockDim.x) Z[i] += bias[i % cols];
}

// attention forward of one decoder step, window only (attention.py:526-575): one BLOCK of four warps per row -- a warp per
// window position for the energies, all threads over the columns for the context.  (A warp per row left a 64-line
// batch on 16 blocks: 106 us per step, 60 % of the whole training step together with its backward twin.)
__global__ void __launch_bounds__(128) tr_attn_fwd_kernel(const float* wq, const float* u, const float* enc, const float* a_prev, const float* v_a, const float* b_v,
                                                          int B, int Ts, int W, int D, int window, float* a_out, float* ctx, int32_t* win) {
    __shared__ float s_red[4];
    __shared__ float s_e[MAXWIN];
    __shared__ int s_win[2];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float part = 0.f;
    for (int j = tid; j < Ts; j += 128) part = fmaf(a_prev[(size_t)b * Ts + j], (float)j, part);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if (lane == 0) s_red[warp] = part;
    __syncthreads();
    if (tid == 0) {
        const float centre = ((s_red[0] + s_red[1]) + (s_red[2] + s_red[3])) + 1.0f;
        int j0 = 0, nj = 0;
        if (window > 0) {
            const int j_lo = max(0, (int)ceilf(centre - (float)window) - 1), j_hi = min(Ts - 1, (int)floorf(centre + (float)window) + 1);
            for (int j = j_lo; j <= j_hi && nj < MAXWIN; ++j)
                if (fabsf(centre - (float)j) <= (float)window) { if (nj == 0) j0 = j; ++nj; }
        } else { j0 = 0; nj = min(Ts, MAXWIN); }
        s_win[0] = j0; s_win[1] = nj;
        win[2 * b] = j0; win[2 * b + 1] = nj;
    }
    __syncthreads();
    const int j0 = s_win[0], nj = s_win[1];
    for (int q = warp; q < nj; q += 4) {
        const float* uj = u + ((size_t)(j0 + q) * B + b) * W;
        float s = 0.f;
        for (int k = lane; k < W; k += 32) s = fmaf(tanhf(wq[(size_t)b * W + k] + uj[k]), v_a[k], s);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) s_e[q] = expf(s + b_v[0]);
    }
    __syncthreads();
    float esum = 0.f;
    for (int q = 0; q < nj; ++q) esum += s_e[q];
    for (int j = tid; j < Ts; j += 128) {
        const int q = j - j0;
        a_out[(size_t)b * Ts + j] = (q >= 0 && q < nj) ? s_e[q] / esum : 0.f;
    }
    for (int k = tid; k < D; k += 128) {
        float c = 0.f;
        for (int q = 0; q < nj; ++q) c = fmaf(s_e[q] / esum, enc[((size_t)(j0 + q) * B + b) * D + k], c);
        ctx[(size_t)b * D + k] = c;
    }
}
// attention backward of one decoder step: d_ctx -> d_enc, d_u (accumulated), d_wq, partial dv_a / db_v of the row; block per row
__global__ void __launch_bounds__(128) tr_attn_bwd_kernel(const float* wq, const float* u, const float* enc, const float* a, const float* v_a, const float* d_ctx,
                                                          const int32_t* win, int B, int Ts, int W, int D, float* d_enc, float* d_u, float* d_wq, float* dv_part, float* dbv_part) {
    __shared__ float s_da[MAXWIN], s_aw[MAXWIN];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int j0 = win[2 * b], nj = win[2 * b + 1];
    for (int q = warp; q < nj; q += 4) {
        const float* ej = enc + ((size_t)(j0 + q) * B + b) * D;
        float s = 0.f;
        for (int k = lane; k < D; k += 32) s = fmaf(d_ctx[(size_t)b * D + k], ej[k], s);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) { s_da[q] = s; s_aw[q] = a[(size_t)b * Ts + j0 + q]; }
    }
    __syncthreads();
    float dot = 0.f;
    for (int q = 0; q < nj; ++q) dot = fmaf(s_aw[q], s_da[q], dot);
    float ds[MAXWIN];
    float dbv = 0.f;
#pragma unroll
    for (int q = 0; q < MAXWIN; ++q) { ds[q] = q < nj ? s_aw[q] * (s_da[q] - dot) : 0.f; dbv += ds[q]; }     // gradient wrt the log-energies
    for (int k = tid; k < D; k += 128) {
        const float dc = d_ctx[(size_t)b * D + k];
        for (int q = 0; q < nj; ++q) {
            float* dej = d_enc + ((size_t)(j0 + q) * B + b) * D + k;
            *dej = fmaf(s_aw[q], dc, *dej);
        }
    }
    for (int k = tid; k < W; k += 128) {
        const float w = wq[(size_t)b * W + k], v = v_a[k];
        float dw = 0.f, dv = 0.f;
#pragma unroll
        for (int q = 0; q < MAXWIN; ++q) {
            if (q < nj) {
                const size_t o = ((size_t)(j0 + q) * B + b) * W + k;
                const float th = tanhf(w + u[o]);
                const float dt = ds[q] * v * (1.0f - th * th);
                dv = fmaf(ds[q], th, dv);
                dw += dt;
                d_u[o] += dt;
            }
        }
        d_wq[(size_t)b * W + k] = dw;
        dv_part[(size_t)b * W + k] = dv;
    }
    if (tid == 0) dbv_part[b] = dbv;
}

// softmax + weighted cross-entropy of one row (t, b): P in place (logits -> d_logits = w * (p - y)), loss sums
__global__ void tr_softmax_ce_kernel(float* P, const int32_t* target, int B, int Td, int V, double* scal) {
    const int r = blockIdx.x;                 // row of the time-major [Td][B] enumeration
    const int t = r / B, b = r % B;
    float* z = P + (size_t)r * V;
    __shared__ float red[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    float m = -INFINITY;
    for (int v = threadIdx.x; v < V; v += blockDim.x) m = fmaxf(m, z[v]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) red[warp] = m;
    __syncthreads();
    m = red[0];
    for (int k = 1; k < nw; ++k) m = fmaxf(m, red[k]);
    __syncthreads();
    float s = 0.f;
    for (int v = threadIdx.x; v < V; v += blockDim.x) s += expf(z[v] - m);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) red[warp] = s;
    __syncthreads();
    s = 0.f;
    for (int k = 0; k < nw; ++k) s += red[k];
    const int y = target[(size_t)b * Td + t];
    // Keras: renormalise (a no-op on a softmax), clip to [1e-7, 1 - 1e-7], -log; a clipped probability has no gradient
    float py = 0.f; bool clipped = false;
    if (y >= 0) {
        py = expf(z[y] - m) / s;
        clipped = py < 1e-7f || py > 1.0f - 1e-7f;
    }
    __syncthreads();
    for (int v = threadIdx.x; v < V; v += blockDim.x) {
        const float p = expf(z[v] - m) / s;
        z[v] = (y >= 0 && !clipped) ? (p - (v == y ? 1.0f : 0.0f)) : 0.0f;
    }
    if (threadIdx.x == 0 && y >= 0) {
        const float q = fminf(fmaxf(py, 1e-7f), 1.0f - 1e-7f);
        atomicAdd(&scal[0], (double)(-logf(q)));
        atomicAdd(&scal[1], 1.0);
    }
}

// regulariser gradient (seq2seq.py:530-553) added to dE, after the cross-entropy gradient has been scaled: one block per
// embedding row; colsum = column sums over ALL rows (tr_colsum_kernel), so mean_{i>=1} = (colsum - E_0) / (V - 1)
__global__ void __launch_bounds__(128) tr_regulariser_kernel(const float* E, int V, int W, const float* colsum, float* dE, double* scal) {
    __shared__ double s_red[4];
    const int v = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float* row = E + (size_t)v * W;
    double n2 = 0.0;
    for (int u = tid; u < W; u += 128) n2 += (double)row[u] * row[u];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n2 += __shfl_xor_sync(0xffffffffu, n2, o);
    if (lane == 0) s_red[warp] = n2;
    __syncthreads();
    n2 = (s_red[0] + s_red[1]) + (s_red[2] + s_red[3]);
    __syncthreads();
    const float coef = (float)(-0.04 * (1.0 - n2));
    double under = 0.0;
    for (int u = tid; u < W; u += 128) {
        float g = coef * row[u];
        if (v == 0) {
            const float dlt = row[u] - (colsum[u] - row[u]) / (float)(V - 1);
            g += 2.0f * dlt;
            under += (double)dlt * dlt;
        }
        dE[(size_t)v * W + u] += g;
    }
    if (v == 0) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) under += __shfl_xor_sync(0xffffffffu, under, o);
        if (lane == 0) s_red[warp] = under;
        __syncthreads();
        under = (s_red[0] + s_red[1]) + (s_red[2] + s_red[3]);
    }
    if (tid == 0) atomicAdd(&scal[2], 0.01 * (1.0 - n2) * (1.0 - n2) + (v == 0 ? under : 0.0));
}

__global__ void tr_scale_kernel(float* x, size_t n, float s) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) x[i] *= s;
}
__global__ void tr_sumsq_kernel(const float* x, size_t n, double* out) {
    double acc = 0.0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) acc += (double)x[i] * x[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    __shared__ double red[32];
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) { double s = 0.0; for (int k = 0; k < (int)(blockDim.x >> 5); ++k) s += red[k]; atomicAdd(out, s); }
}
// Adam with global-norm clipping (Keras 2.3.1 stand-alone): g *= clip / max(clip, |g|)
__global__ void tr_adam_kernel(float* p, const float* g, float* m, float* v, size_t n, const double* sumsq, float clip, float lr_t, float b1, float b2, float eps) {
    const float gn = (float)sqrt(*sumsq);
    const float scale = clip > 0.f ? clip / fmaxf(clip, gn) : 1.0f;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float gi = g[i] * scale;
        const float mi = b1 * m[i] + (1.0f - b1) * gi;
        const float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
        m[i] = mi; v[i] = vi;
        p[i] -= lr_t * mi / (sqrtf(vi) + eps);
    }
}

inline int grid_for_n(size_t n) { return (int)std::min<size_t>((n + 255) / 256, 148 * 8); }

// C[M,N] (ldc) = alpha * op(A) * op(B) + beta * C, all row-major (see the header comment for the operand order)
int gemm(cor_trainer* t, bool ta, bool tb, int M, int N, int K, float alpha, const float* A, int lda, const float* B, int ldb, float beta, float* C, int ldc) {
    if (M <= 0 || N <= 0) return 0;
    if (K <= 0) return 0;
    COR_BLAS(cublasSgemm(t->cur_blas, tb ? CUBLAS_OP_T : CUBLAS_OP_N, ta ? CUBLAS_OP_T : CUBLAS_OP_N, N, M, K, &alpha, B, ldb, A, lda, &beta, C, ldc));
    return 0;
}
// Work issued between fork() and join() on the returned stream runs beside the main stream (the two directions of the
// bidirectional layer are independent chains of ~T dependent launches each); it has its own cuBLAS handle (workspace).
int fork_side(cor_trainer* t, cudaStream_t st, cudaStream_t* side) {
    COR_CUDA(cudaEventRecord(t->ev_fork, st));
    COR_CUDA(cudaStreamWaitEvent(t->side, t->ev_fork, 0));
    COR_BLAS(cublasSetStream(t->blas2, t->side));
    t->cur_blas = t->blas2;
    *side = t->side;
    return 0;
}
void end_side(cor_trainer* t) { t->cur_blas = t->blas; }      // what follows goes to the main stream again; the side work keeps running
int join_side(cor_trainer* t, cudaStream_t st) {
    COR_CUDA(cudaEventRecord(t->ev_join, t->side));
    COR_CUDA(cudaStreamWaitEvent(st, t->ev_join, 0));
    return 0;
}

int dalloc(cor_trainer* t, size_t bytes, void** out) {
    COR_CUDA(cudaMalloc(out, bytes ? bytes : 4));
    t->allocs.push_back(*out);
    return 0;
}
template <typename T> int take(cor_trainer* t, size_t n, T** out) { return dalloc(t, n * sizeof(T), (void**)out); }

void drop_graphs(cor_trainer* t) {
    for (auto& g : t->graphs) { cudaGraphExecDestroy(g.exec); cudaGraphDestroy(g.graph); }
    t->graphs.clear();
}
void free_store(cor_trainer* t) {
    drop_graphs(t);         // they hold addresses of the store
    for (void* p : t->allocs) cudaFree(p);
    t->allocs.clear();
    t->cap_B = t->cap_Ts = t->cap_Td = t->cap_k = 0;
}

int ensure_store(cor_trainer* t, int B, int Ts, int Td, int k) {
    if (B <= t->cap_B && Ts <= t->cap_Ts && Td <= t->cap_Td && k <= t->cap_k) return 0;
    const int B0 = t->cap_B, Ts0 = t->cap_Ts, Td0 = t->cap_Td, k0 = t->cap_k;
    COR_CUDA(cudaDeviceSynchronize());
    free_store(t);
    B = std::max(B, B0); Ts = std::max(Ts, Ts0); Td = std::max(Td, Td0); k = std::max(k, k0);
    const int W = t->W, d = t->d, D = t->D, V = t->V;
    const size_t BW = (size_t)B * W;
    if (take(t, (size_t)Ts * BW, &t->X0) || take(t, (size_t)Td * BW, &t->D0)) return 1;
    auto act = [&](LstmAct& a, int T) -> int {
        return take(t, (size_t)T * B * 4 * W, &a.G) || take(t, (size_t)(T + 1) * BW, &a.C) || take(t, (size_t)(T + 1) * BW, &a.H);
    };
    t->a_fw.assign(d, LstmAct{}); t->a_bw.assign(1, LstmAct{}); t->a_dec.assign(d, LstmAct{});
    t->enc_in.assign(d, nullptr); t->dec_in.assign(d, nullptr); t->enc_out_raw.assign(d, nullptr);
    t->dh_init.assign(d, nullptr); t->dc_init.assign(d, nullptr);
    for (int n = 0; n < d; ++n) {
        if (act(t->a_fw[n], Ts) || act(t->a_dec[n], Td)) return 1;
        const int Dn = n == 0 ? 2 * W : W;
        if (take(t, (size_t)Ts * B * Dn, &t->enc_out_raw[n])) return 1;
        if (n > 0 && take(t, (size_t)Ts * B * t->enc_fw[n].in, &t->enc_in[n])) return 1;
        if (n > 0 && take(t, (size_t)Td * BW, &t->dec_in[n])) return 1;
        if (take(t, BW, &t->dh_init[n]) || take(t, BW, &t->dc_init[n])) return 1;
    }
    if (act(t->a_bw[0], Ts)) return 1;
    t->enc_in[0] = t->X0; t->dec_in[0] = t->D0;
    if (take(t, (size_t)Ts * B * D, &t->enc_out) || take(t, (size_t)Ts * BW, &t->u)) return 1;
    if (take(t, (size_t)Td * BW, &t->wq) || take(t, (size_t)(Td + 1) * B * Ts, &t->al) || take(t, (size_t)Td * B * D, &t->ctx) ||
        take(t, (size_t)Td * B * D, &t->ctxm) || take(t, (size_t)Td * BW, &t->xm) || take(t, (size_t)Td * B * 2, &t->win)) return 1;
    if (take(t, (size_t)Td * BW, &t->top_out) || take(t, (size_t)Td * B * V, &t->prob)) return 1;
    if (take(t, (size_t)Td * BW, &t->d_top) || take(t, (size_t)Ts * B * D, &t->d_encout) || take(t, (size_t)Ts * BW, &t->d_u) ||
        take(t, (size_t)Td * BW, &t->d_wq) || take(t, (size_t)Td * B * D, &t->d_ctx) || take(t, (size_t)Td * BW, &t->dv_part) ||
        take(t, (size_t)Td * B, &t->dbv_part)) return 1;
    const size_t Tm = std::max(Ts, Td);
    if (take(t, Tm * B * 2 * W, &t->d_seq_a) || take(t, Tm * B * 2 * W, &t->d_seq_b)) return 1;
    if (take(t, BW, &t->dh) || take(t, BW, &t->dc) || take(t, BW, &t->dh_tmp) || take(t, BW, &t->dh_r) || take(t, BW, &t->dc_r)) return 1;
    if (take(t, (size_t)Ts * BW, &t->d_x2) || take(t, BW, &t->dh_s) || take(t, BW, &t->dc_s)) return 1;
    if (take(t, (size_t)Ts * B * 2 * W, &t->e_seq_a) || take(t, (size_t)Ts * B * 2 * W, &t->e_seq_b)) return 1;
    if (take(t, (size_t)B * Ts * k, &t->s_src) || take(t, (size_t)B * Ts * k, &t->s_val) || take(t, (size_t)B * Td, &t->s_din) ||
        take(t, (size_t)B * Td, &t->s_dout) || take(t, (size_t)std::max(d - 1, 1) * 2 * W, &t->s_menc) ||
        take(t, (size_t)std::max(d - 1, 1) * W, &t->s_mdec) || take(t, (size_t)B * (W + D), &t->s_mcell)) return 1;
    t->cap_B = B; t->cap_Ts = Ts; t->cap_Td = Td; t->cap_k = k;
    return 0;
}

// One LSTM scan forward.  x [T][B][in] (all steps), P = kernel / recurrent / bias; h0/c0 initial state or nullptr (zeros).
// reverse: scan step s works on sequence position T-1-s.  Output h by POSITION into out (ld_out, column offset folded in).
int lstm_forward(cor_trainer* t, cudaStream_t st, const LstmP& lp, const float* x, int T, int B, LstmAct& a, const float* h0, const float* c0,
                 bool reverse, float* out, int ld_out) {
    const int W = t->W;
    const size_t BW = (size_t)B * W;
    const float* Pm = t->P;
    // pre-activations of every step at once: G = x . K + b
    if (gemm(t, false, false, T * B, 4 * W, lp.in, 1.f, x, lp.in, Pm + lp.K, 4 * W, 0.f, a.G, 4 * W)) return 1;
    tr_bias_rows_kernel<<<grid_for_n((size_t)T * B * 4 * W), 256, 0, st>>>(a.G, Pm + lp.b, (size_t)T * B, 4 * W);
    // state slots: forward scan -- slot s before step s, slot s + 1 after; backward scan -- slot pos + 1 before the step on
    // position pos, slot pos after (initial state in slot T, final in slot 0): either way "the state before position pos"
    // of all positions is one contiguous block, which the recurrent-weight gradient wants as one GEMM operand
    const size_t s0 = reverse ? (size_t)T * BW : 0;
    if (h0) { COR_CUDA(cudaMemcpyAsync(a.H + s0, h0, BW * 4, cudaMemcpyDeviceToDevice, st)); COR_CUDA(cudaMemcpyAsync(a.C + s0, c0, BW * 4, cudaMemcpyDeviceToDevice, st)); }
    else { COR_CUDA(cudaMemsetAsync(a.H + s0, 0, BW * 4, st)); COR_CUDA(cudaMemsetAsync(a.C + s0, 0, BW * 4, st)); }
    for (int s = 0; s < T; ++s) {
        const int pos = reverse ? T - 1 - s : s;
        const size_t before = (size_t)(reverse ? pos + 1 : pos) * BW, after = (size_t)(reverse ? pos : pos + 1) * BW;
        float* Z = a.G + (size_t)pos * B * 4 * W;
        if (gemm(t, false, false, B, 4 * W, W, 1.f, a.H + before, W, Pm + lp.R, 4 * W, 1.f, Z, 4 * W)) return 1;
        tr_cell_fwd_kernel<<<(int)((BW + 255) / 256), 256, 0, st>>>(Z, a.C + before, a.C + after, a.H + after, W,
                                                                  out + (size_t)pos * B * ld_out, ld_out, B, W);
    }
    COR_CUDA(cudaGetLastError());
    return 0;
}

// One LSTM scan backward.  d_out [T][B][ld] = gradient wrt the output sequence by POSITION (column offset folded in);
// dh_last/dc_last: extra gradient wrt the FINAL state (from a decoder layer that started there) or nullptr.
// Produces dK, dR, db (accumulated into the flat gradient), d_x [T][B][in] (if wanted; beta = accumulate), and leaves
// the gradient wrt the initial state in t->dh / t->dc.
int lstm_backward(cor_trainer* t, cudaStream_t st, const LstmP& lp, const float* x, int T, int B, LstmAct& a, bool reverse,
                  const float* d_out, int ld, const float* dh_last, const float* dc_last, float* d_x, float beta_x, int bufset = -1) {
    const int W = t->W;
    const size_t BW = (size_t)B * W;
    const float* Pm = t->P; float* Gm = t->Gd;
    // scans that run side by side have their own running gradients: 0 forward encoder scans (and the result slot the
    // callers read: t->dh / t->dc), 1 the backward direction, 2 a decoder layer beside an encoder layer
    if (bufset < 0) bufset = reverse ? 1 : 0;
    float* dh = bufset == 2 ? t->dh_s : (bufset == 1 ? t->dh_r : t->dh);
    float* dc = bufset == 2 ? t->dc_s : (bufset == 1 ? t->dc_r : t->dc);
    if (dh_last) { COR_CUDA(cudaMemcpyAsync(dh, dh_last, BW * 4, cudaMemcpyDeviceToDevice, st)); COR_CUDA(cudaMemcpyAsync(dc, dc_last, BW * 4, cudaMemcpyDeviceToDevice, st)); }
    else { COR_CUDA(cudaMemsetAsync(dh, 0, BW * 4, st)); COR_CUDA(cudaMemsetAsync(dc, 0, BW * 4, st)); }
    for (int s = T - 1; s >= 0; --s) {
        const int pos = reverse ? T - 1 - s : s;
        const size_t before = (size_t)(reverse ? pos + 1 : pos) * BW, after = (size_t)(reverse ? pos : pos + 1) * BW;
        float* G = a.G + (size_t)pos * B * 4 * W;
        tr_cell_bwd_kernel<<<(int)((BW + 255) / 256), 256, 0, st>>>(G, a.C + before, a.C + after, dh,
                                                                  d_out ? d_out + (size_t)pos * B * ld : nullptr, ld, dc, B, W);
        // dh_{s-1} = dZ . R^T
        if (gemm(t, false, true, B, W, 4 * W, 1.f, G, 4 * W, Pm + lp.R, 4 * W, 0.f, dh, W)) return 1;
    }
    COR_CUDA(cudaGetLastError());
    // parameter gradients over all steps at once: "state before position pos" is slots 0..T-1 (forward) or 1..T (backward)
    if (gemm(t, true, false, W, 4 * W, T * B, 1.f, a.H + (reverse ? BW : 0), W, a.G, 4 * W, 1.f, Gm + lp.R, 4 * W)) return 1;
    if (gemm(t, true, false, lp.in, 4 * W, T * B, 1.f, x, lp.in, a.G, 4 * W, 1.f, Gm + lp.K, 4 * W)) return 1;
    tr_colsum_kernel<<<(4 * W + 31) / 32, dim3(32, 8), 0, st>>>(a.G, 4 * W, (size_t)T * B, 4 * W, Gm + lp.b, 1);
    if (d_x && gemm(t, false, true, T * B, lp.in, 4 * W, 1.f, a.G, 4 * W, Pm + lp.K, 4 * W, beta_x, d_x, lp.in)) return 1;
    COR_CUDA(cudaGetLastError());
    return 0;
}

int add_slot(cor_trainer* t, const std::string& layer, int rows, int cols, size_t* off) {
    PSlot s{t->n_params, (size_t)rows * std::max(cols, 1), rows, cols};
    *off = s.off;
    t->n_params += (s.n + 63) & ~size_t(63);          // 256-byte aligned slots (cuBLAS / float4)
    t->slots[layer].push_back(s);
    return 0;
}

}  // namespace
}  // namespace cor

extern "C" {

int cor_trainer_create(const cor_config* cfg, int device, cor_trainer** out) {
    COR_CHECK(cfg && out, "null argument");
    COR_CHECK(cfg->width > 0 && cfg->depth >= 1 && cfg->depth <= 8 && cfg->voc_size >= 2, "bad topology");
    COR_CHECK(!cfg->deep_bidirectional_encoder && !cfg->bridge_dense && !cfg->lm_predict,
              "training supports the stacked-forward encoder without bridge and LM loss (config 5 family)");
    COR_CHECK(cfg->window_width >= 0 && cfg->window_width <= 6, "window_width must be in 0..6");
    int ndev = 0;
    COR_CUDA(cudaGetDeviceCount(&ndev));
    COR_CHECK(device >= 0 && device < ndev, "no such CUDA device");
    COR_CUDA(cudaSetDevice(device));
    cor_trainer* t = new cor_trainer();
    t->cfg = *cfg; t->device = device;
    const int W = t->W = cfg->width, d = t->d = cfg->depth, V = t->V = cfg->voc_size;
    const int D = t->D = d == 1 ? 2 * W : W;
    auto lstm = [&](const std::string& layer, int in, LstmP* lp) {
        lp->in = in;
        add_slot(t, layer, in, 4 * W, &lp->K); add_slot(t, layer, W, 4 * W, &lp->R); add_slot(t, layer, 4 * W, 0, &lp->b);
    };
    t->layer_order.push_back("char_input_projection");
    add_slot(t, "char_input_projection", V, W, &t->E);
    t->enc_fw.resize(d); t->enc_bw.resize(1); t->dec.resize(d);
    for (int n = 0; n < d; ++n) {
        const std::string name = "encoder_lstm_" + std::to_string(n + 1);
        t->layer_order.push_back(name);
        lstm(name, n == 0 ? W : (n == 1 ? 2 * W : W), &t->enc_fw[n]);
        if (n == 0) lstm(name, W, &t->enc_bw[0]);
    }
    t->layer_order.push_back("attention_dense");
    add_slot(t, "attention_dense", D, W, &t->U);
    for (int n = 0; n < d; ++n) {
        const std::string name = "decoder_lstm_" + std::to_string(n + 1);
        t->layer_order.push_back(name);
        if (n == d - 1) {
            add_slot(t, name, W, W, &t->Wa); add_slot(t, name, W, 1, &t->va); add_slot(t, name, W, 0, &t->bUW); add_slot(t, name, 1, 0, &t->bv);
            lstm(name, W + D, &t->dec[n]);
        } else {
            lstm(name, W, &t->dec[n]);
        }
    }
    if (cudaMalloc((void**)&t->d_scal, 64) != cudaSuccess) { set_error("out of device memory for the trainer"); delete t; return 1; }
    if (cublasCreate(&t->blas) != CUBLAS_STATUS_SUCCESS || cublasCreate(&t->blas2) != CUBLAS_STATUS_SUCCESS) { set_error("cublasCreate failed"); delete t; return 1; }
    t->cur_blas = t->blas;
    cublasSetMathMode(t->blas2, CUBLAS_DEFAULT_MATH);
    if (cudaStreamCreateWithFlags(&t->side, cudaStreamNonBlocking) != cudaSuccess || cudaEventCreateWithFlags(&t->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&t->ev_join, cudaEventDisableTiming) != cudaSuccess) { set_error("cannot create the side stream"); delete t; return 1; }
    cublasSetMathMode(t->blas, CUBLAS_DEFAULT_MATH);        // plain fp32 FFMA: TF32 tensor ops stay off unless asked for, which nothing here does
    *out = t;
    return 0;
}

void cor_trainer_destroy(cor_trainer* t) {
    if (!t) return;
    cudaSetDevice(t->device);
    free_store(t);
    if (t->cap_stream) cudaStreamDestroy(t->cap_stream);
    if (t->blas) cublasDestroy(t->blas);
    if (t->blas2) cublasDestroy(t->blas2);
    if (t->side) cudaStreamDestroy(t->side);
    if (t->ev_fork) cudaEventDestroy(t->ev_fork);
    if (t->ev_join) cudaEventDestroy(t->ev_join);
    cudaFree(t->d_scal);
    if (t->ones) cudaFree(t->ones);
    delete t;
}

int cor_trainer_param_count(const cor_trainer* t, int64_t* n) {
    COR_CHECK(t && n, "null argument");
    *n = (int64_t)t->n_params;
    return 0;
}

int cor_trainer_slot(const cor_trainer* t, const char* layer, int index, int64_t* offset, int64_t* count) {
    COR_CHECK(t && layer && offset && count, "null argument");
    auto it = t->slots.find(layer);
    COR_CHECK(it != t->slots.end() && index >= 0 && index < (int)it->second.size(), "no such weight");
    *offset = (int64_t)it->second[index].off; *count = (int64_t)it->second[index].n;
    return 0;
}

int cor_trainer_bind(cor_trainer* t, float* params, float* grads, float* adam_m, float* adam_v) {
    COR_CHECK(t && params && grads && adam_m && adam_v, "null argument");
    t->P = params; t->Gd = grads; t->M1 = adam_m; t->M2 = adam_v;
    return 0;
}

int64_t cor_trainer_launch_count(cor_trainer* t, int reset) {
    if (!t) return -1;
    const int64_t v = t->launches;
    if (reset) t->launches = 0;
    return v;
}

int cor_trainer_reset_optimizer(cor_trainer* t) {
    COR_CHECK(t && t->M1, "trainer has no buffers bound");
    COR_CUDA(cudaSetDevice(t->device));
    COR_CUDA(cudaMemset(t->M1, 0, t->n_params * 4)); COR_CUDA(cudaMemset(t->M2, 0, t->n_params * 4));
    t->adam_t = 0;
    return 0;
}

namespace cor { namespace {
// The whole step on fixed addresses (the staged batch `bt`): what gets recorded into a graph.
int step_body(cor_trainer* t, const cor_train_batch* bt, cudaStream_t st) {
    const int B = bt->n_rows, Ts = bt->t_src, Td = bt->t_dec, k = bt->k_alt;
    COR_BLAS(cublasSetStream(t->blas, st));
    const int W = t->W, d = t->d, D = t->D, V = t->V;
    const size_t BW = (size_t)B * W;
    const float* P = t->P; float* G = t->Gd;
    const bool resid = t->cfg.residual_connections != 0;
    COR_CUDA(cudaMemsetAsync(G, 0, t->n_params * 4, st));
    COR_CUDA(cudaMemsetAsync(t->d_scal, 0, 64, st));

    // ================================================================ forward
    tr_embed_kernel<<<dim3(Ts, B), 128, 0, st>>>(bt->src_idx, bt->src_val, B, Ts, k, P + t->E, W, t->X0);
    tr_embed_kernel<<<dim3(Td, B), 128, 0, st>>>(bt->dec_in, nullptr, B, Td, 1, P + t->E, W, t->D0);
    // state handed to decoder layer n: backward final of the bottom layer, forward final otherwise (seq2seq.py:280-283)
    auto enc_final = [&](int n, const float** h, const float** c) {
        LstmAct& a = n == 0 ? t->a_bw[0] : t->a_fw[n];
        const size_t fin = n == 0 ? 0 : (size_t)Ts * BW;          // final state: slot 0 of a backward scan, slot T of a forward one
        *h = a.H + fin; *c = a.C + fin;
    };
    auto dec_layer_forward = [&](int n, cudaStream_t s_) -> int {
        const float *h0, *c0;
        enc_final(n, &h0, &c0);
        float* raw = t->d_seq_a;        // raw output as scratch, then residual + dropout into the next layer's input
        if (lstm_forward(t, s_, t->dec[n], t->dec_in[n], Td, B, t->a_dec[n], h0, c0, false, raw, W)) return 1;
        const float* mask = bt->mask_dec ? bt->mask_dec + (size_t)n * W : nullptr;
        tr_combine_kernel<<<grid_for_n((size_t)Td * BW), 256, 0, s_>>>(raw, W, (n > 0 && resid) ? t->dec_in[n] : nullptr, W, mask, 1, B,
                                                                     t->dec_in[n + 1], W, (size_t)Td * B, W);
        return 0;
    };
    // ---- encoder, with decoder layer n - 1 (which needs encoder layer n - 1's final state only) beside encoder layer n
    for (int n = 0; n < d; ++n) {
        const float* x = t->enc_in[n];
        const int Dn = n == 0 ? 2 * W : W;
        float* raw = t->enc_out_raw[n];
        cudaStream_t side;
        if (n == 0) {       // the backward direction beside the forward one
            if (fork_side(t, st, &side) || lstm_forward(t, side, t->enc_bw[0], x, Ts, B, t->a_bw[0], nullptr, nullptr, true, raw + W, Dn)) return 1;
            end_side(t);
        } else {
            if (fork_side(t, st, &side) || dec_layer_forward(n - 1, side)) return 1;
            end_side(t);
        }
        if (lstm_forward(t, st, t->enc_fw[n], x, Ts, B, t->a_fw[n], nullptr, nullptr, false, raw, Dn)) return 1;
        // what the next layer (or the attention) sees: residual from the third forward layer up, then the time-constant dropout
        // (the bottom layer's backward half must be there: join first)
        if (n == 0 && join_side(t, st)) return 1;
        const bool rs = resid && n >= 2;
        const float* mask = (n < d - 1 && bt->mask_enc) ? bt->mask_enc + (size_t)n * 2 * W : nullptr;
        float* dst = n < d - 1 ? t->enc_in[n + 1] : t->enc_out;
        tr_combine_kernel<<<grid_for_n((size_t)Ts * B * Dn), 256, 0, st>>>(raw, Dn, rs ? x : nullptr, Dn, mask, 1, B, dst, Dn, (size_t)Ts * B, Dn);
        if (n > 0 && join_side(t, st)) return 1;
    }
    if (gemm(t, false, false, Ts * B, W, D, 1.f, t->enc_out, D, P + t->U, W, 0.f, t->u, W)) return 1;
    // ---- attention cell
    {
        const int top = d - 1;
        const LstmP& lp = t->dec[top];
        LstmAct& a = t->a_dec[top];
        const float* x = t->dec_in[top];
        const float *h0, *c0;
        enc_final(top, &h0, &c0);
        const float* mx = bt->mask_cell;                 // [B][W + D]
        // x-part of the cell input for all steps: G = (x * m_x) . Kx + b
        const float* xin = x;
        if (mx) {       // the x columns of the per-row cell mask as a dense [B][W] block, then x * m_x[b] for every step
            COR_CUDA(cudaMemcpy2DAsync(t->dh_tmp, (size_t)W * 4, mx, (size_t)(W + D) * 4, (size_t)W * 4, B, cudaMemcpyDeviceToDevice, st));
            tr_combine_kernel<<<grid_for_n((size_t)Td * BW), 256, 0, st>>>(x, W, nullptr, 0, t->dh_tmp, B, B, t->xm, W, (size_t)Td * B, W);
            xin = t->xm;
        }
        if (gemm(t, false, false, Td * B, 4 * W, W, 1.f, xin, W, P + lp.K, 4 * W, 0.f, a.G, 4 * W)) return 1;
        tr_bias_rows_kernel<<<grid_for_n((size_t)Td * B * 4 * W), 256, 0, st>>>(a.G, P + lp.b, (size_t)Td * B, 4 * W);
        COR_CUDA(cudaMemcpyAsync(a.H, h0, BW * 4, cudaMemcpyDeviceToDevice, st));
        COR_CUDA(cudaMemcpyAsync(a.C, c0, BW * 4, cudaMemcpyDeviceToDevice, st));
        COR_CUDA(cudaMemsetAsync(t->al, 0, (size_t)B * Ts * 4, st));
        float* cmask = nullptr;                          // context part of the cell mask as a dense [B][D] block
        if (mx) {
            cmask = t->d_ctx;                            // scratch until the backward pass
            COR_CUDA(cudaMemcpy2DAsync(cmask, (size_t)D * 4, mx + W, (size_t)(W + D) * 4, (size_t)D * 4, B, cudaMemcpyDeviceToDevice, st));
        }
        for (int s = 0; s < Td; ++s) {
            const float* hp = a.H + (size_t)s * BW;
            float* wq = t->wq + (size_t)s * BW;
            if (gemm(t, false, false, B, W, W, 1.f, hp, W, P + t->Wa, W, 0.f, wq, W)) return 1;
            tr_bias_rows_kernel<<<grid_for_n(BW), 256, 0, st>>>(wq, P + t->bUW, B, W);
            float* ctx = t->ctx + (size_t)s * B * D;
            tr_attn_fwd_kernel<<<B, 128, 0, st>>>(wq, t->u, t->enc_out, t->al + (size_t)s * B * Ts, P + t->va, P + t->bv, B, Ts, W, D,
                                                          t->cfg.window_width, t->al + (size_t)(s + 1) * B * Ts, ctx, t->win + (size_t)s * B * 2);
            const float* cin = ctx;
            if (cmask) {
                float* cm = t->ctxm + (size_t)s * B * D;
                tr_combine_kernel<<<grid_for_n((size_t)B * D), 256, 0, st>>>(ctx, D, nullptr, 0, cmask, B, B, cm, D, B, D);
                cin = cm;
            }
            float* Z = a.G + (size_t)s * B * 4 * W;
            if (gemm(t, false, false, B, 4 * W, D, 1.f, cin, D, P + lp.K + (size_t)W * 4 * W, 4 * W, 1.f, Z, 4 * W)) return 1;
            if (gemm(t, false, false, B, 4 * W, W, 1.f, hp, W, P + lp.R, 4 * W, 1.f, Z, 4 * W)) return 1;
            tr_cell_fwd_kernel<<<(int)((BW + 255) / 256), 256, 0, st>>>(Z, a.C + (size_t)s * BW, a.C + (size_t)(s + 1) * BW, a.H + (size_t)(s + 1) * BW, W,
                                                                      nullptr, 0, B, W);
        }
        // top output (+ residual from the layer input when there is a layer below), logits, softmax + loss
        tr_combine_kernel<<<grid_for_n((size_t)Td * BW), 256, 0, st>>>(a.H + BW, W, (d > 1 && resid) ? x : nullptr, W, nullptr, 1, B, t->top_out, W, (size_t)Td * B, W);
        if (gemm(t, false, true, Td * B, V, W, 1.f, t->top_out, W, P + t->E, W, 0.f, t->prob, V)) return 1;
        tr_softmax_ce_kernel<<<Td * B, 128, 0, st>>>(t->prob, bt->dec_out, B, Td, V, t->d_scal);
        COR_CUDA(cudaGetLastError());
    }

    // ================================================================ backward (gradients of the SUM of weighted cross-entropies)
    // prob now holds d_logits
    if (gemm(t, false, false, Td * B, W, V, 1.f, t->prob, V, P + t->E, W, 0.f, t->d_top, W)) return 1;
    if (gemm(t, true, false, V, W, Td * B, 1.f, t->prob, V, t->top_out, W, 1.f, G + t->E, W)) return 1;
    COR_CUDA(cudaMemsetAsync(t->d_encout, 0, (size_t)Ts * B * D * 4, st));
    COR_CUDA(cudaMemsetAsync(t->d_u, 0, (size_t)Ts * BW * 4, st));
    float* d_in_top = t->d_seq_b;        // gradient wrt the top cell's input sequence x [Td][B][W]
    {
        const int top = d - 1;
        const LstmP& lp = t->dec[top];
        LstmAct& a = t->a_dec[top];
        const float* x = t->dec_in[top];
        const float* mx = bt->mask_cell;
        float* cmask = nullptr;
        if (mx) {        // dense [B][D] context mask again (d_ctx was scratch in the forward pass): keep it in dv_part's tail? use dh_tmp-sized blocks
            cmask = t->d_seq_a;          // [B][D] fits (d_seq_a is at least T*B*2W)
            COR_CUDA(cudaMemcpy2DAsync(cmask, (size_t)D * 4, mx + W, (size_t)(W + D) * 4, (size_t)D * 4, B, cudaMemcpyDeviceToDevice, st));
        }
        COR_CUDA(cudaMemsetAsync(t->dh, 0, BW * 4, st));
        COR_CUDA(cudaMemsetAsync(t->dc, 0, BW * 4, st));
        for (int s = Td - 1; s >= 0; --s) {
            float* Gs = a.G + (size_t)s * B * 4 * W;
            tr_cell_bwd_kernel<<<(int)((BW + 255) / 256), 256, 0, st>>>(Gs, a.C + (size_t)s * BW, a.C + (size_t)(s + 1) * BW, t->dh,
                                                                      t->d_top + (size_t)s * BW, W, t->dc, B, W);
            // d_ctx = (dZ . Kc^T) * m_c ; dh_prev = dZ . R^T
            float* dctx = t->d_ctx + (size_t)s * B * D;
            if (gemm(t, false, true, B, D, 4 * W, 1.f, Gs, 4 * W, P + lp.K + (size_t)W * 4 * W, 4 * W, 0.f, dctx, D)) return 1;
            if (cmask) tr_combine_kernel<<<grid_for_n((size_t)B * D), 256, 0, st>>>(dctx, D, nullptr, 0, cmask, B, B, dctx, D, B, D);
            if (gemm(t, false, true, B, W, 4 * W, 1.f, Gs, 4 * W, P + lp.R, 4 * W, 0.f, t->dh, W)) return 1;
            float* dwq = t->d_wq + (size_t)s * BW;
            tr_attn_bwd_kernel<<<B, 128, 0, st>>>(t->wq + (size_t)s * BW, t->u, t->enc_out, t->al + (size_t)(s + 1) * B * Ts, P + t->va, dctx,
                                                          t->win + (size_t)s * B * 2, B, Ts, W, D, t->d_encout, t->d_u, dwq,
                                                          t->dv_part + (size_t)s * BW, t->dbv_part + (size_t)s * B);
            // dh_prev += d_wq . W_a^T
            if (gemm(t, false, true, B, W, W, 1.f, dwq, W, P + t->Wa, W, 1.f, t->dh, W)) return 1;
        }
        COR_CUDA(cudaGetLastError());
        // gradient wrt the cell's initial state = what is left in dh / dc
        COR_CUDA(cudaMemcpyAsync(t->dh_init[top], t->dh, BW * 4, cudaMemcpyDeviceToDevice, st));
        COR_CUDA(cudaMemcpyAsync(t->dc_init[top], t->dc, BW * 4, cudaMemcpyDeviceToDevice, st));
        // parameter gradients of the cell over all steps
        const float* xin = mx ? t->xm : x;
        const float* cin = mx ? t->ctxm : t->ctx;
        if (gemm(t, true, false, W, 4 * W, Td * B, 1.f, xin, W, a.G, 4 * W, 1.f, G + lp.K, 4 * W)) return 1;
        if (gemm(t, true, false, D, 4 * W, Td * B, 1.f, cin, D, a.G, 4 * W, 1.f, G + lp.K + (size_t)W * 4 * W, 4 * W)) return 1;
        if (gemm(t, true, false, W, 4 * W, Td * B, 1.f, a.H, W, a.G, 4 * W, 1.f, G + lp.R, 4 * W)) return 1;
        tr_colsum_kernel<<<(4 * W + 31) / 32, dim3(32, 8), 0, st>>>(a.G, 4 * W, (size_t)Td * B, 4 * W, G + lp.b, 1);
        // attention parameters: W_a from (h_prev, d_wq), b_UW, v_a, b_v by column sums
        if (gemm(t, true, false, W, W, Td * B, 1.f, a.H, W, t->d_wq, W, 1.f, G + t->Wa, W)) return 1;
        tr_colsum_kernel<<<(W + 31) / 32, dim3(32, 8), 0, st>>>(t->d_wq, W, (size_t)Td * B, W, G + t->bUW, 1);
        tr_colsum_kernel<<<(W + 31) / 32, dim3(32, 8), 0, st>>>(t->dv_part, W, (size_t)Td * B, W, G + t->va, 1);
        tr_colsum_kernel<<<1, dim3(32, 8), 0, st>>>(t->dbv_part, 1, (size_t)Td * B, 1, G + t->bv, 1);
        // gradient wrt the cell's input sequence: (dZ . Kx^T) * m_x, plus the residual path of the top output
        if (gemm(t, false, true, Td * B, W, 4 * W, 1.f, a.G, 4 * W, P + lp.K, 4 * W, 0.f, d_in_top, W)) return 1;
        if (mx) {
            COR_CUDA(cudaMemcpy2DAsync(t->dh_tmp, (size_t)W * 4, mx, (size_t)(W + D) * 4, (size_t)W * 4, B, cudaMemcpyDeviceToDevice, st));
            tr_combine_kernel<<<grid_for_n((size_t)Td * BW), 256, 0, st>>>(d_in_top, W, nullptr, 0, t->dh_tmp, B, B, d_in_top, W, (size_t)Td * B, W);
        }
        if (d > 1 && resid) tr_accum_kernel<<<grid_for_n((size_t)Td * BW), 256, 0, st>>>(d_in_top, W, t->d_top, W, nullptr, 1, B, (size_t)Td * B, W);
        COR_CUDA(cudaGetLastError());
    }
    // attention keys: u = enc_out . U
    if (gemm(t, true, false, D, W, Ts * B, 1.f, t->enc_out, D, t->d_u, W, 1.f, G + t->U, W)) return 1;
    if (gemm(t, false, true, Ts * B, D, W, 1.f, t->d_u, W, P + t->U, W, 1.f, t->d_encout, D)) return 1;
    // ---- below the top: decoder layer m = d-2-k beside encoder layer n = d-1-k (stage k).  Encoder layer n needs the gradient
    // wrt its final state from decoder layer n, which ran one stage earlier; the decoder chain only needs itself.
    float* d_cur = d_in_top; float* d_other = t->d_seq_a;       // d_cur = gradient wrt the output side of the next decoder layer down
    if (d == 1) tr_embed_bwd_kernel<<<dim3(Td, B), 128, 0, st>>>(bt->dec_in, nullptr, B, Td, 1, d_cur, W, G + t->E);
    float* d_enc = t->d_encout;                                 // gradient wrt the output side of the next encoder layer down
    float* ebufs[2] = {t->e_seq_a, t->e_seq_b};
    for (int stage = 0, pp = 0; stage < d; ++stage, pp ^= 1) {
        const int n = d - 1 - stage, mdec = d - 2 - stage;
        if (mdec >= 0) {
            cudaStream_t side;
            if (fork_side(t, st, &side)) return 1;
            // output side: out = (lstm + (m > 0 && resid ? in : 0)) * mask
            const float* mask = bt->mask_dec ? bt->mask_dec + (size_t)mdec * W : nullptr;
            if (mask) tr_combine_kernel<<<grid_for_n((size_t)Td * BW), 256, 0, side>>>(d_cur, W, nullptr, 0, mask, 1, B, d_cur, W, (size_t)Td * B, W);
            if (lstm_backward(t, side, t->dec[mdec], t->dec_in[mdec], Td, B, t->a_dec[mdec], false, d_cur, W, nullptr, nullptr, d_other, 0.f, 2)) return 1;
            COR_CUDA(cudaMemcpyAsync(t->dh_init[mdec], t->dh_s, BW * 4, cudaMemcpyDeviceToDevice, side));
            COR_CUDA(cudaMemcpyAsync(t->dc_init[mdec], t->dc_s, BW * 4, cudaMemcpyDeviceToDevice, side));
            if (mdec > 0 && resid) tr_accum_kernel<<<grid_for_n((size_t)Td * BW), 256, 0, side>>>(d_other, W, d_cur, W, nullptr, 1, B, (size_t)Td * B, W);
            std::swap(d_cur, d_other);
            // d_cur = gradient wrt the embedded decoder input D0 once the bottom decoder layer is through
            if (mdec == 0) tr_embed_bwd_kernel<<<dim3(Td, B), 128, 0, side>>>(bt->dec_in, nullptr, B, Td, 1, d_cur, W, G + t->E);
            end_side(t);
        }
        const int Dn = n == 0 ? 2 * W : W;
        const bool rs = resid && n >= 2;
        const float* mask = (n < d - 1 && bt->mask_enc) ? bt->mask_enc + (size_t)n * 2 * W : nullptr;
        if (mask) tr_combine_kernel<<<grid_for_n((size_t)Ts * B * Dn), 256, 0, st>>>(d_enc, Dn, nullptr, 0, mask, 1, B, d_enc, Dn, (size_t)Ts * B, Dn);
        const float* x = t->enc_in[n];
        const int in = t->enc_fw[n].in;
        float* d_x = ebufs[pp];                      // gradient wrt the layer's input sequence [Ts][B][in]
        // decoder layer n started from this layer's final state: backward final for the bottom layer, forward final otherwise
        if (n == 0) {
            cudaStream_t side;
            if (fork_side(t, st, &side) ||
                lstm_backward(t, side, t->enc_bw[0], x, Ts, B, t->a_bw[0], true, d_enc + W, Dn, t->dh_init[0], t->dc_init[0], t->d_x2, 0.f)) return 1;
            end_side(t);
            if (lstm_backward(t, st, t->enc_fw[0], x, Ts, B, t->a_fw[0], false, d_enc, Dn, nullptr, nullptr, d_x, 0.f) || join_side(t, st)) return 1;
            tr_accum_kernel<<<grid_for_n((size_t)Ts * B * W), 256, 0, st>>>(d_x, W, t->d_x2, W, nullptr, 1, B, (size_t)Ts * B, W);
        } else {
            if (lstm_backward(t, st, t->enc_fw[n], x, Ts, B, t->a_fw[n], false, d_enc, Dn, t->dh_init[n], t->dc_init[n], d_x, 0.f)) return 1;
            if (rs) tr_accum_kernel<<<grid_for_n((size_t)Ts * B * in), 256, 0, st>>>(d_x, in, d_enc, Dn, nullptr, 1, B, (size_t)Ts * B, in);
        }
        if (mdec >= 0 && join_side(t, st)) return 1;
        d_enc = d_x;
    }
    tr_embed_bwd_kernel<<<dim3(Ts, B), 128, 0, st>>>(bt->src_idx, bt->src_val, B, Ts, k, d_enc, W, G + t->E);
    COR_CUDA(cudaGetLastError());
    return 0;
}
} }  // namespace cor::<anon>

int cor_train_forward_backward(cor_trainer* t, const cor_train_batch* bt, double* loss_out, void* stream) {
    COR_CHECK(t && bt && loss_out, "null argument");
    COR_CHECK(t->P && t->Gd, "trainer has no buffers bound (cor_trainer_bind)");
    cudaStream_t st = (cudaStream_t)stream;
    COR_CUDA(cudaSetDevice(t->device));
    const int B = bt->n_rows, Ts = bt->t_src, Td = bt->t_dec, k = bt->k_alt;
    COR_CHECK(B > 0 && Ts > 0 && Td > 0 && k > 0, "empty batch geometry");
    COR_CHECK(bt->src_idx && bt->dec_in && bt->dec_out, "null batch arrays");
    if (ensure_store(t, B, Ts, Td, k)) return 1;
    // stage the batch: a recorded step reads these addresses
    const int W = t->W, d = t->d, D = t->D;
    cor_train_batch sb = *bt;
    auto stage = [&](const void* src, void* dst, size_t bytes, const void** field) -> int {
        if (!src) { *field = nullptr; return 0; }
        COR_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, st));
        *field = dst;
        return 0;
    };
    if (stage(bt->src_idx, t->s_src, (size_t)B * Ts * k * 4, (const void**)&sb.src_idx) || stage(bt->src_val, t->s_val, (size_t)B * Ts * k * 4, (const void**)&sb.src_val) ||
        stage(bt->dec_in, t->s_din, (size_t)B * Td * 4, (const void**)&sb.dec_in) || stage(bt->dec_out, t->s_dout, (size_t)B * Td * 4, (const void**)&sb.dec_out) ||
        stage(d > 1 ? bt->mask_enc : nullptr, t->s_menc, (size_t)(d - 1) * 2 * W * 4, (const void**)&sb.mask_enc) ||
        stage(d > 1 ? bt->mask_dec : nullptr, t->s_mdec, (size_t)(d - 1) * W * 4, (const void**)&sb.mask_dec) ||
        stage(bt->mask_cell, t->s_mcell, (size_t)B * (W + D) * 4, (const void**)&sb.mask_cell)) return 1;
    const char* off = getenv("COR_NO_GRAPH");
    if (off && off[0] && off[0] != '0') {
        if (step_body(t, &sb, st)) return 1;
    } else {
        const std::vector<int> key = {B, Ts, Td, k, sb.src_val != nullptr, sb.mask_enc != nullptr, sb.mask_dec != nullptr, sb.mask_cell != nullptr};
        cudaGraphExec_t exec = nullptr;
        long long nodes = 0;
        for (auto& g : t->graphs) if (g.key == key) { exec = g.exec; nodes = g.nodes; }
        if (!exec) {
            if (!t->cap_stream) COR_CUDA(cudaStreamCreateWithFlags(&t->cap_stream, cudaStreamNonBlocking));
            COR_CUDA(cudaStreamSynchronize(st));                 // the staging copies above must not race the warm-up run below
            // cuBLAS picks its algorithms (and may allocate) on first use: run once for real, then record
            if (step_body(t, &sb, t->cap_stream)) return 1;
            COR_CUDA(cudaStreamSynchronize(t->cap_stream));
            COR_CUDA(cudaStreamBeginCapture(t->cap_stream, cudaStreamCaptureModeRelaxed));
            const int rc = step_body(t, &sb, t->cap_stream);
            cudaGraph_t graph = nullptr;
            const cudaError_t ce = cudaStreamEndCapture(t->cap_stream, &graph);
            if (rc) return 1;
            COR_CUDA(ce);
            COR_CUDA(cudaGraphInstantiate(&exec, graph, 0));
            size_t n_nodes = 0;
            COR_CUDA(cudaGraphGetNodes(graph, nullptr, &n_nodes));
            nodes = (long long)n_nodes;
            if (t->graphs.size() >= 16) drop_graphs(t);
            t->graphs.push_back(cor_trainer::StepGraph{key, graph, exec, nodes});
        }
        COR_CUDA(cudaGraphLaunch(exec, st));
        t->launches += nodes;
    }
    double scal[3];
    COR_CUDA(cudaMemcpyAsync(scal, t->d_scal, sizeof(scal), cudaMemcpyDeviceToHost, st));
    COR_CUDA(cudaStreamSynchronize(st));
    loss_out[0] = scal[0]; loss_out[1] = scal[1];
    return 0;
}

int cor_train_apply(cor_trainer* t, double n_target_steps, int with_regulariser, float lr, float clipnorm,
                    double* stats_out, void* stream) {
    COR_CHECK(t && t->P && t->Gd && t->M1 && t->M2, "trainer has no buffers bound (cor_trainer_bind)");
    cudaStream_t st = (cudaStream_t)stream;
    COR_CUDA(cudaSetDevice(t->device));
    const int W = t->W, V = t->V;
    COR_CUDA(cudaMemsetAsync(t->d_scal + 2, 0, 16, st));
    const float inv = n_target_steps > 0 ? (float)(1.0 / n_target_steps) : 0.f;
    tr_scale_kernel<<<grid_for_n(t->n_params), 256, 0, st>>>(t->Gd, t->n_params, inv);
    if (with_regulariser) {
        COR_CHECK(t->dh_tmp, "cor_train_apply before any cor_train_forward_backward");
        tr_colsum_kernel<<<(W + 31) / 32, dim3(32, 8), 0, st>>>(t->P + t->E, W, (size_t)V, W, t->dh_tmp, 0);
        tr_regulariser_kernel<<<V, 128, 0, st>>>(t->P + t->E, V, W, t->dh_tmp, t->Gd + t->E, t->d_scal);
    }
    tr_sumsq_kernel<<<grid_for_n(t->n_params), 256, 0, st>>>(t->Gd, t->n_params, t->d_scal + 3);
    t->launches += 3 + (with_regulariser ? 2 : 0);
    if (lr > 0.f) {
        t->adam_t += 1;
        const double b1 = 0.9, b2 = 0.999;
        const float lr_t = (float)((double)lr * sqrt(1.0 - pow(b2, (double)t->adam_t)) / (1.0 - pow(b1, (double)t->adam_t)));
        tr_adam_kernel<<<grid_for_n(t->n_params), 256, 0, st>>>(t->P, t->Gd, t->M1, t->M2, t->n_params, t->d_scal + 3, clipnorm, lr_t, 0.9f, 0.999f, 1e-7f);
    }
    COR_CUDA(cudaGetLastError());
    double scal[4];
    COR_CUDA(cudaMemcpyAsync(scal, t->d_scal, sizeof(scal), cudaMemcpyDeviceToHost, st));
    COR_CUDA(cudaStreamSynchronize(st));
    if (stats_out) { stats_out[0] = scal[2]; stats_out[1] = sqrt(scal[3]); }     // regulariser value, global gradient norm (before clipping)
    return 0;
}

}  // extern "C"
