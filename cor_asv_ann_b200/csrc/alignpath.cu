// alignpath.cu — hard re-alignment path from the decoders' soft alignment rows, on the device.
//
// Replaces the per-line Python loop `_alignment2path(alignment, input_line_len, len(output_line),
// 1/voc_size)` of wrapper/transcode.py:279-349 (called at :126), the only consumer of the
// `alignments` that correct_lines returns: with the path made here only (T+1) ints and one double
// per line cross to the host instead of S*T floats.
//
// One CTA per line.  The (i_max x j_max) forward table lives in shared memory, laid out [i][j]
// like the reference's numpy array.  Phase 1 copies the scores a[j][i] into it (coalesced over i).
// Phase 2 is the forward pass as an anti-diagonal wavefront: cell (i,j) needs (i-1,j), (i,j-1),
// (i-1,j-1), all on the two previous diagonals, and the value is the same fp32 add / max the
// reference does in its row-major walk, so the table is bit-identical.  Cells whose score is not
// above min_score are never visited by the reference and stay 0 (except (0,0)).  Phase 3, one
// thread: start cell (argmax over the clipped slice when the table is not square) and trace-back
// with the reference's negative-index wrap-around and strict comparisons; visited cells are logged.
// Phase 4: (1 - a) of the visited cells is fetched in parallel and summed by one thread in visit
// order in fp64 (the reference's numpy<1.24 promotion), so `dist` is reproducible.
#include "engine.h"

namespace cor {
namespace {

struct PathArgs {
    const float* align;        // a(row, j, i) = align[j*stride_j + row*stride_row + i]
    long long stride_j, stride_row;
    const int32_t* i_max;      // [n_rows] input length
    const int32_t* j_max;      // [n_rows] output length; < 0: identity matrix of size -j_max (beam fallback, seq2seq.py:835)
    double min_score;          // compared in fp64 like np.float32 > python float under numpy<1.24
    int32_t* path;             // [n_rows][path_ld], -1 = no entry
    int path_ld;
    double* dist;              // [n_rows]
    int t_cap, s_cap;          // shared-memory capacity: i_max <= t_cap, |j_max| <= s_cap
};

__device__ __forceinline__ float score_at(const PathArgs& p, const float* base, int ident, int i, int j) {
    if (ident) return (i == j && i < ident) ? 1.0f : 0.0f;
    return base[(long long)j * p.stride_j + i];
}

__global__ void __launch_bounds__(128) alignment_path_kernel(PathArgs p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int row = blockIdx.x;
    const int I = p.i_max[row];
    int J = p.j_max[row];
    int ident = 0;
    if (J < 0) { ident = -J; J = -J; }
    int32_t* path = p.path + (long long)row * p.path_ld;
    for (int i = threadIdx.x; i < p.path_ld; i += blockDim.x) path[i] = -1;
    if (I < 1 || J < 1 || I > p.t_cap || J > p.s_cap) {      // nothing to align (the reference would raise)
        __syncthreads();
        if (threadIdx.x == 0) {
            if (I >= 0 && I < p.path_ld) path[I] = J;
            path[0] = 0;
            p.dist[row] = 0.0;
        }
        return;
    }
    const int ld = J | 1;                                      // odd row stride: fewer bank conflicts on [i][j]
    float* tab = reinterpret_cast<float*>(smem_raw);            // [I][ld]
    int32_t* cells = reinterpret_cast<int32_t*>(tab + (((size_t)p.t_cap * (p.s_cap | 1) + 1) & ~(size_t)1));   // visit log, <= I+J+1 entries
    double* terms = reinterpret_cast<double*>(cells + ((p.t_cap + p.s_cap + 2 + 1) & ~1));
    const float* base = p.align + (long long)row * p.stride_row;

    // phase 1: raw scores into the table
    for (int e = threadIdx.x; e < I * J; e += blockDim.x) {
        int j = e / I, i = e - j * I;
        tab[i * ld + j] = score_at(p, base, ident, i, j);
    }
    __syncthreads();

    // phase 2: forward wavefront, in place
    for (int d = 0; d <= I + J - 2; ++d) {
        int i_lo = max(0, d - (J - 1)), i_hi = min(I - 1, d);
        for (int i = i_lo + threadIdx.x; i <= i_hi; i += blockDim.x) {
            int j = d - i;
            float a = tab[i * ld + j];
            float v = 0.0f;
            if ((double)a > p.min_score || d == 0) {
                float best = 0.0f;
                if (i > 0) best = fmaxf(best, tab[(i - 1) * ld + j]);
                if (j > 0) best = fmaxf(best, tab[i * ld + j - 1]);
                if (i > 0 && j > 0) best = fmaxf(best, tab[(i - 1) * ld + j - 1]);
                v = a + best;
            }
            tab[i * ld + j] = v;
        }
        __syncthreads();
    }

    // phase 3: start cell and trace-back (one thread)
    __shared__ int n_cells;
    if (threadIdx.x == 0) {
        int i, j;
        if (I <= J) {
            i = I - 1;
        } else {
            int s = max(0, J - 2), arg = 0;
            float best = tab[s * ld + J - 1];
            for (int k = s + 1; k < I; ++k) {
                float v = tab[k * ld + J - 1];
                if (v > best) { best = v; arg = k - s; }
            }
            i = J - 2 + arg;
        }
        if (J <= I) {
            j = J - 1;
        } else {
            int s = max(0, I - 2), arg = 0;
            float best = tab[(I - 1) * ld + s];
            for (int k = s + 1; k < J; ++k) {
                float v = tab[(I - 1) * ld + k];
                if (v > best) { best = v; arg = k - s; }
            }
            j = I - 2 + arg;
        }
        path[I] = J;
        int n = 0;
        while (i >= 0 && j >= 0) {
            cells[n++] = (i << 16) | j;
            path[i] = j;
            int iw = i > 0 ? i - 1 : I - 1, jw = j > 0 ? j - 1 : J - 1;     // numpy's negative index
            float up = tab[iw * ld + j], left = tab[i * ld + jw], diag = tab[iw * ld + jw];
            if (up > left) {
                if (up > diag) { i -= 1; } else { i -= 1; j -= 1; }
            } else if (left > diag) {
                j -= 1;
            } else {
                i -= 1; j -= 1;
            }
        }
        path[0] = 0;
        n_cells = n;
    }
    __syncthreads();

    // phase 4: dist = sum over the visited cells, in visit order, of 1.0 - a (fp64)
    const int n = n_cells;
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
        int i = cells[e] >> 16, j = cells[e] & 0xffff;
        terms[e] = 1.0 - (double)score_at(p, base, ident, i, j);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int e = 0; e < n; ++e) s += terms[e];
        p.dist[row] = s;
    }
}

}  // namespace

int launch_alignment_paths(cudaStream_t st, const float* align, long long stride_j, long long stride_row,
                           int n_rows, int t_cap, int s_cap, const int32_t* i_max, const int32_t* j_max,
                           double min_score, int32_t* path, int path_ld, double* dist) {
    COR_CHECK(n_rows >= 0 && t_cap >= 1 && s_cap >= 1 && t_cap < 32768 && s_cap < 32768, "bad alignment-path geometry");
    COR_CHECK(path_ld >= t_cap + 1, "path rows must hold t_max + 1 entries");
    if (n_rows == 0) return 0;
    size_t n_log = (size_t)((t_cap + s_cap + 2 + 1) & ~1);
    size_t smem = (((size_t)t_cap * (s_cap | 1) + 1) & ~(size_t)1) * sizeof(float) + n_log * sizeof(int32_t) + n_log * sizeof(double);
    smem = (smem + 15) & ~(size_t)15;
    COR_CHECK(smem <= 227 * 1024, "alignment table does not fit in shared memory (t_max * s_max too large)");
    if (smem > 48 * 1024)
        COR_CUDA(cudaFuncSetAttribute(alignment_path_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    PathArgs p{align, stride_j, stride_row, i_max, j_max, min_score, path, path_ld, dist, t_cap, s_cap};
    alignment_path_kernel<<<n_rows, 128, smem, st>>>(p);
    ++ctx().launches;
    COR_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace cor
