// engine.h — internal interfaces shared by the .cu files of libcor_b200.so
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stddef.h>
#include <string>

namespace cor {

// ------------------------------------------------------------------ errors
void set_error(const std::string& msg);
#define COR_CUDA(expr)                                                                   \
    do {                                                                                 \
        cudaError_t _e = (expr);                                                         \
        if (_e != cudaSuccess) {                                                         \
            cor::set_error(std::string(#expr) + ": " + cudaGetErrorString(_e) + " at " + \
                           __FILE__ + ":" + std::to_string(__LINE__));                   \
            return 1;                                                                    \
        }                                                                                \
    } while (0)
#define COR_CHECK(cond, msg)                                                   \
    do {                                                                       \
        if (!(cond)) {                                                         \
            cor::set_error(std::string(msg) + " (" #cond ") at " + __FILE__ + \
                           ":" + std::to_string(__LINE__));                    \
            return 1;                                                          \
        }                                                                      \
    } while (0)

// ------------------------------------------------------------------ per-handle engine state
// Counters, the tensor-map cache and the test switches belong to ONE cor_model: two handles in a process (two
// models, or one model per GPU) never share them.  A C-ABI entry point installs its handle's Ctx for the duration
// of the call (CtxScope, thread-local); handle-free test hooks (cor_debug_gemm) run on a default Ctx.
struct Ctx {
    int64_t launches = 0;          // kernels launched (bench.py's gpu_launches), graph nodes included
    int64_t in_graph = 0;          // ... of which ran as nodes of a loop graph
    int64_t graph_launches = 0;    // cudaGraphLaunch calls
    // tcgen05 launches by [epilogue: 0 store, 1 LSTM][tile: 0 = 128x32, 1 = 128x256, 2 = CTA pair 256x256], then FFMA [store, LSTM]
    int64_t tc_kind[2][3] = {};
    int64_t ffma_kind[2] = {};
    int force_bn = 0;              // test hook (cor_debug_gemm mode 3): force the narrow tile
    // tile shape of the tcgen05 launches regardless of M: 0 = picked from M and N, 1 = 128x32, 2 = 128x256, 3 = CTA pair 256x256.
    // Set while the branches of a loop graph's SWITCH node are captured (the row count lives on the device there).
    int force_shape = 0;
    void* tc_maps = nullptr;       // tensor-map cache (gemm_tc.cu owns the type)
    void* prof = nullptr;          // per-launch event bracketing (api.cu owns the type)
};
Ctx& ctx();
struct CtxScope {
    Ctx* prev;
    explicit CtxScope(Ctx* c);
    ~CtxScope();
};
void tc_ctx_release(Ctx* c);       // frees the tensor-map cache

// ------------------------------------------------------------------ GEMM
// C[M,N] = sum_s A_s[M,K_s] * B_s[N,K_s]^T   (A row-major, B "K-major": one row per output column)
// Every B operand is a weight prepared at load time; `B_lo` is the fp16 residual of B
// (B = hi + lo with hi = B rounded to fp16; both stored as packed halves behind the float* type), used only by the tcgen05 3-pass path, which also
// wants A split the same way (A = A_hi + A_lo, produced by the epilogue that wrote A).
struct GemmSeg {
    const float* A;  int lda;      // fp32 (FFMA path) or fp16 head (tensor path)
    const float* A_lo;             // fp16 residual or nullptr
    const float* B;  int ldb;
    const float* B_lo;
    int K;
};
constexpr int kMaxSegs = 3;

// plain epilogue: C = act(acc + bias)
struct StoreEpi {
    float* C; int ldc;
    float* C_lo;               // if set: C gets the fp16 head and C_lo the residual
    const float* bias;         // [N] or nullptr
    int act;                   // 0 none, 1 tanh
    // optional second destination: columns >= n_split (a multiple of 32) go to C2[row][col - n_split]
    // as plain fp32 with bias2 (two products of the same A in one launch, e.g. logits | attention query)
    int n_split;               // 0 = off
    float* C2; int ldc2;
    const float* bias2;        // [N - n_split] or nullptr
};

// LSTM-cell epilogue.  N = 4W with columns interleaved n = 4*unit + gate (i,f,c~,o).
// Row r at scan step `step`: active iff step < row_T[r] (row_T == nullptr: always active);
// t = reverse ? row_T[r]-1-step : step is the sequence position the output belongs to.
struct LstmEpi {
    const float* bias;         // [4W] interleaved
    const float* c_in;         // [M][W]
    float* c_out;              // [M][W] (may alias c_in)
    const float* h_prev;       // [M][W] (copied through for inactive rows)
    const float* h_prev_lo;
    float* h_out;              // [M][W]
    float* h_out_lo;           // fp16 residual of h_out or nullptr
    const int32_t* row_T;
    int step;
    int reverse;
    int rows_total;            // B: row stride of the time-major sequence buffers
    float* seq_out;  float* seq_out_lo;  int seq_ld; int seq_col;   // seq_out[t][r][seq_col+u]
    // position of (t, r) in seq_out, in rows of seq_ld elements: t * seq_tm + r * seq_rm.  Time-major (tm = B, rm = 1)
    // for layers that feed another layer's per-timestep GEMM; LINE-major (tm = 1, rm = t_max) for the last layer, whose
    // output is only gathered by window (attention): a line's window is then one contiguous block for the bulk copies.
    // Both 0 = time-major with rows_total.
    int seq_tm, seq_rm;
    int seq_T;                 // time steps the sequence buffers hold (extent of the TMA store maps of the tcgen05 epilogue)
    float* seq_rev;  float* seq_rev_lo;                              // seq_rev[T-1-t][r][seq_col+u]
    const float* resid; const float* resid_lo; int resid_ld;         // residual add into seq_out
    int cross_sum;             // seq value = h[2k]+h[2k+1] for both members of a pair
    int W;
    // optional input term taken from a table instead of a K-segment (first encoder layer): the layer
    // input is x_t = sum_k val_k E[idx_k] (seq2seq.py:237-244), so x_t . Kx = sum_k val_k (E.Kx)[idx_k]
    // with xtab = E.Kx [V][4W] prepared at load time.  idx/val are the call's lines [row][x_tmax][x_k].
    const float* xtab;
    const int32_t* x_idx; const float* x_val; int x_k; int x_tmax;
};

// m_dev != nullptr: the row count is read on the device (beam), M is then the capacity.
int gemm_store_ffma(cudaStream_t st, int M, const int* m_dev, int N, const GemmSeg* segs, int nseg, const StoreEpi& epi);
int gemm_lstm_ffma(cudaStream_t st, int M, const int* m_dev, int N, const GemmSeg* segs, int nseg, const LstmEpi& epi);
// passes: 3 = hi/lo three-pass product (fp32 grade), 1 = fp16 heads only (one pass, ~5e-4 relative per product)
int gemm_store_tc(cudaStream_t st, int M, const int* m_dev, int N, const GemmSeg* segs, int nseg, const StoreEpi& epi, int passes = 3);
int gemm_lstm_tc(cudaStream_t st, int M, const int* m_dev, int N, const GemmSeg* segs, int nseg, const LstmEpi& epi, int passes = 3);
bool tc_available();
void tc_force_narrow(bool on);   // test hook: 128x32 tiles regardless of shape (current Ctx)
// Row counts up to which pick_shape takes the 128x32 tiles / the single-CTA 128x256 tiles for N output columns.
void tc_shape_thresholds(int N, int* narrow_max_rows, int* single_max_rows);

// ------------------------------------------------------------------ persistent recurrence (recur.cu)
// One launch = all T timesteps of one LSTM layer (both directions side by side) for a page-sized batch:
// recurrent weights resident in shared memory, grid-wide barrier between timesteps.
constexpr int kRecurMaxRows = 64;        // lines per page-sized call (processed in blocks of 32 rows)
struct RecurArgs {
    int ndir, W, B, T;
    const float* R_hi[2]; const float* R_lo[2]; int ldr;   // recurrent kernels [4W][ldr] (fp16 head / residual)
    LstmEpi epi[2];              // per direction; step and the h pointers are filled in by the kernel
    float* h[2][2]; float* h_lo[2][2];                     // [dir][ping-pong]: step s reads [s & 1], writes [(s & 1) ^ 1]
    const float* zx[2];          // non-recurrent input term [T][B][4W] (fp32) or nullptr (table gather / none)
    unsigned* barrier;           // one counter in global memory (zeroed by the launcher)
};
bool lstm_recur_supported(int W, int B, int ndir);
int launch_lstm_recur(cudaStream_t st, const RecurArgs& a);

// ------------------------------------------------------------------ persistent page decoder (decpage.cu)
// The whole fast batch-greedy decode of a page (<= 64 lines, depth 2) in one launch; weights resident in
// shared memory, K split over the 8 CTAs of a cluster, partial sums reduced through distributed shared memory.
struct PageDecArgs {
    int B, T, S, W, D, V, Vp, window, eos;
    float b_v;
    const __half *low_hi, *low_lo;       // [4W][Vp+W]   = [E.Kx | R] of the lower cell
    const __half *top_hi, *top_lo;       // [4W][2W+D]   = [Kx | Kc | R] of the attention cell
    const __half *ewa_hi, *ewa_lo; int ewa_ld;   // [Vp+W][W] = [E ; W_a]
    const float *bias_low, *bias_top, *b_UW, *v_a;
    __half *lowA_hi[2], *lowA_lo[2];     // [32][Vp+W]   operand rows [p | h1], ping-pong
    __half *topA_hi[2], *topA_lo[2];     // [32][2W+D]   operand rows [h1' | ctx | h2]
    const __half *h0_hi[2], *h0_lo[2];   // initial h of the two cells [B][W]
    float* c[2];                         // cell states [B][W]
    float *logits, *wq;                  // [B][Vp], [B][W] (wq of step 0 prepared by the caller)
    float* a[2];                         // alignment rows [B][T], ping-pong; a[0] zeroed by the caller
    float* o_align;                      // optional alignment history [S][B][T]
    const float *u, *enc, *enc_lo;       // [B][T][W], [B][T][D] (hi/lo), line-major
    const int32_t* line_T;
    const uint8_t* live; uint8_t* done;
    int32_t* out_chars; float* out_probs; double* score_sum; int32_t* out_len;
    int32_t* out_ties;                   // optional [B] near-tie counter (see GreedyArgs)
    unsigned* barrier;
    unsigned long long* timing;          // optional [S][4] phase stamps (COR_PD_TIMING), else nullptr
};
bool page_decoder_supported(int W, int D, int Vp, int B);
int launch_page_decoder(cudaStream_t st, const PageDecArgs& a);   // 0 ok, 1 error, 2 cooperative launch refused (nothing ran: use the per-timestep path)

// ------------------------------------------------------------------ small kernels (kernels.cu)
// X[t][r][:] = sum_k val*E[idx]; Xr[T_r-1-t][r][:] = same (reversed copy for backward scans)
int launch_embed(cudaStream_t st, const int32_t* idx, const float* val, const int32_t* row_T,
                 int n_rows, int t_max, int k_alt, const float* E, int W, int V,
                 float* X, float* X_lo, float* Xr, float* Xr_lo);

struct AttnArgs {
    const float* wq;           // [R][W]  = h_prev*W_a + b_UW
    const float* u;            // [L][t_max][W]   line-major: a line's window rows are contiguous
    const float* enc;          // [L][t_max][D]  (hi part when enc_lo set)
    const float* enc_lo;
    const float* a_prev;       // [R][t_max]
    const int32_t* row_line;   // [R] or nullptr
    const int32_t* line_T;     // [L]
    const float* v_a; float b_v;
    float* a_out;              // [R][t_max]
    float* ctx; float* ctx_lo; // [R][D]
    float* align_log;          // optional [R][t_max] extra copy (alignment history)
    // timestep on the device (loop graphs): the history row block of this step is align_log + (*step_dev + step_off) * align_stride
    const int* step_dev; int step_off; long long align_stride;
    int align_by_line;         // the history row of a hypothesis row is its LINE's (rows are compacted while lines keep their place)
    int R; const int* r_dev; int L; int t_max; int W; int D; int window; int inclusive;
};
int launch_attention(cudaStream_t st, const AttnArgs& a);
int launch_attention_side(cudaStream_t st, const AttnArgs& a);   // register-only variant that co-resides with the persistent GEMM (overlap mode)

struct GreedyArgs {
    const float* logits;       // [R][V]
    float* p; float* p_lo;     // [R][V] softmax rows (feedback)
    const int32_t* row_T;
    const uint8_t* live;       // [R] line has any input
    int32_t* out_chars; float* out_probs; double* score_sum; int32_t* out_len; uint8_t* done;
    int32_t* out_ties;         // [R] or nullptr: emitted characters whose runner-up is within kTieMargin (relative)
    int R; int V; int step; int s_max; int eos;
    // loop graphs: the timestep is *step_dev + step; *done_count counts the lines that have finished
    const int* step_dev; int32_t* done_count;
    // compacted rows (loop graph with row compaction): row -> line for every per-line array above; live row count on the device
    const int32_t* row_line; const int* r_dev;
};
// An argmax whose runner-up lies within this relative margin is logged as a near-tie: fp32-grade implementations
// (this one, Keras/TF, the numpy oracle) may legitimately pick either, so parity tests report such lines separately.
constexpr float kTieMargin = 1e-6f;
// the beam's threshold cuts (relative beam threshold, rejection threshold) compare probabilities that agree with the oracle's to ~1e-6 ABSOLUTE (small
// probabilities: far more than 1e-6 relative), so its near-tie log adds an absolute term
constexpr float kTieAbs = 2e-6f;
int launch_softmax_greedy(cudaStream_t st, const GreedyArgs& a);
int launch_softmax(cudaStream_t st, const float* logits, float* p, float* p_lo, int R, const int* r_dev, int V);
int launch_split(cudaStream_t st, const float* x, float* hi, float* lo, size_t n);
int launch_join(cudaStream_t st, const float* hi, const float* lo, float* out, size_t n);
int launch_join_lt(cudaStream_t st, const float* hi, const float* lo, float* out, int L, int T, int D);   // line-major (hi, lo) -> time-major fp32
int launch_split_tl(cudaStream_t st, const float* in, float* hi, float* lo, int L, int T, int D);         // time-major fp32 -> line-major (hi, lo)
int launch_finish_scores(cudaStream_t st, double* score_sum, const int32_t* len, int R);
int launch_live_rows(cudaStream_t st, const int32_t* idx, const float* val, int n_rows, int t_max, int k_alt, uint8_t* live);


// ------------------------------------------------------------------ alignpath.cu
// Viterbi re-alignment path per line (wrapper/transcode.py:279-349); a(row,j,i) = align[j*stride_j + row*stride_row + i]
int launch_alignment_paths(cudaStream_t st, const float* align, long long stride_j, long long stride_row,
                           int n_rows, int t_cap, int s_cap, const int32_t* i_max, const int32_t* j_max,
                           double min_score, int32_t* path, int path_ld, double* dist);

}  // namespace cor
