// model.h — device-resident model and workspace layout (internal).
#pragma once
#include "engine.h"
#include "../../include/cor_b200.h"
#include <map>
#include <vector>

namespace cor {

// Weight as a GEMM B operand: [rows = output columns][ld = K], K-major.
struct Mat {
    float* hi = nullptr;   // full fp32 when !split, TF32 head when split
    float* lo = nullptr;
    int rows = 0, ld = 0;
};

struct LstmW {
    Mat Kx;                // [4W][in]   (top cell: the x-rows of the kernel, in = W)
    Mat Kc;                // [4W][D]    (top cell only: the context rows)
    Mat R;                 // [4W][W]
    float* bias = nullptr; // [4W] interleaved 4u+g
};

struct HostWeight { std::vector<int64_t> shape; std::vector<float> data; };

// Per-launch timing of the LSTM-GEMM launches of one handle (bench.py's roofline pass): when enabled, every such
// launch is bracketed by two events on the launching stream; cor_prof_read sums the elapsed times after a synchronize.
struct Prof {
    bool on = false;
    std::vector<cudaEvent_t> ev;     // pairs
    size_t used = 0;
    double flops = 0;                // launches whose row count is known on the host
    int64_t launches = 0;
    // launches whose row count lives on the device (beam): the count is copied into pinned memory
    // right behind the kernel, and the flops are added up in cor_prof_read
    int32_t* rows = nullptr;         // pinned [kMaxDeferred]
    std::vector<double> per_row;     // 2*N*K of the deferred launch
};

}  // namespace cor

struct cor_model {
    cor::Ctx ctx;             // counters, tensor-map cache, test switches of THIS handle
    cor::Prof prof;
    cor_config cfg{};
    int device = 0;
    int W = 0, d = 0, V = 0, Vp = 0, D = 0;
    bool split = false;       // activations and weights kept as TF32 (hi, lo) pairs for the tcgen05 path
    int tc_min_rows = 1;
    int passes = 3;           // tcgen05 passes per product: 3 = fp32 grade; 1 = fp16 heads only (gemm_mode 3)
    bool finalized = false;
    std::map<std::string, std::vector<cor::HostWeight>> host;   // until finalize
    std::vector<void*> allocs;

    float* E_full = nullptr;  // [V][W] for the gather
    cor::Mat E;               // [Vp][W]  logits = h . E^T
    cor::Mat Et;              // [W][Vp]  x = p . E
    std::vector<cor::LstmW> enc_fw, enc_bw;
    std::vector<cor::Mat> bridge_h, bridge_c;
    std::vector<float*> bridge_hb, bridge_cb;
    cor::Mat U;               // [W][D]
    std::vector<cor::LstmW> dec;   // d entries; the last one is the attention cell's LSTMCell
    cor::Mat Wa;              // [W][W]
    cor::Mat EWa;             // [Vp+W][W]  rows of E then rows of W_a: logits | next query in one launch
    cor::Mat PE;              // [4W][Vp]   (E . Kx) of the first decoder cell: z_x = p . PE without the round trip through x
    float* xtab[2] = {nullptr, nullptr};   // [V][4W] (E . Kx) of the first encoder layer, forward / backward scan
    bool use_xtab = false;
    bool fuse_pe = false;     // PE exists (3*Vp <= 4*W: the fused product is the cheaper one)
    cudaStream_t side = nullptr;          // attention runs here, beside the lower-cell GEMMs
    // beam_correct: the previous call's lines nearly all needed (more than half of) the full record quota -- the next call
    // starts with full-quota waves instead of probing every line with a share of the workspace that is sure to run out
    bool beam_full_hint = false;
    bool side_attention = false;          // ... as the register-only kernel that fits on an SM next to the GEMM's CTA
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cor::Mat LowCat, TopCat;  // page decoder (decpage.cu): [E.Kx | R] and [Kx | Kc | R] concatenated along K (depth 2 only)
    unsigned* grid_bar = nullptr;         // grid barrier counter of the persistent recurrence (recur.cu)
    void* loop_graphs = nullptr;          // executable graphs of the decoder loops (graphs.h)
    int64_t* dev_counters = nullptr;      // device: [0] fast-greedy timesteps run by loop graphs since the last cor_launch_count
    int greedy_step_kernels = 0;          // kernels of one fast-greedy timestep (for that count)
    // cor_set_encoding: encoder results handed in by the caller (decode_sequence_*(encoder_outputs=...),
    // decoder_model.predict_on_batch); while set, run_encoder imports them instead of running the encoder graph
    struct { const float* enc_out = nullptr; const float* states = nullptr; int n_rows = 0, t_max = 0; } preset;
    float* b_UW = nullptr;    // [W]
    float* v_a = nullptr;     // [W]
    float b_v = 0.f;

    bool layer_bidi(int n) const { return n == 0 || cfg.deep_bidirectional_encoder; }
    int enc_in_dim(int n) const { return n == 0 ? W : ((cfg.deep_bidirectional_encoder || n == 1) ? 2 * W : W); }
    int enc_out_dim(int n) const { return layer_bidi(n) ? 2 * W : W; }
};

namespace cor {

// Carve of the caller's workspace for (n_rows, t_max, k_alt).  Pure function of its arguments,
// so cor_encode / cor_decode_step / cor_decode_greedy_batch agree on where things live.
struct Ws {
    int B = 0, T = 0, k = 0;
    size_t bytes = 0;
    int32_t* row_T = nullptr; int32_t* idx = nullptr; float* val = nullptr; uint8_t* live = nullptr; uint8_t* done = nullptr;
    float *X = nullptr, *X_lo = nullptr, *Xr = nullptr, *Xr_lo = nullptr;
    std::vector<float*> seq, seq_lo, seqr, seqr_lo;     // per encoder layer [T][B][Dn]; the LAST layer (enc_out) is line-major [B][T][D]
    float* u = nullptr;                                  // [B][T][W] (line-major, like enc_out)
    float* pgA[8] = {};                                  // page decoder operand rows: lowA hi/lo x2, topA hi/lo x2
    float* zx[2] = {nullptr, nullptr};                   // page-sized calls: input term of a layer for all timesteps [T][B][4W]
    float *eh[2][2] = {}, *eh_lo[2][2] = {}, *ec[2] = {};
    std::vector<float*> st_h, st_c;                      // encoder final states per layer [B][W]
    std::vector<float*> dh[2], dh_lo[2], dc;             // decoder state ping-pong
    float* a[2] = {};
    float *x = nullptr, *x_lo = nullptr, *wq = nullptr, *ctx = nullptr, *ctx_lo = nullptr;
    float *logits = nullptr, *p = nullptr, *p_lo = nullptr;
    float *h_lm = nullptr, *h_lm_lo = nullptr, *c_lm = nullptr, *logits_lm = nullptr, *p_lm = nullptr;
    int32_t* o_chars = nullptr; float* o_probs = nullptr; double* o_score = nullptr; int32_t* o_len = nullptr;
    int32_t* o_ties = nullptr;
    float* o_align = nullptr;
    int32_t* loop_state = nullptr;                       // [8] fast-greedy loop graph: [0] timestep, [1] lines finished, [2] live rows, [3] rows before the last compaction, [4] compacted?
    int32_t *g_rl = nullptr, *g_rl2 = nullptr, *g_new = nullptr;      // row -> line, its compacted copy, new row index (or -1) per old row
    uint8_t* g_scratch = nullptr; size_t g_scratch_row = 0;           // compaction scratch: one copy of the per-row decoder state
    float* enc_out() const { return seq.back(); }
    float* enc_out_lo() const { return seq_lo.back(); }
};
int carve(const cor_model* m, int B, int T, int k, int want_align, void* base, Ws* ws);

struct StepIO {
    int R = 0; const int* r_dev = nullptr;
    const int32_t* row_line = nullptr;
    bool first = false;                      // p == 0: x = 0 without the projection
    const float *p = nullptr, *p_lo = nullptr;
    float *x = nullptr, *x_lo = nullptr;
    const float* h_in[8] = {}; const float* h_in_lo[8] = {};
    float* h_out[8] = {}; float* h_out_lo[8] = {};
    const float* c_in[8] = {}; float* c_out[8] = {};
    const float* a_in = nullptr; float* a_out = nullptr; float* align_log = nullptr;
    const int* step_dev = nullptr; int step_off = 0; long long align_stride = 0;   // loop graphs: see AttnArgs
    int align_by_line = 0;
    float *wq = nullptr, *ctx = nullptr, *ctx_lo = nullptr, *logits = nullptr;
    bool wq_ready = false;                   // wq already holds h_in[top].W_a + b (written by the previous step)
    float* wq_next = nullptr;                // if set: the logits launch also writes h_out[top].W_a + b here
    float *h_lm = nullptr, *h_lm_lo = nullptr, *c_lm = nullptr, *logits_lm = nullptr;
};
// One decoder timestep up to the logits (seq2seq.py:416-470); enc_out/u come from `ws`.
int decoder_core(cor_model* m, const Ws& ws, const StepIO& io, cudaStream_t st);
int run_encoder(cor_model* m, const Ws& ws, cudaStream_t st);

}  // namespace cor
