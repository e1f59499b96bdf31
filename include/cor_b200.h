/* cor_b200.h — C ABI of the B200-native cor-asv-ann neural hot path.
 *
 * The reference has no FFI of its own: its boundary to the numeric engine is
 * Keras `Model.predict_on_batch` on three graphs built by
 * `Sequence2Sequence.configure` (ocrd_cor_asv_ann/lib/seq2seq.py:389,403,477).
 * Each entry point below names the reference interface it replaces.
 *
 * Conventions: C linkage; every call returns 0 on success, non-zero on error with
 * the message in cor_last_error(); no exceptions cross the boundary.  The caller
 * owns every host and device buffer it passes; the model handle owns only the
 * repacked weights.  One handle per device, not thread-safe per handle.
 * `stream` is a cudaStream_t passed as void* (0 = legacy default stream).
 *
 * Device tensor layouts (fp32 unless stated):
 *   sequences are TIME-MAJOR: seq[t][row][feature];
 *   per-row state is row-major: h[row][W];
 *   `row_T[row]` is the padded length T of the reference batch the row belongs to
 *   (padding is NOT masked in the reference, SURVEY.md A.6 item 1, so T is part of
 *   the input); rows of several reference batches may share one call.
 */
#ifndef COR_B200_H
#define COR_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cor_model cor_model;

/* Topology switches — seq2seq.py:108-179 (width, depth, voc_size, residual_connections,
 * deep_bidirectional_encoder, bridge_dense, lm_predict) and attention.py:515 (window_width). */
typedef struct cor_config {
    int32_t width;
    int32_t depth;
    int32_t voc_size;
    int32_t residual_connections;
    int32_t deep_bidirectional_encoder;
    int32_t bridge_dense;
    int32_t lm_predict;
    int32_t window_width;      /* 5 in the reference (seq2seq.py:347,452) */
    int32_t eos_index;         /* index of '\n' in the mapping (1 when built by map_files, seq2seq.py:580-586) */
    int32_t gemm_mode;         /* 0 auto, 1 force FFMA fp32, 2 force tcgen05 3-pass fp16 hi/lo (fp32 grade), 3 tcgen05 single-pass fp16
                                * (the north star's reduced-precision variant: looser stated tolerance, never picked by 0) */
} cor_config;

/* Input lines of one call — replaces the dense (B,T,V) array of
 * vectorize_lines (seq2seq.py:1020-1119) by its non-zeros.
 * idx/val are [n_rows][t_max][k_alt]; idx < 0 marks "no entry" (true-zero padding);
 * val == NULL means all ones (plain strings).  Host or device pointers as the call says. */
typedef struct cor_lines {
    int32_t n_rows;
    int32_t t_max;
    int32_t k_alt;
    const int32_t* row_T;      /* [n_rows] */
    const int32_t* idx;
    const float*   val;
} cor_lines;

/* Beam parameters — seq2seq.py:111,162-169. */
typedef struct cor_beam_params {
    int32_t batch_size;          /* 256 */
    int32_t beam_width_in;       /* 15  */
    int32_t beam_width_out;      /* 16  */
    int32_t n_best;              /* final hypotheses returned per line, best first, 0 or 1 = the best one only, at most
                                  * COR_BEAM_NBEST_MAX.  decode_sequence_beam is a generator over ALL finals
                                  * (seq2seq.py:1538-1544); its callers draw one.  With n_best = N > 1 every per-line output
                                  * of the beam entry points has N slots: out_chars/out_probs [n_rows][N][2*t_max],
                                  * out_scores/out_len [n_rows][N], out_align [n_rows][N][2*t_max][t_max]; out_len = 0
                                  * marks a slot beyond the number of finals found. */
    double  beam_threshold_in;   /* 0.2 */
    double  rejection_threshold; /* 0.3 class default / 0.5 CLI default; 0 disables */
} cor_beam_params;

const char* cor_last_error(void);
int cor_abi_version(void);

/* ---- model lifetime: replaces configure()+load_weights()+_resync_decoder()
 *      (seq2seq.py:190-489, 1164-1174, 526-528). */
int cor_model_create(const cor_config* cfg, int device, cor_model** out);
/* One Keras weight by layer name and position inside the layer (SURVEY.md A.5 order), in the
 * Keras (in, out) layout; the Keras -> kernel repack (transposition to K-major, gate
 * interleave, fp16 hi/lo split) happens here.  LSTM layers must be in LSTM (not CuDNNLSTM)
 * format: the host class converts CuDNN-format layers (bias 8W) the way Keras does on load. */
int cor_model_set_weight(cor_model* m, const char* keras_layer, int index,
                         const float* host, const int64_t* shape, int rank);
int cor_model_finalize(cor_model* m);
void cor_model_destroy(cor_model* m);

/* ---- workspace: one caller-allocated device buffer per call group. */
int cor_workspace_bytes(const cor_model* m, int n_rows, int t_max, int k_alt,
                        int want_alignments, size_t* bytes);

/* ---- encoder: replaces encoder_model.predict_on_batch (seq2seq.py:811,1231,1306,1383).
 * `lines` holds DEVICE pointers.  Results stay inside `workspace`; the optional outputs
 * receive copies: enc_out [t_max][n_rows][D], states [2*depth][n_rows][W] (h1,c1,...,hd,cd).
 * want_alignments != 0 reserves the alignment history of a following cor_decode_greedy_batch
 * (workspace size from cor_workspace_bytes with the same flag); it is the caller's explicit choice,
 * never inferred from the size of the buffer. */
int cor_encode(cor_model* m, const cor_lines* lines, int want_alignments, void* workspace, size_t workspace_bytes,
               float* enc_out, float* states, void* stream);

/* ---- encoder results from the caller: the reference lets every decoder take `encoder_outputs` instead of the
 * input (decode_sequence_greedy/beam(encoder_outputs=...), seq2seq.py:1288, 1356; correct_lines itself encodes a
 * batch once and hands slices to the per-line decoders, :811-825) and its decoder model takes the attended
 * sequence as an input (:477-480).  While a preset is installed, every entry point that would run the encoder
 * (cor_encode, cor_correct_greedy_host, cor_correct_beam_host, cor_decode_beam_batch) imports these DEVICE arrays
 * instead: enc_out [t_max][n_rows][D], states [2*depth][n_rows][W], the layouts cor_encode returns; the call's
 * geometry must match, its lines still provide the rejection candidates of the beam and the empty-line flags.
 * NULL, NULL removes the preset.  The arrays must stay valid until then. */
int cor_set_encoding(cor_model* m, int n_rows, int t_max, const float* enc_out, const float* states);

/* ---- one decoder timestep: replaces decoder_model.predict_on_batch (seq2seq.py:1245,1321,1428).
 * Rows are hypotheses; row_line[r] (NULL = identity) names the encoded line in `workspace`
 * each hypothesis attends to (the reference broadcasts a batch-1 enc_out, seq2seq.py:1428).
 * p_in [R][V] previous softmax rows (zeros at the first step); h_in/c_in [depth][R][W];
 * a_in [R][t_max].  Outputs p_out [R][V], lm_out [R][V] (only with lm_predict, else NULL),
 * h_out/c_out [depth][R][W], a_out [R][t_max].  In and out must not alias. */
int cor_decode_step(cor_model* m, void* workspace, size_t workspace_bytes,
                    int n_hyp, const int32_t* row_line,
                    const float* p_in, const float* h_in, const float* c_in, const float* a_in,
                    float* p_out, float* lm_out, float* h_out, float* c_out, float* a_out,
                    void* stream);

/* ---- fast mode: replaces decode_batch_greedy (seq2seq.py:1215-1286) for rows already
 * encoded into `workspace` by cor_encode.  Runs 2*row_T[row] steps per row.
 * out_chars int32 [n_rows][2*t_max], out_probs [n_rows][2*t_max], out_scores double [n_rows]
 * (mean -log p), out_len int32 [n_rows], out_align [2*t_max][n_rows][t_max] or NULL. All DEVICE.
 * out_ties int32 [n_rows] or NULL: per line, the number of emitted characters whose runner-up probability
 * lies within 1e-6 relative of the winner's -- the near-tie argmaxes the north star wants logged (an
 * fp32-grade implementation may legitimately pick either; parity reports such lines separately). */
int cor_decode_greedy_batch(cor_model* m, void* workspace, size_t workspace_bytes,
                            int32_t* out_chars, float* out_probs, double* out_scores,
                            int32_t* out_len, int32_t* out_ties, float* out_align, void* stream);

/* ---- the call a host program makes: HOST lines in, HOST results out; host<->device copies
 * happen inside, on `stream`, and the call returns after the results have landed.
 * Replaces correct_lines(fast=True) (seq2seq.py:782-808). */
int cor_correct_greedy_host(cor_model* m, const cor_lines* host_lines,
                            void* workspace, size_t workspace_bytes,
                            int32_t* out_chars, float* out_probs, double* out_scores,
                            int32_t* out_len, int32_t* out_ties, float* out_align, void* stream);

/* ---- beam mode: replaces the per-line loop of correct_lines(fast=False, greedy=False)
 * over decode_sequence_beam (seq2seq.py:823-836, 1356-1544) for all rows at once.
 * Workspace size comes from cor_beam_workspace_bytes.  Per row: the best final hypothesis
 * (out_len = 0 when the search yields none, i.e. the reference's StopIteration fallback),
 * out_scores = cum_cost/(len-1), out_align compact [n_rows][2*t_max][t_max] or NULL;
 * stats int64 [n_rows][8] (COR_BEAM_STATS) = [0] iterations, [1] hypothesis-steps, [2] finals, [3] open nodes left,
 * [4] expanded hypotheses whose two best scores are EXACTLY equal (the argsort tie the reference leaves to numpy),
 * [5] near-tie ordering decisions (candidate order, beam-threshold cut or pop boundary within 1e-6 relative): where
 *     an fp32-grade implementation may legitimately search differently -- the ties the north star wants logged,
 * [6] checksum of which hypotheses (length, character) were expanded in which iteration (uint64 bit pattern),
 * [7] peak fringe length.  [0] is -1 for a line whose arena overflowed (internal, retried) and -2 for an empty line.
 * All HOST. */
#define COR_BEAM_STATS 8
#define COR_BEAM_NBEST_MAX 64
int cor_beam_workspace_bytes(const cor_model* m, int n_rows, int t_max, int k_alt,
                             const cor_beam_params* bp, size_t budget_bytes, size_t* bytes);
int cor_correct_beam_host(cor_model* m, const cor_lines* host_lines, const cor_beam_params* bp,
                          void* workspace, size_t workspace_bytes,
                          int32_t* out_chars, float* out_probs, double* out_scores,
                          int32_t* out_len, float* out_align, int64_t* stats, void* stream);

/* Same search with the line arrays and every result array in DEVICE memory (inputs resident in HBM);
 * only the per-line overflow flags cross to the host between waves. */
int cor_decode_beam_batch(cor_model* m, const cor_lines* device_lines, const cor_beam_params* bp,
                          void* workspace, size_t workspace_bytes,
                          int32_t* out_chars, float* out_probs, double* out_scores,
                          int32_t* out_len, float* out_align, int64_t* stats, void* stream);

/* ---- hard re-alignment path per line: replaces the per-line Python call
 * `_alignment2path(alignment, input_line_len, len(output_line), 1./voc_size)` of the OCR-D wrapper
 * (wrapper/transcode.py:279-349, called at :126) — the only consumer of the decoders' alignments —
 * so that (t_max+1) ints and one double per line cross to the host instead of S*T floats.
 * `align` is the DEVICE alignment array of cor_decode_greedy_batch (stride_step = n_rows*t_max,
 * stride_row = t_max) or of cor_decode_beam_batch (stride_step = t_max, stride_row = 2*t_max*t_max):
 * a(row, step j, position i) = align[j*stride_step + row*stride_row + i].
 * in_len[row] = input_line_len (<= t_max), out_len[row] = len(output_line) (<= s_max); a negative
 * out_len means "identity matrix of that size" (the beam fallback of correct_lines, seq2seq.py:835).
 * min_score is compared in double, like np.float32 > python float under the reference's numpy.
 * out_path int32 [n_rows][path_ld], path_ld >= t_max+1: out_path[row][i] = output position of input
 * position i, -1 where the reference's dict has no key; entry [in_len] = out_len, entry [0] = 0.
 * out_dist double [n_rows] = sum of (1 - score) along the path.  All pointers DEVICE. */
int cor_alignment_paths(const float* align, int64_t stride_step, int64_t stride_row,
                        int n_rows, int t_max, int s_max, const int32_t* in_len, const int32_t* out_len,
                        double min_score, int32_t* out_path, int path_ld, double* out_dist, void* stream);

/* ---- row T (BASELINE.json config 5): one data-parallel training step.  Replaces encoder_decoder_model.train_on_batch
 * (lib/keras_train.py:195) on the training graph of seq2seq.py:316-390 compiled by _recompile (:491-497): teacher
 * forcing, Keras categorical cross-entropy with temporal weights, embedding regulariser (:530-553), Adam(clipnorm=5).
 * A trainer works on fp32 master weights, gradients and Adam moments as FLAT device vectors; cor_trainer_slot says where a
 * Keras weight (layer name, index in Keras order, Keras (in, out) layout) lives in them.  Topologies: any depth and
 * width, residual_connections; not deep_bidirectional_encoder / bridge_dense / LM loss. */
typedef struct cor_trainer cor_trainer;
typedef struct cor_train_batch {
    int32_t n_rows, t_src, k_alt, t_dec;
    const int32_t* src_idx;    /* DEVICE [n_rows][t_src][k_alt] source non-zeros like cor_lines (idx < 0: none) */
    const float*   src_val;    /* DEVICE same shape, or NULL = ones */
    const int32_t* dec_in;     /* DEVICE [n_rows][t_dec] decoder input index per step, -1 = zero vector (start symbol, padding) */
    const int32_t* dec_out;    /* DEVICE [n_rows][t_dec] target index per step, -1 = padding (temporal weight 0) */
    const float*   mask_enc;   /* DEVICE [depth-1][2*width] dropout mask (already scaled by 1/keep) on the output of encoder layer n, or NULL */
    const float*   mask_dec;   /* DEVICE [depth-1][width] the same for the decoder layers below the top, or NULL */
    const float*   mask_cell;  /* DEVICE [n_rows][width + D] input dropout mask of the attention cell's LSTMCell (seq2seq.py:345), or NULL */
} cor_train_batch;
int  cor_trainer_create(const cor_config* cfg, int device, cor_trainer** out);
void cor_trainer_destroy(cor_trainer* t);
int  cor_trainer_param_count(const cor_trainer* t, int64_t* n_floats);
int  cor_trainer_slot(const cor_trainer* t, const char* layer, int index, int64_t* offset, int64_t* count);
/* the four flat DEVICE vectors (n_floats each, caller-owned: the host side allocates them so that it can all-reduce the
 * gradient and read / write the weights): parameters, gradient, Adam first and second moments */
int  cor_trainer_bind(cor_trainer* t, float* params, float* grads, float* adam_m, float* adam_v);
int  cor_trainer_reset_optimizer(cor_trainer* t);
int64_t cor_trainer_launch_count(cor_trainer* t, int reset);     /* graph nodes (kernels, copies) the steps ran since the last reset */
/* Forward + backward of this rank's batch.  Leaves in the gradient vector the gradient of the SUM of the weighted
 * cross-entropies (not divided by the number of target steps) and returns loss_out[0] = that sum, loss_out[1] = the
 * number of target steps: the caller all-reduces the vector and the pair (NCCL), then every rank applies. */
int  cor_train_forward_backward(cor_trainer* t, const cor_train_batch* batch, double* loss_out, void* stream);
/* gradient /= n_target_steps (the GLOBAL count), + regulariser gradient, global-norm clip, Adam (lr <= 0: no update).
 * stats_out[0] = regulariser value, stats_out[1] = global gradient norm before clipping. */
int  cor_train_apply(cor_trainer* t, double n_target_steps, int with_regulariser, float lr, float clipnorm,
                     double* stats_out, void* stream);

/* ---- instrumentation for bench.py and the tests, all PER HANDLE: kernels launched since the last
 * reset; launches by GEMM kernel variant; device time accumulated by the LSTM-GEMM launches between
 * cor_prof_enable(m, 1) and cor_prof_enable(m, 0). */
int64_t cor_launch_count(cor_model* m, int reset);
/* The decoder loops run as device-side WHILE graphs (one graph launch = all iterations): cor_launch_count counts the
 * KERNELS that ran (graph nodes included, counted on the device), this one the launch calls the host made (kernel
 * launches + graph launches) -- the number that costs launch latency.  Same reset semantics. */
int64_t cor_host_launch_count(cor_model* m, int reset);
/* out[8]: tcgen05 launches with the plain epilogue on 128x32 / 128x256 / CTA-pair 256x256 tiles, the same three
 * with the fused LSTM epilogue, then FFMA launches plain / LSTM.  Lets a parity test assert WHICH kernel it checked. */
int cor_debug_counters(cor_model* m, int64_t* out, int reset);
/* test hook: C[M,N] = A[M,K] . B[N,K]^T (device fp32) through one GEMM kernel: 1 = FFMA, 2 = tcgen05 3-pass fp16 hi/lo, 3 = the same with the page-sized 128x32 tiles forced */
int cor_debug_gemm(int mode, int M, int N, int K, const float* A, const float* B, float* C, void* stream);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  C[M,N] = A[M,K] . B[N,K]^T (device
 * fp32, K a multiple of 64) through the tcgen05 3-pass kernel like cor_debug_gemm(2, ...), but the fp16 (head, residual)
 * copies of A and B go into a 16 MiB file-scope array in device memory (csrc/noalloc.cu) instead of cudaMalloc'ed
 * buffers: no cudaMalloc / cudaFree / allocating library call.  4 * (M + N) * K bytes (each of the four parts rounded
 * up to 256) must fit in cor_gemm_noalloc_max_bytes() = 16 MiB, else the call fails with an error and launches nothing.
 * Stream-ordered, does not synchronise.  Every call on a device shares the one array: calls that overlap on the GPU
 * (different streams without a dependency) race; calls on one stream are safe. */
int cor_gemm_noalloc(int M, int N, int K, const float* A, const float* B, float* C, void* stream);
size_t cor_gemm_noalloc_max_bytes(void);
int cor_prof_enable(cor_model* m, int on);   /* on: reset and start bracketing every LSTM-GEMM launch with events */
int cor_prof_read(cor_model* m, double* total_ms, int64_t* launches, double* flops);   /* synchronizes the device */

#ifdef __cplusplus
}
#endif
#endif /* COR_B200_H */
