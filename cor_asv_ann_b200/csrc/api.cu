// api.cu — C ABI (include/cor_b200.h): model lifetime, weight repack, workspace carve,
// encoder scan, one decoder timestep, greedy driver.
#include "model.h"
#include "epilogue.cuh"
#include "graphs.h"
#include <cuda_fp16.h>
#include <string.h>
#include <stdlib.h>
#include <math.h>
#include <algorithm>
#include <thread>

namespace cor {

static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }

// ------------------------------------------------------------------ per-handle state (engine.h)
static Ctx g_default_ctx;                       // handle-free test hooks only
static thread_local Ctx* t_ctx = nullptr;
Ctx& ctx() { return t_ctx ? *t_ctx : g_default_ctx; }
CtxScope::CtxScope(Ctx* c) : prev(t_ctx) { t_ctx = c; }
CtxScope::~CtxScope() { t_ctx = prev; }

// ------------------------------------------------------------------ per-launch timing (bench.py roofline pass)
constexpr size_t kMaxDeferred = 65536;
static Prof g_no_prof;
static Prof& prof() { return ctx().prof ? *static_cast<Prof*>(ctx().prof) : g_no_prof; }

static void prof_begin(cudaStream_t st) {
    if (!prof().on) return;
    if (prof().used + 2 > prof().ev.size()) {
        if (prof().ev.size() >= 60000) return;
        for (int i = 0; i < 2; ++i) { cudaEvent_t e; cudaEventCreate(&e); prof().ev.push_back(e); }
    }
    cudaEventRecord(prof().ev[prof().used], st);
}
static void prof_end(cudaStream_t st, int M, const int* m_dev, int N, const GemmSeg* s, int n, int k_table) {
    if (!prof().on || prof().used + 2 > prof().ev.size()) return;
    cudaEventRecord(prof().ev[prof().used + 1], st);
    prof().used += 2;
    // algorithmic K: the segments plus the input term a table gather stands in for (SURVEY.md 8d counts x.Kx)
    double k = k_table;
    for (int i = 0; i < n; ++i) k += s[i].K;
    if (m_dev && prof().rows && prof().per_row.size() < kMaxDeferred) {
        cudaMemcpyAsync(prof().rows + prof().per_row.size(), m_dev, 4, cudaMemcpyDeviceToHost, st);
        prof().per_row.push_back(2.0 * N * k);
    } else {
        prof().flops += 2.0 * M * N * k;
    }
    prof().launches += 1;
}

// GEMM dispatch: tcgen05 3xFP16 whenever the model keeps split operands and the shape is TMA-aligned
// (gemm_tc.cu picks pair / single / narrow tiles from M and N), the exact FFMA kernel otherwise.
// Both consume the same operand representation.
static bool tc_shape_ok(const GemmSeg* s, int n) {
    for (int i = 0; i < n; ++i) if (s[i].K % 64 != 0 || s[i].lda % 8 != 0 || s[i].ldb % 8 != 0) return false;
    return true;
}
static int gemm_store(cor_model* m, cudaStream_t st, int M, const int* m_dev, int N, const GemmSeg* s, int n, const StoreEpi& e) {
    if (m->split && M >= m->tc_min_rows && tc_shape_ok(s, n) && tc_available()) return gemm_store_tc(st, M, m_dev, N, s, n, e, m->passes);
    return gemm_store_ffma(st, M, m_dev, N, s, n, e);
}
static int gemm_lstm(cor_model* m, cudaStream_t st, int M, const int* m_dev, int N, const GemmSeg* s, int n, const LstmEpi& e) {
    prof_begin(st);
    int rc;
    if (m->split && M >= m->tc_min_rows && tc_shape_ok(s, n) && tc_available()) rc = gemm_lstm_tc(st, M, m_dev, N, s, n, e, m->passes);
    else rc = gemm_lstm_ffma(st, M, m_dev, N, s, n, e);
    prof_end(st, M, m_dev, N, s, n, e.xtab ? e.W : 0);
    return rc;
}

// ------------------------------------------------------------------ loop graphs (graphs.h)
struct LoopGraphCache {
    std::vector<LoopGraph> entries;
    cudaStream_t capture[2] = {nullptr, nullptr};
    uint64_t tick = 0;
};
constexpr size_t kGraphCacheSize = 16;
static LoopGraphCache* graph_cache(cor_model* m) {
    if (!m->loop_graphs) m->loop_graphs = new LoopGraphCache();
    return static_cast<LoopGraphCache*>(m->loop_graphs);
}
static void drop_graph(LoopGraph& g) {
    if (g.exec) cudaGraphExecDestroy(g.exec);
    if (g.graph) cudaGraphDestroy(g.graph);
    g.exec = nullptr; g.graph = nullptr;
}
cudaStream_t loop_capture_stream(cor_model* m, int which) {
    LoopGraphCache* c = graph_cache(m);
    if (!c->capture[which] && cudaStreamCreateWithFlags(&c->capture[which], cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    return c->capture[which];
}
int loop_graph_switch(cudaStream_t cs, int n, cudaGraphConditionalHandle sw, cudaGraph_t* branches) {
    cudaStreamCaptureStatus status; cudaGraph_t cur = nullptr; const cudaGraphNode_t* deps = nullptr; size_t ndeps = 0;
    COR_CUDA(cudaStreamGetCaptureInfo(cs, &status, nullptr, &cur, &deps, &ndeps));
    COR_CHECK(status == cudaStreamCaptureStatusActive && cur, "stream is not capturing");
    cudaGraphNodeParams cp{};
    cp.type = cudaGraphNodeTypeConditional;
    cp.conditional.handle = sw; cp.conditional.type = n == 0 ? cudaGraphCondTypeIf : cudaGraphCondTypeSwitch; cp.conditional.size = (unsigned)std::max(n, 1);
    cudaGraphNode_t node;
    COR_CUDA(cudaGraphAddNode(&node, cur, deps, ndeps, &cp));
    for (int i = 0; i < std::max(n, 1); ++i) branches[i] = cp.conditional.phGraph_out[i];
    COR_CUDA(cudaStreamUpdateCaptureDependencies(cs, &node, 1, cudaStreamSetCaptureDependencies));
    return 0;
}
bool loop_graphs_enabled() {
    const char* off = getenv("COR_NO_GRAPH");
    return !prof().on && !(off && off[0] && off[0] != '0');
}
int loop_graph_lookup(cor_model* m, const std::vector<uint64_t>& key, LoopGraph** out) {
    LoopGraphCache* c = graph_cache(m);
    ++c->tick;
    for (auto& e : c->entries)
        if (e.key == key) { e.used = c->tick; *out = &e; return 0; }
    if (c->entries.size() < kGraphCacheSize) {
        c->entries.reserve(kGraphCacheSize);
        c->entries.emplace_back();
        *out = &c->entries.back();
    } else {
        LoopGraph* lru = &c->entries[0];
        for (auto& e : c->entries) if (e.used < lru->used) lru = &e;
        // an executable graph may still be running on a caller's stream: drain the device before destroying one
        COR_CUDA(cudaDeviceSynchronize());
        drop_graph(*lru);
        *out = lru;
    }
    (*out)->key = key; (*out)->used = c->tick; (*out)->body_kernels = 0;
    return 0;
}
int loop_graph_begin(LoopGraph* g, const cudaKernelNodeParams& first, cudaGraphConditionalHandle* cond, cudaGraph_t* body) {
    // the handle is an argument of `first` (through its kernelParams, which point at *cond): create it before the node
    drop_graph(*g);
    COR_CUDA(cudaGraphCreate(&g->graph, 0));
    COR_CUDA(cudaGraphConditionalHandleCreate(cond, g->graph, 0, 0));
    cudaGraphNode_t n_first, n_while;
    COR_CUDA(cudaGraphAddKernelNode(&n_first, g->graph, nullptr, 0, &first));
    cudaGraphNodeParams cp{};
    cp.type = cudaGraphNodeTypeConditional;
    cp.conditional.handle = *cond; cp.conditional.type = cudaGraphCondTypeWhile; cp.conditional.size = 1;
    COR_CUDA(cudaGraphAddNode(&n_while, g->graph, &n_first, 1, &cp));
    *body = cp.conditional.phGraph_out[0];
    return 0;
}
void loop_graphs_release(cor_model* m) {
    if (!m || !m->loop_graphs) return;
    LoopGraphCache* c = static_cast<LoopGraphCache*>(m->loop_graphs);
    for (auto& e : c->entries) drop_graph(e);
    for (int i = 0; i < 2; ++i) if (c->capture[i]) cudaStreamDestroy(c->capture[i]);
    delete c;
    m->loop_graphs = nullptr;
}

// ------------------------------------------------------------------ weights
static int dev_alloc(cor_model* m, size_t bytes, void** out) {
    COR_CUDA(cudaMalloc(out, bytes ? bytes : 4));
    m->allocs.push_back(*out);
    return 0;
}

// fp16 (hi, lo) split of a weight matrix on the host (same arithmetic as f16_split on the device)
static int host_split(const std::vector<float>& x, std::vector<__half>& hi, std::vector<__half>& lo) {
    hi.resize(x.size()); lo.resize(x.size());
    for (size_t i = 0; i < x.size(); ++i) {
        COR_CHECK(fabsf(x[i]) < 6.0e4f, "weight outside the fp16 operand range: use gemm_mode 1 (exact FFMA) for this model");
        hi[i] = __float2half_rn(x[i]);
        lo[i] = __float2half_rn(x[i] - __half2float(hi[i]));
    }
    return 0;
}

// upload a K-major matrix [rows][ld]
static int upload_mat(cor_model* m, const std::vector<float>& data, int rows, int ld, Mat* out) {
    out->rows = rows; out->ld = ld;
    const size_t bytes = data.size() * sizeof(float);
    if (m->split) {
        std::vector<__half> hi, lo;
        if (host_split(data, hi, lo)) return 1;
        const size_t hb = data.size() * sizeof(__half);
        if (dev_alloc(m, hb, (void**)&out->hi) || dev_alloc(m, hb, (void**)&out->lo)) return 1;
        COR_CUDA(cudaMemcpy(out->hi, hi.data(), hb, cudaMemcpyHostToDevice));
        COR_CUDA(cudaMemcpy(out->lo, lo.data(), hb, cudaMemcpyHostToDevice));
    } else {
        if (dev_alloc(m, bytes, (void**)&out->hi)) return 1;
        COR_CUDA(cudaMemcpy(out->hi, data.data(), bytes, cudaMemcpyHostToDevice));
    }
    return 0;
}
static int upload_vec(cor_model* m, const std::vector<float>& data, float** out) {
    if (dev_alloc(m, data.size() * sizeof(float), (void**)out)) return 1;
    COR_CUDA(cudaMemcpy(*out, data.data(), data.size() * sizeof(float), cudaMemcpyHostToDevice));
    return 0;
}

// Keras (in, out) kernel rows [r0, r1) -> K-major [out][r1-r0]
static std::vector<float> transpose_rows(const HostWeight& w, int r0, int r1) {
    const int out = (int)w.shape[1], k = r1 - r0;
    std::vector<float> t((size_t)out * k);
    for (int r = r0; r < r1; ++r)
        for (int c = 0; c < out; ++c) t[(size_t)c * k + (r - r0)] = w.data[(size_t)r * out + c];
    return t;
}
// Keras LSTM kernel (in, 4W), columns g*W+u -> K-major [4u+g][in rows r0..r1)
static std::vector<float> lstm_rows(const HostWeight& w, int W, int r0, int r1) {
    const int k = r1 - r0;
    std::vector<float> t((size_t)4 * W * k);
    for (int r = r0; r < r1; ++r)
        for (int g = 0; g < 4; ++g)
            for (int u = 0; u < W; ++u) t[(size_t)(4 * u + g) * k + (r - r0)] = w.data[(size_t)r * 4 * W + g * W + u];
    return t;
}
static std::vector<float> lstm_bias(const HostWeight& b, int W) {
    std::vector<float> t(4 * W);
    for (int g = 0; g < 4; ++g)
        for (int u = 0; u < W; ++u) t[4 * u + g] = b.data[g * W + u];
    return t;
}

// out[a][b] = sum_u A[a][u] * B[b][u]  (fp64 accumulation, one rounding), rows of A spread over host threads
static std::vector<float> product_f64(const float* A, int na, const float* B, int nb, int K, int ld_out) {
    std::vector<float> out((size_t)na * ld_out, 0.f);
    unsigned nt = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    if ((size_t)na * nb * K < (1u << 22)) nt = 1;
    auto work = [&](int a0, int a1) {
        for (int a = a0; a < a1; ++a) {
            const float* ar = A + (size_t)a * K;
            for (int b = 0; b < nb; ++b) {
                const float* br = B + (size_t)b * K;
                double s = 0.0;
                for (int u = 0; u < K; ++u) s += (double)ar[u] * (double)br[u];
                out[(size_t)a * ld_out + b] = (float)s;
            }
        }
    };
    std::vector<std::thread> th;
    const int per = (na + (int)nt - 1) / (int)nt;
    for (unsigned t = 0; t < nt; ++t) {
        const int a0 = (int)t * per, a1 = std::min(na, a0 + per);
        if (a0 < a1) th.emplace_back(work, a0, a1);
    }
    for (auto& t : th) t.join();
    return out;
}

static int need(cor_model* m, const std::string& layer, size_t count, const std::vector<HostWeight>** out) {
    auto it = m->host.find(layer);
    COR_CHECK(it != m->host.end(), ("missing weights for layer " + layer).c_str());
    COR_CHECK(it->second.size() == count, ("wrong number of weights for layer " + layer).c_str());
    for (auto& w : it->second) COR_CHECK(!w.data.empty(), ("unset weight in layer " + layer).c_str());
    *out = &it->second;
    return 0;
}
static int check_shape(const HostWeight& w, int64_t a, int64_t b, const std::string& what) {
    bool ok = b < 0 ? (w.shape.size() == 1 && w.shape[0] == a) : (w.shape.size() == 2 && w.shape[0] == a && w.shape[1] == b);
    COR_CHECK(ok, ("bad shape for " + what).c_str());
    return 0;
}
static int build_lstm(cor_model* m, const HostWeight* w3, int in_dim, int x_rows, LstmW* out, const std::string& what) {
    const int W = m->W;
    if (check_shape(w3[0], in_dim, 4 * W, what + " kernel") || check_shape(w3[1], W, 4 * W, what + " recurrent") ||
        check_shape(w3[2], 4 * W, -1, what + " bias")) return 1;
    if (upload_mat(m, lstm_rows(w3[0], W, 0, x_rows), 4 * W, x_rows, &out->Kx)) return 1;
    if (x_rows < in_dim && upload_mat(m, lstm_rows(w3[0], W, x_rows, in_dim), 4 * W, in_dim - x_rows, &out->Kc)) return 1;
    if (upload_mat(m, lstm_rows(w3[1], W, 0, W), 4 * W, W, &out->R)) return 1;
    return upload_vec(m, lstm_bias(w3[2], W), &out->bias);
}

// ------------------------------------------------------------------ workspace
struct Carver {
    char* base; size_t off = 256;   // the first 256 bytes hold the WsHeader
    template <typename T> T* take(size_t n) {
        off = (off + 255) & ~size_t(255);
        T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
        off += n * sizeof(T);
        return p;
    }
};

// Page-sized calls reserve room for the persistent recurrence.  Must be a pure function of (model, B): the
// workspace size is computed without a device (and before the first launch).
static bool recur_layout(const cor_model* m, int B) {
    return m->split && m->passes == 3 && B <= kRecurMaxRows && m->W % 16 == 0 && m->W <= 768 && !getenv("COR_NO_RECUR");
}

int carve(const cor_model* m, int B, int T, int k, int want_align, void* base, Ws* ws) {
    COR_CHECK(B > 0 && T > 0 && k > 0, "empty call geometry");
    Carver c{(char*)base};
    const int W = m->W, d = m->d, D = m->D, Vp = m->Vp;
    const bool sp = m->split;
    const size_t BT = (size_t)B * T;
    ws->B = B; ws->T = T; ws->k = k;
    ws->row_T = c.take<int32_t>(B);
    ws->idx = c.take<int32_t>(BT * k);
    ws->val = c.take<float>(BT * k);
    ws->live = c.take<uint8_t>(B);
    ws->done = c.take<uint8_t>(B);
    if (!m->use_xtab) {      // projected input sequences (the table path gathers E.Kx in the epilogue instead)
        ws->X = c.take<float>(BT * W);   ws->X_lo = sp ? c.take<float>(BT * W) : nullptr;
        ws->Xr = c.take<float>(BT * W);  ws->Xr_lo = sp ? c.take<float>(BT * W) : nullptr;
    }
    ws->seq.assign(d, nullptr); ws->seq_lo.assign(d, nullptr); ws->seqr.assign(d, nullptr); ws->seqr_lo.assign(d, nullptr);
    for (int n = 0; n < d; ++n) {
        const size_t e = BT * m->enc_out_dim(n);
        ws->seq[n] = c.take<float>(e);
        ws->seq_lo[n] = sp ? c.take<float>(e) : nullptr;
        if (n + 1 < d && m->layer_bidi(n + 1)) {       // the next layer scans backwards too
            ws->seqr[n] = c.take<float>(e);
            ws->seqr_lo[n] = sp ? c.take<float>(e) : nullptr;
        }
    }
    ws->u = c.take<float>(BT * W);
    if (recur_layout(m, B)) {    // persistent recurrence (recur.cu): the upper layers' input term for all timesteps
        ws->zx[0] = c.take<float>(BT * 4 * W);
        ws->zx[1] = (m->cfg.deep_bidirectional_encoder || !m->use_xtab) ? c.take<float>(BT * 4 * W) : nullptr;
    } else { ws->zx[0] = ws->zx[1] = nullptr; }
    for (int q = 0; q < 8; ++q) ws->pgA[q] = nullptr;
    if (recur_layout(m, B) && m->LowCat.hi) {     // operand rows of the persistent page decoder (fp16: 2 bytes per element)
        for (int q = 0; q < 4; ++q) ws->pgA[q] = reinterpret_cast<float*>(c.take<uint16_t>((size_t)kRecurMaxRows * (m->Vp + W)));
        for (int q = 4; q < 8; ++q) ws->pgA[q] = reinterpret_cast<float*>(c.take<uint16_t>((size_t)kRecurMaxRows * (2 * W + D)));
    }
    const size_t BW = (size_t)B * W;
    for (int dir = 0; dir < 2; ++dir) {
        for (int q = 0; q < 2; ++q) {
            ws->eh[dir][q] = c.take<float>(BW);
            ws->eh_lo[dir][q] = sp ? c.take<float>(BW) : nullptr;
        }
        ws->ec[dir] = c.take<float>(BW);
    }
    ws->st_h.assign(d, nullptr); ws->st_c.assign(d, nullptr); ws->dc.assign(d, nullptr);
    for (int q = 0; q < 2; ++q) { ws->dh[q].assign(d, nullptr); ws->dh_lo[q].assign(d, nullptr); }
    for (int n = 0; n < d; ++n) {
        ws->st_h[n] = c.take<float>(BW); ws->st_c[n] = c.take<float>(BW);
        ws->dc[n] = c.take<float>(BW);
        for (int q = 0; q < 2; ++q) {
            ws->dh[q][n] = c.take<float>(BW);
            ws->dh_lo[q][n] = sp ? c.take<float>(BW) : nullptr;
        }
    }
    ws->a[0] = c.take<float>(BT); ws->a[1] = c.take<float>(BT);
    ws->x = c.take<float>(BW);  ws->x_lo = sp ? c.take<float>(BW) : nullptr;
    ws->wq = c.take<float>(BW);
    ws->ctx = c.take<float>((size_t)B * D); ws->ctx_lo = sp ? c.take<float>((size_t)B * D) : nullptr;
    ws->logits = c.take<float>((size_t)B * Vp);
    ws->p = c.take<float>((size_t)B * Vp); ws->p_lo = sp ? c.take<float>((size_t)B * Vp) : nullptr;
    if (m->cfg.lm_predict) {
        ws->h_lm = c.take<float>(BW); ws->h_lm_lo = sp ? c.take<float>(BW) : nullptr;
        ws->c_lm = c.take<float>(BW);
        ws->logits_lm = c.take<float>((size_t)B * Vp);
        ws->p_lm = c.take<float>((size_t)B * Vp);
    }
    const size_t S = 2 * (size_t)T;
    ws->o_chars = c.take<int32_t>(B * S);
    ws->o_probs = c.take<float>(B * S);
    ws->o_score = c.take<double>(B);
    ws->o_len = c.take<int32_t>(B);
    ws->o_ties = c.take<int32_t>(B);
    ws->loop_state = c.take<int32_t>(8);
    // row compaction of the fast-greedy loop graph: row -> line maps, new row indices, one scratch copy of the per-row state
    ws->g_rl = c.take<int32_t>(B); ws->g_rl2 = c.take<int32_t>(B); ws->g_new = c.take<int32_t>(B);
    // per row: h (hi + lo, or fp32) and c of every layer, the alignment row, the feedback row (hi + lo, or fp32), the query;
    // every entry rounded up to 16 bytes (greedy_loop's MoveList)
    ws->g_scratch_row = (size_t)d * 8 * W + (size_t)T * 4 + (size_t)4 * Vp + 4 * W + 16 * (3 * (size_t)d + 5);
    ws->g_scratch_row = (ws->g_scratch_row + 15) & ~size_t(15);
    ws->g_scratch = c.take<uint8_t>(ws->g_scratch_row * B);
    ws->o_align = want_align ? c.take<float>(S * BT) : nullptr;
    ws->bytes = (c.off + 255) & ~size_t(255);
    return 0;
}

// ------------------------------------------------------------------ encoder
// seq2seq.py:237-314.  Per layer and direction: one GEMM+LSTM launch per timestep over all rows.
// Encoder results handed in through cor_set_encoding: enc_out [T][B][D] and states [2d][B][W] (the layouts
// cor_encode returns) become the workspace's line-major (hi, lo) enc_out, the per-layer initial decoder states
// and u = enc_out . U -- everything the decoders read, without running the encoder graph.
static int import_encoding(cor_model* m, const Ws& ws, cudaStream_t st) {
    const int B = ws.B, T = ws.T, W = m->W, D = m->D;
    COR_CHECK(m->preset.n_rows == B && m->preset.t_max == T, "the preset encoding has another geometry than this call");
    if (launch_split_tl(st, m->preset.enc_out, ws.enc_out(), ws.enc_out_lo(), B, T, D)) return 1;
    const size_t BW = (size_t)B * W;
    for (int n = 0; n < m->d; ++n) {
        COR_CUDA(cudaMemcpyAsync(ws.st_h[n], m->preset.states + (size_t)(2 * n) * BW, BW * 4, cudaMemcpyDeviceToDevice, st));
        COR_CUDA(cudaMemcpyAsync(ws.st_c[n], m->preset.states + (size_t)(2 * n + 1) * BW, BW * 4, cudaMemcpyDeviceToDevice, st));
    }
    GemmSeg su = {ws.enc_out(), D, ws.enc_out_lo(), m->U.hi, m->U.ld, m->U.lo, D};
    StoreEpi eu{ws.u, W, nullptr, nullptr, 0};
    return gemm_store(m, st, T * B, nullptr, W, &su, 1, eu);
}

int run_encoder(cor_model* m, const Ws& ws, cudaStream_t st) {
    const int W = m->W, d = m->d, B = ws.B, T = ws.T;
    const bool deep = m->cfg.deep_bidirectional_encoder != 0;
    if (launch_live_rows(st, ws.idx, ws.val, B, T, ws.k, ws.live)) return 1;
    if (m->preset.enc_out) return import_encoding(m, ws, st);
    if (!m->use_xtab && launch_embed(st, ws.idx, ws.val, ws.row_T, B, T, ws.k, m->E_full, W, m->V, ws.X, ws.X_lo, ws.Xr, ws.Xr_lo)) return 1;
    const size_t BW = (size_t)B * W;
    for (int n = 0; n < d; ++n) {
        const bool bidi = m->layer_bidi(n);
        const int in_dim = m->enc_in_dim(n), out_dim = m->enc_out_dim(n);
        const float* in[2] = {n == 0 ? ws.X : ws.seq[n - 1], n == 0 ? ws.Xr : ws.seqr[n - 1]};
        const float* in_lo[2] = {n == 0 ? ws.X_lo : ws.seq_lo[n - 1], n == 0 ? ws.Xr_lo : ws.seqr_lo[n - 1]};
        const int ndir = bidi ? 2 : 1;
        for (int dir = 0; dir < ndir; ++dir) {
            COR_CUDA(cudaMemsetAsync(ws.eh[dir][0], 0, BW * 4, st));
            if (ws.eh_lo[dir][0]) COR_CUDA(cudaMemsetAsync(ws.eh_lo[dir][0], 0, BW * 4, st));
            COR_CUDA(cudaMemsetAsync(ws.ec[dir], 0, BW * 4, st));
        }
        int cur = 0;
        const bool tab0 = n == 0 && m->use_xtab;
        // a bidirectional layer too wide for one grid runs its directions one after the other
        const int per_launch = (ws.zx[0] && lstm_recur_supported(W, B, ndir)) ? ndir : 1;
        if (ws.zx[0] && (tab0 || ws.zx[ndir - 1]) && lstm_recur_supported(W, B, per_launch)) {
            // ---- page-sized batch: one persistent launch per layer (recur.cu)
            for (int d0 = 0; d0 < ndir; d0 += per_launch) {
                RecurArgs ra{};
                ra.ndir = per_launch; ra.W = W; ra.B = B; ra.T = T; ra.barrier = m->grid_bar;
                for (int q = 0; q < per_launch; ++q) {
                    const int dir = d0 + q;
                    const LstmW& lw = dir == 0 ? m->enc_fw[n] : m->enc_bw[n];
                    if (!tab0) {    // input term of every timestep from ONE GEMM over the (t,row) pairs
                        GemmSeg sx = {in[dir], in_dim, in_lo[dir], lw.Kx.hi, lw.Kx.ld, lw.Kx.lo, in_dim};
                        StoreEpi ex{ws.zx[dir], 4 * W, nullptr, nullptr, 0};
                        if (gemm_store(m, st, T * B, nullptr, 4 * W, &sx, 1, ex)) return 1;
                    }
                    ra.R_hi[q] = lw.R.hi; ra.R_lo[q] = lw.R.lo; ra.ldr = lw.R.ld;
                    ra.zx[q] = tab0 ? nullptr : ws.zx[dir];
                    for (int pp = 0; pp < 2; ++pp) { ra.h[q][pp] = ws.eh[dir][pp]; ra.h_lo[q][pp] = ws.eh_lo[dir][pp]; }
                    LstmEpi& e = ra.epi[q];
                    if (tab0) { e.xtab = m->xtab[dir]; e.x_idx = ws.idx; e.x_val = ws.val; e.x_k = ws.k; e.x_tmax = T; }
                    e.bias = lw.bias; e.c_in = ws.ec[dir]; e.c_out = ws.ec[dir];
                    e.row_T = ws.row_T; e.reverse = dir; e.rows_total = B;
                    e.seq_out = ws.seq[n]; e.seq_out_lo = ws.seq_lo[n]; e.seq_ld = out_dim; e.seq_col = dir * W; e.seq_T = T;
                    if (n == d - 1) { e.seq_tm = 1; e.seq_rm = T; }      // enc_out is line-major (attention gathers windows)
                    e.seq_rev = ws.seqr[n]; e.seq_rev_lo = ws.seqr_lo[n];
                    if (m->cfg.residual_connections && !bidi && n >= 2) {
                        e.resid = ws.seq[n - 1]; e.resid_lo = ws.seq_lo[n - 1]; e.resid_ld = m->enc_out_dim(n - 1);
                    }
                    e.cross_sum = deep && n + 1 < d;
                    e.W = W;
                }
                if (launch_lstm_recur(st, ra)) return 1;
            }
            cur = T & 1;
        } else
        for (int s = 0; s < T; ++s) {
            for (int dir = 0; dir < ndir; ++dir) {
                const LstmW& lw = dir == 0 ? m->enc_fw[n] : m->enc_bw[n];
                const size_t slab = (size_t)s * B * in_dim;
                // split operands are __half arrays behind float* (epilogue.cuh): offset in bytes
                const size_t sb = slab * (in_lo[dir] ? sizeof(__half) : sizeof(float));
                const float* a_hi = reinterpret_cast<const float*>(reinterpret_cast<const char*>(in[dir]) + sb);
                const float* a_lo = in_lo[dir] ? reinterpret_cast<const float*>(reinterpret_cast<const char*>(in_lo[dir]) + sb) : nullptr;
                const bool tab = n == 0 && m->use_xtab;
                GemmSeg segs[2]; int ns = 0;
                if (!tab) segs[ns++] = GemmSeg{a_hi, in_dim, a_lo, lw.Kx.hi, lw.Kx.ld, lw.Kx.lo, in_dim};
                segs[ns++] = GemmSeg{ws.eh[dir][cur], W, ws.eh_lo[dir][cur], lw.R.hi, lw.R.ld, lw.R.lo, W};
                LstmEpi e{};
                if (tab) { e.xtab = m->xtab[dir]; e.x_idx = ws.idx; e.x_val = ws.val; e.x_k = ws.k; e.x_tmax = T; }
                e.bias = lw.bias; e.c_in = ws.ec[dir]; e.c_out = ws.ec[dir];
                e.h_prev = ws.eh[dir][cur]; e.h_prev_lo = ws.eh_lo[dir][cur];
                e.h_out = ws.eh[dir][cur ^ 1]; e.h_out_lo = ws.eh_lo[dir][cur ^ 1];
                e.row_T = ws.row_T; e.step = s; e.reverse = dir; e.rows_total = B;
                e.seq_out = ws.seq[n]; e.seq_out_lo = ws.seq_lo[n]; e.seq_ld = out_dim; e.seq_col = dir * W; e.seq_T = T;
                if (n == d - 1) { e.seq_tm = 1; e.seq_rm = T; }          // enc_out is line-major (attention gathers windows)
                e.seq_rev = ws.seqr[n]; e.seq_rev_lo = ws.seqr_lo[n];
                if (m->cfg.residual_connections && !bidi && n >= 2) {
                    e.resid = ws.seq[n - 1]; e.resid_lo = ws.seq_lo[n - 1]; e.resid_ld = m->enc_out_dim(n - 1);
                }
                e.cross_sum = deep && n + 1 < d;
                e.W = W;
                if (gemm_lstm(m, st, B, nullptr, 4 * W, segs, ns, e)) return 1;
            }
            cur ^= 1;
        }
        // state handed to decoder layer n: backward final for bidirectional layers (seq2seq.py:280-281),
        // forward final (after the padding) otherwise (:283)
        const int sd = bidi ? 1 : 0;
        const float* fh = ws.eh[sd][cur]; const float* fh_lo = ws.eh_lo[sd][cur];
        const float* fc = ws.ec[sd];
        if (m->cfg.bridge_dense) {
            GemmSeg sh = {fh, W, fh_lo, m->bridge_h[n].hi, m->bridge_h[n].ld, m->bridge_h[n].lo, W};
            StoreEpi eh_{ws.st_h[n], W, nullptr, m->bridge_hb[n], 1};
            if (gemm_store_ffma(st, B, nullptr, W, &sh, 1, eh_)) return 1;
            GemmSeg sc = {fc, W, nullptr, m->bridge_c[n].hi, m->bridge_c[n].ld, m->bridge_c[n].lo, W};
            StoreEpi ec_{ws.st_c[n], W, nullptr, m->bridge_cb[n], 1};
            if (gemm_store_ffma(st, B, nullptr, W, &sc, 1, ec_)) return 1;
        } else {
            if (fh_lo) { if (launch_join(st, fh, fh_lo, ws.st_h[n], BW)) return 1; }
            else COR_CUDA(cudaMemcpyAsync(ws.st_h[n], fh, BW * 4, cudaMemcpyDeviceToDevice, st));
            COR_CUDA(cudaMemcpyAsync(ws.st_c[n], fc, BW * 4, cudaMemcpyDeviceToDevice, st));
        }
    }
    // u = enc_out . U for every (t,row) at once — hoisted out of the step loop (the reference
    // re-evaluates attention_dense inside every decoder call, seq2seq.py:459-460)
    GemmSeg su = {ws.enc_out(), m->D, ws.enc_out_lo(), m->U.hi, m->U.ld, m->U.lo, m->D};
    StoreEpi eu{ws.u, W, nullptr, nullptr, 0};
    const long long rows = (long long)T * B;
    COR_CHECK(rows < (1ll << 31), "too many (t,row) pairs for one call");
    return gemm_store(m, st, (int)rows, nullptr, W, &su, 1, eu);
}

// ------------------------------------------------------------------ decoder step
// The feedback projection x = p . E (seq2seq.py:421-423) feeds only the first cell's kernel, so
// with PE = E . Kx prepared at load time its product is one K-segment of that cell's GEMM.
static int x_segment(cor_model* m, const StepIO& io, const LstmW& lw, cudaStream_t st, GemmSeg* seg, int* n) {
    if (m->fuse_pe) {
        if (!io.first) seg[(*n)++] = GemmSeg{io.p, m->Vp, io.p_lo, m->PE.hi, m->PE.ld, m->PE.lo, m->Vp};
        return 0;                                   // first step: p == 0, the segment vanishes
    }
    const int W = m->W;
    if (io.first) {
        COR_CUDA(cudaMemsetAsync(io.x, 0, (size_t)io.R * W * 4, st));
        if (io.x_lo) COR_CUDA(cudaMemsetAsync(io.x_lo, 0, (size_t)io.R * W * 4, st));
    } else {
        GemmSeg s = {io.p, m->Vp, io.p_lo, m->Et.hi, m->Et.ld, m->Et.lo, m->Vp};
        StoreEpi e{io.x, W, io.x_lo, nullptr, 0};
        if (gemm_store(m, st, io.R, io.r_dev, W, &s, 1, e)) return 1;
    }
    seg[(*n)++] = GemmSeg{io.x, W, io.x_lo, lw.Kx.hi, lw.Kx.ld, lw.Kx.lo, W};
    return 0;
}

int decoder_core(cor_model* m, const Ws& ws, const StepIO& io, cudaStream_t st) {
    const int W = m->W, d = m->d, D = m->D, R = io.R;
    const int top = d - 1;
    if (!io.wq_ready) {   // wq = h_prev . W_a + b_UW   (attention.py:539)
        GemmSeg s = {io.h_in[top], W, io.h_in_lo[top], m->Wa.hi, m->Wa.ld, m->Wa.lo, W};
        StoreEpi e{io.wq, W, nullptr, m->b_UW, 0};
        if (gemm_store(m, st, R, io.r_dev, W, &s, 1, e)) return 1;
    }
    AttnArgs a{};
    a.wq = io.wq; a.u = ws.u; a.enc = ws.enc_out(); a.enc_lo = ws.enc_out_lo();
    a.a_prev = io.a_in; a.row_line = io.row_line; a.line_T = ws.row_T;
    a.v_a = m->v_a; a.b_v = m->b_v; a.a_out = io.a_out; a.ctx = io.ctx; a.ctx_lo = io.ctx_lo;
    a.align_log = io.align_log; a.step_dev = io.step_dev; a.step_off = io.step_off; a.align_stride = io.align_stride;
    a.align_by_line = io.align_by_line;
    a.R = R; a.r_dev = io.r_dev; a.L = ws.B; a.t_max = ws.T; a.W = W; a.D = D;
    a.window = m->cfg.window_width; a.inclusive = 1;
    // The attention of this step depends on the previous step only (query, alignment); with lower
    // cells in between it runs on the side stream, its HBM traffic under their tensor work.
    const bool overlap = d >= 2 && m->side != nullptr;
    if (overlap) COR_CUDA(cudaEventRecord(m->ev_fork, st));
    else if (launch_attention(st, a)) return 1;

    GemmSeg xs[1]; int nx = 0;
    if (x_segment(m, io, m->dec[0], st, xs, &nx)) return 1;
    const float* x = nullptr; const float* x_lo = nullptr;
    for (int n = 0; n + 1 < d; ++n) {
        const LstmW& lw = m->dec[n];
        GemmSeg segs[2]; int ns = 0;
        if (n == 0) { if (nx) segs[ns++] = xs[0]; }
        else segs[ns++] = GemmSeg{x, W, x_lo, lw.Kx.hi, lw.Kx.ld, lw.Kx.lo, W};
        segs[ns++] = GemmSeg{io.h_in[n], W, io.h_in_lo[n], lw.R.hi, lw.R.ld, lw.R.lo, W};
        LstmEpi e{};
        e.bias = lw.bias; e.c_in = io.c_in[n]; e.c_out = io.c_out[n];
        e.h_prev = io.h_in[n]; e.h_prev_lo = io.h_in_lo[n]; e.h_out = io.h_out[n]; e.h_out_lo = io.h_out_lo[n];
        e.W = W;
        if (gemm_lstm(m, st, R, io.r_dev, 4 * W, segs, ns, e)) return 1;
        x = io.h_out[n]; x_lo = io.h_out_lo[n];
        if (n == 0 && overlap) {
            // submitted behind the persistent GEMM, so that its CTAs take their SMs first and the
            // attention blocks fill the registers and warp slots left beside them
            COR_CUDA(cudaStreamWaitEvent(m->side, m->ev_fork, 0));
            if (m->side_attention ? launch_attention_side(m->side, a) : launch_attention(m->side, a)) return 1;
            COR_CUDA(cudaEventRecord(m->ev_join, m->side));
        }
    }
    if (overlap) COR_CUDA(cudaStreamWaitEvent(st, m->ev_join, 0));
    const LstmW& tw = m->dec[top];
    GemmSeg tsegs[3]; int nt = 0;          // input segment(s) of the top cell, shared with the LM evaluation
    if (d == 1) { if (nx) tsegs[nt++] = xs[0]; }
    else tsegs[nt++] = GemmSeg{x, W, x_lo, tw.Kx.hi, tw.Kx.ld, tw.Kx.lo, W};
    const int n_in = nt;
    {
        tsegs[nt++] = GemmSeg{io.ctx, D, io.ctx_lo, tw.Kc.hi, tw.Kc.ld, tw.Kc.lo, D};
        tsegs[nt++] = GemmSeg{io.h_in[top], W, io.h_in_lo[top], tw.R.hi, tw.R.ld, tw.R.lo, W};
        LstmEpi e{};
        e.bias = tw.bias; e.c_in = io.c_in[top]; e.c_out = io.c_out[top];
        e.h_prev = io.h_in[top]; e.h_prev_lo = io.h_in_lo[top]; e.h_out = io.h_out[top]; e.h_out_lo = io.h_out_lo[top];
        e.W = W;
        if (gemm_lstm(m, st, R, io.r_dev, 4 * W, tsegs, nt, e)) return 1;
    }
    if (io.wq_next) {   // logits = h . E^T (tied output projection, seq2seq.py:379) | next query = h . W_a + b_UW
        GemmSeg s = {io.h_out[top], W, io.h_out_lo[top], m->EWa.hi, m->EWa.ld, m->EWa.lo, W};
        StoreEpi e{io.logits, m->Vp, nullptr, nullptr, 0, m->Vp, io.wq_next, W, m->b_UW};
        if (gemm_store(m, st, R, io.r_dev, m->Vp + W, &s, 1, e)) return 1;
    } else {
        GemmSeg s = {io.h_out[top], W, io.h_out_lo[top], m->E.hi, m->E.ld, m->E.lo, W};
        StoreEpi e{io.logits, m->Vp, nullptr, nullptr, 0};
        if (gemm_store(m, st, R, io.r_dev, m->Vp, &s, 1, e)) return 1;
    }
    if (m->cfg.lm_predict && io.logits_lm) {
        // second evaluation of the top cell with attended == u == 0: context is exactly zero
        // (seq2seq.py:464-469); its states are discarded.
        GemmSeg segs[2]; int ns = 0;
        if (n_in) segs[ns++] = tsegs[0];
        segs[ns++] = GemmSeg{io.h_in[top], W, io.h_in_lo[top], tw.R.hi, tw.R.ld, tw.R.lo, W};
        LstmEpi e{};
        e.bias = tw.bias; e.c_in = io.c_in[top]; e.c_out = io.c_lm;
        e.h_prev = io.h_in[top]; e.h_prev_lo = io.h_in_lo[top]; e.h_out = io.h_lm; e.h_out_lo = io.h_lm_lo;
        e.W = W;
        if (gemm_lstm(m, st, R, io.r_dev, 4 * W, segs, ns, e)) return 1;
        GemmSeg s = {io.h_lm, W, io.h_lm_lo, m->E.hi, m->E.ld, m->E.lo, W};
        StoreEpi e2{io.logits_lm, m->Vp, nullptr, nullptr, 0};
        if (gemm_store(m, st, R, io.r_dev, m->Vp, &s, 1, e2)) return 1;
    }
    return 0;
}

static int copy_in(cudaStream_t st, const float* src, float* hi, float* lo, size_t n) {
    if (lo) return launch_split(st, src, hi, lo, n);
    COR_CUDA(cudaMemcpyAsync(hi, src, n * 4, cudaMemcpyDeviceToDevice, st));
    return 0;
}
static int copy_out(cudaStream_t st, const float* hi, const float* lo, float* dst, size_t n) {
    if (lo) return launch_join(st, hi, lo, dst, n);
    COR_CUDA(cudaMemcpyAsync(dst, hi, n * 4, cudaMemcpyDeviceToDevice, st));
    return 0;
}

// ---- fast-greedy loop graph: device-side loop state (Ws::loop_state) -----------------------------------------------
//   [0] timestep the next body starts at      [2] live rows R (the GEMMs' m_dev)      [3] rows before the last compaction
//   [4] 1 if the last scan decided to compact
constexpr int kCompactEvery = 8;         // timesteps between looks at the finished rows
struct GreedyCtl {
    int32_t* state; int64_t* total;     // total[0] timesteps run, total[1] compactions done (launch accounting)
    int s_max, n_rows;
    cudaGraphConditionalHandle loop;    // WHILE: another body?
    cudaGraphConditionalHandle compact; // IF: look at the finished rows after this body?
};
__global__ void greedy_init_kernel(GreedyCtl c, int32_t* rl) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < c.n_rows) rl[r] = r;
    if (r == 0) {
        c.state[0] = 0; c.state[2] = c.n_rows; c.state[3] = c.n_rows; c.state[4] = 0;
        cudaGraphSetConditional(c.loop, (c.s_max > 0 && c.n_rows > 0) ? 1u : 0u);
    }
}
// end of a body (two timesteps): advance the step, decide about another body and about a look at the finished rows
__global__ void greedy_cond_kernel(GreedyCtl c) {
    const int step = c.state[0] + 2;
    c.state[0] = step;
    if (c.total) c.total[0] += 2;
    const bool more = step < c.s_max && c.state[2] > 0;
    cudaGraphSetConditional(c.loop, more ? 1u : 0u);
    cudaGraphSetConditional(c.compact, (more && step % kCompactEvery == 0) ? 1u : 0u);
}

// Row compaction: a line that has emitted its newline (or used up its 2T steps) only burns
// GEMM rows from then on -- the reference keeps stepping it and ignores the result (seq2seq.py:1255-1256).  Every
// kCompactEvery steps one block scans the live rows; if at least 1/16 of them have finished (or all), the survivors move
// to the front (through a scratch copy: a row may land where another still has to be read), the row -> line map with
// them, and the row count the GEMMs read shrinks.  When none are left the loop ends.
struct CompactArgs {
    GreedyCtl ctl;
    const int32_t* rl; int32_t* rl2; int32_t* g_new;
    const uint8_t* done; const uint8_t* live; const int32_t* row_T;
};
__global__ void __launch_bounds__(1024) greedy_compact_scan_kernel(CompactArgs c) {
    __shared__ int s_wsum[32];
    __shared__ int s_total;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int R = c.ctl.state[2], step = c.ctl.state[0];
    const int per = (R + 1023) / 1024;
    const int r0 = min(R, tid * per), r1 = min(R, r0 + per);
    auto keep = [&](int r) {
        const int line = c.rl[r];
        return !c.done[line] && c.live[line] && step < 2 * c.row_T[line];
    };
    int mine = 0;
    for (int r = r0; r < r1; ++r) mine += keep(r) ? 1 : 0;
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
    if (lane == 31) s_wsum[warp] = incl;
    __syncthreads();
    int off = 0, total = 0;
    for (int w = 0; w < 32; ++w) { if (w < warp) off += s_wsum[w]; total += s_wsum[w]; }
    const bool compact = total == 0 || R - total >= max(1, R >> 4);
    if (compact) {
        int j = off + incl - mine;
        for (int r = r0; r < r1; ++r) {
            if (keep(r)) { c.g_new[r] = j; c.rl2[j] = c.rl[r]; ++j; } else c.g_new[r] = -1;
        }
    }
    if (tid == 0) {
        c.ctl.state[4] = compact ? 1 : 0;
        if (compact) {
            c.ctl.state[3] = R; c.ctl.state[2] = total;
            if (c.ctl.total) c.ctl.total[1] += 1;
            if (total == 0) cudaGraphSetConditional(c.ctl.loop, 0u);
        }
    }
}
struct MoveEntry { char* ptr; int bytes; int off; };
struct MoveList { MoveEntry e[32]; int n; };
__global__ void __launch_bounds__(128) greedy_compact_move_kernel(MoveList ml, const int32_t* state, const int32_t* g_new, int32_t* rl, const int32_t* rl2,
                                                                  uint8_t* scratch, size_t scratch_row, int back) {
    if (!state[4]) return;
    const int rows = back ? state[2] : state[3];
    for (int r = blockIdx.x; r < rows; r += gridDim.x) {
        const int j = back ? r : g_new[r];
        if (j < 0) continue;
        uint8_t* srow = scratch + (size_t)j * scratch_row;
        for (int k = 0; k < ml.n; ++k) {
            const MoveEntry e = ml.e[k];
            char* g = e.ptr + (size_t)(back ? j : r) * e.bytes;
            char* sc = reinterpret_cast<char*>(srow) + e.off;
            const char* src = back ? sc : g; char* dst = back ? g : sc;
            if ((e.bytes & 15) == 0) {
                for (int i = threadIdx.x; i < (e.bytes >> 4); i += blockDim.x) reinterpret_cast<uint4*>(dst)[i] = reinterpret_cast<const uint4*>(src)[i];
            } else {
                for (int i = threadIdx.x; i < (e.bytes >> 2); i += blockDim.x) reinterpret_cast<uint32_t*>(dst)[i] = reinterpret_cast<const uint32_t*>(src)[i];
            }
        }
        if (back && threadIdx.x == 0) rl[j] = rl2[j];
    }
}

static int greedy_loop(cor_model* m, const Ws& ws, cudaStream_t st) {
    const int B = ws.B, T = ws.T, W = m->W, d = m->d;
    const size_t BW = (size_t)B * W;
    for (int n = 0; n < d; ++n) {
        if (copy_in(st, ws.st_h[n], ws.dh[0][n], ws.dh_lo[0][n], BW)) return 1;
        COR_CUDA(cudaMemcpyAsync(ws.dc[n], ws.st_c[n], BW * 4, cudaMemcpyDeviceToDevice, st));
    }
    COR_CUDA(cudaMemsetAsync(ws.a[0], 0, (size_t)B * T * 4, st));
    COR_CUDA(cudaMemsetAsync(ws.done, 0, B, st));
    COR_CUDA(cudaMemsetAsync(ws.o_len, 0, (size_t)B * 4, st));
    COR_CUDA(cudaMemsetAsync(ws.o_ties, 0, (size_t)B * 4, st));
    COR_CUDA(cudaMemsetAsync(ws.o_score, 0, (size_t)B * 8, st));
    COR_CUDA(cudaMemsetAsync(ws.o_chars, 0, (size_t)B * 2 * T * 4, st));
    COR_CUDA(cudaMemsetAsync(ws.o_probs, 0, (size_t)B * 2 * T * 4, st));
    int cur = 0;
    const int S = 2 * T;
    if (ws.pgA[0] && page_decoder_supported(W, m->D, m->Vp, B)) {
        // ---- page-sized batch: the whole loop is one persistent launch (decpage.cu)
        const int top = d - 1;
        {   // query of step 0: wq = h_0 . W_a + b_UW  (attention.py:539)
            GemmSeg sq = {ws.dh[0][top], W, ws.dh_lo[0][top], m->Wa.hi, m->Wa.ld, m->Wa.lo, W};
            StoreEpi eq{ws.wq, W, nullptr, m->b_UW, 0};
            if (gemm_store(m, st, B, nullptr, W, &sq, 1, eq)) return 1;
        }
        PageDecArgs pa{};
        pa.B = B; pa.T = T; pa.S = S; pa.W = W; pa.D = m->D; pa.V = m->V; pa.Vp = m->Vp;
        pa.window = m->cfg.window_width; pa.eos = m->cfg.eos_index; pa.b_v = m->b_v;
        pa.low_hi = (const __half*)m->LowCat.hi; pa.low_lo = (const __half*)m->LowCat.lo;
        pa.top_hi = (const __half*)m->TopCat.hi; pa.top_lo = (const __half*)m->TopCat.lo;
        pa.ewa_hi = (const __half*)m->EWa.hi; pa.ewa_lo = (const __half*)m->EWa.lo; pa.ewa_ld = m->EWa.ld;
        pa.bias_low = m->dec[0].bias; pa.bias_top = m->dec[1].bias; pa.b_UW = m->b_UW; pa.v_a = m->v_a;
        for (int q = 0; q < 2; ++q) {
            pa.lowA_hi[q] = (__half*)ws.pgA[2 * q]; pa.lowA_lo[q] = (__half*)ws.pgA[2 * q + 1];
            pa.topA_hi[q] = (__half*)ws.pgA[4 + 2 * q]; pa.topA_lo[q] = (__half*)ws.pgA[4 + 2 * q + 1];
            pa.h0_hi[q] = (const __half*)ws.dh[0][q]; pa.h0_lo[q] = (const __half*)ws.dh_lo[0][q];
            pa.c[q] = ws.dc[q]; pa.a[q] = ws.a[q];
        }
        pa.logits = ws.logits; pa.wq = ws.wq; pa.o_align = ws.o_align;
        pa.u = ws.u; pa.enc = ws.enc_out(); pa.enc_lo = ws.enc_out_lo();
        pa.line_T = ws.row_T; pa.live = ws.live; pa.done = ws.done;
        pa.out_chars = ws.o_chars; pa.out_probs = ws.o_probs; pa.score_sum = ws.o_score; pa.out_len = ws.o_len;
        pa.out_ties = ws.o_ties;
        pa.barrier = m->grid_bar;
        const int rc = launch_page_decoder(st, pa);
        if (rc == 1) return 1;
        if (rc == 0) return launch_finish_scores(st, ws.o_score, ws.o_len, B);
        // rc == 2: the cooperative launch was refused and nothing ran -- per-timestep launches below
    }
    // one timestep: decoder cells, attention, logits, softmax + greedy bookkeeping.  `s` < 0: the timestep is read on the
    // device (loop graph; `off` = position of the step inside the graph body)
    auto timestep = [&](cudaStream_t s_, int s, int off, int cur_) -> int {
        StepIO io{};
        io.R = B; io.first = false;           // step 0 multiplies a zero feedback row (below), like every other step multiplies its own
        io.p = ws.p; io.p_lo = ws.p_lo; io.x = ws.x; io.x_lo = ws.x_lo;
        for (int n = 0; n < d; ++n) {
            io.h_in[n] = ws.dh[cur_][n]; io.h_in_lo[n] = ws.dh_lo[cur_][n];
            io.h_out[n] = ws.dh[cur_ ^ 1][n]; io.h_out_lo[n] = ws.dh_lo[cur_ ^ 1][n];
            io.c_in[n] = ws.dc[n]; io.c_out[n] = ws.dc[n];
        }
        io.a_in = ws.a[cur_]; io.a_out = ws.a[cur_ ^ 1];
        if (s >= 0) io.align_log = ws.o_align ? ws.o_align + (size_t)s * B * T : nullptr;
        else {
            io.align_log = ws.o_align; io.step_dev = ws.loop_state; io.step_off = off; io.align_stride = (long long)B * T;
            io.r_dev = ws.loop_state + 2; io.row_line = ws.g_rl; io.align_by_line = 1;      // rows are compacted, lines keep their place
        }
        io.wq = ws.wq; io.ctx = ws.ctx; io.ctx_lo = ws.ctx_lo; io.logits = ws.logits;
        io.wq_ready = (s != 0); io.wq_next = ws.wq;
        if (decoder_core(m, ws, io, s_)) return 1;
        GreedyArgs g{};
        g.logits = ws.logits; g.p = ws.p; g.p_lo = ws.p_lo; g.row_T = ws.row_T; g.live = ws.live;
        g.out_chars = ws.o_chars; g.out_probs = ws.o_probs; g.score_sum = ws.o_score; g.out_len = ws.o_len; g.done = ws.done;
        g.out_ties = ws.o_ties;
        g.R = B; g.V = m->V; g.step = s >= 0 ? s : off; g.s_max = S; g.eos = m->cfg.eos_index;
        if (s < 0) { g.step_dev = ws.loop_state; g.row_line = ws.g_rl; g.r_dev = ws.loop_state + 2; }
        return launch_softmax_greedy(s_, g);
    };
    if (loop_graphs_enabled()) {
        // ---- all timesteps as ONE graph launch: WHILE { timestep, timestep, condition } with the step counter and the
        // "some line still wants a step" flag on the device -- the loop ends when every line has emitted its newline
        // (the reference steps on to 2T and ignores finished lines, seq2seq.py:1255-1256: same results).
        // Step 0 is made uniform with the rest: zero feedback row instead of a skipped K-segment (adds exact zeros
        // to the accumulator), its query from its own small GEMM.
        const int top = d - 1;
        COR_CUDA(cudaMemsetAsync(ws.p, 0, (size_t)B * m->Vp * 4, st));
        if (ws.p_lo) COR_CUDA(cudaMemsetAsync(ws.p_lo, 0, (size_t)B * m->Vp * 4, st));
        COR_CUDA(cudaMemsetAsync(ws.loop_state, 0, 16, st));
        {
            GemmSeg sq = {ws.dh[0][top], W, ws.dh_lo[0][top], m->Wa.hi, m->Wa.ld, m->Wa.lo, W};
            StoreEpi eq{ws.wq, W, nullptr, m->b_UW, 0};
            if (gemm_store(m, st, B, nullptr, W, &sq, 1, eq)) return 1;
        }
        LoopGraph* g = nullptr;
        const std::vector<uint64_t> key = {0x63EEDull, (uint64_t)(uintptr_t)ws.row_T, (uint64_t)B, (uint64_t)T, (uint64_t)ws.k,
                                           (uint64_t)(ws.o_align != nullptr), (uint64_t)(ws.val != nullptr), (uint64_t)m->passes,
                                           (uint64_t)(m->side != nullptr), (uint64_t)ws.bytes};
        if (loop_graph_lookup(m, key, &g)) return 1;
        if (!g->exec) {
            GreedyCtl ctl{};
            ctl.state = ws.loop_state; ctl.total = m->dev_counters; ctl.s_max = S; ctl.n_rows = B;
            cudaGraph_t body;
            // the first node needs both handles, which exist only once the graph does: create the graph and the handles by
            // hand, then the node (loop_graph_begin would create the WHILE handle after reading the node's arguments)
            if (g->graph) { cudaGraphExecDestroy(g->exec); cudaGraphDestroy(g->graph); g->exec = nullptr; g->graph = nullptr; }
            COR_CUDA(cudaGraphCreate(&g->graph, 0));
            COR_CUDA(cudaGraphConditionalHandleCreate(&ctl.loop, g->graph, 0, 0));
            COR_CUDA(cudaGraphConditionalHandleCreate(&ctl.compact, g->graph, 0, 0));
            {
                int32_t* rl = ws.g_rl;
                void* kargs[2] = {(void*)&ctl, (void*)&rl};
                cudaKernelNodeParams kp{};
                kp.func = (void*)greedy_init_kernel; kp.gridDim = dim3((B + 255) / 256); kp.blockDim = dim3(256); kp.kernelParams = kargs;
                cudaGraphNode_t n_first, n_while;
                COR_CUDA(cudaGraphAddKernelNode(&n_first, g->graph, nullptr, 0, &kp));
                cudaGraphNodeParams cp{};
                cp.type = cudaGraphNodeTypeConditional;
                cp.conditional.handle = ctl.loop; cp.conditional.type = cudaGraphCondTypeWhile; cp.conditional.size = 1;
                COR_CUDA(cudaGraphAddNode(&n_while, g->graph, &n_first, 1, &cp));
                body = cp.conditional.phGraph_out[0];
            }
            // per-row decoder state that moves with a row
            MoveList ml{};
            size_t off = 0;
            auto add = [&](const void* ptr, size_t bytes) {
                if (!ptr) return;
                ml.e[ml.n++] = MoveEntry{(char*)ptr, (int)bytes, (int)off};
                off = (off + bytes + 15) & ~size_t(15);
            };
            const bool sp = m->split;
            for (int n = 0; n < d; ++n) {
                add(ws.dh[0][n], (size_t)W * (sp ? 2 : 4)); add(ws.dh_lo[0][n], (size_t)W * 2);
                add(ws.dc[n], (size_t)W * 4);
            }
            add(ws.a[0], (size_t)T * 4);
            add(ws.p, (size_t)m->Vp * (sp ? 2 : 4)); add(ws.p_lo, (size_t)m->Vp * 2);
            add(ws.wq, (size_t)W * 4);
            COR_CHECK(off <= ws.g_scratch_row && ml.n <= 32, "compaction scratch row too small");
            cudaStream_t cs = loop_capture_stream(m, 0), cs2 = loop_capture_stream(m, 1);
            COR_CHECK(cs && cs2, "no capture stream");
            const int64_t before = ctx().launches;
            COR_CUDA(cudaStreamBeginCaptureToGraph(cs, body, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
            int rc = timestep(cs, -1, 0, 0) || timestep(cs, -1, 1, 1);
            if (!rc) { greedy_cond_kernel<<<1, 1, 0, cs>>>(ctl); rc = cudaGetLastError() != cudaSuccess; }
            const int step_kernels = (int)(ctx().launches - before);
            if (!rc) {
                cudaGraph_t branch[1];
                rc = loop_graph_switch(cs, 0, ctl.compact, branch);
                if (!rc) {
                    cudaError_t e = cudaStreamBeginCaptureToGraph(cs2, branch[0], nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed);
                    if (e != cudaSuccess) { set_error(std::string("capturing the compaction branch: ") + cudaGetErrorString(e)); rc = 1; }
                    else {
                        CompactArgs ca{ctl, ws.g_rl, ws.g_rl2, ws.g_new, ws.done, ws.live, ws.row_T};
                        const int mgrid = std::min(B, 148 * 8);
                        greedy_compact_scan_kernel<<<1, 1024, 0, cs2>>>(ca);
                        greedy_compact_move_kernel<<<mgrid, 128, 0, cs2>>>(ml, ws.loop_state, ws.g_new, ws.g_rl, ws.g_rl2, ws.g_scratch, ws.g_scratch_row, 0);
                        greedy_compact_move_kernel<<<mgrid, 128, 0, cs2>>>(ml, ws.loop_state, ws.g_new, ws.g_rl, ws.g_rl2, ws.g_scratch, ws.g_scratch_row, 1);
                        rc = cudaGetLastError() != cudaSuccess;
                        cudaGraph_t got = nullptr;
                        e = cudaStreamEndCapture(cs2, &got);
                        if (!rc && e != cudaSuccess) { set_error(std::string("capturing the compaction branch: ") + cudaGetErrorString(e)); rc = 1; }
                    }
                }
            }
            cudaGraph_t captured = nullptr;
            const cudaError_t ce = cudaStreamEndCapture(cs, &captured);
            g->body_kernels = step_kernels;
            ctx().launches = before;            // captured, not run
            COR_CHECK(!rc, "capturing the greedy timestep failed");
            COR_CUDA(ce);
            COR_CUDA(cudaGraphInstantiate(&g->exec, g->graph, 0));
        }
        m->greedy_step_kernels = g->body_kernels / 2;
        COR_CUDA(cudaGraphLaunch(g->exec, st));
        ctx().launches += 1; ctx().in_graph += 1; ctx().graph_launches += 1;     // the graph's first node
        return launch_finish_scores(st, ws.o_score, ws.o_len, B);
    }
    // host-driven loop: the same K-segments in every step as in the graph (a GEMM whose k-blocks are shared out among several
    // accumulators -- the 128x32-tile kernel -- rounds differently when a segment of zeros is left out)
    COR_CUDA(cudaMemsetAsync(ws.p, 0, (size_t)B * m->Vp * 4, st));
    if (ws.p_lo) COR_CUDA(cudaMemsetAsync(ws.p_lo, 0, (size_t)B * m->Vp * 4, st));
    for (int s = 0; s < S; ++s) {
        if (timestep(st, s, 0, cur)) return 1;
        cur ^= 1;
    }
    return launch_finish_scores(st, ws.o_score, ws.o_len, B);
}

}  // namespace cor

using namespace cor;

extern "C" {

const char* cor_last_error(void) { return g_error.c_str(); }
int cor_abi_version(void) { return 2; }
int64_t cor_launch_count(cor_model* m, int reset) {
    if (!m) return -1;
    if (m->dev_counters && m->greedy_step_kernels) {    // timesteps the fast-greedy loop graphs ran (counted on the device)
        int64_t steps = 0;
        if (cudaSetDevice(m->device) == cudaSuccess && cudaDeviceSynchronize() == cudaSuccess &&
            cudaMemcpy(&steps, m->dev_counters, 8, cudaMemcpyDeviceToHost) == cudaSuccess) {
            int64_t compactions = 0;
            cudaMemcpy(&compactions, m->dev_counters + 1, 8, cudaMemcpyDeviceToHost);
            const int64_t ran = steps * m->greedy_step_kernels + steps / 2 + 3 * compactions;
            m->ctx.launches += ran;
            m->ctx.in_graph += ran;
            cudaMemset(m->dev_counters, 0, 16);
        }
    }
    const int64_t v = m->ctx.launches;
    if (reset) m->ctx.launches = 0;
    return v;
}
int64_t cor_host_launch_count(cor_model* m, int reset) {
    if (!m) return -1;
    const int64_t kernels = cor_launch_count(m, 0);       // folds the device-side counts in
    const int64_t v = kernels - m->ctx.in_graph + m->ctx.graph_launches;
    if (reset) { m->ctx.launches = 0; m->ctx.in_graph = 0; m->ctx.graph_launches = 0; }
    return v;
}
int cor_debug_counters(cor_model* m, int64_t* out, int reset) {
    COR_CHECK(m && out, "null argument");
    for (int e = 0; e < 2; ++e) for (int k = 0; k < 3; ++k) out[3 * e + k] = m->ctx.tc_kind[e][k];
    out[6] = m->ctx.ffma_kind[0]; out[7] = m->ctx.ffma_kind[1];
    if (reset) { memset(m->ctx.tc_kind, 0, sizeof(m->ctx.tc_kind)); memset(m->ctx.ffma_kind, 0, sizeof(m->ctx.ffma_kind)); }
    return 0;
}

// Test hook: C[M,N] = A[M,K] . B[N,K]^T through one of the two GEMM kernels (1 = FFMA, 2 = tcgen05 3xFP16).
int cor_debug_gemm(int mode, int M, int N, int K, const float* A, const float* B, float* C, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    COR_CHECK(mode >= 1 && mode <= 4, "mode must be 1 (FFMA), 2 (tcgen05, tile shape picked from M and N), 3 (tcgen05, 128x32 tiles forced) or 4 (tcgen05, fp16 heads only)");
    StoreEpi e{C, N, nullptr, nullptr, 0};
    if (mode == 1) {
        GemmSeg s{A, K, nullptr, B, K, nullptr, K};
        return gemm_store_ffma(st, M, nullptr, N, &s, 1, e);
    }
    float *a_hi = nullptr, *a_lo = nullptr, *b_hi = nullptr, *b_lo = nullptr;
    const size_t na = (size_t)M * K, nb = (size_t)N * K;
    COR_CUDA(cudaMalloc(&a_hi, na * 4)); COR_CUDA(cudaMalloc(&a_lo, na * 4));
    COR_CUDA(cudaMalloc(&b_hi, nb * 4)); COR_CUDA(cudaMalloc(&b_lo, nb * 4));
    int rc = launch_split(st, A, a_hi, a_lo, na) || launch_split(st, B, b_hi, b_lo, nb);
    if (!rc) {
        GemmSeg s{a_hi, K, a_lo, b_hi, K, b_lo, K};
        tc_force_narrow(mode == 3);
        rc = gemm_store_tc(st, M, nullptr, N, &s, 1, e, mode == 4 ? 1 : 3);
        tc_force_narrow(false);
    }
    cudaError_t err = cudaStreamSynchronize(st);
    cudaFree(a_hi); cudaFree(a_lo); cudaFree(b_hi); cudaFree(b_lo);
    if (!rc && err != cudaSuccess) { set_error(std::string("debug gemm: ") + cudaGetErrorString(err)); rc = 1; }
    return rc;
}

int cor_prof_enable(cor_model* m, int on) {
    COR_CHECK(m, "null argument");
    CtxScope scope(&m->ctx);
    COR_CUDA(cudaSetDevice(m->device));
    prof().on = on != 0;
    if (on) {
        prof().used = 0; prof().flops = 0; prof().launches = 0; prof().per_row.clear();
        if (!prof().rows) COR_CUDA(cudaMallocHost((void**)&prof().rows, kMaxDeferred * 4));
    }
    return 0;
}
int cor_prof_read(cor_model* m, double* total_ms, int64_t* launches, double* flops) {
    COR_CHECK(m, "null argument");
    CtxScope scope(&m->ctx);
    COR_CUDA(cudaSetDevice(m->device));
    COR_CUDA(cudaDeviceSynchronize());
    double ms = 0;
    for (size_t i = 0; i + 1 < prof().used; i += 2) {
        float t = 0;
        COR_CUDA(cudaEventElapsedTime(&t, prof().ev[i], prof().ev[i + 1]));
        ms += t;
    }
    double fl = prof().flops;
    for (size_t i = 0; i < prof().per_row.size(); ++i) fl += prof().per_row[i] * (double)prof().rows[i];
    if (total_ms) *total_ms = ms;
    if (launches) *launches = prof().launches;
    if (flops) *flops = fl;
    return 0;
}

int cor_model_create(const cor_config* cfg, int device, cor_model** out) {
    COR_CHECK(cfg && out, "null argument");
    COR_CHECK(cfg->width > 0 && cfg->width % 32 == 0, "width must be a positive multiple of 32");
    COR_CHECK(cfg->depth >= 1 && cfg->depth <= 8, "depth must be in 1..8");
    COR_CHECK(cfg->voc_size >= 2, "voc_size must be at least 2");
    COR_CHECK(cfg->window_width >= 0 && cfg->window_width <= 6, "window_width must be in 0..6");
    int ndev = 0;
    COR_CUDA(cudaGetDeviceCount(&ndev));
    COR_CHECK(device >= 0 && device < ndev, "no such CUDA device");
    COR_CUDA(cudaSetDevice(device));
    cor_model* m = new cor_model();
    m->ctx.prof = &m->prof;
    m->cfg = *cfg; m->device = device;
    m->W = cfg->width; m->d = cfg->depth; m->V = cfg->voc_size; m->Vp = (cfg->voc_size + 127) & ~127;
    m->D = (cfg->depth == 1 || cfg->deep_bidirectional_encoder) ? 2 * cfg->width : cfg->width;
    COR_CHECK(cfg->gemm_mode >= 0 && cfg->gemm_mode <= 3, "gemm_mode must be 0..3");
    m->split = cfg->gemm_mode >= 2 || (cfg->gemm_mode == 0 && tc_available());
    m->passes = cfg->gemm_mode == 3 ? 1 : 3;
    m->tc_min_rows = 1;     // page-sized batches run the narrow-tile tcgen05 kernel (8 rows: 19 ms per page call against 110 ms on the FFMA kernel)
    if (const char* v = getenv("COR_TC_MIN_ROWS")) m->tc_min_rows = atoi(v);
    *out = m;
    return 0;
}

int cor_model_set_weight(cor_model* m, const char* layer, int index, const float* host, const int64_t* shape, int rank) {
    COR_CHECK(m && layer && host && shape, "null argument");
    COR_CHECK(!m->finalized, "model already finalized");
    COR_CHECK(index >= 0 && index < 16 && rank >= 1 && rank <= 2, "bad weight index or rank");
    auto& vec = m->host[layer];
    if ((int)vec.size() <= index) vec.resize(index + 1);
    HostWeight& w = vec[index];
    w.shape.assign(shape, shape + rank);
    size_t n = 1;
    for (int i = 0; i < rank; ++i) n *= (size_t)shape[i];
    w.data.assign(host, host + n);
    return 0;
}

int cor_model_finalize(cor_model* m) {
    COR_CHECK(m && !m->finalized, "bad model state");
    CtxScope scope(&m->ctx);
    COR_CUDA(cudaSetDevice(m->device));
    const int W = m->W, d = m->d, V = m->V, Vp = m->Vp, D = m->D;
    const std::vector<HostWeight>* w;
    if (need(m, "char_input_projection", 1, &w) || check_shape((*w)[0], V, W, "char_input_projection")) return 1;
    {
        const HostWeight& E = (*w)[0];
        if (upload_vec(m, E.data, &m->E_full)) return 1;
        std::vector<float> e((size_t)Vp * W, 0.f), et((size_t)W * Vp, 0.f);
        for (int v = 0; v < V; ++v)
            for (int u = 0; u < W; ++u) { e[(size_t)v * W + u] = E.data[(size_t)v * W + u]; et[(size_t)u * Vp + v] = E.data[(size_t)v * W + u]; }
        if (upload_mat(m, e, Vp, W, &m->E) || upload_mat(m, et, W, Vp, &m->Et)) return 1;
    }
    m->enc_fw.resize(d); m->enc_bw.resize(d); m->dec.resize(d);
    m->bridge_h.resize(d); m->bridge_c.resize(d); m->bridge_hb.assign(d, nullptr); m->bridge_cb.assign(d, nullptr);
    for (int n = 0; n < d; ++n) {
        const std::string name = "encoder_lstm_" + std::to_string(n + 1);
        const int in_dim = m->enc_in_dim(n);
        if (need(m, name, m->layer_bidi(n) ? 6 : 3, &w)) return 1;
        if (build_lstm(m, &(*w)[0], in_dim, in_dim, &m->enc_fw[n], name)) return 1;
        if (m->layer_bidi(n) && build_lstm(m, &(*w)[3], in_dim, in_dim, &m->enc_bw[n], name + " (backward)")) return 1;
        if (m->cfg.bridge_dense) {
            for (int hc = 0; hc < 2; ++hc) {
                const std::string bn = std::string(hc ? "bridge_c_" : "bridge_h_") + std::to_string(n + 1);
                if (need(m, bn, 2, &w) || check_shape((*w)[0], W, W, bn) || check_shape((*w)[1], W, -1, bn + " bias")) return 1;
                if (upload_mat(m, transpose_rows((*w)[0], 0, W), W, W, hc ? &m->bridge_c[n] : &m->bridge_h[n])) return 1;
                if (upload_vec(m, (*w)[1].data, hc ? &m->bridge_cb[n] : &m->bridge_hb[n])) return 1;
            }
        }
    }
    if (need(m, "attention_dense", 1, &w) || check_shape((*w)[0], D, W, "attention_dense")) return 1;
    if (upload_mat(m, transpose_rows((*w)[0], 0, D), W, D, &m->U)) return 1;
    for (int n = 0; n + 1 < d; ++n) {
        const std::string name = "decoder_lstm_" + std::to_string(n + 1);
        if (need(m, name, 3, &w) || build_lstm(m, &(*w)[0], W, W, &m->dec[n], name)) return 1;
    }
    {
        const std::string name = "decoder_lstm_" + std::to_string(d);
        if (need(m, name, 7, &w)) return 1;
        if (check_shape((*w)[0], W, W, name + " W_a") || check_shape((*w)[1], W, 1, name + " v_a") ||
            check_shape((*w)[2], W, -1, name + " b_UW") || check_shape((*w)[3], 1, -1, name + " b_v")) return 1;
        if (upload_mat(m, transpose_rows((*w)[0], 0, W), W, W, &m->Wa)) return 1;
        if (upload_vec(m, (*w)[1].data, &m->v_a) || upload_vec(m, (*w)[2].data, &m->b_UW)) return 1;
        m->b_v = (*w)[3].data[0];
        if (build_lstm(m, &(*w)[4], W + D, W, &m->dec[d - 1], name + " cell")) return 1;
        // E rows then W_a rows (both K-major over W): logits and the next step's query from one launch
        const std::vector<float> wa = transpose_rows((*w)[0], 0, W);
        const HostWeight& E = m->host["char_input_projection"][0];
        std::vector<float> ewa((size_t)(Vp + W) * W, 0.f);
        std::copy(E.data.begin(), E.data.begin() + (size_t)V * W, ewa.begin());
        std::copy(wa.begin(), wa.end(), ewa.begin() + (size_t)Vp * W);
        if (upload_mat(m, ewa, Vp + W, W, &m->EWa)) return 1;
    }
    m->fuse_pe = 3 * Vp <= 4 * W;
    if (m->fuse_pe) {
        // PE[n][v] = sum_u Kx[n][u] * E[v][u] for the first decoder cell (fp64 accumulation, one rounding)
        const HostWeight& E = m->host["char_input_projection"][0];
        const std::string name = "decoder_lstm_1";
        const HostWeight& K = m->host[name][d == 1 ? 4 : 0];
        const std::vector<float> kx = lstm_rows(K, W, 0, W);      // [4W][W]
        const std::vector<float> pe = product_f64(kx.data(), 4 * W, E.data.data(), V, W, Vp);
        if (upload_mat(m, pe, 4 * W, Vp, &m->PE)) return 1;
    }
    m->use_xtab = V <= 4096 && !getenv("COR_NO_XTAB");
    if (m->use_xtab) {
        // first encoder layer: x_t . Kx as a gather from (E . Kx) [V][4W], one table per scan direction
        const HostWeight& E = m->host["char_input_projection"][0];
        const std::vector<HostWeight>& l1 = m->host["encoder_lstm_1"];
        for (int dir = 0; dir < 2; ++dir) {
            const std::vector<float> kx = lstm_rows(l1[3 * dir], W, 0, W);     // [4W][W], gate-interleaved rows
            const std::vector<float> tab = product_f64(E.data.data(), V, kx.data(), 4 * W, W, 4 * W);
            if (upload_vec(m, tab, &m->xtab[dir])) return 1;
        }
    }
    m->side_attention = getenv("COR_ATTN_SIDE") != nullptr;
    if (!m->side && getenv("COR_OVERLAP")) {     // experiment switch: attention beside the lower-cell GEMM (no gain measured: DESIGN.md)
        COR_CUDA(cudaStreamCreateWithFlags(&m->side, cudaStreamNonBlocking));
        COR_CUDA(cudaEventCreateWithFlags(&m->ev_fork, cudaEventDisableTiming));
        COR_CUDA(cudaEventCreateWithFlags(&m->ev_join, cudaEventDisableTiming));
    }
    { void* gb = nullptr; if (dev_alloc(m, 256, &gb)) return 1; m->grid_bar = (unsigned*)gb; }
    { void* dc = nullptr; if (dev_alloc(m, 64, &dc)) return 1; m->dev_counters = (int64_t*)dc; COR_CUDA(cudaMemset(dc, 0, 64)); }
    if (m->split && m->passes == 3 && d == 2 && m->fuse_pe && !m->cfg.lm_predict) {
        // page decoder: the two cells' kernels concatenated along K, so that a CTA's K-slice is one contiguous range
        const LstmW& lo_w = m->dec[0]; const LstmW& tw = m->dec[1];
        const int K_low = Vp + W, K_top = 2 * W + D;
        m->LowCat.rows = 4 * W; m->LowCat.ld = K_low; m->TopCat.rows = 4 * W; m->TopCat.ld = K_top;
        if (dev_alloc(m, (size_t)4 * W * K_low * 2, (void**)&m->LowCat.hi) || dev_alloc(m, (size_t)4 * W * K_low * 2, (void**)&m->LowCat.lo) ||
            dev_alloc(m, (size_t)4 * W * K_top * 2, (void**)&m->TopCat.hi) || dev_alloc(m, (size_t)4 * W * K_top * 2, (void**)&m->TopCat.lo)) return 1;
        auto put = [&](const Mat& dst, int k0, const Mat& src, int k) -> int {
            for (int part = 0; part < 2; ++part) {
                char* d_ = reinterpret_cast<char*>(part ? dst.lo : dst.hi) + (size_t)k0 * 2;
                const char* s_ = reinterpret_cast<const char*>(part ? src.lo : src.hi);
                COR_CUDA(cudaMemcpy2D(d_, (size_t)dst.ld * 2, s_, (size_t)src.ld * 2, (size_t)k * 2, 4 * W, cudaMemcpyDeviceToDevice));
            }
            return 0;
        };
        if (put(m->LowCat, 0, m->PE, Vp) || put(m->LowCat, Vp, lo_w.R, W) ||
            put(m->TopCat, 0, tw.Kx, W) || put(m->TopCat, W, tw.Kc, D) || put(m->TopCat, W + D, tw.R, W)) return 1;
    }
    m->host.clear();
    m->finalized = true;
    return 0;
}

void cor_model_destroy(cor_model* m) {
    if (!m) return;
    cudaSetDevice(m->device);
    loop_graphs_release(m);
    if (m->side) { cudaStreamSynchronize(m->side); cudaStreamDestroy(m->side); }
    if (m->ev_fork) cudaEventDestroy(m->ev_fork);
    if (m->ev_join) cudaEventDestroy(m->ev_join);
    for (void* p : m->allocs) cudaFree(p);
    for (cudaEvent_t e : m->prof.ev) cudaEventDestroy(e);
    if (m->prof.rows) cudaFreeHost(m->prof.rows);
    tc_ctx_release(&m->ctx);
    delete m;
}

int cor_set_encoding(cor_model* m, int n_rows, int t_max, const float* enc_out, const float* states) {
    COR_CHECK(m && m->finalized, "model not finalized");
    COR_CHECK((enc_out == nullptr) == (states == nullptr), "enc_out and states must both be given or both be NULL");
    COR_CHECK(!enc_out || (n_rows > 0 && t_max > 0), "empty geometry");
    m->preset.enc_out = enc_out; m->preset.states = states; m->preset.n_rows = n_rows; m->preset.t_max = t_max;
    return 0;
}

int cor_workspace_bytes(const cor_model* m, int n_rows, int t_max, int k_alt, int want_alignments, size_t* bytes) {
    COR_CHECK(m && bytes, "null argument");
    Ws ws;
    if (carve(m, n_rows, t_max, k_alt, want_alignments, nullptr, &ws)) return 1;
    *bytes = ws.bytes;
    return 0;
}

static int bind(cor_model* m, const cor_lines* l, void* workspace, size_t bytes, int want_align, Ws* ws) {
    COR_CHECK(m && m->finalized, "model not finalized");
    COR_CHECK(l && workspace, "null argument");
    COR_CUDA(cudaSetDevice(m->device));
    if (carve(m, l->n_rows, l->t_max, l->k_alt, want_align, workspace, ws)) return 1;
    COR_CHECK(ws->bytes <= bytes, "workspace too small");
    return 0;
}

static int stage_lines(const Ws& ws, const cor_lines* l, cudaMemcpyKind kind, cudaStream_t st) {
    const size_t n = (size_t)l->n_rows * l->t_max * l->k_alt;
    COR_CUDA(cudaMemcpyAsync(ws.row_T, l->row_T, (size_t)l->n_rows * 4, kind, st));
    COR_CUDA(cudaMemcpyAsync(ws.idx, l->idx, n * 4, kind, st));
    if (l->val) COR_CUDA(cudaMemcpyAsync(ws.val, l->val, n * 4, kind, st));
    return 0;
}

// cor_encode records the call geometry in the first bytes of the workspace; the step and greedy
// entry points re-derive the carve from it.
struct WsHeader { int32_t B, T, k, align; uint32_t magic; };
static int write_header(void* workspace, size_t bytes, const Ws& ws, int align, cudaStream_t st) {
    COR_CHECK(ws.bytes <= bytes, "workspace too small");
    WsHeader h{ws.B, ws.T, ws.k, align, 0xC0A5B200u};
    COR_CUDA(cudaMemcpyAsync(workspace, &h, sizeof(h), cudaMemcpyHostToDevice, st));
    COR_CUDA(cudaStreamSynchronize(st));
    return 0;
}
static int read_header(cor_model* m, void* workspace, size_t bytes, Ws* ws, int* align, cudaStream_t st) {
    COR_CHECK(m && m->finalized && workspace && bytes >= 256, "bad arguments");
    COR_CUDA(cudaSetDevice(m->device));
    WsHeader h{};
    COR_CUDA(cudaMemcpyAsync(&h, workspace, sizeof(h), cudaMemcpyDeviceToHost, st));
    COR_CUDA(cudaStreamSynchronize(st));
    COR_CHECK(h.magic == 0xC0A5B200u, "workspace holds no encoded lines (call cor_encode first)");
    if (carve(m, h.B, h.T, h.k, h.align, workspace, ws)) return 1;
    COR_CHECK(ws->bytes <= bytes, "workspace smaller than at encode time");
    if (align) *align = h.align;
    return 0;
}

int cor_encode(cor_model* m, const cor_lines* lines, int want_alignments, void* workspace, size_t workspace_bytes,
               float* enc_out, float* states, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    Ws ws;
    if (!m || !lines) { set_error("null argument"); return 1; }
    CtxScope scope(&m->ctx);
    // alignments are decided at encode time, by the caller: the header records whether room was reserved
    const int align = want_alignments != 0;
    if (bind(m, lines, workspace, workspace_bytes, align, &ws)) return 1;
    if (stage_lines(ws, lines, cudaMemcpyDeviceToDevice, st)) return 1;
    if (!lines->val) ws.val = nullptr;
    if (run_encoder(m, ws, st)) return 1;
    const size_t BW = (size_t)ws.B * m->W;
    // the internal enc_out is line-major; the C ABI returns the reference's (T, B, D) enumeration
    if (enc_out && launch_join_lt(st, ws.enc_out(), ws.enc_out_lo(), enc_out, ws.B, ws.T, m->D)) return 1;
    if (states)
        for (int n = 0; n < m->d; ++n) {
            COR_CUDA(cudaMemcpyAsync(states + (2 * n) * BW, ws.st_h[n], BW * 4, cudaMemcpyDeviceToDevice, st));
            COR_CUDA(cudaMemcpyAsync(states + (2 * n + 1) * BW, ws.st_c[n], BW * 4, cudaMemcpyDeviceToDevice, st));
        }
    return write_header(workspace, workspace_bytes, ws, align, st);
}

int cor_decode_step(cor_model* m, void* workspace, size_t workspace_bytes, int n_hyp, const int32_t* row_line,
                    const float* p_in, const float* h_in, const float* c_in, const float* a_in,
                    float* p_out, float* lm_out, float* h_out, float* c_out, float* a_out, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    Ws ws;
    COR_CHECK(m, "null argument");
    CtxScope scope(&m->ctx);
    if (read_header(m, workspace, workspace_bytes, &ws, nullptr, st)) return 1;
    COR_CHECK(n_hyp > 0 && n_hyp <= ws.B, "hypothesis count must be in 1..n_rows of the encoded call");
    COR_CHECK(p_in && h_in && c_in && a_in && p_out && h_out && c_out && a_out, "null argument");
    const int R = n_hyp, W = m->W, d = m->d;
    const size_t RW = (size_t)R * W;
    // stage the caller's rows into the step buffers of the workspace (stride V -> Vp for p)
    // (staged through the logits buffer: the fp16 split must not run in place)
    float* p_stage = ws.p_lo ? ws.logits : ws.p;
    COR_CUDA(cudaMemsetAsync(p_stage, 0, (size_t)R * m->Vp * 4, st));
    COR_CUDA(cudaMemcpy2DAsync(p_stage, (size_t)m->Vp * 4, p_in, (size_t)m->V * 4, (size_t)m->V * 4, R, cudaMemcpyDeviceToDevice, st));
    if (ws.p_lo && launch_split(st, p_stage, ws.p, ws.p_lo, (size_t)R * m->Vp)) return 1;
    StepIO io{};
    io.R = R; io.row_line = row_line; io.first = false;
    io.p = ws.p; io.p_lo = ws.p_lo; io.x = ws.x; io.x_lo = ws.x_lo;
    for (int n = 0; n < d; ++n) {
        if (copy_in(st, h_in + n * RW, ws.dh[0][n], ws.dh_lo[0][n], RW)) return 1;
        io.h_in[n] = ws.dh[0][n]; io.h_in_lo[n] = ws.dh_lo[0][n];
        io.h_out[n] = ws.dh[1][n]; io.h_out_lo[n] = ws.dh_lo[1][n];
        io.c_in[n] = c_in + n * RW; io.c_out[n] = c_out + n * RW;
    }
    io.a_in = a_in; io.a_out = a_out;
    io.wq = ws.wq; io.ctx = ws.ctx; io.ctx_lo = ws.ctx_lo; io.logits = ws.logits;
    if (m->cfg.lm_predict && lm_out) {
        io.h_lm = ws.h_lm; io.h_lm_lo = ws.h_lm_lo; io.c_lm = ws.c_lm; io.logits_lm = ws.logits_lm;
    }
    if (decoder_core(m, ws, io, st)) return 1;
    if (launch_softmax(st, ws.logits, ws.logits, nullptr, R, nullptr, m->V)) return 1;
    COR_CUDA(cudaMemcpy2DAsync(p_out, (size_t)m->V * 4, ws.logits, (size_t)m->Vp * 4, (size_t)m->V * 4, R, cudaMemcpyDeviceToDevice, st));
    if (io.logits_lm) {
        if (launch_softmax(st, ws.logits_lm, ws.p_lm, nullptr, R, nullptr, m->V)) return 1;
        COR_CUDA(cudaMemcpy2DAsync(lm_out, (size_t)m->V * 4, ws.p_lm, (size_t)m->Vp * 4, (size_t)m->V * 4, R, cudaMemcpyDeviceToDevice, st));
    }
    for (int n = 0; n < d; ++n)
        if (copy_out(st, ws.dh[1][n], ws.dh_lo[1][n], h_out + n * RW, RW)) return 1;
    return 0;
}

static int greedy_results(const Ws& ws, int32_t* out_chars, float* out_probs, double* out_scores, int32_t* out_len,
                          int32_t* out_ties, float* out_align, cudaMemcpyKind kind, cudaStream_t st) {
    const size_t BS = (size_t)ws.B * 2 * ws.T;
    if (out_chars) COR_CUDA(cudaMemcpyAsync(out_chars, ws.o_chars, BS * 4, kind, st));
    if (out_probs) COR_CUDA(cudaMemcpyAsync(out_probs, ws.o_probs, BS * 4, kind, st));
    if (out_scores) COR_CUDA(cudaMemcpyAsync(out_scores, ws.o_score, (size_t)ws.B * 8, kind, st));
    if (out_len) COR_CUDA(cudaMemcpyAsync(out_len, ws.o_len, (size_t)ws.B * 4, kind, st));
    if (out_ties) COR_CUDA(cudaMemcpyAsync(out_ties, ws.o_ties, (size_t)ws.B * 4, kind, st));
    if (out_align) {
        COR_CHECK(ws.o_align, "alignments were not reserved in this workspace");
        COR_CUDA(cudaMemcpyAsync(out_align, ws.o_align, BS * ws.T * 4, kind, st));
    }
    return 0;
}

int cor_decode_greedy_batch(cor_model* m, void* workspace, size_t workspace_bytes,
                            int32_t* out_chars, float* out_probs, double* out_scores,
                            int32_t* out_len, int32_t* out_ties, float* out_align, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    Ws ws;
    COR_CHECK(m, "null argument");
    CtxScope scope(&m->ctx);
    if (read_header(m, workspace, workspace_bytes, &ws, nullptr, st)) return 1;
    COR_CHECK(!m->cfg.lm_predict, "fast mode with lm_predict is undefined in the reference (seq2seq.py:1248)");
    if (greedy_loop(m, ws, st)) return 1;
    return greedy_results(ws, out_chars, out_probs, out_scores, out_len, out_ties, out_align, cudaMemcpyDeviceToDevice, st);
}

int cor_correct_greedy_host(cor_model* m, const cor_lines* host_lines, void* workspace, size_t workspace_bytes,
                            int32_t* out_chars, float* out_probs, double* out_scores,
                            int32_t* out_len, int32_t* out_ties, float* out_align, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    Ws ws;
    COR_CHECK(m, "null argument");
    CtxScope scope(&m->ctx);
    if (bind(m, host_lines, workspace, workspace_bytes, out_align != nullptr, &ws)) return 1;
    COR_CHECK(!m->cfg.lm_predict, "fast mode with lm_predict is undefined in the reference (seq2seq.py:1248)");
    if (stage_lines(ws, host_lines, cudaMemcpyHostToDevice, st)) return 1;
    if (!host_lines->val) ws.val = nullptr;
    if (run_encoder(m, ws, st)) return 1;
    if (greedy_loop(m, ws, st)) return 1;
    if (greedy_results(ws, out_chars, out_probs, out_scores, out_len, out_ties, out_align, cudaMemcpyDeviceToHost, st)) return 1;
    COR_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int cor_alignment_paths(const float* align, int64_t stride_step, int64_t stride_row,
                        int n_rows, int t_max, int s_max, const int32_t* in_len, const int32_t* out_len,
                        double min_score, int32_t* out_path, int path_ld, double* out_dist, void* stream) {
    COR_CHECK(align && in_len && out_len && out_path && out_dist, "null pointer");
    return launch_alignment_paths((cudaStream_t)stream, align, stride_step, stride_row, n_rows, t_max, s_max,
                                  in_len, out_len, min_score, out_path, path_ld, out_dist);
}

}  // extern "C"
