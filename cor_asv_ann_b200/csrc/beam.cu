// beam.cu — GPU beam search: decode_sequence_beam (ocrd_cor_asv_ann/lib/seq2seq.py:1356-1544) and
// its Node trie (:1546-1608) for many lines at once, entirely on the device.
//
// The reference runs, per line, a best-first (A*-like) search: a global sorted list of open
// hypotheses, from which every iteration pops the <= batch_size best, runs ONE decoder step on them,
// and inserts up to beam_width_in (+1 rejection) children per hypothesis back into the list.
// Here a "wave" of L lines runs those iterations in lock step; every iteration is the kernel sequence
//     pop -> scan -> rowmap -> gather -> [decoder step kernels on R rows] -> expand -> push
// with R (the number of hypothesis rows over all lines) known only on the device (r_dev).
//
// Data layout in HBM (per wave):
//   * expansion records, one per hypothesis that went through a decoder step, line-local index xid
//     in a fixed per-line arena of quota_x records: the output softmax row after the rejection
//     overwrite (V floats), the new LSTM states (2*depth*W), the alignment row (T), and <= 16 children
//     (character, probability, fp64 cumulative cost, rejection flag).
//   * a hypothesis (trie node) is the pair (xid of its parent's record, child slot k), numbered
//     id = 1 + 16*xid + k; id 0 is the root.  Ids grow in insertion order, which is the reference's
//     tie-break (insort_left + pop(): first inserted, first popped among equal keys).
//   * the feedback vector of child k is the record's softmax row with children 0..k-1 zeroed
//     (seq2seq.py:1506-1514: `scores1 = np.copy(scores); scores[idx] = 0`), rebuilt at gather time.
//   * per line a sorted fringe (fp64 key, id), double-buffered; push = bitonic sort of the new
//     children in shared memory + rank-merge with the old fringe, truncated to 2T*batch_size (:1531).
// A line whose arena (or the wave's row capacity) overflows is reported back and re-run by the host
// loop in a smaller wave with a larger arena; results do not depend on the wave a line ran in.
#include "model.h"
#include "epilogue.cuh"
#include "graphs.h"
#include <math.h>
#include <float.h>
#include <string.h>
#include <stdio.h>
#include <stdlib.h>
#include <algorithm>
#include <vector>

namespace cor {

namespace {

constexpr int KMAX = 16;             // children per expansion: beam_width_in (<= 15) + rejection candidate
constexpr int MAX_NEW = 4096;        // batch_size (<= 256) * KMAX
constexpr int ST_ACTIVE = 0, ST_DONE = 1, ST_OVERFLOW = 2, ST_EMPTY = 3;
constexpr uint32_t FINAL_BIT = 0x80000000u;
constexpr int kStats = 8;            // int64 statistics per line (include/cor_b200.h)
constexpr int FMAX = 64;             // final hypotheses kept per line for the n-best read-out (COR_BEAM_NBEST_MAX)

struct BeamDev {
    // geometry and parameters
    int L, T, V, Vp, W, d, D, k_alt;
    int batch, width_in, width_out, eos, lm;
    double thr_in, rej_thr;
    int quota_x, cap, Rcap;
    // encoder products (views into the encoder workspace)
    const int32_t* row_T; const uint8_t* live; const int32_t* idx; const float* val;
    const float* st_h[8]; const float* st_c[8];
    // per line
    int32_t *status, *iter, *x_count, *x_base, *beam_count, *row_base, *fr_len, *fr_off, *fr_buf;
    int32_t *finals_count, *best_final, *open_left;
    double* best_key; long long* hyp_steps;
    // diagnostics per line (stats[4..7]): exact top-2 score ties, near-tie ordering decisions, order-sensitive checksum of
    // the popped hypotheses, peak fringe length
    long long *score_ties, *near_ties; unsigned long long* pop_sum; int32_t* fr_peak;
    // Reach truncation (see reach_keep): the fringe in memory is a PREFIX of the reference's list; fr_gone = entries of that
    // list beyond the prefix (dropped because no remaining iteration can pop them), fin_in = upper bound of the finals in the prefix
    int32_t *fr_gone, *fin_in; int reach_trunc;
    int32_t* push_done;                    // parts of beam_push_kernel that have finished a line's merge
    // final hypotheses in pop order, the first FMAX of them (decode_sequence_beam yields every final, best first)
    uint32_t* fin_id; double* fin_key; int n_best;
    int32_t* r_chars; float* r_probs; double* r_score; int32_t* r_len;      // result staging [L][n_best][...]
    double* fr_key; uint32_t* fr_id;       // [2][L][cap]
    uint32_t* beam_nodes;                  // [L][batch]
    int32_t* rej_table;                    // [L][T]
    int32_t* counters;                     // [0] rows this iteration (r_dev), [1] active lines, [2] lines on the heavy push list
    int32_t* heavy;                        // [L] lines whose new children do not fit one warp
    // per record g = line*quota_x + xid
    float *rec_p, *rec_h, *rec_c, *rec_a;  // [G][Vp], [G][d][W], [G][d][W], [G][T]
    float* rec_wq;                         // [G][W] attention query of the children (h_top . W_a + b)
    float* wq;                             // [Rcap][W] query rows of the current iteration
    double* rec_cur; uint8_t* rec_amax1; int32_t *rec_pos, *rec_len, *rec_nacc; uint32_t* rec_parent;
    int32_t* ch_v; float* ch_p; double* ch_cum; uint8_t* ch_rej;   // [G][KMAX]
    // per row of the current iteration
    int32_t *row_line, *row_xid; uint32_t* row_node;
    long long* o_stats; float* o_align;    // result staging: [L][4], [L][2T][T] (optional)
    float *p, *p_lo, *logits, *logits_lm, *a_in, *a_out;
    float *h_in[8], *h_in_lo[8], *h_out[8], *h_out_lo[8], *c_in[8], *c_out[8];
};

__device__ __forceinline__ bool key_less(double ka, uint32_t ia, double kb, uint32_t ib) {
    return ka < kb || (ka == kb && (ia & ~FINAL_BIT) < (ib & ~FINAL_BIT));
}

// ------------------------------------------------------------------ init
// Root node (seq2seq.py:1389-1394) and the per-position rejection candidates
// (np.any(source_scores) / np.nanargmax(source_scores), :1458-1462).
__global__ void beam_init_kernel(BeamDev b) {
    const int line = blockIdx.x;
    const int Tl = b.row_T[line];
    for (int t = threadIdx.x; t < b.T; t += blockDim.x) {
        int best = -1; float bv = 0.f;
        if (t < Tl) {
            for (int a = 0; a < b.k_alt; ++a) {
                const size_t o = ((size_t)line * b.T + t) * b.k_alt + a;
                const int v = b.idx[o];
                if (v < 0) continue;
                const float x = b.val ? b.val[o] : 1.0f;
                // the dense row holds zeros elsewhere, so only a positive entry can be the (nan)argmax;
                // first index among equal maxima
                if (!(x > 0.0f)) continue;
                if (best < 0 || x > bv || (x == bv && v < best)) { best = v; bv = x; }
            }
        }
        b.rej_table[(size_t)line * b.T + t] = best;
    }
    if (threadIdx.x == 0) {
        b.status[line] = b.live[line] ? ST_ACTIVE : ST_EMPTY;
        b.iter[line] = 0; b.x_count[line] = 0; b.x_base[line] = 0; b.beam_count[line] = 0; b.row_base[line] = 0;
        b.fr_len[line] = 1; b.fr_off[line] = 0; b.fr_buf[line] = 0;
        b.finals_count[line] = 0; b.best_final[line] = -1; b.open_left[line] = 0;
        b.best_key[line] = INFINITY; b.hyp_steps[line] = 0;
        b.score_ties[line] = 0; b.near_ties[line] = 0; b.pop_sum[line] = 0ull; b.fr_peak[line] = 1;
        b.fr_gone[line] = 0; b.fin_in[line] = 0; b.push_done[line] = 0;
        b.fr_key[(size_t)line * b.cap] = 3.0 * fabs(1.0 - (double)Tl);
        b.fr_id[(size_t)line * b.cap] = 0u;
    }
}

// ------------------------------------------------------------------ pop (seq2seq.py:1398-1420)
// One warp per line: the fringe is sorted, so the pop is a prefix of it — walked 32 entries at a time
// until batch_size open nodes have been seen (a few chunks at most, whatever the fringe length).
// `bump`: also reserve the expansion records of the popped nodes (x_base / x_count), which beam_rowmap_kernel does on the
// three-kernel path.
__device__ void pop_line(const BeamDev& b, int line, int lane, bool bump = false) {
    if (b.status[line] != ST_ACTIVE) { if (lane == 0) b.beam_count[line] = 0; return; }
    const int Tl = b.row_T[line];
    const int len = b.fr_len[line];
    if (b.iter[line] >= 2 * Tl) {                    // max_batches reached (:1397-1398)
        if (lane == 0) { b.status[line] = ST_DONE; b.beam_count[line] = 0; b.open_left[line] = len + b.fr_gone[line]; }
        return;
    }
    const int cur = b.fr_buf[line];
    const double* keys = b.fr_key + ((size_t)cur * b.L + line) * b.cap;
    const uint32_t* ids = b.fr_id + ((size_t)cur * b.L + line) * b.cap;
    int first_final = 0x7fffffff, first_open = 0x7fffffff;
    int open_total = 0, consumed = 0, nfin = 0;
    // checksum of WHICH hypotheses are expanded in WHICH iteration: sum over pops of (length << 16 | char + 1) * odd(iteration)
    // -- what the oracle computes from its own beam lists.  The order inside one batch is left out on purpose: two
    // hypotheses whose keys differ in the last bits may swap places there without any effect on the search.
    const unsigned long long it_odd = 2ull * (unsigned long long)b.iter[line] + 1ull;
    unsigned long long csum = 0ull;
    const int fc0 = b.finals_count[line];
    // Rounds of PC chunks of 32 entries: the ids of a round are fetched together, and so are the record reads of the
    // checksum (one chunk after the other made every chunk two dependent memory round trips: 8+ chunks per batch of 256).
    constexpr int PC = 4;
    for (int base = 0; base < len && open_total < b.batch; base += 32 * PC) {
        uint32_t idv[PC];
        int slot[PC];                                   // >= 0: a popped open node and its place in the batch
#pragma unroll
        for (int c = 0; c < PC; ++c) { const int i = base + 32 * c + lane; idv[c] = i < len ? ids[i] : 0u; }
#pragma unroll
        for (int c = 0; c < PC; ++c) {
            slot[c] = -1;
            const int cb = base + 32 * c;
            if (cb < len && open_total < b.batch) {     // (the same for every lane)
                const bool valid = cb + lane < len;
                const bool fin = valid && (idv[c] & FINAL_BIT);
                const bool open = valid && !fin;
                const unsigned bal = __ballot_sync(0xffffffffu, open);
                const int excl = open_total + __popc(bal & ((1u << lane) - 1u));
                // an entry is popped iff fewer than batch_size open nodes precede it (:1400-1415)
                const bool take = valid && excl < b.batch;
                if (take && open) slot[c] = excl;
                const unsigned t_open = __ballot_sync(0xffffffffu, take && open);
                const unsigned t_fin = __ballot_sync(0xffffffffu, take && fin);
                if (t_open && first_open == 0x7fffffff) first_open = cb + __ffs(t_open) - 1;
                if (t_fin && first_final == 0x7fffffff) first_final = cb + __ffs(t_fin) - 1;
                consumed += __popc(t_open | t_fin);
                nfin += __popc(t_fin);
                open_total += __popc(bal);
            }
        }
#pragma unroll
        for (int c = 0; c < PC; ++c) {
            if (slot[c] >= 0) {
                const uint32_t id = idv[c];
                unsigned long long x = (1ull << 16) + 1ull;            // root: length 1, no character
                if (id != 0u) {
                    const size_t pg = (size_t)line * b.quota_x + (id - 1u) / KMAX;
                    x = ((unsigned long long)(b.rec_len[pg] + 1) << 16) + (unsigned long long)b.ch_v[pg * KMAX + (id - 1u) % KMAX] + 1ull;
                }
                csum += x * it_odd * 0x9E3779B97F4A7C15ull;
            }
        }
#pragma unroll
        for (int c = 0; c < PC; ++c)
            if (slot[c] >= 0) b.beam_nodes[(size_t)line * b.batch + slot[c]] = idv[c];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) csum += __shfl_xor_sync(0xffffffffu, csum, o);
    if (lane != 0) return;
    const int count = min(open_total, b.batch);
    int fc = fc0 + nfin;
    double bk = b.best_key[line];
    if (nfin > 0) {
        const double kf = keys[first_final];
        if (kf < bk) { bk = kf; b.best_key[line] = kf; b.best_final[line] = (int32_t)(ids[first_final] & ~FINAL_BIT); }
    }
    if (b.n_best > 1 && nfin > 0) {
        // n-best read-out: keep the FMAX best finals seen so far (unsorted; beam_finish_kernel sorts them).  Finals are
        // rare events, so one lane walks the consumed prefix.
        uint32_t* fid = b.fin_id + (size_t)line * FMAX; double* fk = b.fin_key + (size_t)line * FMAX;
        int have = min(fc0, FMAX);
        for (int i = 0; i < consumed; ++i) {
            const uint32_t id = ids[i];
            if (!(id & FINAL_BIT)) continue;
            const uint32_t nid_ = id & ~FINAL_BIT; const double k = keys[i];
            if (have < FMAX) { fid[have] = nid_; fk[have] = k; ++have; continue; }
            int worst = 0;
            for (int q = 1; q < FMAX; ++q) if (key_less(fk[worst], fid[worst], fk[q], fid[q])) worst = q;
            if (key_less(k, nid_, fk[worst], fid[worst])) { fid[worst] = nid_; fk[worst] = k; }
        }
    }
    b.finals_count[line] = fc;
    b.fr_off[line] = consumed;
    b.fin_in[line] -= nfin;
    const int gone = b.fr_gone[line];
    int out = count;
    if (count == 0) {
        b.status[line] = ST_DONE; b.open_left[line] = len + gone - consumed;
    } else if (fc > b.width_out && bk < keys[first_open]) {       // :1418-1420
        b.status[line] = ST_DONE; b.open_left[line] = len + gone - consumed; out = 0;
    } else if (b.x_count[line] + count > b.quota_x) {
        b.status[line] = ST_OVERFLOW; out = 0;
    } else {
        b.iter[line] += 1;
        b.hyp_steps[line] += count;
        b.pop_sum[line] += csum;
        // near-tie at the pop boundary: which of two hypotheses enters this batch hangs on the last bits of their keys
        if (consumed < len && consumed > 0 && fabs(keys[consumed] - keys[consumed - 1]) <= 1e-6 * fmax(1.0, fabs(keys[consumed]))) b.near_ties[line] += 1;
    }
    b.beam_count[line] = out;
    if (bump && out > 0) { const int xb = b.x_count[line]; b.x_base[line] = xb; b.x_count[line] = xb + out; }
}
__global__ void __launch_bounds__(256) beam_pop_kernel(BeamDev b) {
    const int line = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (line < b.L) pop_line(b, line, threadIdx.x & 31);
}

// ------------------------------------------------------------------ scan: row offsets, row total, active lines
// One block; thread t owns the consecutive lines [t*per, (t+1)*per): all loads are issued at once,
// one block-wide scan of the per-thread sums, then the row bases.  Row order = line order.
// `set_cond`: the launch is part of the iteration graph (run_wave): the kernel counts the body and decides whether the
// WHILE node runs another one (cudaGraphSetConditional is only valid inside the graph that owns the handle).
struct LoopCtl {
    cudaGraphConditionalHandle cond;    // WHILE: another iteration?
    cudaGraphConditionalHandle shape;   // SWITCH: tile shape class of the next iteration's GEMMs, from its row count
    int has_shape, r_narrow, r_single;
};
__device__ __forceinline__ void set_loop(const LoopCtl& lc, int rows, bool more) {
    if (lc.has_shape) cudaGraphSetConditional(lc.shape, rows <= lc.r_narrow ? 0u : (rows <= lc.r_single ? 1u : 2u));
    cudaGraphSetConditional(lc.cond, more ? 1u : 0u);
}
__global__ void __launch_bounds__(1024) beam_scan_kernel(BeamDev b, LoopCtl lc, int set_cond) {
    __shared__ int s_wsum[32];
    __shared__ int s_total, s_active;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) { s_total = 0; s_active = 0; b.counters[2] = 0; }
    const int per = (b.L + 1023) / 1024;
    const int l0 = tid * per, l1 = min(b.L, l0 + per);
    int mine = 0;
    for (int line = l0; line < l1; ++line) mine += b.beam_count[line];
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
    if (lane == 31) s_wsum[warp] = incl;
    __syncthreads();
    int off = 0;
    for (int w = 0; w < warp; ++w) off += s_wsum[w];
    int rb = off + incl - mine;
    int total = 0, active = 0;
    for (int line = l0; line < l1; ++line) {
        int c = b.beam_count[line];
        if (c > 0 && rb + c > b.Rcap) { b.status[line] = ST_OVERFLOW; b.beam_count[line] = 0; }
        else if (c > 0) total = rb + c;
        b.row_base[line] = rb;
        rb += c;
        if (b.status[line] == ST_ACTIVE) ++active;
    }
    if (total) atomicMax(&s_total, total);
    if (active) atomicAdd(&s_active, active);
    __syncthreads();
    if (tid == 0) {
        b.counters[0] = s_total; b.counters[1] = s_active;
        if (set_cond) {
            // a line runs at most 2T iterations (beam_pop_kernel), so the bound below never cuts a search short; it keeps a
            // protocol bug from looping on the device forever
            const int bodies = ++b.counters[3];
            set_loop(lc, s_total, (s_total > 0 || s_active > 0) && bodies <= 2 * b.T + 2);
        }
    }
}

// First node of the iteration graph: the WHILE condition from the counters the last directly launched scan left.
__global__ void beam_cond_kernel(BeamDev b, const LoopCtl* lc) {
    set_loop(*lc, b.counters[0], b.counters[0] > 0 || b.counters[1] > 0);
}

// One warp per line: the rows of a line are its popped nodes in fringe order.
__device__ __forceinline__ void rowmap_line(const BeamDev& b, int line, int lane, int c, int rb) {
    if (c <= 0) return;
    const int xb = b.x_count[line];
    for (int t = lane; t < c; t += 32) {
        b.row_line[rb + t] = line;
        b.row_node[rb + t] = b.beam_nodes[(size_t)line * b.batch + t];
        b.row_xid[rb + t] = xb + t;
    }
    __syncwarp();
    if (lane == 0) { b.x_base[line] = xb; b.x_count[line] = xb + c; }
}
__global__ void __launch_bounds__(256) beam_rowmap_kernel(BeamDev b) {
    const int line = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (line >= b.L) return;
    rowmap_line(b, line, threadIdx.x & 31, b.beam_count[line], b.row_base[line]);
}

// pop + scan + rowmap of a page of lines (L <= kSelectSmall) in ONE block: a warp per line (two rounds), the scan by
// the first warp, no trip through another launch between them.  Same procedure and same results as the three kernels.
constexpr int kSelectSmall = 64;
__global__ void __launch_bounds__(1024) beam_select_small_kernel(BeamDev b, LoopCtl lc, int set_cond) {
    __shared__ int s_cnt[kSelectSmall], s_base[kSelectSmall];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) b.counters[2] = 0;
    for (int line = warp; line < b.L; line += 32) pop_line(b, line, lane);
    __syncthreads();
    if (warp == 0) {
        int carry = 0, total = 0, active = 0;
        for (int base = 0; base < b.L; base += 32) {
            const int line = base + lane;
            const int c0 = line < b.L ? b.beam_count[line] : 0;
            int incl = c0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
            const int rb = carry + incl - c0;
            int c = c0;
            if (line < b.L) {
                if (c > 0 && rb + c > b.Rcap) { b.status[line] = ST_OVERFLOW; b.beam_count[line] = 0; c = 0; }
                b.row_base[line] = rb;
                s_cnt[line] = c; s_base[line] = rb;
            }
            int t = c > 0 ? rb + c : 0, a = (line < b.L && b.status[line] == ST_ACTIVE) ? 1 : 0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) { t = max(t, __shfl_xor_sync(0xffffffffu, t, o)); a += __shfl_xor_sync(0xffffffffu, a, o); }
            total = max(total, t); active += a;
            carry += __shfl_sync(0xffffffffu, incl, 31);
        }
        if (lane == 0) {
            b.counters[0] = total; b.counters[1] = active;
            if (set_cond) {
                const int bodies = ++b.counters[3];
                set_loop(lc, total, (total > 0 || active > 0) && bodies <= 2 * b.T + 2);
            }
        }
    }
    __syncthreads();
    for (int line = warp; line < b.L; line += 32) rowmap_line(b, line, lane, s_cnt[line], s_base[line]);
}

// ------------------------------------------------------------------ gather: decoder-step inputs of every row
// (seq2seq.py:1423-1429: vstack of node.scores / node.state)
__device__ __forceinline__ void gather_row(const BeamDev& b, int split, int row, int line, uint32_t node) {
    // 128-bit accesses throughout: a record is ~12 KB (p, h, c, query) and is read exactly once
    const int W4 = b.W >> 2, V4 = b.Vp >> 2;
    do {
        float* p_lo = split ? b.p_lo : nullptr;       // split operands: base pointers + element offsets only
        const size_t po = (size_t)row * b.Vp;
        float* a_in = b.a_in + (size_t)row * b.T;
        if (node == 0u) {          // root: zero feedback, encoder final states, zero alignment (query: decoder_core)
            for (int v = threadIdx.x; v < V4; v += blockDim.x) store_split4(b.p, p_lo, po + 4 * v, make_float4(0.f, 0.f, 0.f, 0.f));
            for (int j = threadIdx.x; j < b.T; j += blockDim.x) a_in[j] = 0.f;
            for (int n = 0; n < b.d; ++n) {
                const float4* sh = reinterpret_cast<const float4*>(b.st_h[n] + (size_t)line * b.W);
                const float4* sc = reinterpret_cast<const float4*>(b.st_c[n] + (size_t)line * b.W);
                float4* ci = reinterpret_cast<float4*>(b.c_in[n] + (size_t)row * b.W);
                for (int u = threadIdx.x; u < W4; u += blockDim.x) {
                    store_split4(b.h_in[n], split ? b.h_in_lo[n] : nullptr, (size_t)row * b.W + 4 * u, sh[u]);
                    ci[u] = sc[u];
                }
            }
            break;
        }
        const int k = (int)((node - 1u) % KMAX);
        const size_t g = (size_t)line * b.quota_x + (node - 1u) / KMAX;
        const float4* rp = reinterpret_cast<const float4*>(b.rec_p + g * b.Vp);
        const int32_t* acc = b.ch_v + g * KMAX;
        const float* ra = b.rec_a + g * b.T;
        const float4* rq = reinterpret_cast<const float4*>(b.rec_wq + g * b.W);
        float4* wq = reinterpret_cast<float4*>(b.wq + (size_t)row * b.W);
        const int nl = min(b.d, 2);
        // every load of a 128-wide block of the record (scores, alignment, query, h and c of the first two layers)
        // is issued before the first store: one memory round trip per block instead of one per array
        const int span = max(max(V4, W4), b.T);
        for (int base = 0; base < span; base += blockDim.x) {
            const int i = base + threadIdx.x;
            float4 x = make_float4(0.f, 0.f, 0.f, 0.f), tq = x, th[2], tc[2];
            float ta = 0.f;
            if (i < V4) x = rp[i];
            if (i < b.T) ta = ra[i];
            if (i < W4) {
                tq = rq[i];
#pragma unroll
                for (int n = 0; n < 2; ++n)
                    if (n < nl) {
                        th[n] = reinterpret_cast<const float4*>(b.rec_h + (g * b.d + n) * b.W)[i];
                        tc[n] = reinterpret_cast<const float4*>(b.rec_c + (g * b.d + n) * b.W)[i];
                    }
            }
            if (i < V4) {
                for (int q = 0; q < k; ++q) {                // better siblings were zeroed before the copy
                    const int z = acc[q] - 4 * i;
                    if (z == 0) x.x = 0.f; else if (z == 1) x.y = 0.f; else if (z == 2) x.z = 0.f; else if (z == 3) x.w = 0.f;
                }
                store_split4(b.p, p_lo, po + 4 * i, x);
            }
            if (i < b.T) a_in[i] = ta;
            if (i < W4) {
                wq[i] = tq;
#pragma unroll
                for (int n = 0; n < 2; ++n)
                    if (n < nl) {
                        store_split4(b.h_in[n], split ? b.h_in_lo[n] : nullptr, (size_t)row * b.W + 4 * i, th[n]);
                        reinterpret_cast<float4*>(b.c_in[n] + (size_t)row * b.W)[i] = tc[n];
                    }
            }
        }
        for (int n = 2; n < b.d; ++n) {
            const float4* rh = reinterpret_cast<const float4*>(b.rec_h + (g * b.d + n) * b.W);
            const float4* rc = reinterpret_cast<const float4*>(b.rec_c + (g * b.d + n) * b.W);
            float4* ci = reinterpret_cast<float4*>(b.c_in[n] + (size_t)row * b.W);
            for (int u = threadIdx.x; u < W4; u += blockDim.x) {
                store_split4(b.h_in[n], split ? b.h_in_lo[n] : nullptr, (size_t)row * b.W + 4 * u, rh[u]);
                ci[u] = rc[u];
            }
        }
    } while (0);
}
__global__ void __launch_bounds__(128) beam_gather_kernel(BeamDev b, int split) {
    const int R = b.counters[0];
    for (int row = blockIdx.x; row < R; row += gridDim.x) gather_row(b, split, row, b.row_line[row], b.row_node[row]);
}

// Page of lines (L <= kSelectSmall), line-kernel path: the scan over the lines' pop counts is so small that EVERY block of
// the gather redoes it in shared memory instead of waiting for another launch; block 0 publishes the row count for the
// GEMMs, marks overflowing lines and sets the loop controls.  A row finds its line by bisection over <= 64 bases.
__global__ void __launch_bounds__(128) beam_gather_scan_kernel(BeamDev b, int split, LoopCtl lc, int set_cond) {
    __shared__ int s_base[kSelectSmall], s_cnt[kSelectSmall];
    __shared__ int s_total;
    const int lane = threadIdx.x & 31;
    if (threadIdx.x < 32) {
        int carry = 0, total = 0, active = 0;
        for (int base = 0; base < b.L; base += 32) {
            const int line = base + lane;
            const bool live = line < b.L && b.status[line] == ST_ACTIVE;
            const int c0 = live ? b.beam_count[line] : 0;
            int incl = c0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
            const int rb = carry + incl - c0;
            const bool over = c0 > 0 && rb + c0 > b.Rcap;
            const int c = over ? 0 : c0;
            if (line < b.L) {
                s_cnt[line] = c; s_base[line] = rb;
                if (blockIdx.x == 0) { b.row_base[line] = rb; if (over) b.status[line] = ST_OVERFLOW; }
            }
            int t = c > 0 ? rb + c : 0, a = (live && !over) ? 1 : 0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) { t = max(t, __shfl_xor_sync(0xffffffffu, t, o)); a += __shfl_xor_sync(0xffffffffu, a, o); }
            total = max(total, t); active += a;
            carry += __shfl_sync(0xffffffffu, incl, 31);
        }
        if (lane == 0) {
            s_total = total;
            if (blockIdx.x == 0) {
                b.counters[0] = total; b.counters[1] = active; b.counters[2] = 0;
                if (set_cond) {
                    const int bodies = ++b.counters[3];
                    set_loop(lc, total, (total > 0 || active > 0) && bodies <= 2 * b.T + 2);
                }
            }
        }
    }
    __syncthreads();
    const int R = s_total;
    for (int row = blockIdx.x; row < R; row += gridDim.x) {
        int lo = 0, hi = b.L - 1;                  // last line whose base is <= row
        while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (s_base[mid] <= row) lo = mid; else hi = mid - 1; }
        const int line = lo, j = row - s_base[lo];
        const uint32_t node = b.beam_nodes[(size_t)line * b.batch + j];
        if (threadIdx.x == 0) { b.row_line[row] = line; b.row_node[row] = node; b.row_xid[row] = b.x_base[line] + j; }
        gather_row(b, split, row, line, node);
    }
}

// ------------------------------------------------------------------ expand: softmax, rejection, candidates
// (seq2seq.py:1436-1529), one block per hypothesis row.
struct ArgMax { float v; int i; };
__device__ __forceinline__ ArgMax better(ArgMax a, ArgMax c) {
    // descending walk of an ascending argsort: among equal scores the higher index comes first
    return (c.v > a.v || (c.v == a.v && c.i > a.i)) ? c : a;
}

__global__ void __launch_bounds__(128) beam_expand_kernel(BeamDev b, int split) {
    extern __shared__ float sm[];
    float* sp = sm;                 // [Vp] working copy of the scores (taken entries -> -1)
    float* slm = sm + b.Vp;         // [Vp] LM scores (lm only)
    __shared__ float s_red[4];
    __shared__ double s_dred[4];
    __shared__ ArgMax s_arg[4];
    __shared__ int s_cnt[4];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int R = b.counters[0];
    for (int row = blockIdx.x; row < R; row += gridDim.x) {
        const int line = b.row_line[row];
        const uint32_t node = b.row_node[row];
        const int Tl = b.row_T[line];
        const size_t g = (size_t)line * b.quota_x + b.row_xid[row];
        __syncthreads();
        // ---- softmax of the logits row (and of the LM logits)
        for (int pass = 0; pass < (b.lm ? 2 : 1); ++pass) {
            const float* z = (pass ? b.logits_lm : b.logits) + (size_t)row * b.Vp;
            float* dst = pass ? slm : sp;
            float m = -INFINITY;
            for (int v = tid; v < b.V; v += 128) m = fmaxf(m, z[v]);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
            if (lane == 0) s_red[warp] = m;
            __syncthreads();
            m = fmaxf(fmaxf(s_red[0], s_red[1]), fmaxf(s_red[2], s_red[3]));
            __syncthreads();
            float s = 0.f;
            for (int v = tid; v < b.V; v += 128) s += expf(z[v] - m);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == 0) s_red[warp] = s;
            __syncthreads();
            s = (s_red[0] + s_red[1]) + (s_red[2] + s_red[3]);
            for (int v = tid; v < b.Vp; v += 128) dst[v] = v < b.V ? expf(z[v] - m) / s : 0.f;
            __syncthreads();
        }
        // ---- alignment: expected source position in fp64, max == 1.0 test (:1443-1452)
        const float* al = b.a_out + (size_t)row * b.T;
        double cur = 0.0; float amax = 0.f;
        for (int j = tid; j < Tl; j += 128) { const float a = al[j]; cur += (double)a * (double)j; amax = fmaxf(amax, a); }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            cur += __shfl_xor_sync(0xffffffffu, cur, o);
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
        }
        if (lane == 0) { s_dred[warp] = cur; s_red[warp] = amax; }
        __syncthreads();
        cur = (s_dred[0] + s_dred[1]) + (s_dred[2] + s_dred[3]);
        amax = fmaxf(fmaxf(s_red[0], s_red[1]), fmaxf(s_red[2], s_red[3]));
        __syncthreads();
        // ---- the expanded node
        int len_node = 1; double pcum = 0.0, prev = 0.0; bool onehot = false;
        if (node != 0u) {
            const int pk = (int)((node - 1u) % KMAX);
            const size_t pg = (size_t)line * b.quota_x + (node - 1u) / KMAX;
            len_node = b.rec_len[pg] + 1;
            pcum = b.ch_cum[pg * KMAX + pk];
            const bool was_rej = b.ch_rej[pg * KMAX + pk] != 0;
            onehot = was_rej || b.rec_amax1[pg] != 0;
            prev = was_rej ? (double)b.rec_pos[pg] : b.rec_cur[pg];
        }
        int pos = 0; double mis = 0.0;
        if (len_node > 1) {
            mis = fabs(cur - prev - 1.0);
            pos = onehot ? (int)prev + 1 : (int)rint(cur);
        }
        // ---- rejection candidate (:1454-1466); pos == T raises IndexError in the reference: no candidate here
        int rej = -1;
        if (b.rej_thr > 0.0 && (mis < 0.1 || onehot) && pos >= 0 && pos < b.T) {
            const int r = b.rej_table[(size_t)line * b.T + pos];
            if (r >= 0) rej = r;
        }
        bool near_pos = len_node > 1 && ((!onehot && fabs(fabs(cur - floor(cur)) - 0.5) <= 1e-5) || fabs(mis - 0.1) <= 1e-6);
        if (rej >= 0 && fabs((double)sp[rej] - b.rej_thr) <= (double)kTieMargin * b.rej_thr + (double)kTieAbs) near_pos = true;
        __syncthreads();
        if (rej >= 0 && tid == 0 && (double)sp[rej] < b.rej_thr) sp[rej] = (float)b.rej_thr;
        __syncthreads();
        // ---- record: scores after the overwrite, states, alignment
        const int W4 = b.W >> 2;
        float4* rp = reinterpret_cast<float4*>(b.rec_p + g * b.Vp);
        for (int v = tid; v < (b.Vp >> 2); v += 128) rp[v] = reinterpret_cast<const float4*>(sp)[v];
        float* ra = b.rec_a + g * b.T;
        for (int j = tid; j < b.T; j += 128) ra[j] = al[j];
        float4* rq = reinterpret_cast<float4*>(b.rec_wq + g * b.W);
        const float4* wq = reinterpret_cast<const float4*>(b.wq + (size_t)row * b.W);
        for (int u = tid; u < W4; u += 128) rq[u] = wq[u];
        for (int n = 0; n < b.d; ++n) {
            float4* rh = reinterpret_cast<float4*>(b.rec_h + (g * b.d + n) * b.W);
            float4* rc = reinterpret_cast<float4*>(b.rec_c + (g * b.d + n) * b.W);
            const float4* co = reinterpret_cast<const float4*>(b.c_out[n] + (size_t)row * b.W);
            for (int u = tid; u < W4; u += 128) {
                rh[u] = load_split4(b.h_out[n], split ? b.h_out_lo[n] : nullptr, (size_t)row * b.W + 4 * u);
                rc[u] = co[u];
            }
        }
        // ---- beam width from the relative threshold (:1470-1478)
        float hi = 0.f;
        for (int v = tid; v < b.V; v += 128) hi = fmaxf(hi, sp[v]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        if (lane == 0) s_red[warp] = hi;
        __syncthreads();
        hi = fmaxf(fmaxf(s_red[0], s_red[1]), fmaxf(s_red[2], s_red[3]));
        const double bound = (double)hi * b.thr_in;
        int cnt = 0;
        // packed counters: [0,10) scores >= bound, [10,20) scores equal to the maximum, [20,30) scores within the tie margin of the bound
        for (int v = tid; v < b.V; v += 128) {
            const double x = (double)sp[v];
            cnt += (x >= bound) + ((sp[v] == hi) << 10) + ((fabs(x - bound) <= (double)kTieMargin * bound + (double)kTieAbs) << 20);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        if (lane == 0) s_cnt[warp] = cnt;
        __syncthreads();
        const int cnt_all = s_cnt[0] + s_cnt[1] + s_cnt[2] + s_cnt[3];
        const int beampos = min(cnt_all & 1023, b.width_in);
        const bool score_tie = ((cnt_all >> 10) & 1023) >= 2;
        bool near_tie = ((cnt_all >> 20) & 1023) > 0 || near_pos;
        float walked_prev = -1.f;
        // ---- walk the candidates best first (:1480-1529)
        int nacc = 0, pending = rej;
        const float rej_p = rej >= 0 ? sp[rej] : 0.f;
        for (int rank = 1; rank <= beampos + 1; ++rank) {
            int v; float pv; bool isrej;
            if (rank <= beampos) {
                ArgMax best{-1.f, -1};
                for (int q = tid; q < b.V; q += 128) best = better(best, ArgMax{sp[q], q});
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    ArgMax other{__shfl_xor_sync(0xffffffffu, best.v, o), __shfl_xor_sync(0xffffffffu, best.i, o)};
                    best = better(best, other);
                }
                if (lane == 0) s_arg[warp] = best;
                __syncthreads();
                best = better(better(s_arg[0], s_arg[1]), better(s_arg[2], s_arg[3]));
                __syncthreads();
                if (best.i < 0 || best.v < 0.f) break;
                v = best.i; pv = best.v;
                if (walked_prev >= 0.f && walked_prev - pv <= kTieMargin * walked_prev) near_tie = true;
                walked_prev = pv;
                if (tid == 0) sp[v] = -1.f;            // taken
                __syncthreads();
                isrej = (v == pending);
                if (isrej) pending = -1;
            } else {
                // past the beam width only a still pending rejection candidate enters; `if rej_idx:` (:1498)
                // treats index 0 as not pending
                if (pending <= 0) break;
                v = pending; pv = rej_p; isrej = true; pending = -1;
            }
            if (v == 0) continue;                       // '' (underspecification) is never a hypothesis (:1504-1507)
            const float pc = b.lm ? slm[v] : pv;        // LM rates the beam, the decoder enters it (:1486-1489)
            const float cost = (float)(-log((double)pc));
            if (cost != cost) continue;
            if (tid == 0) {
                b.ch_v[g * KMAX + nacc] = v;
                b.ch_p[g * KMAX + nacc] = pv;
                b.ch_cum[g * KMAX + nacc] = pcum + (double)cost;
                b.ch_rej[g * KMAX + nacc] = isrej ? 1 : 0;
            }
            ++nacc;
        }
        if (walked_prev >= 0.f) {       // the first candidate LEFT OUT against the last one taken (see the warp variant)
            float rest = -1.f;
            for (int q = tid; q < b.V; q += 128) rest = fmaxf(rest, sp[q]);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) rest = fmaxf(rest, __shfl_xor_sync(0xffffffffu, rest, o));
            __syncthreads();
            if (lane == 0) s_red[warp] = rest;
            __syncthreads();
            rest = fmaxf(fmaxf(s_red[0], s_red[1]), fmaxf(s_red[2], s_red[3]));
            if (rest >= 0.f && walked_prev - rest <= kTieMargin * walked_prev) near_tie = true;
        }
        if (tid == 0) {
            b.rec_cur[g] = cur; b.rec_amax1[g] = amax == 1.0f ? 1 : 0; b.rec_pos[g] = pos;
            b.rec_len[g] = len_node; b.rec_nacc[g] = nacc; b.rec_parent[g] = node;
            if (score_tie) atomicAdd((unsigned long long*)&b.score_ties[line], 1ull);
            if (near_tie) atomicAdd((unsigned long long*)&b.near_ties[line], 1ull);
        }
    }
}

// Warp-per-row variant for Vp <= 1024: the whole score row lives in registers (element lane + 32k in
// slot k, the assignment softmax_kernel uses, so both decoders see bit-identical probabilities), every
// reduction is a shuffle, and no block barrier is needed.  Same procedure as beam_expand_kernel.
template <int NVL>
__device__ __forceinline__ void expand_row(const BeamDev& b, int split, int row, int lane) {
    const unsigned FULL = 0xffffffffu;
    {
        const int line = b.row_line[row];
        const uint32_t node = b.row_node[row];
        const int Tl = b.row_T[line];
        const size_t g = (size_t)line * b.quota_x + b.row_xid[row];
        // ---- softmax of the logits row (and of the LM logits)
        float p[NVL], plm[NVL];
        auto softmax_row = [&](const float* z, float (&out)[NVL]) {
            float m = -INFINITY;
#pragma unroll
            for (int k = 0; k < NVL; ++k) {
                const int v = lane + 32 * k;
                out[k] = v < b.V ? z[v] : -INFINITY;
                m = fmaxf(m, out[k]);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(FULL, m, o));
            float s = 0.f;
#pragma unroll
            for (int k = 0; k < NVL; ++k)      // one expf per element: the value is reused for the quotient (same bits)
                if (lane + 32 * k < b.V) { out[k] = expf(out[k] - m); s += out[k]; } else out[k] = 0.f;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(FULL, s, o);
#pragma unroll
            for (int k = 0; k < NVL; ++k) out[k] = lane + 32 * k < b.V ? out[k] / s : 0.f;
        };
        softmax_row(b.logits + (size_t)row * b.Vp, p);
        if (b.lm) softmax_row(b.logits_lm + (size_t)row * b.Vp, plm);
        // value of element v of a register row, delivered to every lane
        auto fetch = [&](const float (&arr)[NVL], int v) {
            float mine = 0.f;
#pragma unroll
            for (int k = 0; k < NVL; ++k) if (k == (v >> 5)) mine = arr[k];
            return __shfl_sync(FULL, mine, v & 31);
        };
        // ---- alignment: expected source position in fp64, max == 1.0 test (:1443-1452)
        const float* al = b.a_out + (size_t)row * b.T;
        double cur = 0.0; float amax = 0.f;
        for (int j = lane; j < Tl; j += 32) { const float a = al[j]; cur += (double)a * (double)j; amax = fmaxf(amax, a); }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            cur += __shfl_xor_sync(FULL, cur, o);
            amax = fmaxf(amax, __shfl_xor_sync(FULL, amax, o));
        }
        // ---- the expanded node
        int len_node = 1; double pcum = 0.0, prev = 0.0; bool onehot = false;
        if (node != 0u) {
            const int pk = (int)((node - 1u) % KMAX);
            const size_t pg = (size_t)line * b.quota_x + (node - 1u) / KMAX;
            len_node = b.rec_len[pg] + 1;
            pcum = b.ch_cum[pg * KMAX + pk];
            const bool was_rej = b.ch_rej[pg * KMAX + pk] != 0;
            onehot = was_rej || b.rec_amax1[pg] != 0;
            prev = was_rej ? (double)b.rec_pos[pg] : b.rec_cur[pg];
        }
        int pos = 0; double mis = 0.0;
        if (len_node > 1) {
            mis = fabs(cur - prev - 1.0);
            pos = onehot ? (int)prev + 1 : (int)rint(cur);
        }
        // ---- rejection candidate (:1454-1466); pos == T raises IndexError in the reference: no candidate here
        int rej = -1;
        if (b.rej_thr > 0.0 && (mis < 0.1 || onehot) && pos >= 0 && pos < b.T) {
            const int r = b.rej_table[(size_t)line * b.T + pos];
            if (r >= 0) rej = r;
        }
        // discrete decisions that hang on the last bits of the alignment / the candidate's score are near-ties too
        bool near_pos = len_node > 1 && ((!onehot && fabs(fabs(cur - floor(cur)) - 0.5) <= 1e-5) || fabs(mis - 0.1) <= 1e-6);
        float rej_p = 0.f;
        if (rej >= 0) {
            rej_p = fetch(p, rej);
            if (fabs((double)rej_p - b.rej_thr) <= (double)kTieMargin * b.rej_thr + (double)kTieAbs) near_pos = true;
            if ((double)rej_p < b.rej_thr) {
                rej_p = (float)b.rej_thr;
#pragma unroll
                for (int k = 0; k < NVL; ++k) if (lane + 32 * k == rej) p[k] = rej_p;
            }
        }
        // ---- record: scores after the overwrite, states, alignment, query of the children
        float* rp = b.rec_p + g * b.Vp;
#pragma unroll
        for (int k = 0; k < NVL; ++k) rp[lane + 32 * k] = p[k];
        float* ra = b.rec_a + g * b.T;
        for (int j = lane; j < b.T; j += 32) ra[j] = al[j];
        // The copy is a chain of memory round trips if every array waits for the previous one's stores (the
        // long-scoreboard stalls of the ncu capture): all loads of a 128-float4 block -- query, h (hi, lo) and c
        // of the first two layers -- are issued together, then stored.
        const int W4 = b.W >> 2;
        float4* rq = reinterpret_cast<float4*>(b.rec_wq + g * b.W);
        const float4* wq = reinterpret_cast<const float4*>(b.wq + (size_t)row * b.W);
        const int nl = min(b.d, 2);
        for (int u0 = lane; u0 < W4; u0 += 128) {
            float4 tq[4], tc[2][4];
            uint2 th[2][4], tl[2][4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int u = u0 + 32 * i;
                if (u < W4) {
                    tq[i] = wq[u];
#pragma unroll
                    for (int n = 0; n < 2; ++n) {
                        if (n < nl) {
                            tc[n][i] = reinterpret_cast<const float4*>(b.c_out[n] + (size_t)row * b.W)[u];
                            if (split) {
                                th[n][i] = *reinterpret_cast<const uint2*>(reinterpret_cast<const __half*>(b.h_out[n]) + (size_t)row * b.W + 4 * u);
                                tl[n][i] = *reinterpret_cast<const uint2*>(reinterpret_cast<const __half*>(b.h_out_lo[n]) + (size_t)row * b.W + 4 * u);
                            } else {
                                const float4 hv = reinterpret_cast<const float4*>(b.h_out[n] + (size_t)row * b.W)[u];
                                th[n][i] = make_uint2(__float_as_uint(hv.x), __float_as_uint(hv.y));
                                tl[n][i] = make_uint2(__float_as_uint(hv.z), __float_as_uint(hv.w));
                            }
                        }
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int u = u0 + 32 * i;
                if (u < W4) {
                    rq[u] = tq[i];
#pragma unroll
                    for (int n = 0; n < 2; ++n) {
                        if (n < nl) {
                            float4 hv;
                            if (split) {
                                const float2 h0 = __half22float2(*reinterpret_cast<const __half2*>(&th[n][i].x)), h1 = __half22float2(*reinterpret_cast<const __half2*>(&th[n][i].y));
                                const float2 l0 = __half22float2(*reinterpret_cast<const __half2*>(&tl[n][i].x)), l1 = __half22float2(*reinterpret_cast<const __half2*>(&tl[n][i].y));
                                hv = make_float4(h0.x + l0.x, h0.y + l0.y, h1.x + l1.x, h1.y + l1.y);
                            } else {
                                hv = make_float4(__uint_as_float(th[n][i].x), __uint_as_float(th[n][i].y), __uint_as_float(tl[n][i].x), __uint_as_float(tl[n][i].y));
                            }
                            reinterpret_cast<float4*>(b.rec_h + (g * b.d + n) * b.W)[u] = hv;
                            reinterpret_cast<float4*>(b.rec_c + (g * b.d + n) * b.W)[u] = tc[n][i];
                        }
                    }
                }
            }
        }
        for (int n = 2; n < b.d; ++n) {
            float4* rh = reinterpret_cast<float4*>(b.rec_h + (g * b.d + n) * b.W);
            float4* rc = reinterpret_cast<float4*>(b.rec_c + (g * b.d + n) * b.W);
            const float4* co = reinterpret_cast<const float4*>(b.c_out[n] + (size_t)row * b.W);
            for (int u = lane; u < W4; u += 32) {
                rh[u] = load_split4(b.h_out[n], split ? b.h_out_lo[n] : nullptr, (size_t)row * b.W + 4 * u);
                rc[u] = co[u];
            }
        }
        // ---- beam width from the relative threshold (:1470-1478)
        float hi = 0.f;
#pragma unroll
        for (int k = 0; k < NVL; ++k) if (lane + 32 * k < b.V) hi = fmaxf(hi, p[k]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) hi = fmaxf(hi, __shfl_xor_sync(FULL, hi, o));
        const double bound = (double)hi * b.thr_in;
        int cnt = 0;
        // packed counters: [0,10) scores >= bound, [10,20) scores equal to the maximum, [20,30) scores within the tie margin of the bound
#pragma unroll
        for (int k = 0; k < NVL; ++k)
            if (lane + 32 * k < b.V) {
                const double x = (double)p[k];
                cnt += (x >= bound) + ((p[k] == hi) << 10) + ((fabs(x - bound) <= (double)kTieMargin * bound + (double)kTieAbs) << 20);
            }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(FULL, cnt, o);
        const int beampos = min(cnt & 1023, b.width_in);
        const bool score_tie = ((cnt >> 10) & 1023) >= 2;
        bool near_tie = ((cnt >> 20) & 1023) > 0 || near_pos;
        float walked_prev = -1.f;
        // ---- walk the candidates best first (:1480-1529)
        int nacc = 0, pending = rej;
        for (int rank = 1; rank <= beampos + 1; ++rank) {
            int v; float pv; bool isrej;
            if (rank <= beampos) {
                ArgMax best{-1.f, -1};
#pragma unroll
                for (int k = 0; k < NVL; ++k) if (lane + 32 * k < b.V) best = better(best, ArgMax{p[k], lane + 32 * k});
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    ArgMax other{__shfl_xor_sync(FULL, best.v, o), __shfl_xor_sync(FULL, best.i, o)};
                    best = better(best, other);
                }
                if (best.i < 0 || best.v < 0.f) break;
                v = best.i; pv = best.v;
                if (walked_prev >= 0.f && walked_prev - pv <= kTieMargin * walked_prev) near_tie = true;
                walked_prev = pv;
#pragma unroll
                for (int k = 0; k < NVL; ++k) if (lane + 32 * k == v) p[k] = -1.f;     // taken
                isrej = (v == pending);
                if (isrej) pending = -1;
            } else {
                // past the beam width only a still pending rejection candidate enters; `if rej_idx:` (:1498)
                // treats index 0 as not pending
                if (pending <= 0) break;
                v = pending; pv = rej_p; isrej = true; pending = -1;
            }
            if (v == 0) continue;                       // '' (underspecification) is never a hypothesis (:1504-1507)
            const float pc = b.lm ? fetch(plm, v) : pv; // LM rates the beam, the decoder enters it (:1486-1489)
            const float cost = (float)(-log((double)pc));
            if (cost != cost) continue;
            if (lane == 0) {
                b.ch_v[g * KMAX + nacc] = v;
                b.ch_p[g * KMAX + nacc] = pv;
                b.ch_cum[g * KMAX + nacc] = pcum + (double)cost;
                b.ch_rej[g * KMAX + nacc] = isrej ? 1 : 0;
            }
            ++nacc;
        }
        if (walked_prev >= 0.f) {       // the first candidate LEFT OUT against the last one taken: a near-tie there changes the children
            float rest = -1.f;
#pragma unroll
            for (int k = 0; k < NVL; ++k) if (lane + 32 * k < b.V) rest = fmaxf(rest, p[k]);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) rest = fmaxf(rest, __shfl_xor_sync(FULL, rest, o));
            if (rest >= 0.f && walked_prev - rest <= kTieMargin * walked_prev) near_tie = true;
        }
        if (lane == 0) {
            b.rec_cur[g] = cur; b.rec_amax1[g] = amax == 1.0f ? 1 : 0; b.rec_pos[g] = pos;
            b.rec_len[g] = len_node; b.rec_nacc[g] = nacc; b.rec_parent[g] = node;
            if (score_tie) atomicAdd((unsigned long long*)&b.score_ties[line], 1ull);
            if (near_tie) atomicAdd((unsigned long long*)&b.near_ties[line], 1ull);
        }
    }
}
template <int NVL>
__global__ void __launch_bounds__(128) beam_expand_warp_kernel(BeamDev b, int split) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int R = b.counters[0];
    for (int row = blockIdx.x * 4 + wid; row < R; row += gridDim.x * 4) expand_row<NVL>(b, split, row, lane);
}

// ------------------------------------------------------------------ push: insort of the new children (:1515-1532)
// Reach truncation.  The reference keeps the best max_batches * batch_size entries (:1531) -- "more than can ever be
// processed within limits" at the START of a line.  Later the same argument cuts deeper: with R pop rounds still to come
// (R = 2T - iterations done) a round consumes a prefix of the list holding at most batch_size open nodes (:1400-1415), and
// insertions only push an entry further back, so an entry with at least R * batch_size open nodes ahead of it is never
// popped.  Such entries need not be stored: the fringe in memory is a prefix of the reference's list, `fr_gone` counts the
// rest (list length, `open nodes left` and the peak statistic stay the reference's).  How long the merged prefix may be:
//   * `keep`, the reference's own truncation;
//   * if entries have been dropped before: up to and including the last OLD entry -- new children behind it may belong
//     behind dropped entries; they are unreachable too (every dropped entry had > R * batch_size open nodes ahead after
//     this round's pop, all of them in the stored prefix, hence ahead of those children);
//   * R * batch_size + (finals in the prefix, upper bound) + 1 entries hold R * batch_size + 1 open nodes: the +1 keeps
//     the entry behind the last pop in memory (the near-tie statistic of pop_line looks at it).
// Results, statistics and checksums are unchanged (tests/test_gpu_beam.py against the oracle's); what changes is the
// length of the merge copy, which dominated the competitive / stress brackets (fringe at its cap of 2T * 256 entries).
__device__ __forceinline__ int reach_keep(const BeamDev& b, int line, int Tl, int keep, int n_old, int n_new,
                                          int new_before_last_old, int new_finals) {
    int q = min(keep, n_old + n_new);
    if (!b.reach_trunc) return q;
    if (b.fr_gone[line] > 0) q = min(q, n_old > 0 ? n_old + new_before_last_old : 0);
    const long long rounds = max(0, 2 * Tl - b.iter[line]);
    const long long reach = rounds * b.batch + (long long)b.fin_in[line] + new_finals + 1;
    return (int)min((long long)q, reach);
}
// Three steps, so that several blocks can share one line's merge (beam_push_kernel):
//   push_sort    the line's new children, sorted by (key, id), in shared memory + what the merge has to know (PushDesc);
//                reads the line's state, changes nothing
//   push_merge   rank-merge of slice `part` of `parts` of the old fringe and of the new children into the other buffer
//   push_commit  the line's new state (one thread, after every part has finished)
struct PushDesc { int n_new, n_old, cur, o_off, keep, keep_ref, new_finals; };

__device__ bool push_sort(const BeamDev& b, int line, double* skey, uint32_t* sid, int* s_wsum, PushDesc& d) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, NT = blockDim.x, NW = NT >> 5;
    __shared__ int s_off[257];                   // first child of row j in the unsorted list (batch_size <= 256 rows per line)
    __shared__ int s_new_finals;
    const int c = b.beam_count[line];
    if (c <= 0) return false;
    const int Tl = b.row_T[line];
    const int xb = b.x_base[line];
    // ---- collect: row j of this line holds record xb+j with rec_nacc children
    const int nj = tid < c ? b.rec_nacc[(size_t)line * b.quota_x + xb + tid] : 0;
    int incl = nj;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
    if (lane == 31) s_wsum[warp] = incl;
    if (tid == 0) s_new_finals = 0;
    __syncthreads();
    int off = 0, n_new = 0;
    for (int w = 0; w < NW; ++w) { if (w < warp) off += s_wsum[w]; n_new += s_wsum[w]; }
    off += incl - nj;
    if (tid < c) s_off[tid] = off;
    __syncthreads();
    // one child per thread and round (a thread per ROW walked its up to 16 children one memory round trip after the other)
    for (int e = tid; e < n_new; e += NT) {
        int lo = 0, hi = c;                       // the last row whose first child is <= e (rows without children share an offset)
        while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (s_off[mid] <= e) lo = mid; else hi = mid; }
        const int q = e - s_off[lo];
        const size_t g = (size_t)line * b.quota_x + xb + lo;
        const int len_child = b.rec_len[g] + 1;
        const double pen = 3.0 * fabs((double)(len_child - Tl));      // cost0 * |length - length0| (:1588-1595)
        skey[e] = b.ch_cum[g * KMAX + q] + pen;
        sid[e] = (1u + (uint32_t)(xb + lo) * KMAX + q) | (b.ch_v[g * KMAX + q] == b.eos ? FINAL_BIT : 0u);
    }
    int npad = 1;
    while (npad < n_new) npad <<= 1;
    for (int i = n_new + tid; i < npad; i += NT) { skey[i] = INFINITY; sid[i] = 0x7fffffffu; }
    __syncthreads();
    // ---- bitonic sort ascending by (key, id).  Thread t owns elements t + m * NT; partners at distance j < 32 sit in the
    // same warp, so those steps run on registers with shuffles (no shared-memory traffic, no block barrier): 21 barriers
    // instead of 55 for 1 024 children.  More than 4 elements per thread (256-thread callers): all steps through shared memory.
    const int M = (npad + NT - 1) / NT;
    if (M <= 4) {
        double kr[4]; uint32_t ir[4];
        auto load = [&]() {
#pragma unroll
            for (int m = 0; m < 4; ++m) { const int i = tid + m * NT; const bool in = m < M && i < npad; kr[m] = in ? skey[i] : INFINITY; ir[m] = in ? sid[i] : 0x7fffffffu; }
        };
        auto store = [&]() {
#pragma unroll
            for (int m = 0; m < 4; ++m) { const int i = tid + m * NT; if (m < M && i < npad) { skey[i] = kr[m]; sid[i] = ir[m]; } }
        };
        auto in_warp = [&](int j, int k2) {
#pragma unroll
            for (int m = 0; m < 4; ++m) {
                if (m < M) {                            // (the same for every thread)
                    const int i = tid + m * NT;
                    const double ko = __shfl_xor_sync(0xffffffffu, kr[m], j);
                    const uint32_t io = __shfl_xor_sync(0xffffffffu, ir[m], j);
                    const bool want_min = ((i & j) == 0) == ((i & k2) == 0);     // lower index of an ascending pair, or upper of a descending one
                    const bool take = want_min ? key_less(ko, io, kr[m], ir[m]) : key_less(kr[m], ir[m], ko, io);
                    if (take) { kr[m] = ko; ir[m] = io; }
                }
            }
        };
        load();
        for (int k2 = 2; k2 <= min(32, npad); k2 <<= 1)
            for (int j = k2 >> 1; j > 0; j >>= 1) in_warp(j, k2);
        store();
        __syncthreads();
        for (int k2 = 64; k2 <= npad; k2 <<= 1) {
            for (int j = k2 >> 1; j >= 32; j >>= 1) {
                for (int i = tid; i < npad; i += NT) {
                    const int ixj = i ^ j;
                    if (ixj > i) {
                        const bool up = (i & k2) == 0;
                        const double ka = skey[i], kb = skey[ixj];
                        const uint32_t ia = sid[i], ib = sid[ixj];
                        const bool a_gt_b = key_less(kb, ib, ka, ia);
                        if (a_gt_b == up) { skey[i] = kb; skey[ixj] = ka; sid[i] = ib; sid[ixj] = ia; }
                    }
                }
                __syncthreads();
            }
            load();
            for (int j = 16; j > 0; j >>= 1) in_warp(j, k2);
            store();
            __syncthreads();
        }
    } else {
        for (int k2 = 2; k2 <= npad; k2 <<= 1) {
            for (int j = k2 >> 1; j > 0; j >>= 1) {
                for (int i = tid; i < npad; i += NT) {
                    const int ixj = i ^ j;
                    if (ixj > i) {
                        const bool up = (i & k2) == 0;
                        const double ka = skey[i], kb = skey[ixj];
                        const uint32_t ia = sid[i], ib = sid[ixj];
                        const bool a_gt_b = key_less(kb, ib, ka, ia);
                        if (a_gt_b == up) { skey[i] = kb; skey[ixj] = ka; sid[i] = ib; sid[ixj] = ia; }
                    }
                }
                __syncthreads();
            }
        }
    }
    {
        int nf = 0;
        for (int j = tid; j < n_new; j += NT) nf += (sid[j] & FINAL_BIT) ? 1 : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) nf += __shfl_xor_sync(0xffffffffu, nf, o);
        if (lane == 0 && nf) atomicAdd(&s_new_finals, nf);
    }
    __syncthreads();
    d.n_new = n_new;
    d.new_finals = s_new_finals;
    d.cur = b.fr_buf[line];
    d.o_off = b.fr_off[line];
    d.n_old = b.fr_len[line] - d.o_off;
    const double* okey = b.fr_key + ((size_t)d.cur * b.L + line) * b.cap + d.o_off;
    const uint32_t* oid = b.fr_id + ((size_t)d.cur * b.L + line) * b.cap + d.o_off;
    int lo_last = 0;                              // new entries that precede the LAST old entry (same search in every thread)
    if (d.n_old > 0) {
        const double k = okey[d.n_old - 1]; const uint32_t id = oid[d.n_old - 1];
        int hi = n_new;
        while (lo_last < hi) { const int mid = (lo_last + hi) >> 1; if (key_less(skey[mid], sid[mid], k, id)) lo_last = mid + 1; else hi = mid; }
    }
    // truncation to the best max_batches * batch_size = 2T * batch_size of THIS line's reference batch (seq2seq.py:1397, 1531) ...
    d.keep_ref = (int)min((long long)b.cap, (long long)2 * Tl * b.batch);
    d.keep = reach_keep(b, line, Tl, d.keep_ref, d.n_old, n_new, lo_last, d.new_finals);      // ... and to what can still be popped
    return true;
}

__device__ void push_merge(const BeamDev& b, int line, const PushDesc& d, const double* skey, const uint32_t* sid, int part, int parts) {
    const int tid = threadIdx.x, NT = blockDim.x;
    const double* okey = b.fr_key + ((size_t)d.cur * b.L + line) * b.cap + d.o_off;
    const uint32_t* oid = b.fr_id + ((size_t)d.cur * b.L + line) * b.cap + d.o_off;
    double* nkey = b.fr_key + ((size_t)(d.cur ^ 1) * b.L + line) * b.cap;
    uint32_t* nid = b.fr_id + ((size_t)(d.cur ^ 1) * b.L + line) * b.cap;
    const int keep = d.keep, n_new = d.n_new;
    const int m_old = min(d.n_old, keep);        // an old entry never moves forward: pos >= i
    const int per_o = (m_old + parts - 1) / parts, o1 = min(m_old, (part + 1) * per_o);
    for (int i = part * per_o + tid; i < o1; i += NT) {
        const double k = okey[i]; const uint32_t id = oid[i];
        int lo = 0, hi = n_new;                   // new entries that precede this old one
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (key_less(skey[mid], sid[mid], k, id)) lo = mid + 1; else hi = mid; }
        const int pos = i + lo;
        if (pos < keep) { nkey[pos] = k; nid[pos] = id; }
    }
    const int per_n = (n_new + parts - 1) / parts, n1 = min(n_new, (part + 1) * per_n);
    for (int j = part * per_n + tid; j < n1; j += NT) {
        const double k = skey[j]; const uint32_t id = sid[j];
        int lo = 0, hi = m_old;                   // old entries that precede this new one (past m_old: pos >= keep anyway)
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (key_less(okey[mid], oid[mid], k, id)) lo = mid + 1; else hi = mid; }
        const int pos = j + lo;
        if (pos < keep) { nkey[pos] = k; nid[pos] = id; }
    }
}

__device__ void push_commit(const BeamDev& b, int line, const PushDesc& d) {
    const int listed = (int)min((long long)d.n_old + b.fr_gone[line] + d.n_new, (long long)d.keep_ref);    // the reference's list
    b.fr_len[line] = d.keep;
    b.fr_gone[line] = listed - d.keep;
    b.fin_in[line] += d.new_finals;
    b.fr_off[line] = 0;
    b.fr_buf[line] = d.cur ^ 1;
    if (listed > b.fr_peak[line]) b.fr_peak[line] = listed;
}

__device__ void push_line(const BeamDev& b, int line, double* skey, uint32_t* sid, int* s_wsum) {
    PushDesc d;
    if (!push_sort(b, line, skey, sid, s_wsum, d)) return;
    push_merge(b, line, d, skey, sid, 0, 1);
    __syncthreads();
    if (threadIdx.x == 0) push_commit(b, line, d);
}

// Small case, one warp per line: at most 32 new children (the common case once the rejection candidate
// or a narrow beam prunes the expansion).  Lane t holds child t; its slot in the merged fringe is its rank
// among the new children plus the number of old entries that precede it.  Lines with more children go on
// the heavy list for beam_push_kernel.  Same total order as there: (key, id), first inserted first.
// returns true when the line has more than 32 new children (left to push_line)
__device__ bool push_small_line(const BeamDev& b, int line, int lane, double* s_key_w, uint32_t* s_id_w) {
    const int c = b.beam_count[line];
    if (c <= 0) return false;
    const unsigned FULL = 0xffffffffu;
    const int xb = b.x_base[line];
    int n_new = 33;
    int nj = 0, off = 0;
    if (c <= 32) {
        nj = lane < c ? b.rec_nacc[(size_t)line * b.quota_x + xb + lane] : 0;
        int incl = nj;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += y; }
        n_new = __shfl_sync(FULL, incl, 31);
        off = incl - nj;
    }
    if (n_new > 32) return true;
    const int Tl = b.row_T[line];
    if (lane < c) {
        const size_t g = (size_t)line * b.quota_x + xb + lane;
        const int len_child = b.rec_len[g] + 1;
        const double pen = 3.0 * fabs((double)(len_child - Tl));      // cost0 * |length - length0| (:1588-1595)
        for (int q = 0; q < nj; ++q) {
            s_key_w[off + q] = b.ch_cum[g * KMAX + q] + pen;
            s_id_w[off + q] = (1u + (uint32_t)(xb + lane) * KMAX + q) | (b.ch_v[g * KMAX + q] == b.eos ? FINAL_BIT : 0u);
        }
    }
    __syncwarp();
    const int cur = b.fr_buf[line];
    const int o_off = b.fr_off[line];
    const int n_old = b.fr_len[line] - o_off;
    const double* okey = b.fr_key + ((size_t)cur * b.L + line) * b.cap + o_off;
    const uint32_t* oid = b.fr_id + ((size_t)cur * b.L + line) * b.cap + o_off;
    double* nkey = b.fr_key + ((size_t)(cur ^ 1) * b.L + line) * b.cap;
    uint32_t* nid = b.fr_id + ((size_t)(cur ^ 1) * b.L + line) * b.cap;
    const int keep_ref = (int)min((long long)b.cap, (long long)2 * Tl * b.batch);     // see push_line
    const int new_finals = __popc(__ballot_sync(FULL, lane < n_new && (s_id_w[lane < n_new ? lane : 0] & FINAL_BIT)));
    int lo_last = 0;
    if (n_old > 0) {
        const double kz = okey[n_old - 1]; const uint32_t iz = oid[n_old - 1];
        lo_last = __popc(__ballot_sync(FULL, lane < n_new && key_less(s_key_w[lane < n_new ? lane : 0], s_id_w[lane < n_new ? lane : 0], kz, iz)));
    }
    const int keep = reach_keep(b, line, Tl, keep_ref, n_old, n_new, lo_last, new_finals);
    if (lane < n_new) {
        const double k = s_key_w[lane]; const uint32_t id = s_id_w[lane];
        int rank = 0;
        for (int t = 0; t < n_new; ++t) rank += key_less(s_key_w[t], s_id_w[t], k, id);
        int lo = 0, hi = n_old;                   // old entries that precede this new one
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (key_less(okey[mid], oid[mid], k, id)) lo = mid + 1; else hi = mid; }
        const int pos = rank + lo;
        if (pos < keep) { nkey[pos] = k; nid[pos] = id; }
    }
    for (int i = lane; i < min(n_old, keep); i += 32) {
        const double k = okey[i]; const uint32_t id = oid[i];
        int before = 0;                           // new entries that precede this old one
        for (int t = 0; t < n_new; ++t) before += key_less(s_key_w[t], s_id_w[t], k, id);
        const int pos = i + before;
        if (pos < keep) { nkey[pos] = k; nid[pos] = id; }
    }
    __syncwarp();
    if (lane == 0) {
        const int listed = (int)min((long long)n_old + b.fr_gone[line] + n_new, (long long)keep_ref);
        b.fr_len[line] = keep;
        b.fr_gone[line] = listed - keep;
        b.fin_in[line] += new_finals;
        b.fr_off[line] = 0;
        b.fr_buf[line] = cur ^ 1;
        if (listed > b.fr_peak[line]) b.fr_peak[line] = listed;
    }
    return false;
}
__global__ void __launch_bounds__(256) beam_push_small_kernel(BeamDev b) {
    __shared__ double s_key[8][32];
    __shared__ uint32_t s_id[8][32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int line = blockIdx.x * 8 + wid;
    if (line >= b.L) return;
    if (push_small_line(b, line, lane, s_key[wid], s_id[wid]) && lane == 0) b.heavy[atomicAdd(&b.counters[2], 1)] = line;
}

// Lines on the heavy list (more than 32 new children), one block per line at a time.
// 1024 threads per line: the bitonic sort of up to 4096 children and the rank-merge with a fringe of up to 2T*256 entries are
// per-line work that a wave of a few dozen lines cannot spread over more SMs, only over more threads of one.
constexpr int PUSH_THREADS = 1024;
// gridDim.x blocks share a line (a wave of a page's worth of lines leaves most SMs without one): each sorts the children
// for itself -- cheap beside the merge -- and merges its slice; the block that finishes last writes the line's new state,
// which no other block reads any more by then (push_done counts the finished parts; nobody waits for anybody).
__global__ void __launch_bounds__(PUSH_THREADS) beam_push_kernel(BeamDev b) {
    extern __shared__ unsigned char smraw[];
    double* skey = reinterpret_cast<double*>(smraw);                 // [npad]
    uint32_t* sid = reinterpret_cast<uint32_t*>(skey + MAX_NEW);     // [npad]
    __shared__ int s_wsum[PUSH_THREADS / 32];
    const int n = b.counters[2];
    const int parts = gridDim.x;
    for (int i = blockIdx.y; i < n; i += gridDim.y) {
        const int line = b.heavy[i];
        PushDesc d;
        if (push_sort(b, line, skey, sid, s_wsum, d)) {
            push_merge(b, line, d, skey, sid, blockIdx.x, parts);
            __syncthreads();
            if (threadIdx.x == 0) {
                bool last = true;
                if (parts > 1) { __threadfence(); last = atomicAdd(&b.push_done[line], 1) == parts - 1; }
                if (last) { if (parts > 1) b.push_done[line] = 0; push_commit(b, line, d); }
            }
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------ page of lines: one block per line does the rest
// expand the line's rows (a warp per row), insert their children into the line's fringe (warp-level insert, or the
// block-wide sort + merge when there are more than 32), pop the next batch -- everything of an iteration that touches
// one line only, without a launch boundary in between.  What crosses lines (row offsets) is the gather's scan.
template <int NVL>
__global__ void __launch_bounds__(256) beam_line_kernel(BeamDev b, int split, int do_expand) {
    extern __shared__ unsigned char smraw[];
    double* skey = reinterpret_cast<double*>(smraw);
    uint32_t* sid = reinterpret_cast<uint32_t*>(skey + MAX_NEW);
    __shared__ double s_key[32];
    __shared__ uint32_t s_id[32];
    __shared__ int s_wsum[32];
    __shared__ int s_heavy;
    const int line = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c = (do_expand && b.status[line] == ST_ACTIVE) ? b.beam_count[line] : 0;     // rows of this line in the step that just ran
    if (c > 0) {
        const int rb = b.row_base[line];
        for (int j = warp; j < c; j += 8) expand_row<NVL>(b, split, rb + j, lane);
        __syncthreads();            // the records of all rows of the line are complete
        if (warp == 0) {
            const bool heavy = push_small_line(b, line, lane, s_key, s_id);
            if (lane == 0) s_heavy = heavy ? 1 : 0;
        }
        __syncthreads();
        if (s_heavy) push_line(b, line, skey, sid, s_wsum);
        __syncthreads();
    }
    if (warp == 0) pop_line(b, line, lane, true);
}

// ------------------------------------------------------------------ finish: best final hypothesis per line (:1538-1544)
__global__ void __launch_bounds__(128) beam_finish_kernel(BeamDev b, int32_t* o_chars, float* o_probs, double* o_score,
                                                          int32_t* o_len, float* o_align, long long* o_stats) {
    // Slot (line, r) of the outputs holds the r-th best final hypothesis, r < n_best; unused slots have length 0.
    const int line = blockIdx.x, tid = threadIdx.x;
    const int S = 2 * b.T, N = b.n_best;
    __shared__ int s_n[FMAX];
    __shared__ uint32_t s_fin[FMAX];
    const int st = b.status[line];
    if (tid == 0) {
        int nf = 0;
        if (st != ST_OVERFLOW && st != ST_EMPTY && b.best_final[line] > 0) {
            if (N == 1) {
                s_fin[0] = (uint32_t)b.best_final[line]; nf = 1;
            } else {
                // the recorded finals, best first: ascending (key, id) like the reference's final_beam read from its best end
                const int have = min(b.finals_count[line], FMAX);
                uint32_t* fid = b.fin_id + (size_t)line * FMAX; double* fk = b.fin_key + (size_t)line * FMAX;
                for (int i = 1; i < have; ++i) {
                    const double k = fk[i]; const uint32_t id = fid[i];
                    int j = i - 1;
                    while (j >= 0 && key_less(k, id, fk[j], fid[j])) { fk[j + 1] = fk[j]; fid[j + 1] = fid[j]; --j; }
                    fk[j + 1] = k; fid[j + 1] = id;
                }
                nf = min(have, N);
                for (int r = 0; r < nf; ++r) s_fin[r] = fid[r];
                // more than FMAX finals: the best one is tracked on its own and always comes first
                if (nf > 0 && s_fin[0] != (uint32_t)b.best_final[line] && b.finals_count[line] > FMAX) {
                    for (int r = nf - 1; r > 0; --r) s_fin[r] = s_fin[r - 1];
                    s_fin[0] = (uint32_t)b.best_final[line];
                }
            }
        }
        for (int r = 0; r < N; ++r) {
            const size_t slot = (size_t)line * N + r;
            int n = 0;
            o_score[slot] = 0.0;
            if (r < nf) {
                // walk to the root, filling the slot from its end
                uint32_t node = s_fin[r];
                const size_t g0 = (size_t)line * b.quota_x + (node - 1u) / KMAX;
                const int length = b.rec_len[g0] + 1;
                n = length - 1;
                o_score[slot] = b.ch_cum[g0 * KMAX + (node - 1u) % KMAX] / (double)(length - 1);
                int i = n - 1;
                while (node != 0u && i >= 0) {
                    const size_t g = (size_t)line * b.quota_x + (node - 1u) / KMAX;
                    const int k = (int)((node - 1u) % KMAX);
                    o_chars[slot * S + i] = b.ch_v[g * KMAX + k];
                    o_probs[slot * S + i] = b.ch_p[g * KMAX + k];
                    if (o_align)        // temporarily keep the node id in the first cell of its alignment row
                        reinterpret_cast<uint32_t*>(o_align)[(slot * S + i) * b.T] = node;
                    node = b.rec_parent[g];
                    --i;
                }
            }
            o_len[slot] = n;
            s_n[r] = n;
        }
        long long* os = o_stats + (size_t)line * kStats;
        os[0] = st == ST_OVERFLOW ? -1 : (st == ST_EMPTY ? -2 : b.iter[line]);
        os[1] = b.hyp_steps[line];
        os[2] = b.finals_count[line];
        // (a line that ran out of records is retried: its row tells the host how far it came, beam_correct sizes the retry from it)
        os[3] = st == ST_ACTIVE ? b.fr_len[line] + b.fr_gone[line] : (st == ST_OVERFLOW ? b.iter[line] : b.open_left[line]);
        os[4] = b.score_ties[line];
        os[5] = b.near_ties[line];
        os[6] = (long long)b.pop_sum[line];
        os[7] = b.fr_peak[line];
    }
    __syncthreads();
    if (!o_align) return;
    for (int r = 0; r < N; ++r)
        for (int i = 0; i < s_n[r]; ++i) {
            float* row = o_align + (((size_t)line * N + r) * S + i) * b.T;
            const uint32_t node = reinterpret_cast<uint32_t*>(row)[0];
            __syncthreads();
            const size_t g = (size_t)line * b.quota_x + (node - 1u) / KMAX;
            const bool rej = b.ch_rej[g * KMAX + (node - 1u) % KMAX] != 0;
            const int pos = b.rec_pos[g];
            const float* ra = b.rec_a + g * b.T;
            for (int j = tid; j < b.T; j += 128) row[j] = rej ? (j == pos ? 1.0f : 0.0f) : ra[j];   // np.eye(T)[pos] (:1494-1495)
            __syncthreads();
        }
}

// ------------------------------------------------------------------ workspace layout of one wave
struct Carver {
    char* base; size_t off;
    template <typename T> T* take(size_t n) {
        off = (off + 255) & ~size_t(255);
        T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
        off += n * sizeof(T);
        return p;
    }
};

struct Plan { int L, quota_x, cap, Rcap, want_align; size_t bytes; };

size_t row_bytes(const cor_model* m, int T) {
    const size_t W = m->W, Vp = m->Vp, D = m->D, d = m->d;
    const size_t sp = m->split ? 2 : 1;
    size_t b = sp * Vp * 4 /*p*/ + sp * W * 4 /*x*/ + W * 4 /*wq*/ + sp * D * 4 /*ctx*/ + Vp * 4 /*logits*/ + 2 * (size_t)T * 4 /*a*/ +
               d * (2 * sp * W * 4 /*h in,out*/ + 2 * W * 4 /*c in,out*/) + 12;
    if (m->cfg.lm_predict) b += sp * W * 4 + W * 4 + Vp * 4;
    return b + 64;
}
size_t rec_bytes(const cor_model* m, int T) {
    return (size_t)m->Vp * 4 + (2 * (size_t)m->d + 1) * m->W * 4 + (size_t)T * 4 + 8 + 1 + 4 * 4 + KMAX * (4 + 4 + 8 + 1) + 16;
}

int layout(const cor_model* m, int L, int T, int k, const cor_beam_params* bp, int quota_x, int Rcap, int want_align,
           void* base, Ws* ws, BeamDev* b, StepIO* io, size_t* bytes) {
    if (carve(m, L, T, k, 0, base, ws)) return 1;
    Carver c{(char*)base, ws->bytes};
    const int W = m->W, d = m->d, D = m->D, Vp = m->Vp;
    const bool sp = m->split;
    const int cap = (int)std::min<long long>((long long)2 * T * bp->batch_size, (long long)quota_x * KMAX + 1);
    *b = BeamDev{};
    b->L = L; b->T = T; b->V = m->V; b->Vp = Vp; b->W = W; b->d = d; b->D = D; b->k_alt = k;
    b->batch = bp->batch_size; b->width_in = bp->beam_width_in; b->width_out = bp->beam_width_out;
    b->eos = m->cfg.eos_index; b->lm = m->cfg.lm_predict ? 1 : 0;
    b->thr_in = bp->beam_threshold_in; b->rej_thr = bp->rejection_threshold;
    b->quota_x = quota_x; b->cap = cap; b->Rcap = Rcap;
    b->row_T = ws->row_T; b->live = ws->live; b->idx = ws->idx; b->val = ws->val;
    for (int n = 0; n < d; ++n) { b->st_h[n] = ws->st_h[n]; b->st_c[n] = ws->st_c[n]; }
    b->status = c.take<int32_t>(L); b->iter = c.take<int32_t>(L); b->x_count = c.take<int32_t>(L); b->x_base = c.take<int32_t>(L);
    b->beam_count = c.take<int32_t>(L); b->row_base = c.take<int32_t>(L); b->fr_len = c.take<int32_t>(L);
    b->fr_off = c.take<int32_t>(L); b->fr_buf = c.take<int32_t>(L); b->finals_count = c.take<int32_t>(L);
    b->best_final = c.take<int32_t>(L); b->open_left = c.take<int32_t>(L);
    b->best_key = c.take<double>(L); b->hyp_steps = c.take<long long>(L);
    b->score_ties = c.take<long long>(L); b->near_ties = c.take<long long>(L); b->pop_sum = c.take<unsigned long long>(L);
    b->fr_peak = c.take<int32_t>(L);
    b->fr_gone = c.take<int32_t>(L); b->fin_in = c.take<int32_t>(L); b->push_done = c.take<int32_t>(L);
    b->reach_trunc = getenv("COR_NO_REACH_TRUNC") ? 0 : 1;
    b->fr_key = c.take<double>((size_t)2 * L * cap); b->fr_id = c.take<uint32_t>((size_t)2 * L * cap);
    b->beam_nodes = c.take<uint32_t>((size_t)L * bp->batch_size);
    b->rej_table = c.take<int32_t>((size_t)L * T);
    b->counters = c.take<int32_t>(4 + 16);      // [4..]: LoopCtl of the iteration graph (run_wave)
    b->heavy = c.take<int32_t>(L);
    b->o_stats = c.take<long long>((size_t)L * kStats);
    const int NB = std::max(1, std::min(bp->n_best, FMAX));
    b->n_best = NB;
    b->fin_id = c.take<uint32_t>((size_t)L * FMAX); b->fin_key = c.take<double>((size_t)L * FMAX);
    b->r_chars = c.take<int32_t>((size_t)L * NB * 2 * T); b->r_probs = c.take<float>((size_t)L * NB * 2 * T);
    b->r_score = c.take<double>((size_t)L * NB); b->r_len = c.take<int32_t>((size_t)L * NB);
    b->o_align = want_align ? c.take<float>((size_t)L * NB * 2 * T * T) : nullptr;
    const size_t G = (size_t)L * quota_x;
    b->rec_p = c.take<float>(G * Vp); b->rec_h = c.take<float>(G * d * W); b->rec_c = c.take<float>(G * d * W);
    b->rec_a = c.take<float>(G * T);
    b->rec_wq = c.take<float>(G * W);
    b->rec_cur = c.take<double>(G); b->rec_amax1 = c.take<uint8_t>(G); b->rec_pos = c.take<int32_t>(G);
    b->rec_len = c.take<int32_t>(G); b->rec_nacc = c.take<int32_t>(G); b->rec_parent = c.take<uint32_t>(G);
    b->ch_v = c.take<int32_t>(G * KMAX); b->ch_p = c.take<float>(G * KMAX); b->ch_cum = c.take<double>(G * KMAX);
    b->ch_rej = c.take<uint8_t>(G * KMAX);
    const size_t R = Rcap;
    b->row_line = c.take<int32_t>(R); b->row_xid = c.take<int32_t>(R); b->row_node = c.take<uint32_t>(R);
    b->p = c.take<float>(R * Vp); b->p_lo = sp ? c.take<float>(R * Vp) : nullptr;
    b->logits = c.take<float>(R * Vp);
    b->a_in = c.take<float>(R * T); b->a_out = c.take<float>(R * T);
    *io = StepIO{};
    io->R = Rcap; io->r_dev = b->counters; io->row_line = b->row_line; io->first = false;
    io->p = b->p; io->p_lo = b->p_lo;
    io->x = c.take<float>(R * W); io->x_lo = sp ? c.take<float>(R * W) : nullptr;
    io->wq = c.take<float>(R * W);
    b->wq = io->wq; io->wq_next = io->wq;
    io->ctx = c.take<float>(R * D); io->ctx_lo = sp ? c.take<float>(R * D) : nullptr;
    io->logits = b->logits;
    io->a_in = b->a_in; io->a_out = b->a_out;
    for (int n = 0; n < d; ++n) {
        b->h_in[n] = c.take<float>(R * W); b->h_in_lo[n] = sp ? c.take<float>(R * W) : nullptr;
        b->h_out[n] = c.take<float>(R * W); b->h_out_lo[n] = sp ? c.take<float>(R * W) : nullptr;
        b->c_in[n] = c.take<float>(R * W); b->c_out[n] = c.take<float>(R * W);
        io->h_in[n] = b->h_in[n]; io->h_in_lo[n] = b->h_in_lo[n]; io->h_out[n] = b->h_out[n]; io->h_out_lo[n] = b->h_out_lo[n];
        io->c_in[n] = b->c_in[n]; io->c_out[n] = b->c_out[n];
    }
    if (m->cfg.lm_predict) {
        io->h_lm = c.take<float>(R * W); io->h_lm_lo = sp ? c.take<float>(R * W) : nullptr;
        io->c_lm = c.take<float>(R * W);
        io->logits_lm = c.take<float>(R * Vp);
        b->logits_lm = io->logits_lm;
    }
    *bytes = (c.off + 255) & ~size_t(255);
    return 0;
}

// Largest wave (<= n_rows lines) whose layout fits `budget`: rows for at least one hypothesis per line,
// a record arena of at least 2T+2 per line (one hypothesis per iteration), the rest split between them.
int make_plan(const cor_model* m, int n_rows, int T, int k, const cor_beam_params* bp, int want_align, size_t budget, Plan* out) {
    const long long q_full = (long long)2 * T * bp->batch_size;
    for (int L = n_rows; L >= 1; L = L > 1 ? (L + 1) / 2 : 0) {
        Ws ws; BeamDev b; StepIO io; size_t fixed = 0;
        if (layout(m, L, T, k, bp, 1, 1, want_align, nullptr, &ws, &b, &io, &fixed)) return 1;
        const size_t rb = row_bytes(m, T), xb = rec_bytes(m, T) + 2 * KMAX * 12;
        const long long q_min = std::min<long long>(q_full, 2 * T + 2);
        const size_t need = fixed + (size_t)L * rb + (size_t)L * q_min * xb;
        if (need > budget) continue;
        const size_t rest = budget - fixed;
        long long Rcap = std::max<long long>(L, std::min<long long>((long long)L * bp->batch_size, (long long)(rest / 4 / rb)));
        long long quota = std::min<long long>(q_full, (long long)((rest - (size_t)Rcap * rb) / ((size_t)L * xb)));
        // shrink until the exact layout fits (alignment padding)
        for (int tries = 0; tries < 64; ++tries) {
            size_t bytes = 0;
            if (layout(m, L, T, k, bp, (int)quota, (int)Rcap, want_align, nullptr, &ws, &b, &io, &bytes)) return 1;
            if (bytes <= budget) { *out = Plan{L, (int)quota, b.cap, (int)Rcap, want_align, bytes}; return 0; }
            if (quota > q_min) quota = std::max<long long>(q_min, quota - std::max<long long>(1, quota / 32));
            else if (Rcap > L) Rcap = std::max<long long>(L, Rcap - std::max<long long>(1, Rcap / 32));
            else break;
        }
    }
    set_error("beam workspace too small for a single line");
    return 1;
}

int check_params(const cor_model* m, const cor_beam_params* bp) {
    COR_CHECK(m && m->finalized, "model not finalized");
    COR_CHECK(bp, "null beam parameters");
    COR_CHECK(bp->batch_size >= 1 && bp->batch_size <= 256, "batch_size must be in 1..256");
    COR_CHECK(bp->beam_width_in >= 1 && bp->beam_width_in <= KMAX - 1, "beam_width_in must be in 1..15");
    COR_CHECK(bp->beam_width_out >= 0, "beam_width_out must be non-negative");
    COR_CHECK(bp->n_best >= 0 && bp->n_best <= FMAX, "n_best must be in 0..64");
    COR_CHECK(m->d <= 8, "depth above 8 is not supported");
    return 0;
}


std::vector<uint64_t> graph_key(const cor_model* m, int L, int T, int k, const cor_beam_params* bp, const Plan& plan, bool want_align,
                                const void* workspace, bool has_val) {
    uint64_t thr = 0, rej = 0;
    memcpy(&thr, &bp->beam_threshold_in, 8); memcpy(&rej, &bp->rejection_threshold, 8);
    return {0xBEA3ull, (uint64_t)(uintptr_t)workspace, (uint64_t)L, (uint64_t)T, (uint64_t)k, (uint64_t)plan.quota_x, (uint64_t)plan.Rcap,
            (uint64_t)bp->batch_size, (uint64_t)bp->beam_width_in, (uint64_t)bp->beam_width_out, (uint64_t)bp->n_best, thr, rej,
            (uint64_t)want_align, (uint64_t)has_val, (uint64_t)m->passes, (uint64_t)(m->side != nullptr)};
}

// One wave over `L` lines given as HOST arrays; results to HOST arrays (same row order).
int run_wave(cor_model* m, int L, int T, int k, const int32_t* h_row_T, const int32_t* h_idx, const float* h_val,
             const cor_beam_params* bp, const Plan& plan, void* workspace,
             int32_t* o_chars, float* o_probs, double* o_scores, int32_t* o_len, float* o_align, int64_t* o_stats,
             bool on_device, int64_t* host_stats, cudaStream_t st) {
    // on_device: the line arrays and the result arrays are DEVICE memory; host_stats (always HOST, [L][kStats])
    // additionally receives the statistics, which carry the per-line overflow flag the retry loop needs
    const cudaMemcpyKind kin = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    const cudaMemcpyKind kout = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    Ws ws; BeamDev b; StepIO io; size_t bytes = 0;
    if (layout(m, L, T, k, bp, plan.quota_x, plan.Rcap, o_align != nullptr, workspace, &ws, &b, &io, &bytes)) return 1;
    const size_t n = (size_t)L * T * k;
    COR_CUDA(cudaMemcpyAsync(ws.row_T, h_row_T, (size_t)L * 4, kin, st));
    COR_CUDA(cudaMemcpyAsync(ws.idx, h_idx, n * 4, kin, st));
    if (h_val) COR_CUDA(cudaMemcpyAsync(ws.val, h_val, n * 4, kin, st));
    else { ws.val = nullptr; b.val = nullptr; }
    if (run_encoder(m, ws, st)) return 1;
    COR_CUDA(cudaMemsetAsync(b.counters, 0, 16, st));
    beam_init_kernel<<<L, 128, 0, st>>>(b); ++ctx().launches;
    COR_CUDA(cudaGetLastError());
    const int split = m->split ? 1 : 0;
    const int row_grid = std::min(plan.Rcap, 148 * 8);
    const size_t expand_smem = (size_t)(m->cfg.lm_predict ? 2 : 1) * m->Vp * 4;
    const size_t push_smem = (size_t)MAX_NEW * 12;
    // blocks per line of beam_push_kernel: as many as keep 2 x 148 blocks busy when every line is on the heavy list
    const int push_parts = getenv("COR_PUSH_PARTS") ? std::max(1, std::min(16, atoi(getenv("COR_PUSH_PARTS")))) : std::max(1, std::min(8, 148 * 2 / L));
    const int t_hi = T;      // every row_T is <= t_max
    // pop + scan: which hypotheses of every line are expanded next, their rows, and whether anything is left to do
    const bool small_sel = L <= kSelectSmall && !getenv("COR_NO_SELECT_SMALL");
    // Page of lines with the score row in registers: gather (with the scan) -> decoder step -> one kernel per line for the
    // rest (6 kernels per iteration instead of 9).  OPT-IN (COR_LINE_KERNEL=1): measured 8.08 -> 7.67 ms on a 32-line page
    // of one hypothesis per line, but 96 -> 257 ms on a 64-line page with ~116 hypotheses per line -- one block per line
    // serialises what the row-parallel expand and the 1024-thread push spread out -- and which of the two a page is
    // is not known when the graph is recorded.
    const bool line_path = small_sel && (m->Vp == 128 || m->Vp == 256 || m->Vp == 512 || m->Vp == 1024) && getenv("COR_LINE_KERNEL") != nullptr;
    auto line_kernel = [&](cudaStream_t s_, int do_expand) -> int {
        switch (m->Vp) {
            case 128:  beam_line_kernel<4><<<L, 256, push_smem, s_>>>(b, split, do_expand); break;
            case 256:  beam_line_kernel<8><<<L, 256, push_smem, s_>>>(b, split, do_expand); break;
            case 512:  beam_line_kernel<16><<<L, 256, push_smem, s_>>>(b, split, do_expand); break;
            default:   beam_line_kernel<32><<<L, 256, push_smem, s_>>>(b, split, do_expand); break;
        }
        ctx().launches += 1;
        COR_CUDA(cudaGetLastError());
        return 0;
    };
    auto select = [&](cudaStream_t s_, const LoopCtl& lc, int set_cond) -> int {
        if (line_path) return 0;       // the line kernel pops; the scan lives in the gather
        if (small_sel) {        // a page of lines: pop, scan and rowmap in one block
            beam_select_small_kernel<<<1, 1024, 0, s_>>>(b, lc, set_cond);
            ctx().launches += 1;
            COR_CUDA(cudaGetLastError());
            return 0;
        }
        beam_pop_kernel<<<(L + 7) / 8, 256, 0, s_>>>(b);
        beam_scan_kernel<<<1, 1024, 0, s_>>>(b, lc, set_cond);
        ctx().launches += 2;
        COR_CUDA(cudaGetLastError());
        return 0;
    };
    LoopCtl body_ctl{};             // loop controls of the body being captured (the line path sets them in its gather)
    int body_cond = 0;
    auto gather = [&](cudaStream_t s_) -> int {
        if (line_path) {
            beam_gather_scan_kernel<<<row_grid, 128, 0, s_>>>(b, split, body_ctl, body_cond);
            ctx().launches += 1;
            COR_CUDA(cudaGetLastError());
            return 0;
        }
        if (!small_sel) { beam_rowmap_kernel<<<(L + 7) / 8, 256, 0, s_>>>(b); ++ctx().launches; }
        beam_gather_kernel<<<row_grid, 128, 0, s_>>>(b, split);
        ctx().launches += 1;
        COR_CUDA(cudaGetLastError());
        return 0;
    };
    auto insert = [&](cudaStream_t s_) -> int {
        if (line_path) return line_kernel(s_, 1);
        const int warp_grid = std::min((plan.Rcap + 3) / 4, 148 * 8);
        switch (m->Vp) {        // score row in registers when it fits (32 values per lane at most)
            case 128:  beam_expand_warp_kernel<4><<<warp_grid, 128, 0, s_>>>(b, split); break;
            case 256:  beam_expand_warp_kernel<8><<<warp_grid, 128, 0, s_>>>(b, split); break;
            case 512:  beam_expand_warp_kernel<16><<<warp_grid, 128, 0, s_>>>(b, split); break;
            case 1024: beam_expand_warp_kernel<32><<<warp_grid, 128, 0, s_>>>(b, split); break;
            default:   beam_expand_kernel<<<row_grid, 128, expand_smem, s_>>>(b, split);
        }
        beam_push_small_kernel<<<(L + 7) / 8, 256, 0, s_>>>(b);
        beam_push_kernel<<<dim3(push_parts, std::min(L, std::max(1, 148 * 2 / push_parts))), PUSH_THREADS, push_smem, s_>>>(b);
        ctx().launches += 3;
        COR_CUDA(cudaGetLastError());
        return 0;
    };
    auto iteration = [&](cudaStream_t s_, bool roots) -> int {
        io.wq_ready = !roots;     // iteration 0 expands the roots: their query comes from the encoder state
        return gather(s_) || decoder_core(m, ws, io, s_) || insert(s_);
    };
    // Iterations 1.. are ONE graph launch: a WHILE node whose body is {iteration, select} and whose condition the
    // scan kernel sets on the device -- no launch latency between the ~10 kernels of an iteration and no host
    // round trip to find out that every line has finished.  (Per-launch profiling brackets launches with events on
    // the host: it keeps the host loop, as does COR_NO_GRAPH=1.)
    const bool use_graph = loop_graphs_enabled();
    int graph_body_kernels = 0;
    const LoopCtl no_ctl{};
    if (line_path ? line_kernel(st, 0) : select(st, no_ctl, 0)) return 1;
    for (int it = 0; it <= 2 * t_hi; ++it) {
        if (use_graph && it == 1) {
            LoopGraph* g = nullptr;
            if (loop_graph_lookup(m, graph_key(m, L, T, k, bp, plan, o_align != nullptr, workspace, ws.val != nullptr), &g)) return 1;
            if (!g->exec) {
                // The GEMMs' tile shape depends on the row count, which only the device knows: where the capacity
                // (Rcap) would pick another shape than a handful of rows, the decoder step is captured once per
                // shape class as the branches of a SWITCH node that the scan kernel sets from the next row count
                // (one page of lines = one hypothesis row per line most of the time: 128x32 tiles over 64 SMs
                // instead of 8 CTA pairs of a 148-CTA grid).
                LoopCtl lc{};
                int r_narrow = 0, r_single = 0;
                tc_shape_thresholds(4 * m->W, &r_narrow, &r_single);
                const bool tc = m->split && tc_available();
                lc.has_shape = (tc && plan.Rcap > r_narrow && !getenv("COR_NO_SHAPE_SWITCH")) ? 1 : 0;
                lc.r_narrow = r_narrow; lc.r_single = r_single;
                cudaGraph_t body;
                LoopCtl* d_lc = reinterpret_cast<LoopCtl*>(b.counters + 4);      // the first node reads the handles from memory: they exist only after it has been added
                void* kargs[2] = {(void*)&b, (void*)&d_lc};
                cudaKernelNodeParams kp{};
                kp.func = (void*)beam_cond_kernel; kp.gridDim = dim3(1); kp.blockDim = dim3(1); kp.kernelParams = kargs;
                if (loop_graph_begin(g, kp, &lc.cond, &body)) return 1;
                cudaStream_t cs = loop_capture_stream(m, 0), cs2 = loop_capture_stream(m, 1);
                COR_CHECK(cs && cs2, "no capture stream");
                const int64_t before = ctx().launches;
                COR_CUDA(cudaStreamBeginCaptureToGraph(cs, body, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
                io.wq_ready = true;
                // the SWITCH selector is an argument of the kernel that sets it (the gather's scan on the line path), which is
                // captured BEFORE the node exists: create the handle first
                if (lc.has_shape) COR_CUDA(cudaGraphConditionalHandleCreate(&lc.shape, g->graph, 0, 0));
                body_ctl = lc; body_cond = 1;
                int rc = gather(cs);
                body_ctl = LoopCtl{}; body_cond = 0;
                int64_t extra = 0;
                if (!rc && lc.has_shape) {
                    cudaGraph_t branch[3];
                    rc = loop_graph_switch(cs, 3, lc.shape, branch);
                    for (int c = 0; c < 3 && !rc; ++c) {
                        const int64_t b0 = ctx().launches;
                        cudaError_t e = cudaStreamBeginCaptureToGraph(cs2, branch[c], nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed);
                        if (e != cudaSuccess) { set_error(std::string("capturing a shape branch: ") + cudaGetErrorString(e)); rc = 1; break; }
                        ctx().force_shape = c + 1;
                        rc = decoder_core(m, ws, io, cs2);
                        ctx().force_shape = 0;
                        cudaGraph_t got = nullptr;
                        e = cudaStreamEndCapture(cs2, &got);
                        if (!rc && e != cudaSuccess) { set_error(std::string("capturing a shape branch: ") + cudaGetErrorString(e)); rc = 1; }
                        if (c > 0) extra += ctx().launches - b0;     // one branch runs per iteration
                    }
                } else if (!rc) {
                    rc = decoder_core(m, ws, io, cs);
                }
                rc = rc || insert(cs) || select(cs, lc, 1);
                cudaGraph_t captured = nullptr;
                const cudaError_t ce = cudaStreamEndCapture(cs, &captured);
                g->body_kernels = (int)(ctx().launches - before - extra);
                ctx().launches = before;            // captured, not run
                if (rc) return 1;
                COR_CUDA(ce);
                COR_CUDA(cudaGraphInstantiate(&g->exec, g->graph, 0));
                g->ctl.assign(reinterpret_cast<const char*>(&lc), reinterpret_cast<const char*>(&lc) + sizeof(lc));
            }
            // the loop controls (graph-owned handles) for the first node, in this wave's workspace
            COR_CUDA(cudaMemcpyAsync(b.counters + 4, g->ctl.data(), g->ctl.size(), cudaMemcpyHostToDevice, st));
            COR_CUDA(cudaGraphLaunch(g->exec, st));
            ctx().launches += 1; ctx().in_graph += 1; ctx().graph_launches += 1;    // beam_cond_kernel; the bodies are counted after the wave's final sync
            graph_body_kernels = g->body_kernels;
            break;
        }
        if (!use_graph && (it % 8 == 7 || it == 2 * t_hi)) {          // host check: has every line finished?
            int32_t cnt[2] = {0, 0};
            COR_CUDA(cudaMemcpyAsync(cnt, b.counters, 8, cudaMemcpyDeviceToHost, st));
            COR_CUDA(cudaStreamSynchronize(st));
            if (cnt[1] == 0 && cnt[0] == 0) break;
        }
        if (iteration(st, it == 0) || select(st, no_ctl, 0)) return 1;
    }
    // results: staged in the wave's own buffers, n_best slots per line
    const size_t S = 2 * (size_t)T * b.n_best;
    int32_t* d_chars = b.r_chars; float* d_probs = b.r_probs; double* d_score = b.r_score; int32_t* d_len = b.r_len;
    float* d_align = b.o_align;
    COR_CUDA(cudaMemsetAsync(d_chars, 0, (size_t)L * S * 4, st));
    COR_CUDA(cudaMemsetAsync(d_probs, 0, (size_t)L * S * 4, st));
    beam_finish_kernel<<<L, 128, 0, st>>>(b, d_chars, d_probs, d_score, d_len, d_align, b.o_stats);
    ++ctx().launches;
    COR_CUDA(cudaGetLastError());
    COR_CUDA(cudaMemcpyAsync(o_chars, d_chars, (size_t)L * S * 4, kout, st));
    COR_CUDA(cudaMemcpyAsync(o_probs, d_probs, (size_t)L * S * 4, kout, st));
    COR_CUDA(cudaMemcpyAsync(o_scores, d_score, (size_t)L * b.n_best * 8, kout, st));
    COR_CUDA(cudaMemcpyAsync(o_len, d_len, (size_t)L * b.n_best * 4, kout, st));
    COR_CUDA(cudaMemcpyAsync(o_stats, b.o_stats, (size_t)L * kStats * 8, kout, st));
    COR_CUDA(cudaMemcpyAsync(host_stats, b.o_stats, (size_t)L * kStats * 8, cudaMemcpyDeviceToHost, st));
    if (o_align) COR_CUDA(cudaMemcpyAsync(o_align, d_align, (size_t)L * S * T * 4, kout, st));
    int32_t h_cnt[4] = {0, 0, 0, 0};
    if (graph_body_kernels) COR_CUDA(cudaMemcpyAsync(h_cnt, b.counters, 16, cudaMemcpyDeviceToHost, st));
    COR_CUDA(cudaStreamSynchronize(st));
    ctx().launches += (int64_t)h_cnt[3] * graph_body_kernels;     // kernels the WHILE node ran (bodies counted by the scan kernel)
    ctx().in_graph += (int64_t)h_cnt[3] * graph_body_kernels;
    return 0;
}

}  // namespace

}  // namespace cor

using namespace cor;

extern "C" {

int cor_beam_workspace_bytes(const cor_model* m, int n_rows, int t_max, int k_alt, const cor_beam_params* bp,
                             size_t budget_bytes, size_t* bytes) {
    COR_CHECK(bytes, "null argument");
    if (check_params(m, bp)) return 1;
    COR_CHECK(n_rows > 0 && t_max > 0 && k_alt > 0, "empty call geometry");
    Plan plan;
    if (make_plan(m, n_rows, t_max, k_alt, bp, 1, budget_bytes, &plan)) return 1;
    // ask for the whole budget when the full-size arena does not fit, so that retries of heavy lines have room
    const long long q_full = (long long)2 * t_max * bp->batch_size;
    *bytes = (plan.L == n_rows && plan.quota_x >= q_full && plan.Rcap >= (long long)n_rows * bp->batch_size) ? plan.bytes : budget_bytes;
    return 0;
}

// Staging memory of the retry waves: host (malloc) or device (cudaMalloc), freed on scope exit.
struct Stage {
    bool dev; std::vector<void*> ptrs;
    explicit Stage(bool d) : dev(d) {}
    ~Stage() { for (void* p : ptrs) { if (dev) cudaFree(p); else free(p); } }
    void* take(size_t bytes) {
        void* p = nullptr;
        if (!bytes) bytes = 8;
        if (dev) { if (cudaMalloc(&p, bytes) != cudaSuccess) return nullptr; }
        else p = malloc(bytes);
        if (p) ptrs.push_back(p);
        return p;
    }
};

static int beam_correct(cor_model* m, const cor_lines* hl, bool on_device, const cor_beam_params* bp, void* workspace,
                        size_t workspace_bytes, int32_t* out_chars, float* out_probs, double* out_scores, int32_t* out_len,
                        float* out_align, int64_t* stats, cudaStream_t st) {
    if (check_params(m, bp)) return 1;
    CtxScope scope(&m->ctx);
    COR_CHECK(hl && workspace && out_chars && out_probs && out_scores && out_len && stats, "null argument");
    COR_CHECK(hl->n_rows > 0 && hl->t_max > 0 && hl->k_alt > 0, "empty call geometry");
    COR_CUDA(cudaSetDevice(m->device));
    COR_CUDA(cudaFuncSetAttribute(beam_push_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_NEW * 12));
    COR_CUDA(cudaFuncSetAttribute(beam_line_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_NEW * 12));
    COR_CUDA(cudaFuncSetAttribute(beam_line_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_NEW * 12));
    COR_CUDA(cudaFuncSetAttribute(beam_line_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_NEW * 12));
    COR_CUDA(cudaFuncSetAttribute(beam_line_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_NEW * 12));
    const int B = hl->n_rows, T = hl->t_max, k = hl->k_alt;
    const int NB = std::max(1, bp->n_best);                  // result slots per line
    const size_t S = 2 * (size_t)T * NB, tk = (size_t)T * k;
    auto xcopy = [&](void* dst, const void* src, size_t bytes) -> int {
        if (!on_device) { memcpy(dst, src, bytes); return 0; }
        COR_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, st));
        return 0;
    };
    std::vector<int> pending(B);
    for (int j = 0; j < B; ++j) pending[j] = j;
    std::vector<int64_t> hstats;
    // Lines that run out of expansion records are retried with a larger share of the workspace.  How large: a failed
    // line reports the records it used and the iterations it came through (stats[1], stats[3]); 1.5 x that rate over all
    // 2T iterations, at least twice what it had, at most the 2T * batch_size it can ever use.  The retry waves take the
    // lines in the order of that need, each wave as many lines as its hardest line allows (a power of two, so that few
    // distinct wave geometries -- loop graphs -- exist): easy lines share large waves, only the hard ones get small ones.
    const long long q_full = (long long)2 * T * bp->batch_size;
    std::vector<long long> need(B, m->beam_full_hint ? q_full : 0);
    int lines_done = 0, lines_heavy = 0;         // ... of which used more than half of the full quota
    const bool sized_retry = getenv("COR_NO_SIZED_RETRY") == nullptr;
    const double retry_factor = getenv("COR_RETRY_FACTOR") ? atof(getenv("COR_RETRY_FACTOR")) : 1.5;
    for (int level = 0; !pending.empty(); ++level) {
        COR_CHECK(level < 16, "beam search: retry depth exceeded");
        Plan plan;
        if (level > 0) std::stable_sort(pending.begin(), pending.end(), [&](int a, int b2) { return need[a] < need[b2]; });
        std::vector<int> failed;
        for (size_t p0 = 0; p0 < pending.size(); p0 += plan.L) {
            const int rest = (int)(pending.size() - p0);
            if (level == 0 && !(sized_retry && m->beam_full_hint)) {
                if (p0 == 0 && make_plan(m, rest, T, k, bp, out_align != nullptr, workspace_bytes, &plan)) return 1;
            } else {
                // the largest wave (all the rest, or a power of two) whose plan gives its hardest line what it needs
                int best = 0;
                for (int cand = rest; cand >= 1; cand = cand == rest ? (1 << (31 - __builtin_clz((unsigned)std::max(1, rest - 1)))) : cand / 2) {
                    Plan tmp;
                    if (make_plan(m, cand, T, k, bp, out_align != nullptr, workspace_bytes, &tmp)) return 1;
                    if (tmp.L == cand && tmp.quota_x >= need[pending[p0 + cand - 1]]) { best = cand; plan = tmp; break; }
                    if (cand == 1) break;
                }
                if (!best && make_plan(m, 1, T, k, bp, out_align != nullptr, workspace_bytes, &plan)) return 1;     // all there is for one line
            }
            const int L = (int)std::min<size_t>(plan.L, pending.size() - p0);
            const bool direct = level == 0 && L == B;        // first wave over everything: use the caller's buffers
            Stage stage(on_device);
            const int32_t *w_T = hl->row_T, *w_idx = hl->idx; const float* w_val = hl->val;
            int32_t *w_chars = out_chars, *w_len = out_len; float *w_probs = out_probs, *w_align = out_align;
            double* w_scores = out_scores; int64_t* w_stats = stats;
            if (!direct) {
                int32_t* rT = (int32_t*)stage.take((size_t)L * 4);
                int32_t* ix = (int32_t*)stage.take((size_t)L * tk * 4);
                float* vl = hl->val ? (float*)stage.take((size_t)L * tk * 4) : nullptr;
                w_chars = (int32_t*)stage.take((size_t)L * S * 4); w_probs = (float*)stage.take((size_t)L * S * 4);
                w_scores = (double*)stage.take((size_t)L * NB * 8); w_len = (int32_t*)stage.take((size_t)L * NB * 4);
                w_stats = (int64_t*)stage.take((size_t)L * kStats * 8);
                w_align = out_align ? (float*)stage.take((size_t)L * S * T * 4) : nullptr;
                COR_CHECK(rT && ix && (vl || !hl->val) && w_chars && w_probs && w_scores && w_len && w_stats && (w_align || !out_align),
                          "out of memory for beam retry staging");
                for (int j = 0; j < L; ++j) {
                    const int src = pending[p0 + j];
                    if (xcopy(rT + j, hl->row_T + src, 4) || xcopy(ix + (size_t)j * tk, hl->idx + (size_t)src * tk, tk * 4)) return 1;
                    if (hl->val && xcopy(vl + (size_t)j * tk, hl->val + (size_t)src * tk, tk * 4)) return 1;
                }
                w_T = rT; w_idx = ix; w_val = vl;
            }
            Plan wp = plan;
            if (L != plan.L && make_plan(m, L, T, k, bp, out_align != nullptr, workspace_bytes, &wp)) return 1;
            hstats.assign((size_t)L * kStats, 0);
            if (run_wave(m, L, T, k, w_T, w_idx, w_val, bp, wp, workspace, w_chars, w_probs, w_scores, w_len, w_align, w_stats,
                         on_device, hstats.data(), st)) return 1;
            if (getenv("COR_BEAM_TRACE")) {
                int nfail = 0; long long used = 0, its = 0;
                for (int j = 0; j < L; ++j) { nfail += hstats[(size_t)j * kStats] == -1; used += hstats[(size_t)j * kStats + 1]; its = std::max<long long>(its, hstats[(size_t)j * kStats] == -1 ? hstats[(size_t)j * kStats + 3] : hstats[(size_t)j * kStats]); }
                fprintf(stderr, "[beam] level %d wave of %d lines, %d records per line, %d rows per iteration: %d lines out of records, %lld records used, longest line %lld iterations\n",
                        level, L, wp.quota_x, wp.Rcap, nfail, used, its);
            }
            for (int j = 0; j < L; ++j) {
                const int dst = pending[p0 + j];
                if (hstats[(size_t)j * kStats] == -1) {
                    COR_CHECK(wp.L > 1, "beam search: a single line overflows the whole workspace");
                    const long long used = hstats[(size_t)j * kStats + 1], its = hstats[(size_t)j * kStats + 3];
                    const long long est = (sized_retry && its > 0) ? (long long)(retry_factor * (double)used / (double)its * 2.0 * T) + bp->batch_size : q_full;
                    need[dst] = std::min(q_full, std::max(std::max<long long>(2 * wp.quota_x, 8), est));
                    failed.push_back(dst);
                    continue;
                }
                ++lines_done;
                if (2 * hstats[(size_t)j * kStats + 1] > q_full) ++lines_heavy;
                if (direct) continue;
                if (xcopy(out_chars + (size_t)dst * S, w_chars + (size_t)j * S, S * 4) ||
                    xcopy(out_probs + (size_t)dst * S, w_probs + (size_t)j * S, S * 4) ||
                    xcopy(out_scores + (size_t)dst * NB, w_scores + (size_t)j * NB, (size_t)NB * 8) ||
                    xcopy(out_len + (size_t)dst * NB, w_len + (size_t)j * NB, (size_t)NB * 4) ||
                    xcopy(stats + (size_t)dst * kStats, w_stats + (size_t)j * kStats, kStats * 8)) return 1;
                if (out_align && xcopy(out_align + (size_t)dst * S * T, w_align + (size_t)j * S * T, S * T * 4)) return 1;
            }
            if (on_device) COR_CUDA(cudaStreamSynchronize(st));      // staging is freed at the end of this scope
        }
        pending.swap(failed);
    }
    m->beam_full_hint = lines_done >= 8 && 10 * lines_heavy >= 9 * lines_done;
    return 0;
}

int cor_correct_beam_host(cor_model* m, const cor_lines* hl, const cor_beam_params* bp, void* workspace, size_t workspace_bytes,
                          int32_t* out_chars, float* out_probs, double* out_scores, int32_t* out_len, float* out_align,
                          int64_t* stats, void* stream) {
    return beam_correct(m, hl, false, bp, workspace, workspace_bytes, out_chars, out_probs, out_scores, out_len, out_align,
                        stats, (cudaStream_t)stream);
}

int cor_decode_beam_batch(cor_model* m, const cor_lines* dl, const cor_beam_params* bp, void* workspace, size_t workspace_bytes,
                          int32_t* out_chars, float* out_probs, double* out_scores, int32_t* out_len, float* out_align,
                          int64_t* stats, void* stream) {
    return beam_correct(m, dl, true, bp, workspace, workspace_bytes, out_chars, out_probs, out_scores, out_len, out_align,
                        stats, (cudaStream_t)stream);
}

}  // extern "C"
