// decpage.cu — persistent fused decoder for page-sized batches (fast batch-greedy decode, <= 64 lines, in blocks of 32 rows).
//
// North star item (3): "one fused decoder-step kernel that performs the LSTM cell, the local monotonic
// additive attention cell, the tied-weight softmax and the direct softmax feedback".  For a page of lines
// the whole decode_batch_greedy loop (seq2seq.py:1215-1286: 2T timesteps of lower cell, attention cell,
// top cell, tied output projection, softmax, argmax bookkeeping) is ONE cooperative launch.
//
// All decoder weights of the depth-2 topology stay in shared memory for the 2T timesteps, as fp16 head +
// residual pairs (18.9 MB at W = V = 512 -> 176 KB in each of 128 CTAs).  The grid is 4W/64 clusters of
// 4 CTAs (32 clusters at W = 512; a B200 holds 33 such clusters at one CTA per SM, but only 15 clusters of
// 8).  A cluster owns 64 gate columns (16 LSTM units) of both cells and 1/32 of the output columns;
// inside a cluster the CTAs split the reduction dimension K four ways:
//   * CTA r keeps the K-slice r of the cluster's columns of  [E.Kx | R]   (lower cell, K = Vp + W),
//     [Kx | Kc | R]  (attention cell, K = 2W + D)  and  [E ; W_a]  (logits | next query, K = W);
//   * per phase it fetches only ITS K-slice of the activation rows (32 x K/4, cp.async) -- 48 KB instead
//     of the 196 KB a column-split CTA would have to read -- and forms a 32 x 64 partial product with
//     mma.sync m16n8k16 (three-pass hi/lo: fp32-grade, like every GEMM of this library);
//   * after a cluster barrier CTA r sums the four partials of ITS 16 columns through distributed shared
//     memory (ld.shared::cluster, fixed order: deterministic) and runs the LSTM cell / bias epilogue.
// Operand tiles are stored unpadded with a 16-byte-chunk XOR swizzle (chunk ^ (row & 7)): conflict-free
// fragment loads without the padding that would not fit.
//
// A timestep is four phases separated by grid barriers:
//   P0  softmax + greedy bookkeeping of step s-1 (one warp per line, CTAs 0..B-1)  ||
//       attention of step s (one CTA per line: window energies, alignment row, context)
//   P1  lower cell     [p | h1] . [E.Kx | R]^T      -> h1', c1'
//   P2  attention cell [h1' | ctx | h2] . [...]^T   -> h2', c2'
//   P3  [logits | next query] = h2' . [E ; W_a]^T
#include "engine.h"
#include "epilogue.cuh"
#include <cuda_fp16.h>
#include <math.h>
#include <stdlib.h>
#include <algorithm>
#include <stdio.h>
#include <vector>

namespace cor {

namespace {

constexpr int PD_ROWS = 32;              // rows per pass (two m16 tiles); a call has up to kRecurMaxRows rows, taken in blocks
constexpr int PD_CS = 4;                 // CTAs per cluster = K-slices (B200 fits 33 clusters of 4 one-CTA-per-SM blocks, but only 15 of 8)
constexpr int PD_NCOL = 64;              // gate columns per cluster (16 LSTM units): one n-tile per warp
constexpr int PD_THREADS = 256;
constexpr int PART_LD = PD_NCOL + 4;

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ uint32_t ld_acquire(const unsigned* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void grid_sync(unsigned* counter, unsigned target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(counter, 1u);
        const long long t0 = clock64();
        while (ld_acquire(counter) < target) {
            if (clock64() - t0 > 4000000000LL) __trap();     // a protocol bug must fail the launch, not hang the GPU
        }
    }
    __syncthreads();
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// optional phase stamps (COR_PD_TIMING): CTA 0 records the time at which it LEAVES each barrier
__device__ __forceinline__ void stamp(const PageDecArgs& a, int s, int phase) {
    if (a.timing && blockIdx.x == 0 && threadIdx.x == 0) a.timing[(size_t)s * 4 + phase] = globaltimer_ns();
}
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ float4 ld_dsmem4(const float* local_ptr, uint32_t rank) {
    const uint32_t la = (uint32_t)__cvta_generic_to_shared(local_ptr);
    uint32_t ra;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(la), "r"(rank));
    float4 v;
    asm volatile("ld.shared::cluster.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(ra) : "memory");
    return v;
}
__device__ __forceinline__ float ld_dsmem1(const float* local_ptr, uint32_t rank) {
    const uint32_t la = (uint32_t)__cvta_generic_to_shared(local_ptr);
    uint32_t ra;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(la), "r"(rank));
    float v;
    asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(ra) : "memory");
    return v;
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
__device__ __forceinline__ uint32_t lds32(const __half* p) { return *reinterpret_cast<const uint32_t*>(p); }
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

// A [rows][KS] fp16 tile is stored with a 16-byte-chunk swizzle: element (row, k) sits at
// row * KS + (((k >> 3) ^ (row & 7)) << 3) + (k & 7).

// rows [0, nrows) x K-slice [k0, k0+KS) of a row-major fp16 matrix (leading dimension ld) -> swizzled tile
__device__ __forceinline__ void load_tile_async(__half* dst, const __half* src, int ld, int row0, int nrows, int k0, int KS) {
    const int chunks = KS >> 3;
    for (int i = threadIdx.x; i < nrows * chunks; i += PD_THREADS) {
        const int r = i / chunks, c8 = i % chunks;
        cp_async16(dst + r * KS + ((c8 ^ (r & 7)) << 3), src + (size_t)(row0 + r) * ld + k0 + c8 * 8);
    }
}

// part[32][ncol] (this CTA's K-slice) = A[32][KS] . Wt[ncol][KS]^T, three-pass hi/lo; ncol <= 64: one 8-column
// tile per warp.  `part` ALIASES the A tile (there is no room for both): every warp keeps its accumulators in
// registers until all warps are done reading A.
__device__ __forceinline__ void partial_product(float* part, const __half* a_hi, const __half* a_lo, const __half* w_hi, const __half* w_lo,
                                                int KS, int ncol, int B) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, q = lane & 3;
    const int m_tiles = B > 16 ? 2 : 1;
    const int nt = warp;
    float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
    if (nt < (ncol >> 3)) {
        const __half* bh = w_hi + (nt * 8 + g) * KS + 2 * q;
        const __half* bl = w_lo + (nt * 8 + g) * KS + 2 * q;
#pragma unroll 2
        for (int k0 = 0; k0 < KS; k0 += 16) {
            const int c0 = (((k0 >> 3) ^ g) << 3), c1 = ((((k0 >> 3) + 1) ^ g) << 3);     // row & 7 == g for every fragment row
            const uint32_t b_h[2] = {lds32(bh + c0), lds32(bh + c1)};
            const uint32_t b_l[2] = {lds32(bl + c0), lds32(bl + c1)};
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) {
                if (mt < m_tiles) {
                    const __half* ah = a_hi + (mt * 16 + g) * KS + 2 * q;
                    const __half* al = a_lo + (mt * 16 + g) * KS + 2 * q;
                    const uint32_t a_h[4] = {lds32(ah + c0), lds32(ah + 8 * KS + c0), lds32(ah + c1), lds32(ah + 8 * KS + c1)};
                    const uint32_t a_l[4] = {lds32(al + c0), lds32(al + 8 * KS + c0), lds32(al + c1), lds32(al + 8 * KS + c1)};
                    mma16816(acc[mt], a_l, b_h);
                    mma16816(acc[mt], a_h, b_l);
                    mma16816(acc[mt], a_h, b_h);
                }
            }
        }
    }
    __syncthreads();
    if (nt < (ncol >> 3)) {
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
            float* z0 = part + (mt * 16 + g) * PART_LD + nt * 8 + 2 * q;
            z0[0] = acc[mt][0]; z0[1] = acc[mt][1];
            z0[8 * PART_LD] = acc[mt][2]; z0[8 * PART_LD + 1] = acc[mt][3];
        }
    }
}

__device__ __forceinline__ void store_h(__half* hi, __half* lo, size_t off, float v) {
    __half h, l;
    f16_split(v, h, l);
    hi[off] = h; lo[off] = l;
}

// Attention of one line by one CTA (attention.py:526-575 restricted to the window; the arithmetic and its
// order follow attention_kernel in kernels.cu: per-lane fmaf chains over k, butterfly sums, weights in window order).
__device__ void attention_row_cta(const PageDecArgs& a, int row, bool first, float* a_out, float* align_log,
                                  __half* ctx_hi, __half* ctx_lo, float* s_scr) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int T = a.line_T[row];
    // centre = sum_j a_prev[j]*j + 1 (attention.py:551-553).  The previous alignment row was produced by THIS
    // CTA one timestep ago, so the centre is carried in shared memory (s_scr[30], computed at the end of the
    // previous call with the reference kernel's lane assignment and summation order) instead of re-reading the
    // row from global memory; the first step has a_prev == 0.
    const float centre = first ? 1.0f : s_scr[30];
    const int j_lo = max(0, (int)ceilf(centre - (float)a.window) - 1);
    const int j_hi = min(T - 1, (int)floorf(centre + (float)a.window) + 1);
    int j0 = 0, nj = 0;
    for (int j = j_lo; j <= j_hi && nj < 13; ++j) {
        if (fabsf(centre - (float)j) <= (float)a.window) { if (nj == 0) j0 = j; ++nj; }
    }
    // the annotation rows of the window do not depend on the energies: their loads are issued first and wait in
    // registers (one memory round trip for keys and annotations together)
    const int kx = threadIdx.x * 4;
    const bool has_x = kx < a.D && a.D <= PD_THREADS * 4 && a.enc_lo != nullptr;
    uint2 xh[13], xl[13];
    if (has_x) {
        const __half* eh = reinterpret_cast<const __half*>(a.enc) + ((size_t)row * a.T + j0) * a.D + kx;
        const __half* el = reinterpret_cast<const __half*>(a.enc_lo) + ((size_t)row * a.T + j0) * a.D + kx;
#pragma unroll
        for (int qq = 0; qq < 13; ++qq)
            if (qq < nj) {
                xh[qq] = *reinterpret_cast<const uint2*>(eh + (size_t)qq * a.D);
                xl[qq] = *reinterpret_cast<const uint2*>(el + (size_t)qq * a.D);
            }
    }
    const float* wq = a.wq + (size_t)row * a.W;
    for (int qq = warp; qq < nj; qq += PD_THREADS / 32) {      // one warp per window position
        const float* u = a.u + ((size_t)row * a.T + j0 + qq) * a.W;
        float t = 0.f;
        for (int k = lane * 4; k < a.W; k += 128) {
            const float4 w4 = *reinterpret_cast<const float4*>(wq + k);
            const float4 v4 = *reinterpret_cast<const float4*>(a.v_a + k);
            const float4 u4 = *reinterpret_cast<const float4*>(u + k);
            t = fmaf(tanhf_fast(w4.x + u4.x), v4.x, t);
            t = fmaf(tanhf_fast(w4.y + u4.y), v4.y, t);
            t = fmaf(tanhf_fast(w4.z + u4.z), v4.z, t);
            t = fmaf(tanhf_fast(w4.w + u4.w), v4.w, t);
        }
        t = warp_sum(t);
        if (lane == 0) s_scr[qq] = expf(t + a.b_v);
    }
    __syncthreads();
    float esum = 0.f;
    for (int qq = 0; qq < nj; ++qq) esum += s_scr[qq];
    for (int j = threadIdx.x; j < a.T; j += PD_THREADS) {
        const int qq = j - j0;
        const float v = (qq >= 0 && qq < nj) ? s_scr[qq] / esum : 0.f;
        a_out[(size_t)row * a.T + j] = v;
        if (align_log) align_log[(size_t)row * a.T + j] = v;
    }
    float next_centre = 0.f;
    if (warp == 0) {    // lane l owns the positions j = l (mod 32); the window has at most one of them, the rest are zeros
        float part = 0.f;
        for (int j = lane; j < T; j += 32) {
            const int qq = j - j0;
            if (qq >= 0 && qq < nj) part = fmaf(s_scr[qq] / esum, (float)j, part);
        }
        next_centre = warp_sum(part) + 1.0f;
    }
    if (has_x) {
        float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int qq = 0; qq < 13; ++qq)
            if (qq < nj) {
                const float w = s_scr[qq] / esum;
                const float2 h0 = __half22float2(*reinterpret_cast<const __half2*>(&xh[qq].x)), h1 = __half22float2(*reinterpret_cast<const __half2*>(&xh[qq].y));
                const float2 l0 = __half22float2(*reinterpret_cast<const __half2*>(&xl[qq].x)), l1 = __half22float2(*reinterpret_cast<const __half2*>(&xl[qq].y));
                c.x = fmaf(w, h0.x + l0.x, c.x); c.y = fmaf(w, h0.y + l0.y, c.y); c.z = fmaf(w, h1.x + l1.x, c.z); c.w = fmaf(w, h1.y + l1.y, c.w);
            }
        store_h(ctx_hi, ctx_lo, kx + 0, c.x); store_h(ctx_hi, ctx_lo, kx + 1, c.y);
        store_h(ctx_hi, ctx_lo, kx + 2, c.z); store_h(ctx_hi, ctx_lo, kx + 3, c.w);
    } else if (!(a.D <= PD_THREADS * 4 && a.enc_lo != nullptr)) {
        for (int k = threadIdx.x * 4; k < a.D; k += PD_THREADS * 4) {
            float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int qq = 0; qq < nj; ++qq) {
                const float w = s_scr[qq] / esum;
                const float4 x = load_split4(a.enc, a.enc_lo, ((size_t)row * a.T + j0 + qq) * a.D + k);
                c.x = fmaf(w, x.x, c.x); c.y = fmaf(w, x.y, c.y); c.z = fmaf(w, x.z, c.z); c.w = fmaf(w, x.w, c.w);
            }
            store_h(ctx_hi, ctx_lo, k + 0, c.x); store_h(ctx_hi, ctx_lo, k + 1, c.y);
            store_h(ctx_hi, ctx_lo, k + 2, c.z); store_h(ctx_hi, ctx_lo, k + 3, c.w);
        }
    }
    __syncthreads();                                   // everyone is done with the energies in s_scr
    if (warp == 0 && lane == 0) s_scr[30] = next_centre;
}

// softmax of one line's logits + greedy bookkeeping (decode_batch_greedy, seq2seq.py:1250-1260) by one warp;
// same arithmetic as softmax_kernel<true> in kernels.cu.  p goes to the [p | h1] operand rows (hi/lo).
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
__device__ void softmax_greedy_row(const PageDecArgs& a, int row, int step, __half* p_hi, __half* p_lo) {
    // One pass over the logits (kept in registers) and ONE expf per element: softmax_kernel evaluates
    // expf(z - m) twice with identical results, so the bits are the same, but this warp's chain of loads, expf
    // and divisions is the critical path of the phase (~6 us as three passes).  Vp <= 32 * VL (checked on the host).
    constexpr int VL = 16;
    const int lane = threadIdx.x & 31;
    const float* z = a.logits + (size_t)row * a.Vp;
    float e[VL];
    float m = -INFINITY;
#pragma unroll
    for (int k = 0; k < VL; ++k) {
        const int v = lane + 32 * k;
        e[k] = v < a.V ? z[v] : -INFINITY;
    }
#pragma unroll
    for (int k = 0; k < VL; ++k) m = fmaxf(m, e[k]);
    m = warp_max(m);
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < VL; ++k)
        if (lane + 32 * k < a.V) { e[k] = expf(e[k] - m); s += e[k]; } else e[k] = 0.f;
    s = warp_sum(s);
    float best = -INFINITY, second = -INFINITY;
    int best_v = 0x7fffffff;
#pragma unroll
    for (int k = 0; k < VL; ++k) {
        const int v = lane + 32 * k;
        if (v < a.Vp) {
            const float pv = v < a.V ? e[k] / s : 0.f;
            if (v < a.V && v >= 1) {
                if (pv > best) { second = best; best = pv; best_v = v; }
                else second = fmaxf(second, pv);
            }
            store_h(p_hi, p_lo, v, pv);
        }
    }
    top2_merge(best, best_v, second);
    if (lane == 0) {
        const bool alive = !a.done[row] && a.live[row] && step < 2 * a.line_T[row];
        if (alive) {
            const int n = a.out_len[row];
            a.out_chars[(size_t)row * a.S + n] = best_v;
            a.out_probs[(size_t)row * a.S + n] = best;
            a.score_sum[row] += (double)(-logf(best));
            a.out_len[row] = n + 1;
            if (best_v == a.eos) a.done[row] = 1;
            if (a.out_ties && best - second <= kTieMargin * best) a.out_ties[row] += 1;
        }
    }
}

__global__ void __launch_bounds__(PD_THREADS, 1) page_decoder_kernel(const PageDecArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int B = a.B, W = a.W;
    const int K_low = a.Vp + W, K_top = 2 * W + a.D;
    const int KS_low = K_low / PD_CS, KS_top = K_top / PD_CS, KS_e = W / PD_CS;
    const int n_clusters = gridDim.x / PD_CS;
    const int NE = (a.Vp + W) / n_clusters;          // output columns (logits | query) per cluster
    const int KS_max = max(KS_low, max(KS_top, KS_e));
    __half* wl_hi = reinterpret_cast<__half*>(smem);             // [128][KS_low]
    __half* wl_lo = wl_hi + PD_NCOL * KS_low;
    __half* wt_hi = wl_lo + PD_NCOL * KS_low;                     // [128][KS_top]
    __half* wt_lo = wt_hi + PD_NCOL * KS_top;
    __half* we_hi = wt_lo + PD_NCOL * KS_top;                     // [NE][KS_e]
    __half* we_lo = we_hi + NE * KS_e;
    __half* a_hi = we_lo + NE * KS_e;                             // [32][KS_max]
    __half* a_lo = a_hi + PD_ROWS * KS_max;
    float* part = reinterpret_cast<float*>(a_hi);                 // [32][PART_LD] partial sums, aliasing the A tile (see partial_product)
    float* s_scr = reinterpret_cast<float*>(a_lo + PD_ROWS * KS_max);  // 32 floats
    const int tid = threadIdx.x;
    const uint32_t rank = cluster_ctarank();
    const int cl = blockIdx.x / PD_CS;
    const int col0 = cl * PD_NCOL;                   // first gate column of the cluster
    const int ecol0 = cl * NE;                       // first output column of the cluster

    // ---- resident weights: K-slice `rank` of the cluster's columns
    load_tile_async(wl_hi, a.low_hi, K_low, col0, PD_NCOL, rank * KS_low, KS_low);
    load_tile_async(wl_lo, a.low_lo, K_low, col0, PD_NCOL, rank * KS_low, KS_low);
    load_tile_async(wt_hi, a.top_hi, K_top, col0, PD_NCOL, rank * KS_top, KS_top);
    load_tile_async(wt_lo, a.top_lo, K_top, col0, PD_NCOL, rank * KS_top, KS_top);
    load_tile_async(we_hi, a.ewa_hi, a.ewa_ld, ecol0, NE, rank * KS_e, KS_e);
    load_tile_async(we_lo, a.ewa_lo, a.ewa_ld, ecol0, NE, rank * KS_e, KS_e);
    {
        const uint4 zero = make_uint4(0, 0, 0, 0);
        for (int i = tid; i < 2 * PD_ROWS * KS_max / 8; i += PD_THREADS) reinterpret_cast<uint4*>(a_hi)[i] = zero;   // a_hi and a_lo are adjacent
    }
    // ---- operand rows of step 0: [p = 0 | h1_0] and [. | . | h2_0]
    for (int i = blockIdx.x * PD_THREADS + tid; i < B * W; i += gridDim.x * PD_THREADS) {
        const int r = i / W, k = i % W;
        a.lowA_hi[0][(size_t)r * K_low + a.Vp + k] = a.h0_hi[0][i];
        a.lowA_lo[0][(size_t)r * K_low + a.Vp + k] = a.h0_lo[0][i];
        a.topA_hi[0][(size_t)r * K_top + W + a.D + k] = a.h0_hi[1][i];
        a.topA_lo[0][(size_t)r * K_top + W + a.D + k] = a.h0_lo[1][i];
    }
    for (int i = blockIdx.x * PD_THREADS + tid; i < B * a.Vp; i += gridDim.x * PD_THREADS) {
        const int r = i / a.Vp, k = i % a.Vp;
        a.lowA_hi[0][(size_t)r * K_low + k] = __float2half_rn(0.f);
        a.lowA_lo[0][(size_t)r * K_low + k] = __float2half_rn(0.f);
    }
    cp_async_wait_all();
    unsigned epoch = 0;
    grid_sync(a.barrier, gridDim.x * ++epoch);

    for (int s = 0; s <= a.S; ++s) {
        const int cur = s & 1, nxt = cur ^ 1;
        // ================= P0: softmax of step s-1 || attention of step s
        if (s > 0 && (int)blockIdx.x < (B + 7) / 8) {
            const int row = blockIdx.x * 8 + (tid >> 5);
            if (row < B) softmax_greedy_row(a, row, s - 1, a.lowA_hi[cur] + (size_t)row * K_low, a.lowA_lo[cur] + (size_t)row * K_low);
        }
        if (s < a.S) {
            const int row = (int)gridDim.x - 1 - (int)blockIdx.x;       // attention rows on the other end of the grid
            if (row < B)
                attention_row_cta(a, row, s == 0, a.a[nxt], a.o_align ? a.o_align + (size_t)s * B * a.T : nullptr,
                                  a.topA_hi[cur] + (size_t)row * K_top + W, a.topA_lo[cur] + (size_t)row * K_top + W, s_scr);
        }
        grid_sync(a.barrier, gridDim.x * ++epoch);
        if (s == a.S) break;
        stamp(a, s, 0);

        // ================= P1: lower cell  [p | h1] -> h1', c1'   (rows in blocks of 32 over the resident weights)
        for (int r0 = 0; r0 < B; r0 += PD_ROWS) {
            const int Bb = min(PD_ROWS, B - r0);
            if (r0 > 0) cluster_sync();                  // the peers are done with this CTA's partial sums of the previous block
            load_tile_async(a_hi, a.lowA_hi[cur], K_low, r0, Bb, rank * KS_low, KS_low);
            load_tile_async(a_lo, a.lowA_lo[cur], K_low, r0, Bb, rank * KS_low, KS_low);
            cp_async_wait_all();
            __syncthreads();
            partial_product(part, a_hi, a_lo, wl_hi, wl_lo, KS_low, PD_NCOL, Bb);
            cluster_sync();
            if (tid < PD_ROWS * 4) {
                const int rl = tid >> 2, u = tid & 3, row = r0 + rl;
                if (rl < Bb) {
                    const int unit = (col0 >> 2) + 4 * rank + u;
                    float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
                    const float* src = part + rl * PART_LD + 16 * rank + 4 * u;
#pragma unroll
                    for (uint32_t r = 0; r < PD_CS; ++r) { const float4 v = ld_dsmem4(src, r); z.x += v.x; z.y += v.y; z.z += v.z; z.w += v.w; }
                    const float4 b = *reinterpret_cast<const float4*>(a.bias_low + 4 * unit);
                    const float ig = sigmoidf_fast(z.x + b.x), fg = sigmoidf_fast(z.y + b.y), gg = tanhf_fast(z.z + b.z), og = sigmoidf_fast(z.w + b.w);
                    const size_t so = (size_t)row * W + unit;
                    const float c = fg * a.c[0][so] + ig * gg;
                    a.c[0][so] = c;
                    const float h = og * tanhf_fast(c);
                    store_h(a.topA_hi[cur], a.topA_lo[cur], (size_t)row * K_top + unit, h);                 // input of this step's attention cell
                    store_h(a.lowA_hi[nxt], a.lowA_lo[nxt], (size_t)row * K_low + a.Vp + unit, h);          // recurrent input of the next step
                }
            }
        }
        grid_sync(a.barrier, gridDim.x * ++epoch);
        stamp(a, s, 1);

        // ================= P2: attention cell  [h1' | ctx | h2] -> h2', c2'
        for (int r0 = 0; r0 < B; r0 += PD_ROWS) {
            const int Bb = min(PD_ROWS, B - r0);
            if (r0 > 0) cluster_sync();
            load_tile_async(a_hi, a.topA_hi[cur], K_top, r0, Bb, rank * KS_top, KS_top);
            load_tile_async(a_lo, a.topA_lo[cur], K_top, r0, Bb, rank * KS_top, KS_top);
            cp_async_wait_all();
            __syncthreads();
            partial_product(part, a_hi, a_lo, wt_hi, wt_lo, KS_top, PD_NCOL, Bb);
            cluster_sync();
            if (tid < PD_ROWS * 4) {
                const int rl = tid >> 2, u = tid & 3, row = r0 + rl;
                if (rl < Bb) {
                    const int unit = (col0 >> 2) + 4 * rank + u;
                    float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
                    const float* src = part + rl * PART_LD + 16 * rank + 4 * u;
#pragma unroll
                    for (uint32_t r = 0; r < PD_CS; ++r) { const float4 v = ld_dsmem4(src, r); z.x += v.x; z.y += v.y; z.z += v.z; z.w += v.w; }
                    const float4 b = *reinterpret_cast<const float4*>(a.bias_top + 4 * unit);
                    const float ig = sigmoidf_fast(z.x + b.x), fg = sigmoidf_fast(z.y + b.y), gg = tanhf_fast(z.z + b.z), og = sigmoidf_fast(z.w + b.w);
                    const size_t so = (size_t)row * W + unit;
                    const float c = fg * a.c[1][so] + ig * gg;
                    a.c[1][so] = c;
                    const float h = og * tanhf_fast(c);
                    store_h(a.topA_hi[nxt], a.topA_lo[nxt], (size_t)row * K_top + W + a.D + unit, h);       // h2 of the next step (and of P3)
                }
            }
        }
        grid_sync(a.barrier, gridDim.x * ++epoch);
        stamp(a, s, 2);

        // ================= P3: [logits | next query] = h2' . [E ; W_a]^T
        for (int r0 = 0; r0 < B; r0 += PD_ROWS) {
            const int Bb = min(PD_ROWS, B - r0);
            if (r0 > 0) cluster_sync();
            load_tile_async(a_hi, a.topA_hi[nxt] + W + a.D, K_top, r0, Bb, rank * KS_e, KS_e);
            load_tile_async(a_lo, a.topA_lo[nxt] + W + a.D, K_top, r0, Bb, rank * KS_e, KS_e);
            cp_async_wait_all();
            __syncthreads();
            partial_product(part, a_hi, a_lo, we_hi, we_lo, KS_e, NE, Bb);
            cluster_sync();
            const int per = NE / PD_CS;              // output columns finalised by this CTA
            for (int i = tid; i < PD_ROWS * per; i += PD_THREADS) {
                const int rl = i / per, cc = i % per, row = r0 + rl;
                if (rl < Bb) {
                    const float* src = part + rl * PART_LD + per * rank + cc;
                    float z = 0.f;
#pragma unroll
                    for (uint32_t r = 0; r < PD_CS; ++r) z += ld_dsmem1(src, r);
                    const int col = ecol0 + per * rank + cc;
                    if (col < a.Vp) a.logits[(size_t)row * a.Vp + col] = z;
                    else a.wq[(size_t)row * W + (col - a.Vp)] = z + a.b_UW[col - a.Vp];
                }
            }
        }
        grid_sync(a.barrier, gridDim.x * ++epoch);     // logits and query complete before the next P0
        stamp(a, s, 3);
    }
    cluster_sync();     // no CTA leaves while a peer may still read its shared memory
}

size_t pagedec_smem_bytes(int W, int D, int Vp) {
    const int K_low = Vp + W, K_top = 2 * W + D;
    const int KS_low = K_low / PD_CS, KS_top = K_top / PD_CS, KS_e = W / PD_CS;
    const int n_clusters = 4 * W / PD_NCOL;
    const int NE = (Vp + W) / n_clusters;
    const int KS_max = std::max(KS_low, std::max(KS_top, KS_e));
    return (size_t)2 * 2 * (PD_NCOL * KS_low + PD_NCOL * KS_top + NE * KS_e + PD_ROWS * KS_max) + 32 * 4 + 16;
}

}  // namespace

bool page_decoder_supported(int W, int D, int Vp, int B) {
    if (B < 1 || B > kRecurMaxRows || Vp > 512) return false;     // softmax_greedy_row keeps a row in 16 registers per lane
    const int K_low = Vp + W, K_top = 2 * W + D;
    if (K_low % (PD_CS * 64) || K_top % (PD_CS * 64) || W % (PD_CS * 64)) return false;     // swizzle: K-slices in 64-element blocks
    if ((4 * W) % PD_NCOL) return false;
    const int n_clusters = 4 * W / PD_NCOL;
    if ((Vp + W) % (n_clusters * PD_CS * 8)) return false;
    const int NE = (Vp + W) / n_clusters;
    if (NE > PD_NCOL || (size_t)PD_ROWS * PART_LD * 4 > (size_t)2 * 2 * PD_ROWS * (W / PD_CS)) return false;
    const size_t smem = pagedec_smem_bytes(W, D, Vp);
    if (smem > 227 * 1024) return false;
    if (getenv("COR_NO_PAGEDEC")) return false;
    // the grid barrier needs every cluster resident: ask the occupancy calculator (per device and footprint)
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return false;
    static int c_dev = -1, c_clusters = 0; static size_t c_smem = 0;
    if (c_dev != dev || c_smem != smem) {
        c_dev = dev; c_smem = smem; c_clusters = 0;
        if (cudaFuncSetAttribute(page_decoder_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) == cudaSuccess) {
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3(n_clusters * PD_CS); cfg.blockDim = dim3(PD_THREADS); cfg.dynamicSmemBytes = smem;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = PD_CS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            int nc = 0;
            if (cudaOccupancyMaxActiveClusters(&nc, page_decoder_kernel, &cfg) == cudaSuccess) c_clusters = nc;
        }
        (void)cudaGetLastError();
    }
    return n_clusters <= c_clusters;
}

// COR_PD_TIMING=<file>: per-phase times of the LAST call (ns between CTA 0 leaving consecutive grid barriers,
// averaged over the timesteps) are appended to <file>; a measurement aid, off by default.
static void report_timing(const unsigned long long* dev, int S, cudaStream_t st, const char* path, int B) {
    std::vector<unsigned long long> h((size_t)S * 4);
    if (cudaStreamSynchronize(st) != cudaSuccess) return;
    if (cudaMemcpy(h.data(), dev, h.size() * 8, cudaMemcpyDeviceToHost) != cudaSuccess) return;
    double ph[4] = {0, 0, 0, 0};
    for (int s = 1; s < S; ++s) {
        ph[0] += (double)(h[(size_t)s * 4 + 0] - h[(size_t)(s - 1) * 4 + 3]);       // P0: softmax || attention
        for (int k = 1; k < 4; ++k) ph[k] += (double)(h[(size_t)s * 4 + k] - h[(size_t)s * 4 + k - 1]);
    }
    if (FILE* f = fopen(path, "a")) {
        fprintf(f, "{\"rows\": %d, \"steps\": %d, \"ns_P0_softmax_attention\": %.0f, \"ns_P1_lower_cell\": %.0f, \"ns_P2_attention_cell\": %.0f, \"ns_P3_logits_query\": %.0f}\n",
                B, S, ph[0] / (S - 1), ph[1] / (S - 1), ph[2] / (S - 1), ph[3] / (S - 1));
        fclose(f);
    }
}

int launch_page_decoder(cudaStream_t st, const PageDecArgs& a_in) {
    PageDecArgs a = a_in;
    COR_CHECK(page_decoder_supported(a.W, a.D, a.Vp, a.B), "persistent page decoder not available for this shape");
    const char* tpath = getenv("COR_PD_TIMING");
    static unsigned long long* t_buf = nullptr; static int t_cap = 0;
    if (tpath) {
        if (t_cap < a.S) { if (t_buf) cudaFree(t_buf); COR_CUDA(cudaMalloc(&t_buf, (size_t)a.S * 4 * 8)); t_cap = a.S; }
        a.timing = t_buf;
    }
    COR_CUDA(cudaMemsetAsync(a.barrier, 0, sizeof(unsigned), st));
    const int n_clusters = 4 * a.W / PD_NCOL;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(n_clusters * PD_CS); cfg.blockDim = dim3(PD_THREADS);
    cfg.dynamicSmemBytes = pagedec_smem_bytes(a.W, a.D, a.Vp); cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = PD_CS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeCooperative; at[1].val.cooperative = 1;
    // COR_PD_NOCOOP: plain cluster launch (ncu cannot replay the cooperative one); co-residency then rests on the occupancy check alone
    cfg.attrs = at; cfg.numAttrs = getenv("COR_PD_NOCOOP") ? 1 : 2;
    cudaError_t err = cudaLaunchKernelEx(&cfg, page_decoder_kernel, a);
    if (err != cudaSuccess && cfg.numAttrs == 2) {
        // cooperative + cluster launch refused.  The occupancy check does not guarantee co-residency when another kernel
        // holds SMs (side stream, other process, MPS), and the hand-rolled grid barrier would then spin into its trap:
        // tell the caller to take the per-timestep launches instead (return 2).  COR_PD_NOCOOP (above) stays a profiling
        // switch only (ncu cannot replay a cooperative launch) and is unsafe outside an otherwise idle GPU.
        (void)cudaGetLastError();
        return 2;
    }
    COR_CUDA(err);
    ++ctx().launches;
    if (tpath) report_timing(t_buf, a.S, st, tpath, a.B);
    return 0;
}

}  // namespace cor
