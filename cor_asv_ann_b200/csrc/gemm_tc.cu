// gemm_tc.cu — tcgen05 "3xFP16" GEMM for sm_100a with the fused epilogues.
//
//   C[M,N] = sum_s A_s[M,K_s] * B_s[N,K_s]^T,   A = A_hi + A_lo,  B = B_hi + B_lo  (fp16 heads + fp16 residuals)
//   acc   += A_lo*B_hi + A_hi*B_lo + A_hi*B_hi   (fp32 accumulation in TMEM; the dropped lo*lo term is <= 2^-22)
//
// fp32-grade results are a hard requirement of the path (parity 1e-4 abs on probabilities, identical
// strings), which rules out single-pass TF32/BF16/FP16.  An fp16 head carries the same 11 significand
// bits as a TF32 operand, so the three-pass product is as accurate as "3xTF32" (round 1's first
// version of this kernel) at twice the MMA rate and half the operand bytes: the L2->SMEM traffic, which
// sat at the ~6.3 KB/clk LTS cap with 4-byte operands, stays where it was while the math doubles.
// One pipeline stage holds all four tiles {A_hi, A_lo, B_hi, B_lo} of a k-block and the three MMA
// groups are issued from it (4 loads per 3 passes).
//
// Persistent, warp-specialised: warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane),
// warp 2 = TMEM allocator, warps 4..7 = epilogue (TMEM -> registers -> global).  Tile 128 x 256,
// BLOCK_K = 64 fp16 (one 128-byte swizzle atom), 2 stages x 96 KB, accumulator double-buffered in
// TMEM (2 x 256 columns), so the epilogue of tile i overlaps the main loop of tile i+1.
#include "engine.h"
#include "epilogue.cuh"
#include <cuda.h>
#include <stdlib.h>
#include <algorithm>
#include <map>
#include <mutex>
#include <tuple>

namespace cor {

namespace {

constexpr int BM = 128, BK = 64, UMMA_K = 16;
constexpr int BN_WIDE = 256, BN_NARROW = 32;   // tile widths: full tiles / page-sized M (see pick_bn)
constexpr int ELT = 2;                          // bytes per operand element (fp16)
constexpr int A_TILE = BM * BK * ELT;          // 16 KB
constexpr int NSUB_NARROW = 3;                // MMA issuer warps of the 128x32-tile kernel, each with its own accumulator
constexpr int A_BOX_NARROW = 32;               // rows per TMA box of the A operand in the 128x32-tile kernel (a multiple of the 8-row swizzle atom)
// LSTM epilogue staging (per epilogue warp): the results of a 32-row x 8-unit chunk -- c (fp32), h (hi, lo), sequence
// output (hi, lo) -- are written to shared memory as [32][8] tiles and leave through TMA box stores.  The epilogue is
// "one row per lane", so every direct global store of it touches 32 different lines (32 LSU wavefronts whatever its
// width): 160 of the 324 wavefronts per warp and chunk were stores (ncu: LSU data pipe 58 % busy in the K = 512 launches,
// the tensor pipe waiting on it); through the async proxy they cost 24 shared-memory wavefronts.
constexpr int EPI_WARPS_MAX = 8;
constexpr int EPI_STAGE_C = 0, EPI_STAGE_HHI = 1024, EPI_STAGE_HLO = 1536, EPI_STAGE_SHI = 2048, EPI_STAGE_SLO = 2560;
constexpr int EPI_STAGE_WARP = 3072;
constexpr int EPI_STAGE_BYTES = EPI_WARPS_MAX * EPI_STAGE_WARP;
// NCTA = 1: one CTA per 128 x 256 tile, 2 stages of {A_hi, A_lo, B_hi, B_lo} = 96 KB.
// NCTA = 2: a CTA pair (cta_group::2) per 256 x 256 tile; each CTA stages its own 128 rows of A and
// HALF of the B tile (128 of the 256 columns), 64 KB per stage, 3 stages.  The pair reads 128 KB from
// L2 per k-block for twice the math of the single CTA's 96 KB: the kernel sits on the L2->SM path
// (LTS cap ~6.3 KB/clk chip-wide), so this is what lifts the tensor-pipe duty.
// BN = 32 (NCTA = 1 only): the page-sized regime (one OCR-D page per call, M of a few dozen rows).  There a
// launch is bound by how fast ONE SM can pull its share of the weights out of L2, so the columns are
// spread over 8x as many CTAs (64 at N = 2048) with 5 stages of 40 KB in flight per CTA.
template <int NCTA, int BN> struct Geo {
    static constexpr int B_ROWS = BN / NCTA;
    static constexpr int B_TILE = B_ROWS * BK * ELT;
    static constexpr int STAGE_BYTES = 2 * A_TILE + 2 * B_TILE;
    static constexpr int STAGES = NCTA == 2 ? 3 : (BN == BN_WIDE ? 2 : 5);
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/ + EPI_STAGE_BYTES;
    static constexpr uint32_t IDESC = (1u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((BM * NCTA) >> 4) << 24);
};
constexpr int EPI_WARPS = 8;                   // two per TMEM lane quarter: each takes half of the tile's columns
constexpr int NTHREADS = 128 + 32 * EPI_WARPS;
constexpr int NTHREADS_NARROW = NTHREADS + 32;  // the 128x32-tile kernel has a second TMA producer warp (warp 12)
constexpr int TMEM_COLS = 512;

struct TcParams {
    CUtensorMap a_hi[kMaxSegs], a_lo[kMaxSegs], b_hi[kMaxSegs], b_lo[kMaxSegs];
    int kblocks[kMaxSegs];
    int nseg;
    int M; const int* m_dev; int N;
    int tiles_n;
    int passes;     // 3: fp32-grade hi/lo product (the product path); 1: heads only (gemm_mode 3, stated looser tolerance)
    // TMA store maps of the LSTM epilogue (box 8 x 32): cell state, h (hi, lo), sequence output (hi, lo)
    CUtensorMap e_c, e_hhi, e_hlo, e_shi, e_slo;
    int e_use;      // 0: direct stores; 1: c and h through TMA; 2: ... and the sequence output
    int e_seq_mode; // sequence output layout: 0 time-major (row' = t * rows_total + row), 1 line-major (column' = t * seq_ld + column)
};

// ------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
// Bounded wait: a protocol bug must trap (launch error) rather than hang the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait(bar, parity)) {
        if (clock64() - t0 > 4000000000LL) __trap();
    }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
// pair variant: executed by both CTAs, the bytes are accounted on the LEADER's barrier (`bar` is a
// shared::cluster address obtained with mapa)
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, uint32_t src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(map), "r"(src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_f16_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrives on the barrier at this smem offset in BOTH CTAs of the pair when the pair's MMAs retire
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, 128-byte swizzle: rows of 128 B, 8-row groups 1024 B apart (SBO), LBO unused (=1),
// descriptor version 1 (sm_100), layout type 2 = SWIZZLE_128B.   (cute/arch/mma_sm100_desc.hpp)
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// kind::f16, D = f32 (c_format 1 at [4,6)), A/B = f16 (format 0 at [7,10) / [10,13)), both K-major,
// N at [17,23) >> 3, M at [24,29) >> 4
// (Geo<NCTA, BN>::IDESC; M = 256 for the pair)

// 256-bit global accesses (sm_100): the epilogue is "one row per lane", so every warp-wide access touches 32
// different lines and costs 32 L1 wavefronts whatever its width -- the LSU pipe, not the MUFU or the tensor
// pipe, bounded the K = 512 launches (ncu: 129 us against 66 us of MMA time).  32-byte accesses halve the
// instruction count of the 16-byte version.
__device__ __forceinline__ void ld8(const float* p, float* v) {
    asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]) : "l"(p));
}
__device__ __forceinline__ void ld8_nc(const float* p, float* v) {      // read-only data (bias, tables)
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]) : "l"(p));
}
__device__ __forceinline__ void st8(float* p, const float* v) {
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 ::"l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]) : "memory");
}

// ------------------------------------------------------------------ vectorised LSTM epilogue (8 units of one row)
__device__ __forceinline__ void st4(float* p, size_t off, const float* v) {
    *reinterpret_cast<float4*>(p + off) = make_float4(v[0], v[1], v[2], v[3]);
}
__device__ __forceinline__ void ld4(const float* p, size_t off, float* v) {
    const float4 t = *reinterpret_cast<const float4*>(p + off);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
__device__ __forceinline__ void store8_split(float* hi_p, float* lo_p, size_t off, const float* v) {
    if (lo_p) {
        __align__(16) __half hi[8], lo[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) f16_split(v[k], hi[k], lo[k]);
        *reinterpret_cast<uint4*>(reinterpret_cast<__half*>(hi_p) + off) = *reinterpret_cast<const uint4*>(hi);
        *reinterpret_cast<uint4*>(reinterpret_cast<__half*>(lo_p) + off) = *reinterpret_cast<const uint4*>(lo);
    } else if ((((uintptr_t)(hi_p + off)) & 31) == 0) {
        st8(hi_p + off, v);
    } else {
        st4(hi_p, off, v); st4(hi_p, off + 4, v + 4);
    }
}
__device__ __forceinline__ void load8_split(const float* hi_p, const float* lo_p, size_t off, float* v) {
    if (lo_p) {
        const float4 a = load_split4(hi_p, lo_p, off), b = load_split4(hi_p, lo_p, off + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
        ld4(hi_p, off, v); ld4(hi_p, off + 4, v + 4);
    }
}

// Per-row constants of a tile's LSTM epilogue, loaded BEFORE the wait on the accumulator: the padded length,
// the sequence position and the first table index of the row (three dependent global loads that used to sit
// inside every 32-column chunk).
struct RowCtx {
    int T; bool active; int t;
    int idx0; float w0;        // first alternative of the row's input position (xtab); idx0 < 0: none
    size_t xo;                 // offset of the row's alternatives in x_idx / x_val
};
__device__ __forceinline__ RowCtx lstm_row_begin(const LstmEpi& e, int row, bool valid) {
    RowCtx rc;
    rc.T = (valid && e.row_T) ? e.row_T[row] : 0x7fffffff;
    rc.active = e.step < rc.T;
    rc.t = e.reverse ? rc.T - 1 - e.step : e.step;
    rc.idx0 = -1; rc.w0 = 1.0f; rc.xo = 0;
    if (valid && rc.active && e.xtab) {
        rc.xo = ((size_t)row * e.x_tmax + rc.t) * e.x_k;
        rc.idx0 = e.x_idx[rc.xo];
        if (e.x_val) rc.w0 = e.x_val[rc.xo];
    }
    return rc;
}
// Operands of one chunk that do not depend on the accumulator: b = bias (+ the table rows of the input term),
// c = the cell state.  Issued before the TMEM load so that their latency overlaps it.
__device__ __forceinline__ void lstm_chunk_operands(const LstmEpi& e, const RowCtx& rc, int row, int unit0, float (&b)[32], float (&c)[8]) {
    ld8(e.c_in + (size_t)row * e.W + unit0, c);
#pragma unroll
    for (int q = 0; q < 4; ++q) ld8_nc(e.bias + (size_t)4 * unit0 + 8 * q, b + 8 * q);
    if (e.xtab) {
        if (rc.idx0 >= 0) {
            const float* src = e.xtab + (size_t)rc.idx0 * 4 * e.W + 4 * unit0;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                float x[8];
                ld8_nc(src + 8 * q, x);
#pragma unroll
                for (int k = 0; k < 8; ++k) b[8 * q + k] = fmaf(rc.w0, x[k], b[8 * q + k]);
            }
        }
        for (int a = 1; a < e.x_k; ++a) {           // further alternatives (confusion networks): rare
            const int v = e.x_idx[rc.xo + a];
            if (v < 0) continue;
            const float w = e.x_val ? e.x_val[rc.xo + a] : 1.0f;
            const float* src = e.xtab + (size_t)v * 4 * e.W + 4 * unit0;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                float x[8];
                ld8_nc(src + 8 * q, x);
#pragma unroll
                for (int k = 0; k < 8; ++k) b[8 * q + k] = fmaf(w, x[k], b[8 * q + k]);
            }
        }
    }
}

// z = 32 accumulator columns = units [unit0, unit0+8) x gates (i,f,c~,o) of one row
__device__ __forceinline__ void lstm_row8(const LstmEpi& e, const RowCtx& rc, int row, int unit0, const uint32_t (&z)[32],
                                          const float (&b)[32], float (&c)[8]) {
    const int T = rc.T;
    const bool active = rc.active;
    const size_t so = (size_t)row * e.W + unit0;
    float h[8];
    if (active) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float i = sigmoidf_fast(__uint_as_float(z[4 * k + 0]) + b[4 * k + 0]);
            const float f = sigmoidf_fast(__uint_as_float(z[4 * k + 1]) + b[4 * k + 1]);
            const float g = tanhf_fast(__uint_as_float(z[4 * k + 2]) + b[4 * k + 2]);
            const float o = sigmoidf_fast(__uint_as_float(z[4 * k + 3]) + b[4 * k + 3]);
            c[k] = f * c[k] + i * g;
            h[k] = o * tanhf_fast(c[k]);
        }
        st8(e.c_out + so, c);
    } else {
        load8_split(e.h_prev, e.h_prev_lo, so, h);
        if (e.c_out != e.c_in) {
            float cc[8];
            ld8(e.c_in + so, cc);
            st8(e.c_out + so, cc);
        }
    }
    store8_split(e.h_out, e.h_out_lo, so, h);
    if (!active || (!e.seq_out && !e.seq_rev)) return;
    const int t = rc.t;
    float y[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) y[k] = h[k];
    if (e.cross_sum) {
#pragma unroll
        for (int k = 0; k < 8; k += 2) { const float s = h[k] + h[k + 1]; y[k] = s; y[k + 1] = s; }
    }
    if (e.seq_out) {
        float v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = y[k];
        if (e.resid) {
            float r[8];
            load8_split(e.resid, e.resid_lo, ((size_t)t * e.rows_total + row) * e.resid_ld + unit0, r);
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] += r[k];
        }
        store8_split(e.seq_out, e.seq_out_lo, seq_pos(e, t, row) * e.seq_ld + e.seq_col + unit0, v);
    }
    if (e.seq_rev)
        store8_split(e.seq_rev, e.seq_rev_lo, seq_pos(e, T - 1 - t, row) * e.seq_ld + e.seq_col + unit0, y);
}

// The same cell for a warp whose 32 rows are all valid, active and at the same sequence position: results go to the warp's
// staging tiles (row = lane), not to global memory; lstm_stage_flush sends them.
__device__ __forceinline__ void lstm_row8_stage(const LstmEpi& e, const RowCtx& rc, int row, int unit0, const uint32_t (&z)[32],
                                                const float (&b)[32], float (&c)[8], uint8_t* stage, int lane, bool with_seq) {
    float h[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float i = sigmoidf_fast(__uint_as_float(z[4 * k + 0]) + b[4 * k + 0]);
        const float f = sigmoidf_fast(__uint_as_float(z[4 * k + 1]) + b[4 * k + 1]);
        const float g = tanhf_fast(__uint_as_float(z[4 * k + 2]) + b[4 * k + 2]);
        const float o = sigmoidf_fast(__uint_as_float(z[4 * k + 3]) + b[4 * k + 3]);
        c[k] = f * c[k] + i * g;
        h[k] = o * tanhf_fast(c[k]);
    }
    float4* sc = reinterpret_cast<float4*>(stage + EPI_STAGE_C + lane * 32);
    sc[0] = make_float4(c[0], c[1], c[2], c[3]); sc[1] = make_float4(c[4], c[5], c[6], c[7]);
    {
        __align__(16) __half hi[8], lo[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) f16_split(h[k], hi[k], lo[k]);
        *reinterpret_cast<uint4*>(stage + EPI_STAGE_HHI + lane * 16) = *reinterpret_cast<const uint4*>(hi);
        *reinterpret_cast<uint4*>(stage + EPI_STAGE_HLO + lane * 16) = *reinterpret_cast<const uint4*>(lo);
    }
    if (!with_seq) return;
    float y[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) y[k] = h[k];
    if (e.cross_sum) {
#pragma unroll
        for (int k = 0; k < 8; k += 2) { const float s = h[k] + h[k + 1]; y[k] = s; y[k + 1] = s; }
    }
    if (e.resid) {
        float r[8];
        load8_split(e.resid, e.resid_lo, ((size_t)rc.t * e.rows_total + row) * e.resid_ld + unit0, r);
#pragma unroll
        for (int k = 0; k < 8; ++k) y[k] += r[k];
    }
    __align__(16) __half hi[8], lo[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) f16_split(y[k], hi[k], lo[k]);
    *reinterpret_cast<uint4*>(stage + EPI_STAGE_SHI + lane * 16) = *reinterpret_cast<const uint4*>(hi);
    *reinterpret_cast<uint4*>(stage + EPI_STAGE_SLO + lane * 16) = *reinterpret_cast<const uint4*>(lo);
}

__device__ __forceinline__ void store_row32(const StoreEpi& e, int row, int col0, int N, const uint32_t (&z)[32]) {
    if (e.n_split && col0 >= e.n_split && col0 + 32 <= N && (e.ldc2 & 3) == 0) {
        const int c2 = col0 - e.n_split;
        const bool wide = (e.ldc2 & 7) == 0 && ((uintptr_t)e.C2 & 31) == 0;     // 32-byte stores (see ld8/st8)
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = __uint_as_float(z[8 * q + k]) + (e.bias2 ? e.bias2[c2 + 8 * q + k] : 0.0f);
            float* dst = e.C2 + (size_t)row * e.ldc2 + c2 + 8 * q;
            if (wide) st8(dst, v); else { st4(dst, 0, v); st4(dst, 4, v + 4); }
        }
    } else if (col0 + 32 <= N && (e.ldc & 3) == 0 && !(e.n_split && col0 >= e.n_split)) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                v[k] = __uint_as_float(z[8 * q + k]) + (e.bias ? e.bias[col0 + 8 * q + k] : 0.0f);
                if (e.act == 1) v[k] = tanhf(v[k]);
            }
            store8_split(e.C, e.C_lo, (size_t)row * e.ldc + col0 + 8 * q, v);
        }
    } else {
#pragma unroll
        for (int k = 0; k < 32; ++k)
            if (col0 + k < N) store_elem(e, row, col0 + k, __uint_as_float(z[k]));
    }
}

// ------------------------------------------------------------------ the kernel
template <int EPI, int NCTA, int BN, typename Epi>
__global__ void __launch_bounds__(NTHREADS_NARROW, 1) gemm_tc_kernel(const __grid_constant__ TcParams p, const Epi epi) {
    using G = Geo<NCTA, BN>;
    constexpr int STAGES = G::STAGES, STAGE_BYTES = G::STAGE_BYTES, B_TILE = G::B_TILE;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;       // SWIZZLE_128B wants 1024-byte alignment
    const uint32_t bar_base = smem_base + STAGES * STAGE_BYTES;
    // barriers (8 bytes each): full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2]; then the TMEM base slot.
    // Pair: full[] and tmem_empty[] are used in the leader CTA only (both CTAs' loads and both CTAs'
    // epilogues signal there); empty[] and tmem_full[] exist in both and are hit by multicast commits.
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + 2 + a); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
    uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + STAGES * STAGE_BYTES + 8 * (2 * STAGES + 4));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = NCTA == 2 ? cluster_rank() : 0u;
    const bool leader = rank == 0;
    const int unit = blockIdx.x / NCTA, n_units = gridDim.x / NCTA;     // a unit = one CTA or one CTA pair
    const int M = p.m_dev ? min(*p.m_dev, p.M) : p.M;
    const int tiles_m = (M + BM * NCTA - 1) / (BM * NCTA);
    const int num_tiles = tiles_m * p.tiles_n;
    int kb_total = 0;
    for (int s = 0; s < p.nseg; ++s) kb_total += p.kblocks[s];

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < p.nseg; ++s) {
            tma_prefetch_desc(&p.a_hi[s]); tma_prefetch_desc(&p.a_lo[s]);
            tma_prefetch_desc(&p.b_hi[s]); tma_prefetch_desc(&p.b_lo[s]);
        }
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), BN == BN_NARROW ? 2 : 1); mbar_init(empty_bar(s), 1); }   // 128x32 tiles: two producers
        for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), BN == BN_NARROW ? NSUB_NARROW : 1); mbar_init(tempty_bar(a), EPI_WARPS * NCTA); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        if constexpr (NCTA == 1) {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(TMEM_COLS) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(TMEM_COLS) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        }
    }
    tc_fence_before();
    if constexpr (NCTA == 2) cluster_sync_all(); else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;

    if (warp == 0) {
        // ===== TMA producer (one per CTA) =====
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int tile = unit; tile < num_tiles; tile += n_units) {
                const int m0 = (tile / p.tiles_n) * (BM * NCTA) + (int)rank * BM;
                const int n0 = (tile % p.tiles_n) * BN + (int)rank * G::B_ROWS;
                for (int s = 0; s < p.nseg; ++s) {
                    for (int kb = 0; kb < p.kblocks[s]; ++kb) {
                        mbar_wait(empty_bar(stage), phase ^ 1);
                        const uint32_t sa = smem_base + stage * STAGE_BYTES;
                        const bool lo = p.passes == 3;          // heads-only mode leaves the residual tiles out
                        const uint32_t bytes = lo ? STAGE_BYTES : A_TILE + B_TILE;
                        if constexpr (NCTA == 1 && BN == BN_NARROW) {
                            // Page regime: a tile has a few dozen valid rows.  The A operand comes in boxes of A_BOX_NARROW
                            // rows, only as many as hold valid rows (a 128-row box moves 16 KB per operand part and k-block
                            // whatever M is -- zero fill included -- and the k-loop of this kernel runs at the speed one SM
                            // takes its stage in: 40 KB per k-block became 16 KB for a 32-line page).  The other rows of the
                            // tile keep whatever shared memory held: they feed accumulator rows >= M, which nothing reads.
                            // This thread loads A only: a TMA instruction costs its issuing thread ~85 cycles like an MMA does,
                            // and with three MMA issuers the four loads of a k-block were the kernel's bound (366 ns per
                            // k-block).  Warp 12 loads B (below); the full barrier counts both producers.
                            const int nbox = (min(BM, M - m0) + A_BOX_NARROW - 1) / A_BOX_NARROW;
                            constexpr uint32_t BOX_BYTES = A_BOX_NARROW * BK * ELT;
                            mbar_expect_tx(full_bar(stage), (lo ? 2u : 1u) * (nbox * BOX_BYTES));
                            for (int r = 0; r < nbox; ++r) {
                                tma_load_2d(sa + r * BOX_BYTES, &p.a_hi[s], kb * BK, m0 + r * A_BOX_NARROW, full_bar(stage));
                                if (lo) tma_load_2d(sa + A_TILE + r * BOX_BYTES, &p.a_lo[s], kb * BK, m0 + r * A_BOX_NARROW, full_bar(stage));
                            }
                        } else if constexpr (NCTA == 1) {
                            mbar_expect_tx(full_bar(stage), bytes);
                            tma_load_2d(sa, &p.a_hi[s], kb * BK, m0, full_bar(stage));
                            if (lo) tma_load_2d(sa + A_TILE, &p.a_lo[s], kb * BK, m0, full_bar(stage));
                            tma_load_2d(sa + 2 * A_TILE, &p.b_hi[s], kb * BK, n0, full_bar(stage));
                            if (lo) tma_load_2d(sa + 2 * A_TILE + B_TILE, &p.b_lo[s], kb * BK, n0, full_bar(stage));
                        } else {
                            if (leader) mbar_expect_tx(full_bar(stage), 2 * bytes);    // both CTAs' bytes
                            const uint32_t fb = map_to_cta(full_bar(stage), 0);
                            tma_load_2d_pair(sa, &p.a_hi[s], kb * BK, m0, fb);
                            if (lo) tma_load_2d_pair(sa + A_TILE, &p.a_lo[s], kb * BK, m0, fb);
                            tma_load_2d_pair(sa + 2 * A_TILE, &p.b_hi[s], kb * BK, n0, fb);
                            if (lo) tma_load_2d_pair(sa + 2 * A_TILE + B_TILE, &p.b_lo[s], kb * BK, n0, fb);
                        }
                        if (++stage == STAGES) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else if (BN == BN_NARROW && warp >= 1 && warp <= NSUB_NARROW) {
        // ===== MMA issuers of the page regime (BN = 32) =====
        // One thread gets a tcgen05.mma out every ~85 cycles whatever its shape: the 288 small MMAs of a tile at K = 1536
        // were 13 of the kernel's 17 us.  Warps 1..3 therefore share the k-blocks round robin (k-block kb belongs to issuer
        // kb % 3); MMAs of different threads are not ordered against each other, so each issuer owns an accumulator, and the
        // epilogue adds the three up in a fixed order.  Another summation order than the wide tiles' (still fp32 sums of the
        // same products): page-sized and batch-sized calls agree to rounding, not bit for bit.
        if constexpr (BN == BN_NARROW) {
            const int qi = warp - 1;
            if (lane == 0) {
                long long g0 = 0;                       // k-blocks of the tiles before this one (stage and phase follow from the count)
                int acc = 0; uint32_t acc_phase = 0;
                for (int tile = unit; tile < num_tiles; tile += n_units) {
                    mbar_wait(tempty_bar(acc), acc_phase ^ 1);      // epilogue(s) have drained this accumulator
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + acc * (NSUB_NARROW * BN) + qi * BN;
                    bool first = true;
                    for (int kb = qi; kb < kb_total; kb += NSUB_NARROW) {
                        const long long g = g0 + kb;
                        const int stage = (int)(g % STAGES);
                        mbar_wait(full_bar(stage), (uint32_t)((g / STAGES) & 1));
                        tc_fence_after();
                        const uint32_t sa = smem_base + stage * STAGE_BYTES;
                        const uint64_t a_hi = make_smem_desc(sa), a_lo = make_smem_desc(sa + A_TILE);
                        const uint64_t b_hi = make_smem_desc(sa + 2 * A_TILE), b_lo = make_smem_desc(sa + 2 * A_TILE + B_TILE);
                        const bool lo = p.passes == 3;
                        if (lo) {
#pragma unroll
                            for (int k = 0; k < BK / UMMA_K; ++k) {
                                const uint64_t adv = (uint64_t)((k * UMMA_K * ELT) >> 4);
                                umma_f16(d_tmem, a_lo + adv, b_hi + adv, G::IDESC, !(first && k == 0));
                            }
#pragma unroll
                            for (int k = 0; k < BK / UMMA_K; ++k) {
                                const uint64_t adv = (uint64_t)((k * UMMA_K * ELT) >> 4);
                                umma_f16(d_tmem, a_hi + adv, b_lo + adv, G::IDESC, 1);
                            }
                        }
#pragma unroll
                        for (int k = 0; k < BK / UMMA_K; ++k) {
                            const uint64_t adv = (uint64_t)((k * UMMA_K * ELT) >> 4);
                            umma_f16(d_tmem, a_hi + adv, b_hi + adv, G::IDESC, lo || !(first && k == 0));
                        }
                        first = false;
                        umma_commit(empty_bar(stage));      // frees the smem stage when this issuer's MMAs retire
                    }
                    g0 += kb_total;
                    umma_commit(tfull_bar(acc));            // this issuer's part of the accumulator is complete (NSUB_NARROW arrivals)
                    if (++acc == 2) { acc = 0; acc_phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (the leader CTA's, for the pair) =====
        if (lane == 0 && leader) {
            int stage = 0; uint32_t phase = 0;
            int acc = 0; uint32_t acc_phase = 0;
            for (int tile = unit; tile < num_tiles; tile += n_units) {
                mbar_wait(tempty_bar(acc), acc_phase ^ 1);      // epilogue(s) have drained this accumulator
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = 0; kb < kb_total; ++kb) {
                    mbar_wait(full_bar(stage), phase);
                    tc_fence_after();
                    const uint32_t sa = smem_base + stage * STAGE_BYTES;
                    const uint64_t a_hi = make_smem_desc(sa), a_lo = make_smem_desc(sa + A_TILE);
                    const uint64_t b_hi = make_smem_desc(sa + 2 * A_TILE), b_lo = make_smem_desc(sa + 2 * A_TILE + B_TILE);
                    auto mma = [&](uint64_t a, uint64_t b, uint32_t accum) {
                        if constexpr (NCTA == 1) umma_f16(d_tmem, a, b, G::IDESC, accum);
                        else umma_f16_pair(d_tmem, a, b, G::IDESC, accum);
                    };
                    // small terms first: A_lo*B_hi, A_hi*B_lo, then A_hi*B_hi
                    const bool lo = p.passes == 3;
                    if (lo) {
#pragma unroll
                        for (int k = 0; k < BK / UMMA_K; ++k) {
                            const uint64_t adv = (uint64_t)((k * UMMA_K * ELT) >> 4);
                            mma(a_lo + adv, b_hi + adv, (kb | k) != 0);
                        }
#pragma unroll
                        for (int k = 0; k < BK / UMMA_K; ++k) {
                            const uint64_t adv = (uint64_t)((k * UMMA_K * ELT) >> 4);
                            mma(a_hi + adv, b_lo + adv, 1);
                        }
                    }
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) {
                        const uint64_t adv = (uint64_t)((k * UMMA_K * ELT) >> 4);
                        mma(a_hi + adv, b_hi + adv, lo || (kb | k) != 0);
                    }
                    // frees the smem stage (in both CTAs) when these MMAs retire
                    if constexpr (NCTA == 1) umma_commit(empty_bar(stage)); else umma_commit_pair(empty_bar(stage));
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                if constexpr (NCTA == 1) umma_commit(tfull_bar(acc)); else umma_commit_pair(tfull_bar(acc));   // accumulator complete
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (BN == BN_NARROW && warp == 12) {
        // ===== second TMA producer of the page regime: the B (weight) tiles =====
        if constexpr (BN == BN_NARROW) {
            if (lane == 0) {
                int stage = 0; uint32_t phase = 0;
                for (int tile = unit; tile < num_tiles; tile += n_units) {
                    const int n0 = (tile % p.tiles_n) * BN;
                    for (int s = 0; s < p.nseg; ++s) {
                        for (int kb = 0; kb < p.kblocks[s]; ++kb) {
                            mbar_wait(empty_bar(stage), phase ^ 1);
                            const uint32_t sa = smem_base + stage * STAGE_BYTES;
                            const bool lo = p.passes == 3;
                            mbar_expect_tx(full_bar(stage), (lo ? 2u : 1u) * B_TILE);
                            tma_load_2d(sa + 2 * A_TILE, &p.b_hi[s], kb * BK, n0, full_bar(stage));
                            if (lo) tma_load_2d(sa + 2 * A_TILE + B_TILE, &p.b_lo[s], kb * BK, n0, full_bar(stage));
                            if (++stage == STAGES) { stage = 0; phase ^= 1; }
                        }
                    }
                }
            }
        }
    } else if (warp >= 4 && warp < 4 + EPI_WARPS) {
        // ===== epilogue: warp w may touch TMEM lanes 32*(w%4) .. +31 of its own CTA =====
        // the LSTM epilogue is a chain of MUFU and load latencies per row: two warps per scheduler hide them
        const int wq = (warp - 4) & 3, half = (warp - 4) >> 2;
        int acc = 0; uint32_t acc_phase = 0;
        for (int tile = unit; tile < num_tiles; tile += n_units) {
            const int m0 = (tile / p.tiles_n) * (BM * NCTA) + (int)rank * BM, n0 = (tile % p.tiles_n) * BN;
            const int row = m0 + wq * 32 + lane;
            constexpr int ACC_STRIDE = BN == BN_NARROW ? NSUB_NARROW * BN : BN;
            const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16) + acc * ACC_STRIDE;
            // accumulator chunk `ch` as 32 fp32 bit patterns; the 128x32-tile kernel adds up its issuers' accumulators in a fixed
            // order (an issuer without a k-block -- fewer k-blocks than issuers -- has written nothing)
            auto load_acc = [&](int ch, uint32_t (&z)[32]) {
                tmem_ld32(taddr + ch * 32, z);
                tmem_ld_wait();
                if constexpr (BN == BN_NARROW) {
                    const int nsub = min(NSUB_NARROW, kb_total);
#pragma unroll 1
                    for (int q = 1; q < nsub; ++q) {
                        uint32_t y[32];
                        tmem_ld32(taddr + q * BN + ch * 32, y);
                        tmem_ld_wait();
#pragma unroll
                        for (int k = 0; k < 32; ++k) z[k] = __float_as_uint(__uint_as_float(z[k]) + __uint_as_float(y[k]));
                    }
                }
            };
            if constexpr (EPI == 1) {
                // everything that does not depend on the accumulator is fetched before waiting for it
                const RowCtx rc = lstm_row_begin(epi, row, row < M);
                // TMA-staged stores need the warp's 32 rows valid, active and at one sequence position (box store)
                const int t0 = __shfl_sync(0xffffffffu, rc.t, 0);
                const bool staged = p.e_use != 0 && __all_sync(0xffffffffu, row < M && rc.active && rc.t == t0);
                const bool with_seq = p.e_use == 2;
                uint8_t* stage = smem_gen + STAGES * STAGE_BYTES + 256 + (warp - 4) * EPI_STAGE_WARP;
                const uint32_t stage_s = bar_base + 256u + (uint32_t)(warp - 4) * EPI_STAGE_WARP;
                const int row0 = m0 + wq * 32;
                bool waited = false;
#pragma unroll 1
                for (int ch = half; ch < BN / 32; ch += EPI_WARPS / 4) {
                    const int col0 = n0 + ch * 32;
                    const bool live = row < M && col0 < p.N;
                    float b[32], c[8];
                    if (live && rc.active) lstm_chunk_operands(epi, rc, row, col0 >> 2, b, c);
                    if (!waited) { mbar_wait(tfull_bar(acc), acc_phase); tc_fence_after(); waited = true; }
                    uint32_t z[32];
                    load_acc(ch, z);
                    if (staged && col0 < p.N) {
                        const int unit0 = col0 >> 2;
                        if (lane == 0) bulk_wait_read();            // the previous chunk's tiles have been read out
                        __syncwarp();
                        lstm_row8_stage(epi, rc, row, unit0, z, b, c, stage, lane, with_seq);
                        fence_async_smem();
                        __syncwarp();
                        if (lane == 0) {
                            tma_store_2d(&p.e_c, stage_s + EPI_STAGE_C, unit0, row0);
                            tma_store_2d(&p.e_hhi, stage_s + EPI_STAGE_HHI, unit0, row0);
                            tma_store_2d(&p.e_hlo, stage_s + EPI_STAGE_HLO, unit0, row0);
                            if (with_seq) {
                                const int sc0 = p.e_seq_mode ? t0 * epi.seq_ld + epi.seq_col + unit0 : epi.seq_col + unit0;
                                const int sc1 = p.e_seq_mode ? row0 : t0 * epi.rows_total + row0;
                                tma_store_2d(&p.e_shi, stage_s + EPI_STAGE_SHI, sc0, sc1);
                                tma_store_2d(&p.e_slo, stage_s + EPI_STAGE_SLO, sc0, sc1);
                            }
                            bulk_commit();
                        }
                    } else if (live) {
                        lstm_row8(epi, rc, row, col0 >> 2, z, b, c);
                    }
                }
                if (!waited) { mbar_wait(tfull_bar(acc), acc_phase); tc_fence_after(); }
            } else {
                mbar_wait(tfull_bar(acc), acc_phase);
                tc_fence_after();
#pragma unroll 1
                for (int ch = half; ch < BN / 32; ch += EPI_WARPS / 4) {
                    uint32_t z[32];
                    load_acc(ch, z);
                    const int col0 = n0 + ch * 32;
                    if (row < M && col0 < p.N) store_row32(epi, row, col0, p.N, z);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {        // one arrival per epilogue warp, on the (leader's) barrier the MMA issuer waits on
                if constexpr (NCTA == 1) mbar_arrive(tempty_bar(acc));
                else mbar_arrive_cluster(map_to_cta(tempty_bar(acc), 0));
            }
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if constexpr (EPI == 1) { if (lane == 0) bulk_wait_all(); }      // staged stores have landed before the CTA retires
    }
    tc_fence_before();
    // pair: neither CTA may leave (or free TMEM) while its peer can still read its smem or signal its barriers
    if constexpr (NCTA == 2) cluster_sync_all(); else __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        if constexpr (NCTA == 1)
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
        else
            asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
    }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;
int g_tc_state = -1;        // -1 unknown, 0 unavailable, 1 ok
int g_num_sms = 148;
int g_max_pairs = 0;        // resident CTA pairs (0: pair kernel unusable)
std::mutex g_mu;
// tensor maps are cached per handle (Ctx::tc_maps), keyed by pointer and geometry; a handle is used by one thread at a time
using MapCache = std::map<std::tuple<const void*, long long, long long, long long, int>, CUtensorMap>;
MapCache& maps_of(Ctx& c) {
    if (!c.tc_maps) c.tc_maps = new MapCache();
    return *static_cast<MapCache*>(c.tc_maps);
}

bool init_tc() {
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_tc_state >= 0) return g_tc_state == 1;
    g_tc_state = 0;
    int dev = 0;
    cudaDeviceProp prop{};
    if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess) return false;
    if (prop.major != 10) return false;
    g_num_sms = prop.multiProcessorCount;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn) return false;
    g_encode = (EncodeTiledFn)fn;
    auto opt_in = [](auto kern, int bytes) { return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) == cudaSuccess; };
    if (!opt_in(gemm_tc_kernel<0, 1, BN_WIDE, StoreEpi>, Geo<1, BN_WIDE>::SMEM_BYTES) || !opt_in(gemm_tc_kernel<1, 1, BN_WIDE, LstmEpi>, Geo<1, BN_WIDE>::SMEM_BYTES) ||
        !opt_in(gemm_tc_kernel<0, 1, BN_NARROW, StoreEpi>, Geo<1, BN_NARROW>::SMEM_BYTES) || !opt_in(gemm_tc_kernel<1, 1, BN_NARROW, LstmEpi>, Geo<1, BN_NARROW>::SMEM_BYTES) ||
        !opt_in(gemm_tc_kernel<0, 2, BN_WIDE, StoreEpi>, Geo<2, BN_WIDE>::SMEM_BYTES) || !opt_in(gemm_tc_kernel<1, 2, BN_WIDE, LstmEpi>, Geo<2, BN_WIDE>::SMEM_BYTES)) return false;
    {   // how many CTA pairs can be resident at once (a GPC with an odd SM count leaves one SM unpaired):
        // a persistent grid must not exceed it, or the surplus pairs would run as a second wave
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(g_num_sms & ~1); cfg.blockDim = dim3(NTHREADS); cfg.dynamicSmemBytes = Geo<2, BN_WIDE>::SMEM_BYTES;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        int nc = 0;
        if (cudaOccupancyMaxActiveClusters(&nc, gemm_tc_kernel<1, 2, BN_WIDE, LstmEpi>, &cfg) == cudaSuccess && nc > 0) g_max_pairs = nc;
        else { g_max_pairs = 0; (void)cudaGetLastError(); }
        if (getenv("COR_NO_PAIR")) g_max_pairs = 0;     // debug switch: single-CTA kernel everywhere
    }
    g_tc_state = 1;
    return true;
}

// 2-D fp16 row-major [rows][cols] with leading dimension ld (elements); box = 64 columns x box_rows
int get_map(const float* ptr, long long rows, long long cols, long long ld, int box_rows, CUtensorMap* out) {
    auto key = std::make_tuple((const void*)ptr, rows, cols, ld, box_rows);
    MapCache& maps = maps_of(ctx());
    {
        auto it = maps.find(key);
        if (it != maps.end()) { *out = it->second; return 0; }
    }
    COR_CHECK(((uintptr_t)ptr & 15) == 0 && (ld * ELT) % 16 == 0, "TMA operand must be 16-byte aligned with a 16-byte multiple pitch");
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * ELT};
    cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (void*)ptr, dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    COR_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed");
    if (maps.size() > 4096) maps.clear();
    maps[key] = *out;
    return 0;
}

// 2-D row-major [rows][cols] of fp32 or fp16 with leading dimension ld (elements), plain (unswizzled) box: the store maps
// of the LSTM epilogue.  Cached like the operand maps (the key carries the element size and the box).
int get_store_map(const void* ptr, int elt, long long rows, long long cols, long long ld, int box_cols, int box_rows, CUtensorMap* out) {
    auto key = std::make_tuple(ptr, rows, cols, ld, -(elt * 1000000 + box_cols * 1000 + box_rows));
    MapCache& maps = maps_of(ctx());
    {
        auto it = maps.find(key);
        if (it != maps.end()) { *out = it->second; return 0; }
    }
    COR_CHECK(((uintptr_t)ptr & 15) == 0 && (ld * elt) % 16 == 0, "TMA store target must be 16-byte aligned with a 16-byte multiple pitch");
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * elt};
    cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(out, elt == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    COR_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed for an epilogue store map");
    if (maps.size() > 4096) maps.clear();
    maps[key] = *out;
    return 0;
}

// Store maps of the LSTM epilogue, when its outputs are the split (hi, lo) kind the box stores can serve.
int fill_epi_maps(const LstmEpi& e, int M, TcParams* p) {
    p->e_use = 0; p->e_seq_mode = 0;
    if (!e.h_out_lo || !e.h_out || !e.c_out || e.seq_rev || (e.W & 7) || getenv("COR_NO_EPI_TMA")) return 0;
    if (e.seq_out && (!e.seq_out_lo || (e.seq_ld & 7) || ((e.seq_col) & 7) || e.seq_T <= 0)) return 0;
    if (get_store_map(e.c_out, 4, M, e.W, e.W, 8, 32, &p->e_c) || get_store_map(e.h_out, 2, M, e.W, e.W, 8, 32, &p->e_hhi) ||
        get_store_map(e.h_out_lo, 2, M, e.W, e.W, 8, 32, &p->e_hlo)) return 1;
    p->e_use = 1;
    if (e.seq_out) {
        const bool line_major = e.seq_tm != 0 && !(e.seq_tm == e.rows_total && e.seq_rm == 1);
        if (line_major) {
            // element (t, row, col) at (t * seq_tm + row * seq_rm) * seq_ld + col with seq_tm == 1: a row of the map is a line
            if (e.seq_tm != 1) { p->e_use = 0; return 0; }
            const long long ld = (long long)e.seq_rm * e.seq_ld;
            if (get_store_map(e.seq_out, 2, e.rows_total, ld, ld, 8, 32, &p->e_shi) || get_store_map(e.seq_out_lo, 2, e.rows_total, ld, ld, 8, 32, &p->e_slo)) return 1;
            p->e_seq_mode = 1;
        } else {
            const long long rows = (long long)e.seq_T * e.rows_total;
            if (get_store_map(e.seq_out, 2, rows, e.seq_ld, e.seq_ld, 8, 32, &p->e_shi) || get_store_map(e.seq_out_lo, 2, rows, e.seq_ld, e.seq_ld, 8, 32, &p->e_slo)) return 1;
        }
        p->e_use = 2;
    }
    return 0;
}

int fill_params(int ncta, int bn, int M, const int* m_dev, int N, const GemmSeg* segs, int nseg, int passes, TcParams* p) {
    COR_CHECK(nseg >= 1 && nseg <= kMaxSegs, "bad segment count");
    COR_CHECK(passes == 1 || passes == 3, "passes must be 1 or 3");
    p->passes = passes; p->nseg = nseg; p->M = M; p->m_dev = m_dev; p->N = N;
    p->tiles_n = (N + bn - 1) / bn;
    for (int s = 0; s < nseg; ++s) {
        const GemmSeg& g = segs[s];
        COR_CHECK(g.A_lo && g.B_lo, "tcgen05 path needs split (hi, lo) operands");
        COR_CHECK(g.K > 0 && g.K % BK == 0, "K must be a positive multiple of 64");
        p->kblocks[s] = g.K / BK;
        const int a_box = (ncta == 1 && bn == BN_NARROW) ? A_BOX_NARROW : BM;
        if (get_map(g.A, M, g.K, g.lda, a_box, &p->a_hi[s]) || get_map(g.A_lo, M, g.K, g.lda, a_box, &p->a_lo[s]) ||
            get_map(g.B, N, g.K, g.ldb, bn / ncta, &p->b_hi[s]) || get_map(g.B_lo, N, g.K, g.ldb, bn / ncta, &p->b_lo[s])) return 1;
    }
    return 0;
}

// Tile shape.  The pair kernel pays off once there are enough 256-row tiles to fill the resident pairs;
// below that the single-CTA kernel has the finer granularity (128-row tiles on 148 CTAs); and while the
// 32-column tiles still fit in ONE wave (a page of lines: M <= 256 rows at N = 2048 is <= 128 tiles)
// they spread the weight stream over 8x as many SMs as the wide tiles would.
void pick_shape(int M, int N, int* ncta, int* bn) {
    if (ctx().force_shape) {
        const int f = ctx().force_shape;
        *bn = f == 1 ? BN_NARROW : BN_WIDE;
        *ncta = (f == 3 && g_max_pairs > 0) ? 2 : 1;
        return;
    }
    const long long tiles_m = (M + BM - 1) / BM, narrow = tiles_m * ((N + BN_NARROW - 1) / BN_NARROW);
    if (ctx().force_bn == BN_NARROW || (narrow <= g_num_sms && !getenv("COR_NO_NARROW"))) { *ncta = 1; *bn = BN_NARROW; return; }
    *bn = BN_WIDE;
    const long long pair_tiles = (long long)((M + 2 * BM - 1) / (2 * BM)) * ((N + BN_WIDE - 1) / BN_WIDE);
    *ncta = (g_max_pairs > 0 && pair_tiles >= g_max_pairs) ? 2 : 1;
}
int grid_for(int ncta, int bn, int M, int N) {
    const long long tiles = (long long)((M + BM * ncta - 1) / (BM * ncta)) * ((N + bn - 1) / bn);
    if (ncta == 2) return 2 * (int)std::min<long long>(tiles, g_max_pairs);
    return (int)std::min<long long>(tiles, g_num_sms);
}

template <int EPI, int NCTA, int BN, typename Epi>
int launch_tc(cudaStream_t st, int M, int N, const TcParams& p, const Epi& epi) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid_for(NCTA, BN, M, N)); cfg.blockDim = dim3(BN == BN_NARROW ? NTHREADS_NARROW : NTHREADS);
    cfg.dynamicSmemBytes = Geo<NCTA, BN>::SMEM_BYTES; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = NCTA; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    COR_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<EPI, NCTA, BN, Epi>, p, epi));
    ++ctx().launches; ++ctx().tc_kind[EPI][NCTA == 2 ? 2 : (BN == BN_WIDE ? 1 : 0)];
    return 0;
}

template <int EPI, typename Epi>
int dispatch_tc(cudaStream_t st, int M, const int* m_dev, int N, const GemmSeg* segs, int nseg, int passes, const Epi& epi) {
    int ncta, bn;
    pick_shape(M, N, &ncta, &bn);
    TcParams p;
    if (fill_params(ncta, bn, M, m_dev, N, segs, nseg, passes, &p)) return 1;
    p.e_use = 0; p.e_seq_mode = 0;
    if constexpr (EPI == 1) { if (fill_epi_maps(epi, M, &p)) return 1; }
    if (bn == BN_NARROW) return launch_tc<EPI, 1, BN_NARROW, Epi>(st, M, N, p, epi);
    return ncta == 2 ? launch_tc<EPI, 2, BN_WIDE, Epi>(st, M, N, p, epi) : launch_tc<EPI, 1, BN_WIDE, Epi>(st, M, N, p, epi);
}

}  // namespace

bool tc_available() { return init_tc(); }
void tc_shape_thresholds(int N, int* narrow_max_rows, int* single_max_rows) {
    init_tc();
    const int tiles_n32 = (N + BN_NARROW - 1) / BN_NARROW, tiles_n256 = (N + BN_WIDE - 1) / BN_WIDE;
    *narrow_max_rows = getenv("COR_NO_NARROW") ? 0 : (g_num_sms / tiles_n32) * BM;
    // the pair kernel takes over once ceil(M / 256) * tiles_n256 >= g_max_pairs
    *single_max_rows = g_max_pairs > 0 ? ((g_max_pairs + tiles_n256 - 1) / tiles_n256 - 1) * 2 * BM : 0x7fffffff;
    if (*single_max_rows < *narrow_max_rows) *single_max_rows = *narrow_max_rows;
}
void tc_force_narrow(bool on) { ctx().force_bn = on ? BN_NARROW : 0; }
void tc_ctx_release(Ctx* c) {
    if (c && c->tc_maps) { delete static_cast<MapCache*>(c->tc_maps); c->tc_maps = nullptr; }
}

int gemm_store_tc(cudaStream_t st, int M, const int* m_dev, int N, const GemmSeg* segs, int nseg, const StoreEpi& epi, int passes) {
    if (M <= 0) return 0;
    COR_CHECK(init_tc(), "tcgen05 path unavailable on this device");
    return dispatch_tc<0, StoreEpi>(st, M, m_dev, N, segs, nseg, passes, epi);
}

int gemm_lstm_tc(cudaStream_t st, int M, const int* m_dev, int N, const GemmSeg* segs, int nseg, const LstmEpi& epi, int passes) {
    if (M <= 0) return 0;
    COR_CHECK(init_tc(), "tcgen05 path unavailable on this device");
    COR_CHECK(N == 4 * epi.W && epi.W % 8 == 0, "LSTM epilogue needs N == 4W with W a multiple of 8");
    return dispatch_tc<1, LstmEpi>(st, M, m_dev, N, segs, nseg, passes, epi);
}

}  // namespace cor
