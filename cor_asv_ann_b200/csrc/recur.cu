// recur.cu — persistent LSTM recurrence for page-sized batches (one OCR-D page per call, <= 64 lines, in blocks of 32 rows).
//
// North star item (2): "persistent recurrence kernels for the bidirectional bottom encoder layer and the
// stacked forward LSTMs, with recurrent weights held in shared memory across timesteps".  With a few dozen
// rows a timestep is ~0.1 GFLOP: one launch per timestep and direction (gemm_tc.cu, narrow tiles) spends
// its ~15 us on launch, TMEM/barrier set-up and on pulling the 4 MB recurrent kernel out of L2 again.  Here
// ONE launch runs all T timesteps of a layer (both directions of a bidirectional layer side by side):
//
//   * a CTA owns 8 LSTM units (32 gate columns n = 4*unit + gate) of one direction and keeps that slice of
//     the recurrent kernel R (fp16 head + residual, 32 x W x 4 bytes = 64 KB at W = 512) in shared memory
//     for the whole scan; grid = directions x W/8 CTAs (128 at W = 512, bidirectional);
//   * per timestep it copies h_{t-1} of all rows (fp16 hi/lo, B x W x 4 bytes) into shared memory with
//     cp.async, forms its 32 x 32 block of h.R with warp-level tensor-core MMAs (mma.sync m16n8k16, the
//     same three-pass hi/lo product as the tcgen05 kernel: fp32-grade), adds the non-recurrent input term --
//     gathered from the (E.Kx) table for the first layer, read from a [T][B][4W] buffer that ONE big
//     tcgen05 GEMM over all timesteps prepared for the upper layers -- and runs the fused LSTM cell
//     (the same lstm_row_units epilogue as every other kernel), writing h_t (hi/lo) and the sequence outputs;
//   * timesteps are separated by a grid-wide barrier (release/acquire counter in global memory; the grid
//     is launched cooperatively so that all CTAs are co-resident, and the spin is bounded).
//
// The warp tile is 16 rows x 8 columns over the full K = W (8 warps = 2 row tiles x 4 column tiles), so no
// cross-warp reduction is needed; rows beyond the call's B are zero.
#include "engine.h"
#include "epilogue.cuh"
#include <cuda_fp16.h>
#include <stdlib.h>

namespace cor {

namespace {

constexpr int RC_ROWS = 32;              // rows per pass: two m16 tiles (a call has up to kRecurMaxRows rows)
constexpr int RC_COLS = 32;              // gate columns per CTA = 8 units
constexpr int RC_THREADS = 256;
constexpr int ZLD = RC_COLS + 1;

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
// All CTAs of the grid have finished the work before their (target / gridDim.x)-th arrival.
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
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
__device__ __forceinline__ uint32_t lds32(const __half* p) { return *reinterpret_cast<const uint32_t*>(p); }

__global__ void __launch_bounds__(RC_THREADS, 1) lstm_recur_kernel(const RecurArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int W = a.W, LD = W + 8, B = a.B;          // +8 halves: rows land on distinct banks for the fragment loads
    __half* Whi = reinterpret_cast<__half*>(smem);   // [32][LD] this CTA's rows of R
    __half* Wlo = Whi + RC_COLS * LD;
    __half* Ahi = Wlo + RC_COLS * LD;                // [32][LD] h_{t-1}
    __half* Alo = Ahi + RC_ROWS * LD;
    float* zs = reinterpret_cast<float*>(Alo + RC_ROWS * LD);   // [32][33] pre-activations of the CTA's block
    const int tid = threadIdx.x;
    const int per_dir = W / 8;
    const int dir = blockIdx.x / per_dir, cd = blockIdx.x % per_dir;
    const int col0 = cd * RC_COLS;
    const int chunks = W / 8;                        // 16-byte pieces per row

    {   // resident operands: the CTA's slice of R (hi, lo); A rows beyond B stay zero
        const __half* rh = reinterpret_cast<const __half*>(a.R_hi[dir]) + (size_t)col0 * a.ldr;
        const __half* rl = reinterpret_cast<const __half*>(a.R_lo[dir]) + (size_t)col0 * a.ldr;
        for (int i = tid; i < RC_COLS * chunks; i += RC_THREADS) {
            const int r = i / chunks, c8 = i % chunks;
            *reinterpret_cast<uint4*>(Whi + r * LD + c8 * 8) = *reinterpret_cast<const uint4*>(rh + (size_t)r * a.ldr + c8 * 8);
            *reinterpret_cast<uint4*>(Wlo + r * LD + c8 * 8) = *reinterpret_cast<const uint4*>(rl + (size_t)r * a.ldr + c8 * 8);
        }
        const uint4 zero = make_uint4(0, 0, 0, 0);
        for (int i = tid; i < RC_ROWS * LD / 8; i += RC_THREADS) {
            reinterpret_cast<uint4*>(Ahi)[i] = zero;
            reinterpret_cast<uint4*>(Alo)[i] = zero;
        }
    }
    __syncthreads();

    LstmEpi e = a.epi[dir];
    const int warp = tid >> 5, lane = tid & 31, mt = warp & 1, nt = warp >> 1, g = lane >> 2, q = lane & 3;
    const __half* ah = Ahi + (mt * 16 + g) * LD + 2 * q;
    const __half* al = Alo + (mt * 16 + g) * LD + 2 * q;
    const __half* bh = Whi + (nt * 8 + g) * LD + 2 * q;
    const __half* bl = Wlo + (nt * 8 + g) * LD + 2 * q;

    for (int s = 0; s < a.T; ++s) {
        const int cur = s & 1;
        if (s > 0) grid_sync(a.barrier, gridDim.x * (unsigned)s);
        // rows in blocks of 32 (two m16 tiles): a page of up to kRecurMaxRows lines takes one or two passes over the
        // resident weights per timestep, still with ONE grid barrier
        for (int r0 = 0; r0 < B; r0 += RC_ROWS) {
            const int Bb = min(RC_ROWS, B - r0);
            const bool tile_live = mt * 16 < Bb;         // the second row tile is empty for <= 16 rows
            if (r0 > 0) __syncthreads();                 // the previous block's epilogue is done with zs, its MMAs with A
            {
                const __half* hp = reinterpret_cast<const __half*>(a.h[dir][cur]) + (size_t)r0 * W;
                const __half* hl = reinterpret_cast<const __half*>(a.h_lo[dir][cur]) + (size_t)r0 * W;
                for (int i = tid; i < Bb * chunks; i += RC_THREADS) {
                    const int r = i / chunks, c8 = i % chunks;
                    cp_async16(Ahi + r * LD + c8 * 8, hp + (size_t)r * W + c8 * 8);
                    cp_async16(Alo + r * LD + c8 * 8, hl + (size_t)r * W + c8 * 8);
                }
                cp_async_wait_all();
            }
            __syncthreads();
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            if (tile_live) {
#pragma unroll 4
                for (int k0 = 0; k0 < W; k0 += 16) {
                    const uint32_t a_h[4] = {lds32(ah + k0), lds32(ah + 8 * LD + k0), lds32(ah + k0 + 8), lds32(ah + 8 * LD + k0 + 8)};
                    const uint32_t a_l[4] = {lds32(al + k0), lds32(al + 8 * LD + k0), lds32(al + k0 + 8), lds32(al + 8 * LD + k0 + 8)};
                    const uint32_t b_h[2] = {lds32(bh + k0), lds32(bh + k0 + 8)};
                    const uint32_t b_l[2] = {lds32(bl + k0), lds32(bl + k0 + 8)};
                    mma16816(acc, a_l, b_h);             // small terms first, like the tcgen05 kernel
                    mma16816(acc, a_h, b_l);
                    mma16816(acc, a_h, b_h);
                }
            }
            {
                float* z0 = zs + (mt * 16 + g) * ZLD + nt * 8 + 2 * q;
                z0[0] = acc[0]; z0[1] = acc[1];
                z0[8 * ZLD] = acc[2]; z0[8 * ZLD + 1] = acc[3];
            }
            __syncthreads();
            if (tid < RC_ROWS * 4) {                     // (row, pair of units): the cross-sum couples units 2k, 2k+1
                const int rl = tid >> 2, up = tid & 3, row = r0 + rl;
                if (rl < Bb) {
                    float z[2][4];
#pragma unroll
                    for (int k = 0; k < 2; ++k)
#pragma unroll
                        for (int gate = 0; gate < 4; ++gate) z[k][gate] = zs[rl * ZLD + 8 * up + 4 * k + gate];
                    if (a.zx[dir]) {
                        const float4* x = reinterpret_cast<const float4*>(a.zx[dir] + ((size_t)s * B + row) * 4 * W + col0 + 8 * up);
#pragma unroll
                        for (int k = 0; k < 2; ++k) {
                            const float4 v = x[k];
                            z[k][0] += v.x; z[k][1] += v.y; z[k][2] += v.z; z[k][3] += v.w;
                        }
                    }
                    e.step = s;
                    e.h_prev = a.h[dir][cur]; e.h_prev_lo = a.h_lo[dir][cur];
                    e.h_out = a.h[dir][cur ^ 1]; e.h_out_lo = a.h_lo[dir][cur ^ 1];
                    lstm_row_units<2>(e, row, cd * 8 + 2 * up, z);
                }
            }
        }
        // the next step's grid_sync starts with __syncthreads: zs and A are not touched before everyone is here
    }
}

size_t recur_smem_bytes(int W) {
    return (size_t)(2 * RC_COLS + 2 * RC_ROWS) * (W + 8) * sizeof(__half) + (size_t)RC_ROWS * ZLD * sizeof(float);
}

}  // namespace

bool lstm_recur_supported(int W, int B, int ndir) {
    if (getenv("COR_NO_RECUR")) return false;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return false;
    static int c_dev = -1, c_ok = 0, c_sms = 0;
    if (c_dev != dev) {         // per device: cooperative launch support, SM count, shared-memory opt-in
        c_dev = dev; c_ok = 0;
        int coop = 0;
        if (cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev) == cudaSuccess && coop &&
            cudaDeviceGetAttribute(&c_sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess &&
            cudaFuncSetAttribute(lstm_recur_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) == cudaSuccess)
            c_ok = 1;
        (void)cudaGetLastError();
    }
    return c_ok && B >= 1 && B <= kRecurMaxRows && W % 16 == 0 && ndir * (W / 8) <= c_sms && recur_smem_bytes(W) <= 227 * 1024;
}

int launch_lstm_recur(cudaStream_t st, const RecurArgs& a) {
    COR_CHECK(lstm_recur_supported(a.W, a.B, a.ndir), "persistent recurrence not available for this shape");
    COR_CHECK(a.barrier && a.T >= 1, "bad recurrence arguments");
    COR_CUDA(cudaMemsetAsync(a.barrier, 0, sizeof(unsigned), st));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(a.ndir * (a.W / 8)); cfg.blockDim = dim3(RC_THREADS);
    cfg.dynamicSmemBytes = recur_smem_bytes(a.W); cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeCooperative; at[0].val.cooperative = 1;      // co-residency of the whole grid
    cfg.attrs = at; cfg.numAttrs = 1;
    COR_CUDA(cudaLaunchKernelEx(&cfg, lstm_recur_kernel, a));
    ++ctx().launches;
    return 0;
}

}  // namespace cor
