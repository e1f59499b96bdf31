// gemm_ffma.cu — exact-fp32 CUDA-core GEMM with the fused epilogues.
//
// Role: the small-batch path (an OCR-D page is 20-60 lines: fewer rows than one tcgen05
// M=128 tile) and the bring-up / cross-check path for the tensor-core kernel in gemm_tc.cu.
// C[M,N] = sum_s A_s[M,K_s] * B_s[N,K_s]^T; 128x128 tile, 256 threads, 8x8 register tile,
// BK=16 double-buffered through shared memory.  Operands stored as fp16 (hi, lo) pairs are
// re-joined on load (hi + lo is exact in fp32), so this kernel always accumulates in full fp32.
#include "engine.h"
#include "epilogue.cuh"

namespace cor {

namespace {
constexpr int BM = 128, BN = 128, BK = 16, NT = 256;
constexpr int LDS = BM + 4;

struct FfmaParams {
    GemmSeg segs[kMaxSegs];
    int nseg;
    int M;
    const int* m_dev;
    int N;
};

// logical tile column -> shared-memory column: the two float4 halves of a thread's 8 columns
// live 64 floats apart so that the 128-bit reads of a half-warp are conflict-free.
__device__ __forceinline__ int bcol(int n) { return ((n & 4) << 4) + ((n >> 3) << 2) + (n & 3); }

template <int EPI, typename Epi>
__global__ void __launch_bounds__(NT) gemm_ffma_kernel(FfmaParams p, Epi epi) {
    __shared__ __align__(16) float As[2][BK][LDS];
    __shared__ __align__(16) float Bs[2][BK][LDS];
    const int M = p.m_dev ? min(*p.m_dev, p.M) : p.M;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    if (m0 >= M) return;
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int lrow = tid >> 2, lk = (tid & 3) * 4;

    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

    int ntiles = 0;
    for (int s = 0; s < p.nseg; ++s) ntiles += p.segs[s].K / BK;

    float4 ra[2], rb[2];
    auto fetch = [&](int tile) {
        int s = 0, k0 = tile * BK;
        while (k0 >= p.segs[s].K) { k0 -= p.segs[s].K; ++s; }
        const GemmSeg& g = p.segs[s];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            int r = min(m0 + lrow + 64 * h, M - 1);
            size_t off = (size_t)r * g.lda + k0 + lk;
            ra[h] = load_split4(g.A, g.A_lo, off);
            size_t offb = (size_t)min(n0 + lrow + 64 * h, p.N - 1) * g.ldb + k0 + lk;
            rb[h] = load_split4(g.B, g.B_lo, offb);
        }
    };
    auto stash = [&](int buf) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            int r = lrow + 64 * h;
            As[buf][lk + 0][r] = ra[h].x; As[buf][lk + 1][r] = ra[h].y;
            As[buf][lk + 2][r] = ra[h].z; As[buf][lk + 3][r] = ra[h].w;
            int c = bcol(r);
            Bs[buf][lk + 0][c] = rb[h].x; Bs[buf][lk + 1][c] = rb[h].y;
            Bs[buf][lk + 2][c] = rb[h].z; Bs[buf][lk + 3][c] = rb[h].w;
        }
    };

    fetch(0);
    stash(0);
    __syncthreads();
    for (int t = 0; t < ntiles; ++t) {
        const int buf = t & 1;
        if (t + 1 < ntiles) fetch(t + 1);
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            float a[8], b[8];
            *reinterpret_cast<float4*>(a)     = *reinterpret_cast<const float4*>(&As[buf][k][ty * 4]);
            *reinterpret_cast<float4*>(a + 4) = *reinterpret_cast<const float4*>(&As[buf][k][64 + ty * 4]);
            *reinterpret_cast<float4*>(b)     = *reinterpret_cast<const float4*>(&Bs[buf][k][tx * 4]);
            *reinterpret_cast<float4*>(b + 4) = *reinterpret_cast<const float4*>(&Bs[buf][k][64 + tx * 4]);
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        if (t + 1 < ntiles) stash(buf ^ 1);
        __syncthreads();
    }

    // thread's rows: ty*4+i and 64+ty*4+i; logical columns tx*8 .. tx*8+7
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int row = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (row >= M) continue;
        if constexpr (EPI == 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (n0 + tx * 8 + j < p.N) store_elem(epi, row, n0 + tx * 8 + j, acc[i][j]);
        } else {
            float z[2][4];
#pragma unroll
            for (int j = 0; j < 8; ++j) z[j >> 2][j & 3] = acc[i][j];
            lstm_row_units<2>(epi, row, (n0 + tx * 8) >> 2, z);
        }
    }
}

int check_segs(int N, const GemmSeg* segs, int nseg) {
    COR_CHECK(nseg >= 1 && nseg <= kMaxSegs, "bad segment count");
    COR_CHECK(N > 0, "N must be positive");
    for (int s = 0; s < nseg; ++s) {
        COR_CHECK(segs[s].K % BK == 0 && segs[s].K > 0, "K must be a positive multiple of 16");
        COR_CHECK(segs[s].lda % 4 == 0 && segs[s].ldb % 4 == 0, "leading dimensions must be multiples of 4");
    }
    return 0;
}
}  // namespace

int gemm_store_ffma(cudaStream_t st, int M, const int* m_dev, int N, const GemmSeg* segs, int nseg, const StoreEpi& epi) {
    if (M <= 0) return 0;
    if (check_segs(N, segs, nseg)) return 1;
    FfmaParams p{};
    for (int s = 0; s < nseg; ++s) p.segs[s] = segs[s];
    p.nseg = nseg; p.M = M; p.m_dev = m_dev; p.N = N;
    dim3 grid((N + BN - 1) / BN, (M + BM - 1) / BM);
    gemm_ffma_kernel<0, StoreEpi><<<grid, NT, 0, st>>>(p, epi);
    ++ctx().launches; ++ctx().ffma_kind[0];
    COR_CUDA(cudaGetLastError());
    return 0;
}

int gemm_lstm_ffma(cudaStream_t st, int M, const int* m_dev, int N, const GemmSeg* segs, int nseg, const LstmEpi& epi) {
    if (M <= 0) return 0;
    if (check_segs(N, segs, nseg)) return 1;
    COR_CHECK(N == 4 * epi.W && N % BN == 0, "LSTM epilogue needs N == 4W, a multiple of 128");
    FfmaParams p{};
    for (int s = 0; s < nseg; ++s) p.segs[s] = segs[s];
    p.nseg = nseg; p.M = M; p.m_dev = m_dev; p.N = N;
    dim3 grid((N + BN - 1) / BN, (M + BM - 1) / BM);
    gemm_ffma_kernel<1, LstmEpi><<<grid, NT, 0, st>>>(p, epi);
    ++ctx().launches; ++ctx().ffma_kind[1];
    COR_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace cor
