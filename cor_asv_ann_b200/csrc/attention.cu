// attention.cu — local monotonic additive attention (DenseAnnotationAttention.attention_call,
// ocrd_cor_asv_ann/lib/attention.py:526-575) for one decoder step over all hypothesis rows.
//
// Per row the cell needs, for the <= 2w+1 window positions j around the previous alignment centre,
// the hoisted key u_j = enc_j.U (W floats) and the annotation enc_j (D values): ~45 KB at W = D = 512,
// each byte used once, nothing shared between rows — an HBM-bound gather.  One group of warps per row:
//   1. centre and window from the previous alignment row (attention.py:551-557),
//   2. two bursts of bulk async copies (cp.async.bulk, the 1-D TMA path) into the group's shared-memory
//      slab, each counted on its own mbarrier: the query row + the window block of u, and the window
//      blocks of enc (hi, lo) — u and enc are line-major, so a window is one contiguous block per array;
//      ~45 KB in flight per row without holding a register,
//   3. energies e_j = exp(v . tanh(wq + u_j) + b_v) and the context sum_j a_j enc_j from shared memory.
// Positions outside the window have e*0 = 0 exactly in the reference (mask multiply, :557-566), so
// evaluating only the window gives the same sums.
// A row belongs to a GROUP of four warps that split the W (energies) and D (context) columns: the
// arithmetic is a chain of MUFU/FMA latencies, and four slabs of ~46 KB are all an SM can hold, so
// the sixteen warps that hide those latencies have to share rows.  Groups loop over rows (persistent
// grid) and refill each half of the slab for their next row as soon as its last reader is done.
#include "engine.h"
#include "epilogue.cuh"
#include <math.h>
#include <algorithm>

namespace cor {

namespace {


__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}

struct SlabGeo {
    int np;             // window positions a slab holds
    int wq_off, u_off, e_hi_off, e_lo_off, w_off, bytes;   // byte offsets inside a warp's slab
    int enc_elt;        // bytes per enc element in global memory (2: fp16 hi + lo arrays, 4: fp32)
};

constexpr int GW = 4;            // warps per row group

__global__ void __launch_bounds__(512, 1) attention_tma_kernel(AttnArgs a, SlabGeo g, int v_bytes) {
    extern __shared__ __align__(128) uint8_t sm[];
    // [0,128): two mbarriers per group (keys, annotations); then v_a (W floats); then one slab per group
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int gid = warp / GW, sub = warp % GW, groups = blockDim.x / (32 * GW);
    float* s_v = reinterpret_cast<float*>(sm + 128);
    uint8_t* slab = sm + 128 + v_bytes + (size_t)gid * g.bytes;
    const uint32_t slab_s = smem_u32(slab);
    const uint32_t bar_u = smem_u32(sm + 16 * gid), bar_e = bar_u + 8;
    float* s_part = reinterpret_cast<float*>(slab + g.w_off);       // [GW][16] partial energies
    for (int k = threadIdx.x; k < a.W; k += blockDim.x) s_v[k] = a.v_a[k];
    if (sub == 0 && lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_u));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_e));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int R = a.r_dev ? min(*a.r_dev, a.R) : a.R;
    const uint32_t u_row = (uint32_t)a.W * 4u, e_row = (uint32_t)a.D * (uint32_t)g.enc_elt;
    const int stride = gridDim.x * groups;
    auto group_sync = [&]() { asm volatile("bar.sync %0, %1;" ::"r"(1 + gid), "r"(32 * GW) : "memory"); };

    // window of a row from the centre of its previous alignment (attention.py:551-557): all j with
    // |centre - j| <= w (inclusive) or < w — a contiguous range [j0, j0+nj)
    auto window = [&](float centre, int T, int& j0, int& nj) {
        const int j_lo = max(0, (int)ceilf(centre - (float)a.window) - 1);
        const int j_hi = min(T - 1, (int)floorf(centre + (float)a.window) + 1);
        j0 = 0; nj = 0;
        for (int j = j_lo; j <= j_hi && nj < g.np; ++j) {
            const float dist = fabsf(centre - (float)j);
            const bool keep = a.inclusive ? (dist <= (float)a.window) : (dist < (float)a.window);
            if (keep) { if (nj == 0) j0 = j; ++nj; }
        }
    };
    // the two bursts of a row (issued by the group's first warp): query + keys u_j on bar_u,
    // annotations enc_j (hi, lo) on bar_e
    // u and enc are line-major ([L][t_max][.]): the window rows of a line are ONE contiguous block per array,
    // fetched by a single bulk copy (two for hi/lo) instead of one 1-2 KB copy per position
    auto issue_u = [&](int row, int line, int j0, int nj) {
        if (lane == 0) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_u), "r"(u_row * (uint32_t)(1 + nj)) : "memory");
            bulk_g2s(slab_s + g.wq_off, a.wq + (size_t)row * a.W, u_row, bar_u);
            if (nj > 0) bulk_g2s(slab_s + g.u_off, a.u + ((size_t)line * a.t_max + j0) * a.W, u_row * (uint32_t)nj, bar_u);
        }
        __syncwarp();
    };
    auto issue_e = [&](int line, int j0, int nj) {
        if (lane == 0) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_e), "r"((uint32_t)nj * e_row * (a.enc_lo ? 2u : 1u)) : "memory");
            if (nj > 0) {
                const size_t pos = (size_t)line * a.t_max + j0;
                bulk_g2s(slab_s + g.e_hi_off, reinterpret_cast<const uint8_t*>(a.enc) + pos * e_row, e_row * (uint32_t)nj, bar_e);
                if (a.enc_lo)
                    bulk_g2s(slab_s + g.e_lo_off, reinterpret_cast<const uint8_t*>(a.enc_lo) + pos * e_row, e_row * (uint32_t)nj, bar_e);
            }
        }
        __syncwarp();
    };
    auto wait = [&](uint32_t bar, uint32_t parity) {
        if (mbar_try(bar, parity)) return;
        const long long t0 = clock64();
        while (!mbar_try(bar, parity))
            if (clock64() - t0 > 4000000000LL) __trap();       // a protocol bug must not hang the GPU
    };
    // centre = sum_j a_prev[j]*j + 1  (fp32, attention.py:551-553), j ascending per lane
    auto centre_of = [&](const float* ap, int T, const float (&pre)[4]) {
        float part = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) { const int j = lane + 32 * i; if (j < T) part = fmaf(pre[i], (float)j, part); }
        for (int j = lane + 128; j < T; j += 32) part = fmaf(ap[j], (float)j, part);
        return warp_sum(part) + 1.0f;
    };
    auto prefetch_a = [&](const float* ap, int T, float (&pre)[4]) {
#pragma unroll
        for (int i = 0; i < 4; ++i) { const int j = lane + 32 * i; pre[i] = j < T ? ap[j] : 0.f; }
    };

    // every warp of the group derives the windows itself (same arithmetic, same result)
    int row = blockIdx.x * groups + gid;
    int line = 0, j0 = 0, nj = 0;
    if (row < R) {
        line = a.row_line ? a.row_line[row] : row;
        const int T = a.line_T[line];
        const float* ap = a.a_prev + (size_t)row * a.t_max;
        float pre[4];
        prefetch_a(ap, T, pre);
        window(centre_of(ap, T, pre), T, j0, nj);
        if (sub == 0) { issue_u(row, line, j0, nj); issue_e(line, j0, nj); }
    }
    uint32_t parity = 0;
    while (row < R) {
        // the next row of this group: its alignment row is fetched now and consumed after the energies
        const int nrow = row + stride;
        const bool has_next = nrow < R;
        int nline = 0, nT = 0, nj0 = 0, nnj = 0;
        const float* nap = nullptr;
        float pre[4] = {0.f, 0.f, 0.f, 0.f};
        if (has_next) {
            nline = a.row_line ? a.row_line[nrow] : nrow;
            nT = a.line_T[nline];
            nap = a.a_prev + (size_t)nrow * a.t_max;
            prefetch_a(nap, nT, pre);
        }
        // ---- energies e_j = exp(v . tanh(wq + u_j) + b_v): this warp's quarter of the W columns,
        //      four positions at a time with one accumulator per float4 component (16 chains per lane)
        wait(bar_u, parity);
        const float4* s_wq = reinterpret_cast<const float4*>(slab + g.wq_off);
        const float4* s_v4 = reinterpret_cast<const float4*>(s_v);
        const int nk4 = a.W >> 2;
        for (int q = 0; q < nj; q += 4) {
            const float4* su[4];
            float4 acc[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                su[i] = reinterpret_cast<const float4*>(slab + g.u_off + min(q + i, nj - 1) * u_row);
                acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
            for (int k4 = sub * 32 + lane; k4 < nk4; k4 += 32 * GW) {
                const float4 w4 = s_wq[k4], v4 = s_v4[k4];
                float4 x[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) x[i] = su[i][k4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    acc[i].x = fmaf(tanhf_fast(w4.x + x[i].x), v4.x, acc[i].x);
                    acc[i].y = fmaf(tanhf_fast(w4.y + x[i].y), v4.y, acc[i].y);
                    acc[i].z = fmaf(tanhf_fast(w4.z + x[i].z), v4.z, acc[i].z);
                    acc[i].w = fmaf(tanhf_fast(w4.w + x[i].w), v4.w, acc[i].w);
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float p = warp_sum((acc[i].x + acc[i].y) + (acc[i].z + acc[i].w));
                if (lane == 0 && q + i < nj) s_part[sub * 16 + q + i] = p;
            }
        }
        group_sync();      // partial sums are complete, and every warp is done with the key half of the slab
        if (has_next) {
            window(centre_of(nap, nT, pre), nT, nj0, nnj);
            if (sub == 0) issue_u(nrow, nline, nj0, nnj);
        }
        // lane q holds e_q, then a_q = e_q / sum (every warp computes the same values)
        float wgt = 0.f;
        if (lane < nj)
            wgt = expf(((s_part[lane] + s_part[16 + lane]) + (s_part[32 + lane] + s_part[48 + lane])) + a.b_v);
        float esum = 0.f;
        for (int q = 0; q < nj; ++q) esum += __shfl_sync(0xffffffffu, wgt, q);
        wgt = wgt / esum;
        // ---- alignment row: zero everywhere, a_j = e_j / sum in the window
        if (sub == 0) {
            float* ao = a.a_out + (size_t)row * a.t_max;
            float* al = a.align_log ? a.align_log + (a.step_dev ? (size_t)(*a.step_dev + a.step_off) * (size_t)a.align_stride : 0) + (size_t)(a.align_by_line ? line : row) * a.t_max : nullptr;
            for (int jb = 0; jb < a.t_max; jb += 32) {
                const int j = jb + lane, q = j - j0;
                float v = __shfl_sync(0xffffffffu, wgt, min(max(q, 0), 31));
                if (q < 0 || q >= nj) v = 0.f;
                if (j < a.t_max) { ao[j] = v; if (al) al[j] = v; }
            }
        }
        // ---- context = sum_j a_j * enc_j: this warp's quarter of the D columns
        wait(bar_e, parity);
        const int nd4 = a.D >> 2;
        for (int kb = sub * 32; kb < nd4; kb += 32 * GW) {
            const int k4 = min(kb + lane, nd4 - 1);
            float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int q = 0; q < nj; ++q) {
                const float w = __shfl_sync(0xffffffffu, wgt, q);
                float4 x;
                if (a.enc_lo) {
                    const uint2 h = *reinterpret_cast<const uint2*>(slab + g.e_hi_off + q * e_row + k4 * 8);
                    const uint2 l = *reinterpret_cast<const uint2*>(slab + g.e_lo_off + q * e_row + k4 * 8);
                    const float2 h0 = __half22float2(*reinterpret_cast<const __half2*>(&h.x)), h1 = __half22float2(*reinterpret_cast<const __half2*>(&h.y));
                    const float2 l0 = __half22float2(*reinterpret_cast<const __half2*>(&l.x)), l1 = __half22float2(*reinterpret_cast<const __half2*>(&l.y));
                    x = make_float4(h0.x + l0.x, h0.y + l0.y, h1.x + l1.x, h1.y + l1.y);
                } else {
                    x = *reinterpret_cast<const float4*>(slab + g.e_hi_off + q * e_row + k4 * 16);
                }
                c.x = fmaf(w, x.x, c.x); c.y = fmaf(w, x.y, c.y); c.z = fmaf(w, x.z, c.z); c.w = fmaf(w, x.w, c.w);
            }
            if (kb + lane < nd4) store_split4(a.ctx, a.ctx_lo, (size_t)row * a.D + k4 * 4, c);
        }
        group_sync();      // every warp is done with the annotation half of the slab (and the partial sums)
        if (has_next && sub == 0) issue_e(nline, nj0, nnj);
        parity ^= 1;
        row = nrow; line = nline; j0 = nj0; nj = nnj;
    }
}

}  // namespace

int launch_attention_reg(cudaStream_t st, const AttnArgs& a);     // kernels.cu: register-staged variant

int launch_attention(cudaStream_t st, const AttnArgs& a) {
    COR_CHECK(a.window >= 0 && a.window <= 6, "window_width above 6 is not supported");
    COR_CHECK(a.W % 4 == 0 && a.D % 8 == 0, "W must be a multiple of 4 and D of 8");
    if (a.R <= 0) return 0;
    static int n_sms = 0, max_smem = 0;
    if (!n_sms) {
        int dev = 0;
        COR_CUDA(cudaGetDevice(&dev));
        COR_CUDA(cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev));
        COR_CUDA(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        COR_CUDA(cudaFuncSetAttribute(attention_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
    }
    SlabGeo g{};
    g.np = 2 * a.window + 1;
    g.enc_elt = a.enc_lo ? 2 : 4;
    const int u_row = a.W * 4, e_row = a.D * g.enc_elt;
    g.wq_off = 0;
    g.u_off = u_row;
    g.e_hi_off = g.u_off + g.np * u_row;
    g.e_lo_off = g.e_hi_off + g.np * e_row;
    g.w_off = g.e_lo_off + (a.enc_lo ? g.np * e_row : 0);
    g.bytes = (g.w_off + GW * 16 * 4 + 127) & ~127;
    const int v_bytes = (a.W * 4 + 127) & ~127;
    int groups = (max_smem - 128 - v_bytes) / g.bytes;      // row groups (slabs) per block
    if (groups < 1 || (u_row & 15) || (e_row & 15)) return launch_attention_reg(st, a);   // slab does not fit: register variant
    if (groups > 4) groups = 4;
    const size_t smem = 128 + (size_t)v_bytes + (size_t)groups * g.bytes;
    const int blocks = std::min((a.R + groups - 1) / groups, n_sms);
    attention_tma_kernel<<<blocks, groups * GW * 32, smem, st>>>(a, g, v_bytes);
    ++ctx().launches;
    COR_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace cor
