// epilogue.cuh — device-side epilogues shared by the FFMA and the tcgen05 GEMM kernels.
#pragma once
#include "engine.h"
#include <cuda_fp16.h>

namespace cor {

// Split of an fp32 value into an fp16 head and an fp16 residual: x ~= hi + lo with
// |x - hi - lo| <= 2^-22 |x| as long as lo stays a normal fp16 number (|x| >~ 0.12); below that the
// residual sits on the fp16 subnormal grid and the absolute error is <= 2^-25 (3e-8).  Both parts have
// 11-bit significands — what a TF32 operand carries too — but run at the fp16 tensor rate (2x TF32)
// and take half the bytes.  Operand range: |x| < 65504 (activations here are O(1), weights O(0.1)).
//
// Representation convention used by every kernel: an operand is a pair of pointers (hi_p, lo_p), both
// declared float*.  lo_p == nullptr: hi_p holds plain fp32 values.  lo_p != nullptr: BOTH point to
// __half arrays (same element offsets), holding the head and the residual.
__device__ __forceinline__ void f16_split(float x, __half& hi, __half& lo) {
    hi = __float2half_rn(x);
    lo = __float2half_rn(x - __half2float(hi));
}

__device__ __forceinline__ void store_split(float* hi_p, float* lo_p, size_t off, float v) {
    if (lo_p) {
        __half hi, lo;
        f16_split(v, hi, lo);
        reinterpret_cast<__half*>(hi_p)[off] = hi;
        reinterpret_cast<__half*>(lo_p)[off] = lo;
    } else {
        hi_p[off] = v;
    }
}

__device__ __forceinline__ float load_split(const float* hi_p, const float* lo_p, size_t off) {
    if (lo_p) return __half2float(reinterpret_cast<const __half*>(hi_p)[off]) + __half2float(reinterpret_cast<const __half*>(lo_p)[off]);
    return hi_p[off];
}

// 4 consecutive elements (off a multiple of 4)
__device__ __forceinline__ float4 load_split4(const float* hi_p, const float* lo_p, size_t off) {
    if (!lo_p) return *reinterpret_cast<const float4*>(hi_p + off);
    const uint2 h = *reinterpret_cast<const uint2*>(reinterpret_cast<const __half*>(hi_p) + off);
    const uint2 l = *reinterpret_cast<const uint2*>(reinterpret_cast<const __half*>(lo_p) + off);
    const float2 h0 = __half22float2(*reinterpret_cast<const __half2*>(&h.x)), h1 = __half22float2(*reinterpret_cast<const __half2*>(&h.y));
    const float2 l0 = __half22float2(*reinterpret_cast<const __half2*>(&l.x)), l1 = __half22float2(*reinterpret_cast<const __half2*>(&l.y));
    return make_float4(h0.x + l0.x, h0.y + l0.y, h1.x + l1.x, h1.y + l1.y);
}

// Activation functions.  ex2.approx (2^-22 relative) and the fast reciprocal give an absolute error
// below 3e-7 on outputs in [-1, 1] — the same grade as expf/tanhf (2 ulp near 1) and three orders
// below the 1e-4 parity budget — at a fifth of the instructions: the LSTM epilogue evaluates five
// of these per unit and has to hide under the MMA main loop of the next tile.
__device__ __forceinline__ float ex2_approx(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float rcp_approx(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float sigmoidf_fast(float x) { return rcp_approx(1.0f + ex2_approx(-1.4426950408889634f * x)); }
__device__ __forceinline__ float tanhf_fast(float x) {
    // 1 - 2/(e^2x + 1); e = inf for large x gives rcp = 0 and the result 1, e = 0 gives -1
    return fmaf(-2.0f, rcp_approx(ex2_approx(2.8853900817779268f * x) + 1.0f), 1.0f);
}

// row index of sequence position (t, row) in seq_out / seq_rev (see LstmEpi::seq_tm)
__device__ __forceinline__ size_t seq_pos(const LstmEpi& e, int t, int row) {
    return e.seq_tm ? (size_t)t * e.seq_tm + (size_t)row * e.seq_rm : (size_t)t * e.rows_total + row;
}

// One LSTM unit of one row.  z[0..3] = pre-activations i,f,c~,o WITHOUT bias.
// Keras LSTMCell, implementation 2.  Returns the new h.
__device__ __forceinline__ float lstm_unit(const LstmEpi& e, int row, int unit, const float z[4]) {
    const size_t so = (size_t)row * e.W + unit;
    const float zi = z[0] + e.bias[4 * unit + 0];
    const float zf = z[1] + e.bias[4 * unit + 1];
    const float zg = z[2] + e.bias[4 * unit + 2];
    const float zo = z[3] + e.bias[4 * unit + 3];
    const float i = sigmoidf_fast(zi);
    const float f = sigmoidf_fast(zf);
    const float g = tanhf_fast(zg);
    const float o = sigmoidf_fast(zo);
    const float c = f * e.c_in[so] + i * g;
    e.c_out[so] = c;
    return o * tanhf_fast(c);
}

// Write the new h of `n` consecutive units [unit0, unit0+n) of one row (n even when cross_sum).
// Handles the inactive-row copy-through, the recurrent state, and the sequence outputs.
// z[0..n) += sum_k val_k * xtab[idx_k][col0 .. col0+n)  for the row's input position t (n a multiple of 4)
template <int N>
__device__ __forceinline__ void xtab_add(const LstmEpi& e, int row, int t, int col0, float* z) {
    const size_t o = ((size_t)row * e.x_tmax + t) * e.x_k;
    for (int k = 0; k < e.x_k; ++k) {
        const int v = e.x_idx[o + k];
        if (v < 0) continue;
        const float w = e.x_val ? e.x_val[o + k] : 1.0f;
        const float4* src = reinterpret_cast<const float4*>(e.xtab + (size_t)v * 4 * e.W + col0);
#pragma unroll
        for (int q = 0; q < N / 4; ++q) {
            const float4 x = src[q];
            z[4 * q + 0] = fmaf(w, x.x, z[4 * q + 0]); z[4 * q + 1] = fmaf(w, x.y, z[4 * q + 1]);
            z[4 * q + 2] = fmaf(w, x.z, z[4 * q + 2]); z[4 * q + 3] = fmaf(w, x.w, z[4 * q + 3]);
        }
    }
}

template <int NU>
__device__ __forceinline__ void lstm_row_units(const LstmEpi& e, int row, int unit0, const float (*z)[4]) {
    int T = e.row_T ? e.row_T[row] : 0x7fffffff;
    const bool active = e.step < T;
    float h[NU];
    if (active) {
        float zz[NU][4];
#pragma unroll
        for (int k = 0; k < NU; ++k)
#pragma unroll
            for (int g = 0; g < 4; ++g) zz[k][g] = z[k][g];
        if (e.xtab) xtab_add<4 * NU>(e, row, e.reverse ? T - 1 - e.step : e.step, 4 * unit0, &zz[0][0]);
#pragma unroll
        for (int k = 0; k < NU; ++k) h[k] = lstm_unit(e, row, unit0 + k, zz[k]);
    } else {
#pragma unroll
        for (int k = 0; k < NU; ++k) {
            const size_t so = (size_t)row * e.W + unit0 + k;
            h[k] = load_split(e.h_prev, e.h_prev_lo, so);
            if (e.c_out != e.c_in) e.c_out[so] = e.c_in[so];
        }
    }
#pragma unroll
    for (int k = 0; k < NU; ++k) store_split(e.h_out, e.h_out_lo, (size_t)row * e.W + unit0 + k, h[k]);
    if (!active || (!e.seq_out && !e.seq_rev)) return;
    const int t = e.reverse ? T - 1 - e.step : e.step;
    float y[NU];
#pragma unroll
    for (int k = 0; k < NU; ++k) y[k] = h[k];
    if (e.cross_sum) {
#pragma unroll
        for (int k = 0; k + 1 < NU; k += 2) { float s = h[k] + h[k + 1]; y[k] = s; y[k + 1] = s; }
    }
    if (e.seq_out) {
        const size_t base = seq_pos(e, t, row) * e.seq_ld + e.seq_col + unit0;
#pragma unroll
        for (int k = 0; k < NU; ++k) {
            float v = y[k];
            if (e.resid) v += load_split(e.resid, e.resid_lo, ((size_t)t * e.rows_total + row) * e.resid_ld + unit0 + k);
            store_split(e.seq_out, e.seq_out_lo, base + k, v);
        }
    }
    if (e.seq_rev) {
        const size_t base = seq_pos(e, T - 1 - t, row) * e.seq_ld + e.seq_col + unit0;
#pragma unroll
        for (int k = 0; k < NU; ++k) store_split(e.seq_rev, e.seq_rev_lo, base + k, y[k]);
    }
}

// 4 consecutive elements (off a multiple of 4)
__device__ __forceinline__ void store_split4(float* hi_p, float* lo_p, size_t off, float4 v) {
    if (!lo_p) { *reinterpret_cast<float4*>(hi_p + off) = v; return; }
    __half h[4], l[4];
    f16_split(v.x, h[0], l[0]); f16_split(v.y, h[1], l[1]); f16_split(v.z, h[2], l[2]); f16_split(v.w, h[3], l[3]);
    uint2 ph, pl;
    ph.x = (uint32_t)__half_as_ushort(h[0]) | ((uint32_t)__half_as_ushort(h[1]) << 16);
    ph.y = (uint32_t)__half_as_ushort(h[2]) | ((uint32_t)__half_as_ushort(h[3]) << 16);
    pl.x = (uint32_t)__half_as_ushort(l[0]) | ((uint32_t)__half_as_ushort(l[1]) << 16);
    pl.y = (uint32_t)__half_as_ushort(l[2]) | ((uint32_t)__half_as_ushort(l[3]) << 16);
    *reinterpret_cast<uint2*>(reinterpret_cast<__half*>(hi_p) + off) = ph;
    *reinterpret_cast<uint2*>(reinterpret_cast<__half*>(lo_p) + off) = pl;
}

__device__ __forceinline__ void store_elem(const StoreEpi& e, int row, int col, float acc) {
    if (e.n_split && col >= e.n_split) {
        const int c2 = col - e.n_split;
        e.C2[(size_t)row * e.ldc2 + c2] = acc + (e.bias2 ? e.bias2[c2] : 0.0f);
        return;
    }
    float v = acc + (e.bias ? e.bias[col] : 0.0f);
    if (e.act == 1) v = tanhf(v);
    store_split(e.C, e.C_lo, (size_t)row * e.ldc + col, v);
}

}  // namespace cor
