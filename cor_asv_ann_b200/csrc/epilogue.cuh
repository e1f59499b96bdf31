// epilogue.cuh — device-side epilogues shared by the FFMA and the tcgen05 GEMM kernels.
#pragma once
#include "engine.h"

namespace cor {

// Round-to-nearest split of an fp32 value into a TF32-representable head and the exact residual.
// hi has its 13 low mantissa bits clear, so the tensor core's own fp32->tf32 conversion is exact
// on it whatever its rounding mode; lo = x - hi is exact in fp32 (|lo| <= 2^-11 |x|).
__device__ __forceinline__ void tf32_split(float x, float& hi, float& lo) {
    uint32_t b = __float_as_uint(x);
    uint32_t r = (b + 0x1000u) & 0xFFFFE000u;
    // keep inf/nan and values whose rounding would overflow the exponent untouched
    if ((b & 0x7F800000u) == 0x7F800000u || (r & 0x7F800000u) == 0x7F800000u) r = b & 0xFFFFE000u;
    hi = __uint_as_float(r);
    lo = x - hi;
}

__device__ __forceinline__ void store_split(float* hi_p, float* lo_p, size_t off, float v) {
    if (lo_p) {
        float hi, lo;
        tf32_split(v, hi, lo);
        hi_p[off] = hi;
        lo_p[off] = lo;
    } else {
        hi_p[off] = v;
    }
}

__device__ __forceinline__ float load_split(const float* hi_p, const float* lo_p, size_t off) {
    float v = hi_p[off];
    if (lo_p) v += lo_p[off];
    return v;
}

__device__ __forceinline__ float sigmoidf_exact(float x) { return 1.0f / (1.0f + expf(-x)); }

// One LSTM unit of one row.  z[0..3] = pre-activations i,f,c~,o WITHOUT bias.
// Keras LSTMCell, implementation 2.  Returns the new h.
__device__ __forceinline__ float lstm_unit(const LstmEpi& e, int row, int unit, const float z[4]) {
    const size_t so = (size_t)row * e.W + unit;
    const float zi = z[0] + e.bias[4 * unit + 0];
    const float zf = z[1] + e.bias[4 * unit + 1];
    const float zg = z[2] + e.bias[4 * unit + 2];
    const float zo = z[3] + e.bias[4 * unit + 3];
    const float i = sigmoidf_exact(zi);
    const float f = sigmoidf_exact(zf);
    const float g = tanhf(zg);
    const float o = sigmoidf_exact(zo);
    const float c = f * e.c_in[so] + i * g;
    e.c_out[so] = c;
    return o * tanhf(c);
}

// Write the new h of `n` consecutive units [unit0, unit0+n) of one row (n even when cross_sum).
// Handles the inactive-row copy-through, the recurrent state, and the sequence outputs.
template <int NU>
__device__ __forceinline__ void lstm_row_units(const LstmEpi& e, int row, int unit0, const float (*z)[4]) {
    int T = e.row_T ? e.row_T[row] : 0x7fffffff;
    const bool active = e.step < T;
    float h[NU];
    if (active) {
#pragma unroll
        for (int k = 0; k < NU; ++k) h[k] = lstm_unit(e, row, unit0 + k, z[k]);
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
        const size_t base = ((size_t)t * e.rows_total + row) * e.seq_ld + e.seq_col + unit0;
#pragma unroll
        for (int k = 0; k < NU; ++k) {
            float v = y[k];
            if (e.resid) v += load_split(e.resid, e.resid_lo, ((size_t)t * e.rows_total + row) * e.resid_ld + unit0 + k);
            store_split(e.seq_out, e.seq_out_lo, base + k, v);
        }
    }
    if (e.seq_rev) {
        const size_t base = ((size_t)(T - 1 - t) * e.rows_total + row) * e.seq_ld + e.seq_col + unit0;
#pragma unroll
        for (int k = 0; k < NU; ++k) store_split(e.seq_rev, e.seq_rev_lo, base + k, y[k]);
    }
}

__device__ __forceinline__ void store_elem(const StoreEpi& e, int row, int col, float acc) {
    float v = acc + (e.bias ? e.bias[col] : 0.0f);
    if (e.act == 1) v = tanhf(v);
    store_split(e.C, e.C_lo, (size_t)row * e.ldc + col, v);
}

}  // namespace cor
