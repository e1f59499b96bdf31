// noalloc.cu — cor_gemm_noalloc: the tcgen05 3-pass GEMM of cor_debug_gemm(mode 2) WITHOUT any device allocation.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// The tcgen05 kernel reads its operands as fp16 (head, residual) pairs (epilogue.cuh).  cor_debug_gemm makes them in
// four cudaMalloc'ed buffers; here they live in ONE file-scope array in device memory, `g_scratch`, of kScratchBytes
// (16 MiB), which every GPU that loads the library holds until the process exits.  Limits that follow:
//   * 2 * 2 bytes * (M + N) * K (each part rounded up to 256 bytes) must fit: e.g. M = 1024, N = 2048, K = 1024;
//     anything larger is refused with an error, nothing is launched.
//   * every call on a device uses the SAME array: two calls that overlap on the GPU (different streams, no
//     dependency between them) race.  Calls on one stream are ordered and safe.
//   * the array keeps the last call's contents.  Nothing here needs to start at zero: the split kernel writes every
//     element the GEMM will read (the tensor maps cover exactly M x K and N x K; TMA zero-fills past them).
// In host code `g_scratch` is only a stand-in for the symbol, not a device address: the address is taken with
// cudaGetSymbolAddress once per device, with that device current, and that pointer is what the kernels (and the
// tensor maps) get.
//
// The call is stream-ordered and returns without synchronising; it calls no cudaMalloc / cudaFree / allocating
// library routine (first use on a device does one-time set-up: cudaGetSymbolAddress, function attributes).
#include "model.h"
#include "epilogue.cuh"
#include <algorithm>
#include <mutex>

namespace cor {
namespace {

constexpr size_t kScratchBytes = (size_t)16 << 20;
static __device__ __align__(256) unsigned char g_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_scratch_mu;
unsigned char* g_scratch_ptr[kMaxDevices] = {};     // device address of g_scratch per device ordinal

// device address of this device's copy of the array (the current device must be `dev`)
int scratch_base(int dev, unsigned char** out) {
    COR_CHECK(dev >= 0 && dev < kMaxDevices, "device ordinal out of range");
    std::lock_guard<std::mutex> lk(g_scratch_mu);
    if (!g_scratch_ptr[dev]) {
        void* p = nullptr;
        COR_CUDA(cudaGetSymbolAddress(&p, g_scratch));
        COR_CHECK(p != nullptr && ((uintptr_t)p & 255) == 0, "static scratch array has no aligned device address");
        g_scratch_ptr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_scratch_ptr[dev];
    return 0;
}

// Both operands in one launch: element i < na belongs to A, the others to B.
__global__ void __launch_bounds__(256) split_pair_kernel(const float* __restrict__ a, size_t na, __half* a_hi, __half* a_lo,
                                                         const float* __restrict__ b, size_t nb, __half* b_hi, __half* b_lo) {
    const size_t stride = (size_t)gridDim.x * blockDim.x, n = na + nb;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        __half hi, lo;
        if (i < na) { f16_split(a[i], hi, lo); a_hi[i] = hi; a_lo[i] = lo; }
        else        { const size_t j = i - na; f16_split(b[j], hi, lo); b_hi[j] = hi; b_lo[j] = lo; }
    }
}

inline size_t round256(size_t x) { return (x + 255) & ~(size_t)255; }

}  // namespace
}  // namespace cor

using namespace cor;

extern "C" {

size_t cor_gemm_noalloc_max_bytes(void) { return kScratchBytes; }

int cor_gemm_noalloc(int M, int N, int K, const float* A, const float* B, float* C, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    COR_CHECK(A && B && C, "null argument");
    COR_CHECK(M >= 1 && N >= 1 && K >= 64 && K % 64 == 0, "M, N >= 1 and K a positive multiple of 64");
    const size_t na = (size_t)M * K, nb = (size_t)N * K;
    const size_t need = 2 * round256(na * sizeof(__half)) + 2 * round256(nb * sizeof(__half));
    COR_CHECK(need <= kScratchBytes, "cor_gemm_noalloc: 4 * (M + N) * K bytes of split operands exceed the 16 MiB static scratch array");
    int dev = -1;
    COR_CUDA(cudaGetDevice(&dev));
    unsigned char* base = nullptr;
    if (scratch_base(dev, &base)) return 1;
    __half* a_hi = reinterpret_cast<__half*>(base);
    __half* a_lo = reinterpret_cast<__half*>(base + round256(na * sizeof(__half)));
    __half* b_hi = reinterpret_cast<__half*>(base + 2 * round256(na * sizeof(__half)));
    __half* b_lo = reinterpret_cast<__half*>(base + 2 * round256(na * sizeof(__half)) + round256(nb * sizeof(__half)));
    const int blocks = (int)std::min<size_t>((na + nb + 255) / 256, (size_t)148 * 16);
    split_pair_kernel<<<blocks, 256, 0, st>>>(A, na, a_hi, a_lo, B, nb, b_hi, b_lo);
    ++ctx().launches;
    COR_CUDA(cudaGetLastError());
    // (hi, lo) pairs travel behind float* in the engine's interfaces (epilogue.cuh)
    GemmSeg s{reinterpret_cast<const float*>(a_hi), K, reinterpret_cast<const float*>(a_lo),
              reinterpret_cast<const float*>(b_hi), K, reinterpret_cast<const float*>(b_lo), K};
    StoreEpi e{C, N, nullptr, nullptr, 0};
    return gemm_store_tc(st, M, nullptr, N, &s, 1, e, 3);
}

}  // extern "C"
