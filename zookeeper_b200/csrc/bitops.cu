// bitops.cu -- sign + bit-pack (warp ballot, 32 values -> one u32 along K) and the CUDA-core
// XNOR-popcount GEMM over the packed words.  This is the bit-level formulation of larq's
// ste_sign + QuantDense forward (examples/larq_experiment.py:52-57, 95-99); the train step itself
// runs the same products on tcgen05 (gemm_tc.cu) -- DESIGN.md records the measured choice.
#include <cuda_bf16.h>

#include "zkb_host.h"

namespace zkb {

template <typename T>
__device__ __forceinline__ bool nonneg(const T* p, long long i);
template <>
__device__ __forceinline__ bool nonneg<float>(const float* p, long long i) { return p[i] >= 0.0f; }
template <>
__device__ __forceinline__ bool nonneg<uint16_t>(const uint16_t* p, long long i) {
  // bf16: non-negative iff sign bit clear, or the value is -0.0 (sign(-0.0) = +1 like larq)
  const uint16_t h = p[i];
  return (h & 0x8000u) == 0 || (h & 0x7FFFu) == 0;
}

// One warp packs 1024 consecutive elements per iteration: 32 coalesced 128-byte reads, 32
// ballots, one coalesced 128-byte write.  Bit j of word w <-> element 32*w + j (1 = +1).
template <typename T>
__global__ void __launch_bounds__(256)
sign_pack_kernel(const T* __restrict__ x, uint32_t* __restrict__ out, long long nwords) {
  const int lane = threadIdx.x & 31;
  const long long warp = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) >> 5;
  const long long nwarps = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  for (long long w0 = warp * 32; w0 < nwords; w0 += nwarps * 32) {
    uint32_t mine = 0;
#pragma unroll 8
    for (int j = 0; j < 32; ++j) {
      const long long w = w0 + j;
      bool b = false;
      if (w < nwords) b = nonneg<T>(x, w * 32 + lane);
      const uint32_t word = __ballot_sync(0xffffffffu, b);
      if (lane == j) mine = word;
    }
    if (w0 + lane < nwords) out[w0 + lane] = mine;
  }
}

// y[m][n] = 32*Kw - 2 * sum_w popc(a[m][w] ^ b[n][w]).  64x64 output tile per 256-thread block,
// 4x4 outputs per thread, 32-word K slabs staged in shared memory (padded against conflicts).
__global__ void __launch_bounds__(256)
bgemm_popc_kernel(const uint32_t* __restrict__ a, const uint32_t* __restrict__ b,
                  int* __restrict__ y, long long ldy, int M, int N, int Kw) {
  __shared__ uint32_t sa[64][33];
  __shared__ uint32_t sb[64][33];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  int acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0;
  for (int k0 = 0; k0 < Kw; k0 += 32) {
    for (int i = threadIdx.x; i < 64 * 32; i += 256) {
      const int r = i >> 5, c = i & 31;
      const int k = k0 + c;
      sa[r][c] = (m0 + r < M && k < Kw) ? a[static_cast<long long>(m0 + r) * Kw + k] : 0u;
      sb[r][c] = (n0 + r < N && k < Kw) ? b[static_cast<long long>(n0 + r) * Kw + k] : 0u;
    }
    __syncthreads();
#pragma unroll 8
    for (int c = 0; c < 32; ++c) {
      uint32_t av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { av[i] = sa[ty * 4 + i][c]; bv[i] = sb[tx * 4 + i][c]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] += __popc(av[i] ^ bv[j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int m = m0 + ty * 4 + i, n = n0 + tx * 4 + j;
      if (m < M && n < N) y[static_cast<long long>(m) * ldy + n] = 32 * Kw - 2 * acc[i][j];
    }
}

}  // namespace zkb

using namespace zkb;

ZKB_API int zkb_sign_pack(const void* x, void* packed_u32, long long n, int x_is_bf16, void* stream) {
  ZKB_REQUIRE(x && packed_u32, "sign_pack: null pointer");
  ZKB_REQUIRE(n >= 0 && n % 32 == 0, "sign_pack: element count %lld must be a multiple of 32", n);
  if (n == 0) return ZKB_OK;
  const long long nwords = n / 32;
  long long blocks = (nwords + 255) / 256;  // 8 warps x 32 words per block-iteration
  const long long cap = static_cast<long long>(zkb_sm_count()) * 8;
  if (blocks > cap) blocks = cap;
  if (x_is_bf16)
    sign_pack_kernel<uint16_t><<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const uint16_t*>(x), static_cast<uint32_t*>(packed_u32), nwords);
  else
    sign_pack_kernel<float><<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const float*>(x), static_cast<uint32_t*>(packed_u32), nwords);
  ZKB_LAUNCH_CHECK("sign_pack_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_bgemm_popc(const void* a_packed, const void* b_packed, int* y, long long ldy, int M,
                           int N, int k_words, void* stream) {
  ZKB_REQUIRE(a_packed && b_packed && y, "bgemm_popc: null pointer");
  ZKB_REQUIRE(M >= 0 && N >= 0 && k_words > 0 && ldy >= N, "bgemm_popc: bad shape");
  if (M == 0 || N == 0) return ZKB_OK;
  dim3 grid((N + 63) / 64, (M + 63) / 64);
  bgemm_popc_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint32_t*>(a_packed), static_cast<const uint32_t*>(b_packed), y, ldy, M, N,
      k_words);
  ZKB_LAUNCH_CHECK("bgemm_popc_kernel");
  return ZKB_OK;
}
