// gemm_tc.cu -- the two tcgen05 GEMM engines every binary conv / dense layer of the BinaryNet
// train step runs on (forward, dgrad, wgrad).
//
//   kgemm : D[M,N] = sum_t A_t[M,K] * B[N,K]^T       both operands K-major (K contiguous)
//           A is either a plain row-major matrix or the im2col view of an NHWC activation,
//           fetched tap by tap with TMA im2col loads (zero fill outside the image == Larq's
//           zero "same" padding).  t = 1 (forward: +-1 x +-1, exact) or 2 (dgrad: dy = hi + lo
//           bf16 split).  Persistent, warp-specialised: TMA producer / MMA issuer / 4 epilogue
//           warps, smem ring + double-buffered TMEM accumulators.
//   wgemm : D[Ma,N] += sum_t A[P,Ma]^T * B_t[P,N]    both operands MN-major (reduction index P
//           is the slow one: pixels / batch rows).  Split over P across CTAs; partial tiles are
//           reduced into the f32 gradient with red.global.add.
//
// Operands are bf16 (+-1 is exact), accumulation f32 in tensor memory: for +-1 x +-1 with
// K <= 2^24 the result is the exact integer.  Replaces the Eigen/oneDNN float conv + matmul that
// TF runs for larq QuantConv2D / QuantDense (reference call sites:
// examples/larq_experiment.py:71-73,76-82,85-91,95-99).
#include <cuda_bf16.h>

#include "zkb_host.h"
#include "zkb_ptx.cuh"

namespace zkb {

static constexpr int kMaxStages = 8;
static constexpr int kATileBytes = 128 * 128;  // 128 rows x 64 bf16
static constexpr int kSmemBudget = 227 * 1024 - 2048;

struct KGemmParams {
  int M, N;
  int num_kb;    // 64-element k-blocks per A term
  int a_terms;   // 1 or 2
  int block_n;   // 32..256, multiple of 32
  int num_stages;
  int stage_bytes;
  int tiles_m, tiles_n;
  int a_im2col;
  int OH, OW, pad, kw_taps, cblocks;
  int out_kind;  // 0: f32, 1: int16 (exact integers expected)
  void* out;
  long long ldo;
};

struct SharedCtl {
  uint64_t full[kMaxStages];
  uint64_t empty[kMaxStages];
  uint64_t tfull[2];
  uint64_t tempty[2];
  uint32_t tmem_base;
  uint32_t pad_;
};

__device__ __forceinline__ uint8_t* align_1024(uint8_t* p) {
  uintptr_t a = reinterpret_cast<uintptr_t>(p);
  a = (a + 1023) & ~static_cast<uintptr_t>(1023);
  return reinterpret_cast<uint8_t*>(a);
}

__global__ void __launch_bounds__(192, 1)
kgemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                  const __grid_constant__ CUtensorMap tmB, const KGemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_1024(smem_raw);
  SharedCtl* ctl = reinterpret_cast<SharedCtl*>(smem + static_cast<size_t>(p.num_stages) * p.stage_bytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.tiles_m * p.tiles_n;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(smem_u32(&ctl->full[s]), 1);
      mbar_init(smem_u32(&ctl->empty[s]), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(smem_u32(&ctl->tfull[a]), 1);
      mbar_init(smem_u32(&ctl->tempty[a]), 128);
    }
    fence_mbar_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA0);
    if (p.a_terms == 2) tma_prefetch_desc(&tmA1);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) tmem_alloc_512(smem_u32(&ctl->tmem_base));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = ctl->tmem_base;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int tm = tile / p.tiles_n;
        const int tn = tile - tm * p.tiles_n;
        const int m0 = tm * 128;
        const int n0 = tn * p.block_n;
        int img = 0, oh = 0, ow = 0;
        if (p.a_im2col) {
          const int hw = p.OH * p.OW;
          img = m0 / hw;
          const int rem = m0 - img * hw;
          oh = rem / p.OW;
          ow = rem - oh * p.OW;
        }
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(smem_u32(&ctl->empty[stage]), phase ^ 1u);
          const uint32_t full = smem_u32(&ctl->full[stage]);
          mbar_expect_tx(full, static_cast<uint32_t>(p.stage_bytes));
          const uint32_t sA = smem_u32(smem + static_cast<size_t>(stage) * p.stage_bytes);
          const uint32_t sB = sA + static_cast<uint32_t>(p.a_terms) * kATileBytes;
          if (p.a_im2col) {
            const int tap = kb / p.cblocks;
            const int cb = kb - tap * p.cblocks;
            const int kh = tap / p.kw_taps;
            const int kw = tap - kh * p.kw_taps;
            tma_load_im2col_4d(sA, &tmA0, full, cb * 64, ow - p.pad, oh - p.pad, img,
                               static_cast<uint16_t>(kw), static_cast<uint16_t>(kh));
            if (p.a_terms == 2)
              tma_load_im2col_4d(sA + kATileBytes, &tmA1, full, cb * 64, ow - p.pad, oh - p.pad,
                                 img, static_cast<uint16_t>(kw), static_cast<uint16_t>(kh));
          } else {
            tma_load_2d(sA, &tmA0, full, kb * 64, m0);
            if (p.a_terms == 2) tma_load_2d(sA + kATileBytes, &tmA1, full, kb * 64, m0);
          }
          tma_load_2d(sB, &tmB, full, kb * 64, n0);
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(128, p.block_n, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(smem_u32(&ctl->tempty[as]), aphase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(as) * 256u;
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(smem_u32(&ctl->full[stage]), phase);
          tc_fence_after();
          const uint32_t sA = smem_u32(smem + static_cast<size_t>(stage) * p.stage_bytes);
          const uint32_t sB = sA + static_cast<uint32_t>(p.a_terms) * kATileBytes;
          for (int t = 0; t < p.a_terms; ++t) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint64_t ad = umma_smem_desc_sw128(sA + t * kATileBytes + k * 32, 16, 1024);
              const uint64_t bd = umma_smem_desc_sw128(sB + k * 32, 16, 1024);
              umma_f16(d_tmem, ad, bd, idesc, (kb | t | k) != 0 ? 1u : 0u);
            }
          }
          umma_commit(smem_u32(&ctl->empty[stage]));
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        umma_commit(smem_u32(&ctl->tfull[as]));
        as ^= 1;
        if (as == 0) aphase ^= 1u;
      }
    }
  } else {
    // ------------------------------------------------------------ epilogue (warps 2..5)
    const int q = warp & 3;  // TMEM lane quarter this warp may read
    int as = 0;
    uint32_t aphase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int tm = tile / p.tiles_n;
      const int tn = tile - tm * p.tiles_n;
      const long long row = static_cast<long long>(tm) * 128 + q * 32 + lane;
      const int n0 = tn * p.block_n;
      mbar_wait(smem_u32(&ctl->tfull[as]), aphase);
      tc_fence_after();
      const uint32_t tbase =
          tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(as) * 256u;
      for (int c = 0; c < p.block_n / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(tbase + static_cast<uint32_t>(c * 32), r);
        tmem_ld_wait();
        const int col0 = n0 + c * 32;
        if (row < p.M && col0 < p.N) {
          const bool full_cols = (col0 + 32 <= p.N);
          if (p.out_kind == 0) {
            float* o = reinterpret_cast<float*>(p.out) + row * p.ldo + col0;
            if (full_cols) {
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                float4 v = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                       __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
                *reinterpret_cast<float4*>(o + 4 * j) = v;
              }
            } else {
              for (int j = 0; j < 32 && col0 + j < p.N; ++j) o[j] = __uint_as_float(r[j]);
            }
          } else {
            int16_t* o = reinterpret_cast<int16_t*>(p.out) + row * p.ldo + col0;
            if (full_cols) {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                uint32_t w[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const int lo = __float2int_rn(__uint_as_float(r[8 * j + 2 * e]));
                  const int hi = __float2int_rn(__uint_as_float(r[8 * j + 2 * e + 1]));
                  w[e] = (static_cast<uint32_t>(lo) & 0xFFFFu) | (static_cast<uint32_t>(hi) << 16);
                }
                *reinterpret_cast<uint4*>(o + 8 * j) = make_uint4(w[0], w[1], w[2], w[3]);
              }
            } else {
              for (int j = 0; j < 32 && col0 + j < p.N; ++j)
                o[j] = static_cast<int16_t>(__float2int_rn(__uint_as_float(r[j])));
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(smem_u32(&ctl->tempty[as]));
      as ^= 1;
      if (as == 0) aphase ^= 1u;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc_512(tmem_base);
}

// --------------------------------------------------------------------------------------------
struct WGemmParams {
  int Ma;       // rows of D per tap (reduced operand A has Ma channels), multiple of 128
  int N;        // columns of the B operand (multiple of 64)
  int n_valid;  // columns of D actually written (<= N)
  long long P;  // reduction length (pixels / batch rows)
  int b_terms;  // 1 or 2
  int block_n;  // 64, 128 or 256
  int a_im2col;
  int OH, OW, pad, kw_taps, taps;
  int ci_tiles, n_tiles, splits;
  int chunks_per_split;  // 64-row chunks of P
  int num_stages, stage_bytes;
  float* out;
  long long ldo;
};

struct SharedCtlW {
  uint64_t full[kMaxStages];
  uint64_t empty[kMaxStages];
  uint64_t tfull;
  uint32_t tmem_base;
  uint32_t pad_;
};

__global__ void __launch_bounds__(192, 1)
wgemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB0,
                  const __grid_constant__ CUtensorMap tmB1, const WGemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_1024(smem_raw);
  SharedCtlW* ctl =
      reinterpret_cast<SharedCtlW*>(smem + static_cast<size_t>(p.num_stages) * p.stage_bytes);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // work item -> (split, tap, ci_tile, n_tile); n fastest so that neighbours share the A stream.
  int wi = blockIdx.x;
  const int tn = wi % p.n_tiles;
  wi /= p.n_tiles;
  const int tc = wi % p.ci_tiles;
  wi /= p.ci_tiles;
  const int tap = wi % p.taps;
  const int split = wi / p.taps;

  const long long total_chunks = (p.P + 63) / 64;
  const long long chunk_lo = static_cast<long long>(split) * p.chunks_per_split;
  long long chunk_hi = chunk_lo + p.chunks_per_split;
  if (chunk_hi > total_chunks) chunk_hi = total_chunks;
  const int n_chunks = chunk_hi > chunk_lo ? static_cast<int>(chunk_hi - chunk_lo) : 0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(smem_u32(&ctl->full[s]), 1);
      mbar_init(smem_u32(&ctl->empty[s]), 1);
    }
    mbar_init(smem_u32(&ctl->tfull), 1);
    fence_mbar_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB0);
    if (p.b_terms == 2) tma_prefetch_desc(&tmB1);
  }
  if (warp == 1) tmem_alloc_512(smem_u32(&ctl->tmem_base));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = ctl->tmem_base;
  const int nb = p.block_n / 64;                 // 64-channel boxes of B per term
  const int b_term_bytes = nb * 8192;

  if (n_chunks > 0) {
    if (warp == 0) {
      if (lane == 0) {
        const int kh = tap / p.kw_taps;
        const int kw = tap - kh * p.kw_taps;
        int stage = 0;
        uint32_t phase = 0;
        for (int ch = 0; ch < n_chunks; ++ch) {
          const long long pix0 = (chunk_lo + ch) * 64;
          mbar_wait(smem_u32(&ctl->empty[stage]), phase ^ 1u);
          const uint32_t full = smem_u32(&ctl->full[stage]);
          mbar_expect_tx(full, static_cast<uint32_t>(p.stage_bytes));
          const uint32_t sA = smem_u32(smem + static_cast<size_t>(stage) * p.stage_bytes);
          const uint32_t sB = sA + 16384;
          if (p.a_im2col) {
            const int hw = p.OH * p.OW;
            const int img = static_cast<int>(pix0 / hw);
            const int rem = static_cast<int>(pix0 - static_cast<long long>(img) * hw);
            const int oh = rem / p.OW;
            const int ow = rem - oh * p.OW;
            for (int h = 0; h < 2; ++h)
              tma_load_im2col_4d(sA + h * 8192, &tmA, full, tc * 128 + h * 64, ow - p.pad,
                                 oh - p.pad, img, static_cast<uint16_t>(kw),
                                 static_cast<uint16_t>(kh));
          } else {
            for (int h = 0; h < 2; ++h)
              tma_load_2d(sA + h * 8192, &tmA, full, tc * 128 + h * 64, static_cast<int>(pix0));
          }
          for (int j = 0; j < nb; ++j) {
            tma_load_2d(sB + j * 8192, &tmB0, full, tn * p.block_n + j * 64,
                        static_cast<int>(pix0));
            if (p.b_terms == 2)
              tma_load_2d(sB + b_term_bytes + j * 8192, &tmB1, full, tn * p.block_n + j * 64,
                          static_cast<int>(pix0));
          }
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    } else if (warp == 1) {
      if (lane == 0) {
        const uint32_t idesc = umma_idesc_bf16(128, p.block_n, 1, 1);
        int stage = 0;
        uint32_t phase = 0;
        for (int ch = 0; ch < n_chunks; ++ch) {
          mbar_wait(smem_u32(&ctl->full[stage]), phase);
          tc_fence_after();
          const uint32_t sA = smem_u32(smem + static_cast<size_t>(stage) * p.stage_bytes);
          const uint32_t sB = sA + 16384;
          for (int t = 0; t < p.b_terms; ++t) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              // 16 reduction rows per MMA = 2 groups of 8 rows (1024 B each).
              const uint64_t ad = umma_smem_desc_sw128(sA + k * 2048, 8192, 1024);
              const uint64_t bd =
                  umma_smem_desc_sw128(sB + t * b_term_bytes + k * 2048, 8192, 1024);
              umma_f16(tmem_base, ad, bd, idesc, (ch | t | k) != 0 ? 1u : 0u);
            }
          }
          umma_commit(smem_u32(&ctl->empty[stage]));
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        umma_commit(smem_u32(&ctl->tfull));
      }
    } else {
      const int q = warp & 3;
      mbar_wait(smem_u32(&ctl->tfull), 0);
      tc_fence_after();
      const int row_in_tap = tc * 128 + q * 32 + lane;          // channel of A
      const long long orow = static_cast<long long>(tap) * p.Ma + row_in_tap;
      const uint32_t tbase = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
      for (int c = 0; c < p.block_n / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(tbase + static_cast<uint32_t>(c * 32), r);
        tmem_ld_wait();
        const int col0 = tn * p.block_n + c * 32;
        if (row_in_tap < p.Ma) {
          float* o = p.out + orow * p.ldo + col0;
          for (int j = 0; j < 32; ++j)
            if (col0 + j < p.n_valid) atomicAdd(o + j, __uint_as_float(r[j]));
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc_512(tmem_base);
}

}  // namespace zkb

// ============================================================================================
// Host launchers (internal) and C ABI
// ============================================================================================
using namespace zkb;

static int pick_block_n(int N) {
  if (N >= 256) return 256;
  if (N > 128) return 256;
  if (N > 64) return 128;
  if (N > 32) return 64;
  return 32;
}

// A_hi/A_lo: bf16.  im2col != 0: NHWC activation [NB,H,W,C], K = ksize*ksize*C, M = NB*H*W.
// Bmat: bf16 [N][K] row-major.  out: f32 or int16 [M][ldo].
static int launch_kgemm(const void* A_hi, const void* A_lo, const void* Bmat, void* out,
                        int out_kind, long long ldo, long long M, int N, int K, int im2col, int NB,
                        int H, int W, int C, int ksize, int pad, cudaStream_t stream) {
  ZKB_REQUIRE(A_hi && Bmat && out, "kgemm: null pointer");
  ZKB_REQUIRE(M > 0 && N > 0 && K > 0, "kgemm: empty problem M=%lld N=%d K=%d", M, N, K);
  ZKB_REQUIRE(K % 64 == 0, "kgemm: K=%d must be a multiple of 64", K);
  ZKB_REQUIRE(M < (1ll << 31) - 256, "kgemm: M=%lld too large", M);
  ZKB_REQUIRE(out_kind == 0 || out_kind == 1, "kgemm: out_kind must be 0 (f32) or 1 (i16)");
  ZKB_REQUIRE(ldo >= N, "kgemm: ldo=%lld < N=%d", ldo, N);
  ZKB_REQUIRE((ldo * (out_kind == 0 ? 4 : 2)) % 16 == 0 &&
                  (reinterpret_cast<uintptr_t>(out) & 15) == 0,
              "kgemm: output rows must be 16-byte aligned");
  KGemmParams p{};
  p.M = static_cast<int>(M);
  p.N = N;
  p.num_kb = K / 64;
  p.a_terms = A_lo ? 2 : 1;
  p.block_n = pick_block_n(N);
  p.stage_bytes = p.a_terms * kATileBytes + p.block_n * 128;
  p.num_stages = kSmemBudget / p.stage_bytes;
  if (p.num_stages > kMaxStages) p.num_stages = kMaxStages;
  ZKB_REQUIRE(p.num_stages >= 2, "kgemm: tile does not fit in shared memory");
  p.tiles_m = static_cast<int>((M + 127) / 128);
  p.tiles_n = (N + p.block_n - 1) / p.block_n;
  p.a_im2col = im2col;
  p.out_kind = out_kind;
  p.out = out;
  p.ldo = ldo;
  CUtensorMap tmA0, tmA1, tmB;
  int rc;
  if (im2col) {
    ZKB_REQUIRE(C % 64 == 0, "kgemm: channels C=%d must be a multiple of 64", C);
    ZKB_REQUIRE(K == ksize * ksize * C, "kgemm: K=%d != ksize^2*C", K);
    ZKB_REQUIRE(M == static_cast<long long>(NB) * H * W, "kgemm: M != NB*H*W");
    ZKB_REQUIRE(ksize == 2 * pad + 1, "kgemm: only 'same' windows (ksize == 2*pad+1)");
    p.OH = H;
    p.OW = W;
    p.pad = pad;
    p.kw_taps = ksize;
    p.cblocks = C / 64;
    rc = zkb_tmap_im2col_nhwc(&tmA0, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, A_hi, NB, H, W, C, ksize,
                              pad, 64, 128);
    if (rc) return rc;
    tmA1 = tmA0;
    if (A_lo) {
      rc = zkb_tmap_im2col_nhwc(&tmA1, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, A_lo, NB, H, W, C,
                                ksize, pad, 64, 128);
      if (rc) return rc;
    }
  } else {
    rc = zkb_tmap_tiled_2d(&tmA0, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, A_hi, K, M,
                           static_cast<uint64_t>(K) * 2, 64, 128);
    if (rc) return rc;
    tmA1 = tmA0;
    if (A_lo) {
      rc = zkb_tmap_tiled_2d(&tmA1, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, A_lo, K, M,
                             static_cast<uint64_t>(K) * 2, 64, 128);
      if (rc) return rc;
    }
  }
  rc = zkb_tmap_tiled_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, Bmat, K, N,
                         static_cast<uint64_t>(K) * 2, 64, p.block_n);
  if (rc) return rc;

  const size_t smem = static_cast<size_t>(p.num_stages) * p.stage_bytes + sizeof(SharedCtl) + 1024;
  ZKB_CUDA(cudaFuncSetAttribute(kgemm_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                227 * 1024));
  const int tiles = p.tiles_m * p.tiles_n;
  const int sms = zkb_sm_count();
  const int grid = tiles < sms ? tiles : sms;
  kgemm_bf16_kernel<<<grid, 192, smem, stream>>>(tmA0, tmA1, tmB, p);
  ZKB_LAUNCH_CHECK("kgemm_bf16_kernel");
  return ZKB_OK;
}

// A: bf16 [P][Ma] (plain) or NHWC activation [NB,H,W,Ma] viewed through tap (im2col).
// B_hi/B_lo: bf16 [P][N].  out f32 [(taps*Ma)][ldo], accumulated into (caller zeroes it).
static int launch_wgemm(const void* A, const void* B_hi, const void* B_lo, float* out,
                        long long ldo, int Ma, int N, int n_valid, long long P, int im2col, int NB,
                        int H, int W, int ksize, int pad, cudaStream_t stream) {
  ZKB_REQUIRE(A && B_hi && out, "wgemm: null pointer");
  ZKB_REQUIRE(Ma > 0 && N > 0 && P > 0, "wgemm: empty problem");
  ZKB_REQUIRE(Ma % 128 == 0, "wgemm: Ma=%d must be a multiple of 128", Ma);
  ZKB_REQUIRE(N % 64 == 0, "wgemm: N=%d must be a multiple of 64", N);
  ZKB_REQUIRE(P < (1ll << 31) - 256, "wgemm: P too large");
  ZKB_REQUIRE(n_valid > 0 && n_valid <= N && ldo >= n_valid, "wgemm: bad n_valid / ldo");
  WGemmParams p{};
  p.Ma = Ma;
  p.N = N;
  p.n_valid = n_valid;
  p.P = P;
  p.b_terms = B_lo ? 2 : 1;
  p.block_n = (N % 128 == 0) ? 128 : 64;
  p.a_im2col = im2col;
  p.taps = im2col ? ksize * ksize : 1;
  p.kw_taps = im2col ? ksize : 1;
  p.pad = pad;
  p.OH = H;
  p.OW = W;
  p.ci_tiles = Ma / 128;
  p.n_tiles = N / p.block_n;
  p.stage_bytes = 16384 + p.b_terms * (p.block_n / 64) * 8192;
  p.num_stages = kSmemBudget / p.stage_bytes;
  if (p.num_stages > kMaxStages) p.num_stages = kMaxStages;
  p.out = out;
  p.ldo = ldo;
  const long long chunks = (P + 63) / 64;
  const int out_tiles = p.taps * p.ci_tiles * p.n_tiles;
  const int sms = zkb_sm_count();
  long long splits = (2ll * sms + out_tiles - 1) / out_tiles;  // ~2 waves of work items
  if (splits > chunks) splits = chunks;
  if (splits < 1) splits = 1;
  p.chunks_per_split = static_cast<int>((chunks + splits - 1) / splits);
  p.splits = static_cast<int>((chunks + p.chunks_per_split - 1) / p.chunks_per_split);

  CUtensorMap tmA, tmB0, tmB1;
  int rc;
  if (im2col) {
    ZKB_REQUIRE(P == static_cast<long long>(NB) * H * W, "wgemm: P != NB*H*W");
    ZKB_REQUIRE(ksize == 2 * pad + 1, "wgemm: only 'same' windows (ksize == 2*pad+1)");
    rc = zkb_tmap_im2col_nhwc(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, A, NB, H, W, Ma, ksize,
                              pad, 64, 64);
  } else {
    rc = zkb_tmap_tiled_2d(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, A, Ma, P,
                           static_cast<uint64_t>(Ma) * 2, 64, 64);
  }
  if (rc) return rc;
  rc = zkb_tmap_tiled_2d(&tmB0, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, B_hi, N, P,
                         static_cast<uint64_t>(N) * 2, 64, 64);
  if (rc) return rc;
  tmB1 = tmB0;
  if (B_lo) {
    rc = zkb_tmap_tiled_2d(&tmB1, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, B_lo, N, P,
                           static_cast<uint64_t>(N) * 2, 64, 64);
    if (rc) return rc;
  }
  const size_t smem = static_cast<size_t>(p.num_stages) * p.stage_bytes + sizeof(SharedCtlW) + 1024;
  ZKB_CUDA(cudaFuncSetAttribute(wgemm_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                227 * 1024));
  const int grid = out_tiles * p.splits;
  wgemm_bf16_kernel<<<grid, 192, smem, stream>>>(tmA, tmB0, tmB1, p);
  ZKB_LAUNCH_CHECK("wgemm_bf16_kernel");
  return ZKB_OK;
}

// ------------------------------------------------------------------------------ C ABI
// (declared and documented in include/zkb200.h)
ZKB_API int zkb_bconv2d_fwd(const void* act_pm1, const void* wq_oc_k, void* y_i16, int NB, int H,
                            int W, int Cin, int Cout, int ksize, void* stream) {
  ZKB_REQUIRE(ksize % 2 == 1, "bconv2d_fwd: ksize must be odd");
  return launch_kgemm(act_pm1, nullptr, wq_oc_k, y_i16, 1, Cout,
                      static_cast<long long>(NB) * H * W, Cout, ksize * ksize * Cin, 1, NB, H, W,
                      Cin, ksize, ksize / 2, static_cast<cudaStream_t>(stream));
}

ZKB_API int zkb_bdense_fwd(const void* act_pm1, const void* wq_n_k, void* y_i16, long long ldy,
                           int rows, int K, int N, void* stream) {
  return launch_kgemm(act_pm1, nullptr, wq_n_k, y_i16, 1, ldy, rows, N, K, 0, 0, 0, 0, 0, 0, 0,
                      static_cast<cudaStream_t>(stream));
}

ZKB_API int zkb_bconv2d_dgrad(const void* dy_hi, const void* dy_lo, const void* wq_ic_k, void* dx_f32,
                              int NB, int H, int W, int Cin, int Cout, int ksize, void* stream) {
  ZKB_REQUIRE(ksize % 2 == 1, "bconv2d_dgrad: ksize must be odd");
  return launch_kgemm(dy_hi, dy_lo, wq_ic_k, dx_f32, 0, Cin, static_cast<long long>(NB) * H * W, Cin,
                      ksize * ksize * Cout, 1, NB, H, W, Cout, ksize, ksize / 2,
                      static_cast<cudaStream_t>(stream));
}

ZKB_API int zkb_bdense_dgrad(const void* dy_hi, const void* dy_lo, const void* wq_k_n, void* dx_f32,
                             int rows, int K, int N, void* stream) {
  // dx[rows,K] = dy[rows,N] * wq[K,N]^T : reduction over N, "B" operand is wq in its [K][N] layout.
  return launch_kgemm(dy_hi, dy_lo, wq_k_n, dx_f32, 0, K, rows, K, N, 0, 0, 0, 0, 0, 0, 0,
                      static_cast<cudaStream_t>(stream));
}

ZKB_API int zkb_bconv2d_wgrad(const void* act_pm1, const void* dy_hi, const void* dy_lo,
                              float* dw_hwio, int NB, int H, int W, int Cin, int Cout, int ksize,
                              void* stream) {
  ZKB_REQUIRE(ksize % 2 == 1, "bconv2d_wgrad: ksize must be odd");
  return launch_wgemm(act_pm1, dy_hi, dy_lo, dw_hwio, Cout, Cin, Cout, Cout,
                      static_cast<long long>(NB) * H * W, 1, NB, H, W, ksize, ksize / 2,
                      static_cast<cudaStream_t>(stream));
}

ZKB_API int zkb_bdense_wgrad(const void* act_pm1, const void* dy_hi, const void* dy_lo,
                             float* dw_k_n, long long ldw, int rows, int K, int N, int n_out,
                             void* stream) {
  return launch_wgemm(act_pm1, dy_hi, dy_lo, dw_k_n, ldw, K, N, n_out, rows, 0, 0, 0, 0, 1, 0,
                      static_cast<cudaStream_t>(stream));
}
