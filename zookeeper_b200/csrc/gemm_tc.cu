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
// Forward operands are int8 +-1 on kind::i8 with exact int32 accumulation (a bf16 kind::f16 form
// is kept as the cross-check: +-1 x +-1 with K <= 2^24 in f32 is the same integer); backward
// operands are bf16 (dy as a hi + lo pair, +-1 exact) with f32 accumulation in tensor memory.
// Replaces the Eigen/oneDNN float conv + matmul that
// TF runs for larq QuantConv2D / QuantDense (reference call sites:
// examples/larq_experiment.py:71-73,76-82,85-91,95-99).
#include <cuda_bf16.h>
#include <stdlib.h>

#include "zkb_host.h"
#include "zkb_ptx.cuh"

#ifndef ZKB_KGEMM_PLAIN_REGS
#define ZKB_KGEMM_PLAIN_REGS 144   // register cap of the plain kgemm kernel (see kgemm_bf16_kernel)
#endif

// Timing experiments (scripts/bench_gemm.py: drop the MMA issue or the TMA loads to see which side
// bounds a GEMM) exist only in a library built with -DZKB_MEASURE (ZKB_MEASURE=1 python -m
// zookeeper_b200.build); the product build has no such switch.
#ifdef ZKB_MEASURE
#define ZKB_DBG(p, bit) (((p).dbg & (bit)) != 0)
#else
#define ZKB_DBG(p, bit) false
#endif

namespace zkb {

static constexpr int kMaxStages = 8;
static constexpr int kATileBytes = 128 * 128;  // 128 rows x 64 bf16
static constexpr int kSmemBudget = 227 * 1024 - 2048;
static constexpr int kEpiStagingBytes = 8 * 4096;  // one 4 KB TMA-store box per epilogue warp

struct KGemmParams {
  int M, N;
  int kelems;    // K elements per 128-byte row: 64 (bf16) or 128 (int8)
  int num_kb;    // 64-element k-blocks per A term
  int a_terms;   // 1 or 2
  int block_n;   // 32..256, multiple of 32
  int num_stages;
  int stage_bytes;
  int tiles_m, tiles_n;
  int a_im2col;
  int OH, OW, pad, kw_taps, cblocks;   // OH, OW: OUTPUT map; pad: low zero padding along W
  int pad_h, stride;                   // low padding along H; traversal stride (1, or 2 for 'same' stride-2 convs)
  int out_kind;  // 0: f32, 1: int16 (exact integers expected), 2: int8 sign, 3: 2x2 max-pooled int16
  // out_kind 3 (swap mode, int8 operands): a tile is pool_rp whole row PAIRS of the [NB,H,W] output
  // map (H, W even, so row pairs are contiguous across images); the epilogue stages D^T as
  // [pixel][channel] int16 in shared memory, takes the 2x2 / stride-2 maximum with the first-max
  // tie rule of maxpool2_i16_kernel and writes the pooled map (out, row stride ldo), the winning
  // slot (pool_arg) and the BatchNorm statistics of the pooled map (st_sum / st_sumsq): the
  // full-resolution map never reaches memory.
  int pool_w;            // W of the conv output map
  int pool_rp;           // row pairs per tile
  long long pool_rows;   // NB * H / 2 row pairs in total
  uint8_t* pool_arg;
  int is_i8;     // operands are signed int8 (+-1 / 0), accumulators int32 (kind::i8)
  int dbg;       // ZKB_MEASURE builds only: 1 = no MMA issue, 2 = no TMA loads
  // Fused first-layer backward (swap mode, 128 output channels, one input channel, 3x3): the
  // epilogue consumes D^T = ds0 straight from tensor memory -- BatchNorm-backward sums and the
  // weight-gradient sums of the first layer -- and stores nothing.
  int l0_fused;
  int l0_H, l0_W;            // image size (the dgrad output is the (H-2) x (W-2) map)
  const float* l0_x;         // preprocessed images [B, H, W]
  const float* l0_wq;        // sign(W0) [9][128]
  const float* l0_mean;
  const float* l0_rstd;
  const float* l0_beta;
  double* l0_acc;            // [11][128]: sum da, sum da*xhat, S1[k] = sum x_k*da (9 taps)
  // BatchNorm-backward sums of the layer BELOW, fused into the f32 dgrad epilogue (swap mode and
  // CTA pairs): D[row][c] is the gradient w.r.t. that layer's sign output, bn_x its pre-BatchNorm
  // int16 map of the same shape; per channel sum da and sum da*xhat (da = D masked by |a| <= 1)
  // are added into bn_sum_da / bn_sum_dax -- what bn_bwd_reduce_kernel computes from D in HBM.
  // BatchNorm-forward statistics fused into the int16 epilogue (non-swap, one CTA per tile): per
  // output channel sum y and sum y^2 over all rows, exact integers, added into st_sum / st_sumsq
  // (float64) -- what bn_stats_kernel computes from y in a second pass.
  int st_fused;
  double *st_sum, *st_sumsq;
  // out_kind 2 (inference): BatchNorm + sign folded into one integer threshold per output channel;
  // the epilogue writes int8 +1 where the int32 accumulator >= sign_thr[c], else -1.
  const int* sign_thr;
  int bn_fused;
  const int16_t* bn_x;
  const float *bn_mean, *bn_rstd, *bn_beta;
  double *bn_sum_da, *bn_sum_dax;
  int swap_boxes;  // swap mode: pixel boxes per tile and their height (2 x 128)
  int swap_box_px;
  int swap;      // 1: operand roles swapped -- the MMA computes D^T[128 out channels, 256 pixels]
                 //    so that narrow (128-column) layers still issue 256-wide MMAs; f32 output only
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

// One epilogue unit of the fused first-layer backward: r[j] = gradient w.r.t. sign(a0) at pixel
// pix0 + j of this thread's channel.  Lane j works out where pixel pix0 + j's 3x3 window starts in
// the image batch (the div/mod once per unit, SIMD across the warp); the pixel loop then gets that
// offset with one shuffle and loads the nine taps with warp-uniform addresses (L1 broadcasts).
// Pixels at or beyond M have r[j] == 0 exactly (zero-filled operand columns): they add nothing and
// only their address is clamped.
__device__ __forceinline__ void first_layer_unit(const KGemmParams& p, const uint32_t (&r)[32], int pix0,
                                                 int lane, const float (&l0_w)[9], float l0_mu,
                                                 float l0_rs, float l0_be, float (&f)[11]) {
  unsigned pj = static_cast<unsigned>(pix0 + lane);
  if (pj >= static_cast<unsigned>(p.M)) pj = static_cast<unsigned>(p.M) - 1u;
  const unsigned hw = static_cast<unsigned>(p.OH * p.OW);
  const unsigned b = pj / hw;
  const unsigned rem = pj - b * hw;
  const unsigned oh = rem / static_cast<unsigned>(p.OW);
  const unsigned ow = rem - oh * static_cast<unsigned>(p.OW);
  const unsigned my_off = (b * static_cast<unsigned>(p.l0_H) + oh) * static_cast<unsigned>(p.l0_W) + ow;
  // Groups of four pixels: all 36 tap loads of a group are issued before the first FMA that
  // needs one (the FMAs otherwise sit on the loads' scoreboard: 2 epilogue warps per scheduler
  // cannot hide an L1 hit each).
#pragma unroll
  for (int j0 = 0; j0 < 32; j0 += 4) {
    float xv[4][9];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const float* q0 = p.l0_x + __shfl_sync(0xffffffffu, my_off, j0 + u);
      const float* q1 = q0 + p.l0_W;
      const float* q2 = q1 + p.l0_W;
      xv[u][0] = __ldg(q0); xv[u][1] = __ldg(q0 + 1); xv[u][2] = __ldg(q0 + 2);
      xv[u][3] = __ldg(q1); xv[u][4] = __ldg(q1 + 1); xv[u][5] = __ldg(q1 + 2);
      xv[u][6] = __ldg(q2); xv[u][7] = __ldg(q2 + 1); xv[u][8] = __ldg(q2 + 2);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float y = 0.f;
#pragma unroll
      for (int k = 0; k < 9; ++k) y = fmaf(xv[u][k], l0_w[k], y);   // same FMA order as the forward
      const float xh = (y - l0_mu) * l0_rs;
      const float da = (fabsf(xh + l0_be) > 1.0f) ? 0.f : __uint_as_float(r[j0 + u]);
      f[0] += da;
      f[1] = fmaf(da, xh, f[1]);
#pragma unroll
      for (int k = 0; k < 9; ++k) f[2 + k] = fmaf(xv[u][k], da, f[2 + k]);
    }
  }
}

// Column `lane` of the 32 x 32 f32 box a warp has just staged (row j at stage_box + 128 j, 16-byte
// chunks XOR-swizzled by j & 7): the gradient of 32 consecutive rows at one channel.  Adds this
// thread's share of the BatchNorm-backward sums (same arithmetic as bn_bwd_reduce_kernel).
__device__ __forceinline__ void bn_sums_from_box(const KGemmParams& p, uint32_t stage_box, int row0,
                                                 int ch, int lane, float mu, float rs, float be,
                                                 double& acc_da, double& acc_dax) {
  const uint32_t chunk = static_cast<uint32_t>(lane >> 2), byte = static_cast<uint32_t>(lane & 3) * 4u;
  // Rows at or beyond M hold exact zeros in the box (TMA zero-fills the operand rows), so they add
  // nothing; only their x address is clamped.  All 32 loads are issued before the first is used.
  const int last = p.M - 1 - row0;
  const int16_t* xcol = p.bn_x + static_cast<size_t>(row0) * p.ldo + ch;
  int16_t xs[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) xs[j] = __ldg(xcol + static_cast<size_t>(j < last ? j : last) * p.ldo);
  float sa = 0.f, sb = 0.f;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    float g;
    asm volatile("ld.shared.f32 %0, [%1];"
                 : "=f"(g)
                 : "r"(stage_box + static_cast<uint32_t>(j) * 128u +
                       ((chunk ^ static_cast<uint32_t>(j & 7)) << 4) + byte));
    const float xh = (static_cast<float>(xs[j]) - mu) * rs;
    const float da = (fabsf(xh + be) > 1.0f) ? 0.f : g;
    sa += da;
    sb = fmaf(da, xh, sb);
  }
  acc_da += sa;
  acc_dax += sb;
}

// kFirstLayer: the instantiation whose swap epilogue is the fused first-layer backward.  It needs
// 168 registers per thread; the plain instantiation is capped at 144 (no spills; 320 x 144 leaves
// 19 K registers of the SM free, enough for a 256-thread elementwise block of the other stream to
// share the SM with a resident GEMM CTA).
template <bool kFirstLayer>
__device__ __forceinline__ void kgemm_body(const CUtensorMap& tmA0, const CUtensorMap& tmA1,
                                           const CUtensorMap& tmB, const CUtensorMap& tmOut,
                                           const KGemmParams& p) {
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
      mbar_init(smem_u32(&ctl->tempty[a]), 256);
    }
    fence_mbar_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA0);
    if (p.a_terms == 2) tma_prefetch_desc(&tmA1);
    tma_prefetch_desc(&tmB);
    if (!kFirstLayer && p.out_kind != 3) tma_prefetch_desc(&tmOut);
  }
  if (warp == 1) tmem_alloc_512(smem_u32(&ctl->tmem_base));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = ctl->tmem_base;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    // (whole warp, an elected lane issues: coordinates stay in uniform registers)
    if (p.swap) {
      // stage layout: [W tile 128 x 128 B][X term 0: swap_boxes x 64 pixel rows][X term 1: ...]
      const int nbox = p.swap_boxes, box_px = p.swap_box_px;
      const uint32_t box_bytes = static_cast<uint32_t>(box_px) * 128u;
      const uint32_t term_bytes = static_cast<uint32_t>(nbox) * box_bytes;
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int tm = tile / p.tiles_n;
        const int tn = tile - tm * p.tiles_n;
        const int m0 = tm * nbox * box_px;
        const int c0 = tn * 128;
        int img[4] = {0, 0, 0, 0}, oh[4] = {0, 0, 0, 0}, ow[4] = {0, 0, 0, 0};
        int live[4] = {0, 0, 0, 0}, n_live = 0;
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          if (h >= nbox) break;
          const int mh = m0 + h * box_px;
          live[h] = mh < p.M;   // a box entirely beyond the last pixel is not loaded (its columns
                                // of D^T are never stored)
          n_live += live[h];
          if (p.a_im2col && live[h]) {
            const int hw = p.OH * p.OW;
            img[h] = mh / hw;
            const int rem = mh - img[h] * hw;
            oh[h] = rem / p.OW;
            ow[h] = rem - oh[h] * p.OW;
          }
        }
        // (ZKB_MEASURE, dbg bit 8: the second term loads one box instead of n_live -- the bytes an
        // 8-bit residual would need)
        const int lo_boxes = (ZKB_DBG(p, 8) && p.a_terms == 2) ? (n_live > 0 ? 1 : 0) : n_live;
        const uint32_t tx = static_cast<uint32_t>(kATileBytes) +
                            static_cast<uint32_t>((p.a_terms == 2 ? n_live + lo_boxes : n_live)) * box_bytes;
        int cb = 0, kw = 0, kh = 0;   // k-block -> (filter tap, channel block), advanced without divisions
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(smem_u32(&ctl->empty[stage]), phase ^ 1u);
          const uint32_t full = smem_u32(&ctl->full[stage]);
          if (ZKB_DBG(p, 2)) {
            if (elect_one()) mbar_arrive(full);
            __syncwarp();
            if (++stage == p.num_stages) { stage = 0; phase ^= 1u; }
            continue;
          }
          const uint32_t sW = smem_u32(smem + static_cast<size_t>(stage) * p.stage_bytes);
          const uint32_t sX = sW + kATileBytes;
          if (elect_one()) {
          mbar_expect_tx(full, tx);
          tma_load_2d(sW, &tmB, full, kb * p.kelems, c0);
          for (int t = 0; t < p.a_terms; ++t) {
            const CUtensorMap* mp = t == 0 ? &tmA0 : &tmA1;
#pragma unroll
            for (int h = 0; h < 4; ++h) {
              if (h >= nbox || !live[h]) continue;
              if (t == 1 && h >= lo_boxes) continue;
              const uint32_t dst = sX + static_cast<uint32_t>(t) * term_bytes + static_cast<uint32_t>(h) * box_bytes;
              if (p.a_im2col)
                tma_load_im2col_4d(dst, mp, full, cb * p.kelems, ow[h] * p.stride - p.pad, oh[h] * p.stride - p.pad_h, img[h],
                                   static_cast<uint16_t>(kw), static_cast<uint16_t>(kh));
              else
                tma_load_2d(dst, mp, full, kb * p.kelems, m0 + h * box_px);
            }
          }
          }
          __syncwarp();
          if (p.a_im2col && ++cb == p.cblocks) {
            cb = 0;
            if (++kw == p.kw_taps) { kw = 0; ++kh; }
          }
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    } else {
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
        int cb = 0, kw = 0, kh = 0;
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(smem_u32(&ctl->empty[stage]), phase ^ 1u);
          const uint32_t full = smem_u32(&ctl->full[stage]);
          if (ZKB_DBG(p, 2)) {
            if (elect_one()) mbar_arrive(full);
            __syncwarp();
            if (++stage == p.num_stages) { stage = 0; phase ^= 1u; }
            continue;
          }
          const uint32_t sA = smem_u32(smem + static_cast<size_t>(stage) * p.stage_bytes);
          const uint32_t sB = sA + static_cast<uint32_t>(p.a_terms) * kATileBytes;
          if (elect_one()) {
          mbar_expect_tx(full, static_cast<uint32_t>(p.stage_bytes));
          if (p.a_im2col) {
            tma_load_im2col_4d(sA, &tmA0, full, cb * p.kelems, ow * p.stride - p.pad, oh * p.stride - p.pad_h, img,
                               static_cast<uint16_t>(kw), static_cast<uint16_t>(kh));
            if (p.a_terms == 2)
              tma_load_im2col_4d(sA + kATileBytes, &tmA1, full, cb * p.kelems, ow * p.stride - p.pad, oh * p.stride - p.pad_h,
                                 img, static_cast<uint16_t>(kw), static_cast<uint16_t>(kh));
          } else {
            tma_load_2d(sA, &tmA0, full, kb * p.kelems, m0);
            if (p.a_terms == 2) tma_load_2d(sA + kATileBytes, &tmA1, full, kb * p.kelems, m0);
          }
          tma_load_2d(sB, &tmB, full, kb * p.kelems, n0);
          }
          __syncwarp();
          if (p.a_im2col && ++cb == p.cblocks) {
            cb = 0;
            if (++kw == p.kw_taps) { kw = 0; ++kh; }
          }
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    // (the whole warp runs the loop: warp-uniform values stay in uniform registers and each
    // tcgen05.mma is one instruction; an elected lane issues)
    {
      const uint32_t idesc = p.is_i8 ? umma_idesc_s8(128, p.block_n) : umma_idesc_bf16(128, p.block_n, 0, 0);  // swap: block_n = 256 pixels
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
          // Descriptors differ only in their start-address field (bytes >> 4): built once per
          // stage, then one add per MMA -- the single issuing thread has to stay ahead of the
          // tensor pipe.  swap: the weight tile (first 16 KB of the stage) is the M operand and
          // the 256 pixel rows of term t the N operand; otherwise term t's tile is the M operand
          // and the weight tile behind the terms the N operand.
          const uint64_t m_desc = umma_smem_desc_sw128(sA, 16, 1024);
          const uint64_t n_desc = umma_smem_desc_sw128(sA + (p.swap ? 1u : static_cast<uint32_t>(p.a_terms)) * kATileBytes, 16, 1024);
          const uint64_t m_step = p.swap ? 0u : (kATileBytes >> 4);
          const uint64_t n_step = p.swap ? (static_cast<uint32_t>(p.swap_boxes * p.swap_box_px) * 128u >> 4) : 0u;
          if (elect_one()) {
          if (!ZKB_DBG(p, 1)) {
            for (int t = 0; t < p.a_terms; ++t) {
              const uint64_t ad = m_desc + t * m_step, bd = n_desc + t * n_step;
              // a 128-byte row holds 64 bf16 or 128 int8: either way 4 MMAs of 32 bytes of K
              // (ZKB_MEASURE, dbg bit 4: the second term issues HALF its MMAs -- the instruction
              // count an e4m3 residual on kind::f8f6f4 would need; timing only, results are wrong)
              const int nk = (ZKB_DBG(p, 4) && t == 1) ? 2 : 4;
              if (p.is_i8) {
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_i8(d_tmem, ad + 2 * k, bd + 2 * k, idesc, (kb | t | k) != 0 ? 1u : 0u);
              } else {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  if (k < nk) umma_f16(d_tmem, ad + 2 * k, bd + 2 * k, idesc, (kb | t | k) != 0 ? 1u : 0u);
              }
            }
          }
          umma_commit(smem_u32(&ctl->empty[stage]));
          }
          __syncwarp();
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        if (elect_one()) umma_commit(smem_u32(&ctl->tfull[as]));
        __syncwarp();
        as ^= 1;
        if (as == 0) aphase ^= 1u;
      }
    }
  } else {
    // ------------------------------------------------------------ epilogue (warps 2..9)
    // Two warps per TMEM lane quarter, alternating 128-byte-wide column units.  A unit is read
    // from TMEM (tcgen05.ld 32x32b), converted, written to a 4 KB swizzled staging box in shared
    // memory and stored with one TMA store per warp: full 128-byte lines, rows/columns beyond
    // M/N are clipped by the TMA unit.
    const int ew = warp - 2;
    const int q = warp & 3;        // TMEM lane quarter this warp may read
    const int half = ew >> 2;
    const int unit_cols = p.out_kind == 0 ? 32 : (p.out_kind == 1 ? 64 : 128);
    const int units = (p.block_n + unit_cols - 1) / unit_cols;
    const uint32_t stage_box = smem_u32(smem + static_cast<size_t>(p.num_stages) * p.stage_bytes +
                                        1024 + static_cast<size_t>(ew) * 4096);
    const uint32_t my_row = stage_box + static_cast<uint32_t>(lane) * 128u;
    const uint32_t sw = static_cast<uint32_t>(lane & 7);
    // fused first-layer backward: this thread owns output channel l0_c for the whole kernel
    const int l0_c = q * 32 + lane;
    float l0_w[kFirstLayer ? 9 : 1], l0_mu = 0.f, l0_rs = 0.f, l0_be = 0.f;
    double l0_d[kFirstLayer ? 11 : 1];
    float l0_f[kFirstLayer ? 11 : 1];     // float partial sums of one tile (<= 128 pixels), then double
    if constexpr (kFirstLayer) {
#pragma unroll
      for (int k = 0; k < 9; ++k) l0_w[k] = p.l0_wq[k * 128 + l0_c];
#pragma unroll
      for (int i = 0; i < 11; ++i) l0_d[i] = 0.0;
      l0_mu = p.l0_mean[l0_c]; l0_rs = p.l0_rstd[l0_c]; l0_be = p.l0_beta[l0_c];
    }
    // fused BatchNorm-backward sums (swap mode): the grid is a multiple of tiles_n, so this CTA's
    // tn -- and with it this thread's channel -- is the same for every tile
    const int bn_ch = (static_cast<int>(blockIdx.x) % p.tiles_n) * 128 + q * 32 + lane;
    float bn_mu = 0.f, bn_rs = 0.f, bn_be = 0.f;
    double bn_da = 0.0, bn_dax = 0.0;
    if (p.bn_fused) { bn_mu = p.bn_mean[bn_ch]; bn_rs = p.bn_rstd[bn_ch]; bn_be = p.bn_beta[bn_ch]; }
    // fused forward statistics: 2 units of 64 columns per warp and tile, a column pair per lane
    long long st_s[2][2] = {{0, 0}, {0, 0}}, st_q[2][2] = {{0, 0}, {0, 0}};
    long long pool_s[4] = {0, 0, 0, 0}, pool_q[4] = {0, 0, 0, 0};   // fused pooling: statistics of this lane's 4 channels
    int as = 0;
    uint32_t aphase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int tm = tile / p.tiles_n;
      const int tn = tile - tm * p.tiles_n;
      const int row0 = tm * 128 + q * 32;
      const int n0 = tn * p.block_n;
      mbar_wait(smem_u32(&ctl->tfull[as]), aphase);
      tc_fence_after();
      const uint32_t tbase =
          tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(as) * 256u;
      if (p.swap) {
        // D^T: TMEM lane = output channel, column = pixel.  A unit is 32 pixels x this warp's 32
        // channels, transposed through the staging box (row = pixel, 4 bytes per lane).
        const int ch0 = tn * 128 + q * 32;
        if (!kFirstLayer && p.out_kind == 3) {
          // ---- fused 2x2 max-pool.  Slab of this lane quarter: [block_n pixels][32 channels] int16.
          const uint32_t slab = smem_u32(smem + static_cast<size_t>(p.num_stages) * p.stage_bytes + 1024) +
                                static_cast<uint32_t>(q) * static_cast<uint32_t>(p.block_n) * 64u;
          const int n_units = (p.block_n + 31) >> 5;
          for (int u = half; u < n_units; u += 2) {
            if (tm * p.block_n + u * 32 >= p.M) break;
            uint32_t r[32];
            tmem_ld_32x32(tbase + static_cast<uint32_t>(u * 32), r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              if (u * 32 + j < p.block_n)
                asm volatile("st.shared.b16 [%0], %1;" ::"r"(slab + static_cast<uint32_t>(u * 32 + j) * 64u +
                                                             static_cast<uint32_t>(lane) * 2u),
                             "h"(static_cast<short>(static_cast<int>(r[j])))
                             : "memory");
            }
          }
          tc_fence_before();
          mbar_arrive(smem_u32(&ctl->tempty[as]));              // the MMAs of the next tile may start
          named_bar_sync(1u + static_cast<uint32_t>(q), 64u);   // both warps of the quarter have staged
          // Pooling: a lane owns 4 channels (8 bytes of a slab row) of one of the 4 pooled pixels a
          // warp takes per iteration; the two warps of the quarter interleave groups of 4 pixels.
          // Maximum and first-max slot are computed on int16 pairs with the SIMD video instructions.
          const int W2 = p.pool_w, PW = W2 >> 1;
          const int n_pp = p.pool_rp * PW;                      // pooled pixels of a tile
          const int cg = lane & 7, sub = lane >> 3;
          int s_acc[4] = {0, 0, 0, 0}, q_acc[4] = {0, 0, 0, 0};
          for (int pp0 = half * 4; pp0 < n_pp; pp0 += 8) {
            const int pp = pp0 + sub;
            const int u2 = pp / PW, j2 = pp - u2 * PW;
            const long long grp = static_cast<long long>(tm) * p.pool_rp + u2;   // global row pair
            if (pp >= n_pp || grp >= p.pool_rows) continue;
            const uint32_t a = slab + static_cast<uint32_t>(u2 * 2 * W2 + 2 * j2) * 64u + static_cast<uint32_t>(cg) * 8u;
            uint32_t v[4][2];
            asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v[0][0]), "=r"(v[0][1]) : "r"(a));
            asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v[1][0]), "=r"(v[1][1]) : "r"(a + 64u));
            asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v[2][0]), "=r"(v[2][1]) : "r"(a + static_cast<uint32_t>(W2) * 64u));
            asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v[3][0]), "=r"(v[3][1]) : "r"(a + static_cast<uint32_t>(W2) * 64u + 64u));
            uint32_t best[2], slot[2];
#pragma unroll
            for (int w2 = 0; w2 < 2; ++w2) {
              uint32_t bv = v[0][w2], kv = 0u;
#pragma unroll
              for (int k = 1; k < 4; ++k) {
                const uint32_t m = __vcmpgts2(v[k][w2], bv);      // 0xFFFF where window element k is strictly greater
                bv = (v[k][w2] & m) | (bv & ~m);
                kv = ((static_cast<uint32_t>(k) * 0x00010001u) & m) | (kv & ~m);
              }
              best[w2] = bv;
              slot[w2] = kv;
            }
            const long long o = (grp * PW + j2) * p.ldo + ch0 + cg * 4;
            *reinterpret_cast<uint2*>(static_cast<int16_t*>(p.out) + o) = make_uint2(best[0], best[1]);
            *reinterpret_cast<uint32_t*>(p.pool_arg + o) = (slot[0] & 0xFFu) | ((slot[0] >> 16) << 8) |
                                                           ((slot[1] & 0xFFu) << 16) | ((slot[1] >> 16) << 24);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              const int val = static_cast<int16_t>((best[c >> 1] >> (16 * (c & 1))) & 0xFFFFu);
              s_acc[c] += val;
              q_acc[c] += val * val;    // <= 8 pooled pixels x 4608^2 per lane and tile: fits int32
            }
          }
#pragma unroll
          for (int c = 0; c < 4; ++c) { pool_s[c] += s_acc[c]; pool_q[c] += q_acc[c]; }
          named_bar_sync(1u + static_cast<uint32_t>(q), 64u);   // slab free for the next tile
          as ^= 1;
          if (as == 0) aphase ^= 1u;
          continue;
        }
        if constexpr (kFirstLayer) {
#pragma unroll
          for (int i = 0; i < 11; ++i) l0_f[i] = 0.f;
        }
        for (int u = half; u < p.block_n / 32; u += 2) {
          const int pix0 = tm * p.block_n + u * 32;
          if (pix0 >= p.M) break;
          uint32_t r[32];
          tmem_ld_32x32(tbase + static_cast<uint32_t>(u * 32), r);
          tmem_ld_wait();
          if constexpr (kFirstLayer) {
            // r[j] = gradient w.r.t. sign(a0) at pixel pix0 + j, channel l0_c.  Recompute the
            // first conv at that pixel (9 taps, same FMA order as the forward kernels), apply the
            // ste mask and accumulate everything the first layer's backward needs.
            first_layer_unit(p, r, pix0, lane, l0_w, l0_mu, l0_rs, l0_be, l0_f);
            continue;
          }
          if (p.bn_fused) {   // thread = channel: r[j] is the gradient at pixel pix0 + j
            // (pixels at or beyond M: r[j] is an exact zero, only the x address is clamped)
            const int last = p.M - 1 - pix0;
            const int16_t* xcol = p.bn_x + static_cast<size_t>(pix0) * p.ldo + ch0 + lane;
            int16_t xs[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) xs[j] = __ldg(xcol + static_cast<size_t>(j < last ? j : last) * p.ldo);
            float sa = 0.f, sb = 0.f;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float xh = (static_cast<float>(xs[j]) - bn_mu) * bn_rs;
              const float da = (fabsf(xh + bn_be) > 1.0f) ? 0.f : __uint_as_float(r[j]);
              sa += da;
              sb = fmaf(da, xh, sb);
            }
            bn_da += sa;
            bn_dax += sb;
          }
          if (lane == 0) tma_store_wait_read_all();
          __syncwarp();
          if (p.out_kind == 1) {
            // int16 output (the int8 forward of a 128-channel layer): rows of 32 channels = 64
            // bytes, no swizzle; for a given pixel the 32 lanes write one contiguous row
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int v = p.is_i8 ? static_cast<int>(r[j]) : __float2int_rn(__uint_as_float(r[j]));
              asm volatile("st.shared.b16 [%0], %1;" ::"r"(stage_box + static_cast<uint32_t>(j) * 64u +
                                                           static_cast<uint32_t>(lane) * 2u),
                           "h"(static_cast<short>(v))
                           : "memory");
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
              tma_store_2d(&tmOut, stage_box, ch0, pix0);
              tma_store_commit();
            }
            continue;
          }
          const uint32_t lane_chunk = static_cast<uint32_t>(lane >> 2);
          const uint32_t lane_byte = static_cast<uint32_t>(lane & 3) * 4u;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const uint32_t addr = stage_box + static_cast<uint32_t>(j) * 128u +
                                  ((lane_chunk ^ static_cast<uint32_t>(j & 7)) << 4) + lane_byte;
            asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(r[j]) : "memory");
          }
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tmOut, stage_box, ch0, pix0);
            tma_store_commit();
          }
        }
        tc_fence_before();
        mbar_arrive(smem_u32(&ctl->tempty[as]));
        if constexpr (kFirstLayer) {   // (after the accumulator stage is released)
#pragma unroll
          for (int i = 0; i < 11; ++i) l0_d[i] += l0_f[i];
        }
        as ^= 1;
        if (as == 0) aphase ^= 1u;
        continue;
      }
#pragma unroll
      for (int sl = 0; sl < 4; ++sl) {
        const int u = half + 2 * sl;
        if (u >= units) break;
        const int c0 = u * unit_cols;
        uint32_t r[32];
        tmem_ld_32x32(tbase + static_cast<uint32_t>(c0), r);
        tmem_ld_wait();
        if (lane == 0) tma_store_wait_read_all();   // previous box of this warp has left smem
        __syncwarp();
        if (p.out_kind == 2) {
          // 128 int8 signs per row: four accumulator loads of 32 columns, two 16-byte chunks each
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            if (h > 0) {
              if (c0 + 32 * h >= p.block_n) break;
              tmem_ld_32x32(tbase + static_cast<uint32_t>(c0 + 32 * h), r);
              tmem_ld_wait();
            }
            const int4* thr = reinterpret_cast<const int4*>(p.sign_thr + n0 + c0 + 32 * h);
            uint32_t w[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const int4 t = __ldg(thr + e);    // same address in every lane: one broadcast load
              w[e] = ((static_cast<int>(r[4 * e]) >= t.x) ? 0x01u : 0xFFu) |
                     ((static_cast<int>(r[4 * e + 1]) >= t.y) ? 0x0100u : 0xFF00u) |
                     ((static_cast<int>(r[4 * e + 2]) >= t.z) ? 0x010000u : 0xFF0000u) |
                     ((static_cast<int>(r[4 * e + 3]) >= t.w) ? 0x01000000u : 0xFF000000u);
            }
            st_shared_v4(my_row + ((static_cast<uint32_t>(2 * h) ^ sw) << 4), w[0], w[1], w[2], w[3]);
            st_shared_v4(my_row + ((static_cast<uint32_t>(2 * h + 1) ^ sw) << 4), w[4], w[5], w[6], w[7]);
          }
        } else if (p.out_kind == 0) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            st_shared_v4(my_row + ((static_cast<uint32_t>(j) ^ sw) << 4), r[4 * j], r[4 * j + 1],
                         r[4 * j + 2], r[4 * j + 3]);
        } else {
          uint32_t w[16];
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            const int lo = p.is_i8 ? static_cast<int>(r[2 * e]) : __float2int_rn(__uint_as_float(r[2 * e]));
            const int hi = p.is_i8 ? static_cast<int>(r[2 * e + 1])
                                   : __float2int_rn(__uint_as_float(r[2 * e + 1]));
            w[e] = (static_cast<uint32_t>(lo) & 0xFFFFu) | (static_cast<uint32_t>(hi) << 16);
          }
#pragma unroll
          for (int j = 0; j < 4; ++j)
            st_shared_v4(my_row + ((static_cast<uint32_t>(j) ^ sw) << 4), w[4 * j], w[4 * j + 1],
                         w[4 * j + 2], w[4 * j + 3]);
          if (c0 + 32 < p.block_n) {   // second half of the 64-column int16 unit
            tmem_ld_32x32(tbase + static_cast<uint32_t>(c0 + 32), r);
            tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 16; ++e) {
              const int lo = p.is_i8 ? static_cast<int>(r[2 * e]) : __float2int_rn(__uint_as_float(r[2 * e]));
              const int hi = p.is_i8 ? static_cast<int>(r[2 * e + 1])
                                     : __float2int_rn(__uint_as_float(r[2 * e + 1]));
              w[e] = (static_cast<uint32_t>(lo) & 0xFFFFu) | (static_cast<uint32_t>(hi) << 16);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j)
              st_shared_v4(my_row + ((static_cast<uint32_t>(j + 4) ^ sw) << 4), w[4 * j], w[4 * j + 1],
                           w[4 * j + 2], w[4 * j + 3]);
          }
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&tmOut, stage_box, n0 + c0, row0);
          tma_store_commit();
        }
        if (p.st_fused && sl < 2) {
          // columns 2*lane, 2*lane+1 of the staged 32 x 64 int16 box (rows >= M are exact zeros)
          const uint32_t col = stage_box + ((static_cast<uint32_t>(lane >> 2)) << 4) +
                               static_cast<uint32_t>(lane & 3) * 4u;
          int s0 = 0, s1 = 0, q0 = 0, q1 = 0;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            uint32_t w2;
            asm volatile("ld.shared.b32 %0, [%1];"
                         : "=r"(w2)
                         : "r"((col ^ (static_cast<uint32_t>(j & 7) << 4)) + static_cast<uint32_t>(j) * 128u));
            const int a = static_cast<int16_t>(w2 & 0xFFFFu), b = static_cast<int>(w2) >> 16;
            s0 += a; s1 += b;
            q0 += a * a; q1 += b * b;      // 32 rows x 4608^2 < 2^31
          }
          st_s[sl][0] += s0; st_s[sl][1] += s1;
          st_q[sl][0] += q0; st_q[sl][1] += q1;
        }
      }
      tc_fence_before();
      mbar_arrive(smem_u32(&ctl->tempty[as]));
      as ^= 1;
      if (as == 0) aphase ^= 1u;
    }
    if (p.st_fused) {
      const int n0c = (static_cast<int>(blockIdx.x) % p.tiles_n) * p.block_n;   // this CTA's tn is fixed
#pragma unroll
      for (int sl = 0; sl < 2; ++sl) {
        const int ch = n0c + (half + 2 * sl) * 64 + 2 * lane;
        if ((half + 2 * sl) * 64 < p.block_n && ch < p.N) {
          atomicAdd(p.st_sum + ch, static_cast<double>(st_s[sl][0]));
          atomicAdd(p.st_sumsq + ch, static_cast<double>(st_q[sl][0]));
          if (ch + 1 < p.N) {
            atomicAdd(p.st_sum + ch + 1, static_cast<double>(st_s[sl][1]));
            atomicAdd(p.st_sumsq + ch + 1, static_cast<double>(st_q[sl][1]));
          }
        }
      }
    }
    if (!kFirstLayer && p.out_kind == 3) {     // this CTA's tn is fixed (grid is a multiple of tiles_n)
      // the four pixel lanes (lane >> 3) of a channel group are combined by shuffles first
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        long long a = pool_s[c], b2 = pool_q[c];
        a += __shfl_xor_sync(0xffffffffu, a, 8);  b2 += __shfl_xor_sync(0xffffffffu, b2, 8);
        a += __shfl_xor_sync(0xffffffffu, a, 16); b2 += __shfl_xor_sync(0xffffffffu, b2, 16);
        if (lane < 8) {
          const int ch = (static_cast<int>(blockIdx.x) % p.tiles_n) * 128 + q * 32 + lane * 4 + c;
          atomicAdd(p.st_sum + ch, static_cast<double>(a));
          atomicAdd(p.st_sumsq + ch, static_cast<double>(b2));
        }
      }
    }
    if constexpr (kFirstLayer) {
#pragma unroll
      for (int i = 0; i < 11; ++i) atomicAdd(p.l0_acc + i * 128 + l0_c, l0_d[i]);
    }
    if (p.bn_fused) {
      atomicAdd(p.bn_sum_da + bn_ch, bn_da);
      atomicAdd(p.bn_sum_dax + bn_ch, bn_dax);
    }
    if (lane == 0) tma_store_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc_512(tmem_base);
}

__global__ void __maxnreg__(ZKB_KGEMM_PLAIN_REGS)
kgemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                  const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmOut,
                  const KGemmParams p) {
  kgemm_body<false>(tmA0, tmA1, tmB, tmOut, p);
}

__global__ void __launch_bounds__(320, 1)
kgemm_first_bf16_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                        const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmOut,
                        const KGemmParams p) {
  kgemm_body<true>(tmA0, tmA1, tmB, tmOut, p);
}

// Per channel c and tap k: dbeta = sum da; dW0[k][c] = rstd_c * (S1 - mean(da) * S2[k] -
// mean(da*xhat) * S3[k][c]) from the sums the fused epilogue left.  S2[k] = sum x_k and
// S3[k][c] = sum x_k * xhat_c = rstd_c * (sum_k' w_k'c G[k][k'] - mean_c S2[k]) do not depend on the
// gradient: they come from the Gram matrix of the input taps (gram = G[9][9] then S2[9]).
__global__ void conv_first_fused_finalize_kernel(const double* __restrict__ acc,
                                                 const double* __restrict__ gram,
                                                 const float* __restrict__ wq,
                                                 const float* __restrict__ mean,
                                                 const float* __restrict__ rstd, double count,
                                                 double dbeta_scale, float* __restrict__ dbeta,
                                                 float* __restrict__ dw) {
  const int c = threadIdx.x;   // 128 threads
  const double mean_da = acc[c] / count;
  const double mean_dax = acc[128 + c] / count;
  const double rs = rstd[c], mu = mean[c];
  dbeta[c] = static_cast<float>(acc[c] * dbeta_scale);
  for (int k = 0; k < 9; ++k) {
    double wg = 0.0;
    for (int k2 = 0; k2 < 9; ++k2) wg += static_cast<double>(wq[k2 * 128 + c]) * gram[k * 9 + k2];
    const double s2 = gram[81 + k];
    const double s3 = rs * (wg - mu * s2);
    const double s1 = acc[(2 + k) * 128 + c];
    dw[k * 128 + c] += static_cast<float>(rs * (s1 - mean_da * s2 - mean_dax * s3));
  }
}

// --------------------------------------------------------------------------------------------
// kgemm2: the same K-major GEMM on CTA PAIRS (cta_group::2).  A cluster of two CTAs owns a
// 256 x 256 output tile: each CTA loads the A rows of its own 128-row half and ONE HALF of the
// B tile (128 of the 256 columns); a single tcgen05.mma.cta_group::2 issued by the leader CTA
// drives both SMs' tensor cores, each reading its A tile and both B halves.  Per SM that halves
// the shared-memory traffic of the B operand and the bytes that have to be in flight.
// Barriers: the producers of both CTAs count their TMA bytes on the LEADER's full barrier; the
// leader's commits are multicast to the empty / accumulator-full barriers of both CTAs; both
// epilogues release the accumulator on the leader's barrier.
// --------------------------------------------------------------------------------------------
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(320, 1)
kgemm2_bf16_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                   const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmOut,
                   const KGemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_1024(smem_raw);
  SharedCtl* ctl = reinterpret_cast<SharedCtl*>(smem + static_cast<size_t>(p.num_stages) * p.stage_bytes);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int num_tiles = p.tiles_m * p.tiles_n;          // 256 x 256 tiles
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(smem_u32(&ctl->full[s]), 1);
      mbar_init(smem_u32(&ctl->empty[s]), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(smem_u32(&ctl->tfull[a]), 1);
      mbar_init(smem_u32(&ctl->tempty[a]), 512);          // 256 epilogue threads of each CTA
    }
    fence_mbar_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA0);
    if (p.a_terms == 2) tma_prefetch_desc(&tmA1);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmOut);
  }
  if (warp == 1) tmem2_alloc_512(smem_u32(&ctl->tmem_base));
  tc_fence_before();
  cluster_sync_all();          // barriers of BOTH CTAs are initialised before anyone signals them
  tc_fence_after();
  const uint32_t tmem_base = ctl->tmem_base;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer (both CTAs)
    {   // whole warp, an elected lane issues
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        const int tm = tile / p.tiles_n;
        const int tn = tile - tm * p.tiles_n;
        const int m0 = tm * 256 + static_cast<int>(rank) * 128;   // this CTA's rows
        const int nb0 = tn * 256 + static_cast<int>(rank) * 128;  // this CTA's half of the B tile
        const bool live = m0 < p.M;
        int img = 0, oh = 0, ow = 0;
        if (p.a_im2col && live) {
          const int hw = p.OH * p.OW;
          img = m0 / hw;
          const int rem = m0 - img * hw;
          oh = rem / p.OW;
          ow = rem - oh * p.OW;
        }
        int cb = 0, kw = 0, kh = 0;
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(smem_u32(&ctl->empty[stage]), phase ^ 1u);
          const uint32_t full_local = smem_u32(&ctl->full[stage]);
          const uint32_t full_leader = mapa_shared(full_local, 0);
          if (elect_one()) {
          // the leader expects the bytes of both CTAs; a CTA whose rows lie entirely beyond M
          // still loads (zero-filled / clipped by TMA) so that the count is the same every stage
          if (leader) mbar_expect_tx(full_local, 2u * static_cast<uint32_t>(p.stage_bytes));
          const uint32_t sA = smem_u32(smem + static_cast<size_t>(stage) * p.stage_bytes);
          const uint32_t sB = sA + static_cast<uint32_t>(p.a_terms) * kATileBytes;
          if (p.a_im2col && live) {
            tma2_load_im2col_4d(sA, &tmA0, full_leader, cb * p.kelems, ow * p.stride - p.pad, oh * p.stride - p.pad_h, img,
                                static_cast<uint16_t>(kw), static_cast<uint16_t>(kh));
            if (p.a_terms == 2)
              tma2_load_im2col_4d(sA + kATileBytes, &tmA1, full_leader, cb * p.kelems, ow * p.stride - p.pad,
                                  oh * p.stride - p.pad_h, img, static_cast<uint16_t>(kw), static_cast<uint16_t>(kh));
          } else {
            // plain rows, or an im2col tile entirely past the end: a 2D load whose rows are out
            // of bounds delivers zeros and still completes the expected bytes
            const CUtensorMap* m0p = p.a_im2col ? &tmB : &tmA0;
            const CUtensorMap* m1p = p.a_im2col ? &tmB : &tmA1;
            const int row = p.a_im2col ? 0x3FFFFFFF : m0;
            tma2_load_2d(sA, m0p, full_leader, kb * p.kelems, row);
            if (p.a_terms == 2) tma2_load_2d(sA + kATileBytes, m1p, full_leader, kb * p.kelems, row);
          }
          tma2_load_2d(sB, &tmB, full_leader, kb * p.kelems, nb0);
          }
          __syncwarp();
          if (p.a_im2col && ++cb == p.cblocks) {
            cb = 0;
            if (++kw == p.kw_taps) { kw = 0; ++kh; }
          }
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (leader CTA only)
    if (leader) {   // whole warp, an elected lane issues (see kgemm_body)
      const uint32_t idesc = p.is_i8 ? umma_idesc_s8(256, 256) : umma_idesc_bf16(256, 256, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        mbar_wait(smem_u32(&ctl->tempty[as]), aphase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(as) * 256u;
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(smem_u32(&ctl->full[stage]), phase);
          tc_fence_after();
          const uint32_t sA = smem_u32(smem + static_cast<size_t>(stage) * p.stage_bytes);
          const uint64_t a_desc = umma_smem_desc_sw128(sA, 16, 1024);
          const uint64_t bd = umma_smem_desc_sw128(sA + static_cast<uint32_t>(p.a_terms) * kATileBytes, 16, 1024);
          if (elect_one()) {
          for (int t = 0; t < p.a_terms; ++t) {
            const uint64_t ad = a_desc + static_cast<uint64_t>(t) * (kATileBytes >> 4);
            const int nk = (ZKB_DBG(p, 4) && t == 1) ? 2 : 4;     // see kgemm_body
            if (p.is_i8) {
#pragma unroll
              for (int k = 0; k < 4; ++k) umma2_i8(d_tmem, ad + 2 * k, bd + 2 * k, idesc, (kb | t | k) != 0 ? 1u : 0u);
            } else {
#pragma unroll
              for (int k = 0; k < 4; ++k)
                if (k < nk) umma2_f16(d_tmem, ad + 2 * k, bd + 2 * k, idesc, (kb | t | k) != 0 ? 1u : 0u);
            }
          }
          umma2_commit(smem_u32(&ctl->empty[stage]));   // frees the slot in both CTAs
          }
          __syncwarp();
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        if (elect_one()) umma2_commit(smem_u32(&ctl->tfull[as]));          // accumulator ready in both CTAs
        __syncwarp();
        as ^= 1;
        if (as == 0) aphase ^= 1u;
      }
    }
  } else {
    // ------------------------------------------------------------ epilogue (warps 2..9, both CTAs)
    const int ew = warp - 2;
    const int q = warp & 3;
    const int half = ew >> 2;
    const int unit_cols = p.out_kind == 0 ? 32 : 64;
    const int units = 256 / unit_cols;
    const uint32_t stage_box = smem_u32(smem + static_cast<size_t>(p.num_stages) * p.stage_bytes +
                                        1024 + static_cast<size_t>(ew) * 4096);
    const uint32_t my_row = stage_box + static_cast<uint32_t>(lane) * 128u;
    const uint32_t sw = static_cast<uint32_t>(lane & 7);
    // fused BatchNorm-backward sums: the number of clusters is a multiple of tiles_n, so tn is the
    // same for every tile of this cluster and the 4 units of this warp are 4 fixed channel groups
    const int bn_n0 = (cluster_id % p.tiles_n) * 256;
    float bn_mu[4], bn_rs[4], bn_be[4];
    double bn_da[4], bn_dax[4];
#pragma unroll
    for (int sl = 0; sl < 4; ++sl) {
      bn_da[sl] = bn_dax[sl] = 0.0;
      bn_mu[sl] = bn_rs[sl] = bn_be[sl] = 0.f;
      if (p.bn_fused) {
        const int ch = bn_n0 + (half + 2 * sl) * 32 + lane;
        bn_mu[sl] = p.bn_mean[ch]; bn_rs[sl] = p.bn_rstd[ch]; bn_be[sl] = p.bn_beta[ch];
      }
    }
    int as = 0;
    uint32_t aphase = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      const int tm = tile / p.tiles_n;
      const int tn = tile - tm * p.tiles_n;
      const int row0 = tm * 256 + static_cast<int>(rank) * 128 + q * 32;
      const int n0 = tn * 256;
      mbar_wait(smem_u32(&ctl->tfull[as]), aphase);
      tc_fence_after();
      const uint32_t tbase =
          tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(as) * 256u;
#pragma unroll
      for (int sl = 0; sl < 4; ++sl) {   // (unrolled: the fused-sum accumulators stay in registers)
        const int u = half + 2 * sl;
        if (u >= units) break;
        const int c0 = u * unit_cols;
        if (row0 >= p.M || n0 + c0 >= p.N) continue;   // nothing of this box would be stored
        uint32_t r[32];
        tmem_ld_32x32(tbase + static_cast<uint32_t>(c0), r);
        tmem_ld_wait();
        if (lane == 0) tma_store_wait_read_all();
        __syncwarp();
        if (p.out_kind == 0) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            st_shared_v4(my_row + ((static_cast<uint32_t>(j) ^ sw) << 4), r[4 * j], r[4 * j + 1],
                         r[4 * j + 2], r[4 * j + 3]);
        } else {
#pragma unroll
          for (int hseg = 0; hseg < 2; ++hseg) {
            if (hseg == 1) {
              tmem_ld_32x32(tbase + static_cast<uint32_t>(c0 + 32), r);
              tmem_ld_wait();
            }
            uint32_t w[16];
#pragma unroll
            for (int e = 0; e < 16; ++e) {
              const int lo = p.is_i8 ? static_cast<int>(r[2 * e]) : __float2int_rn(__uint_as_float(r[2 * e]));
              const int hi = p.is_i8 ? static_cast<int>(r[2 * e + 1])
                                     : __float2int_rn(__uint_as_float(r[2 * e + 1]));
              w[e] = (static_cast<uint32_t>(lo) & 0xFFFFu) | (static_cast<uint32_t>(hi) << 16);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j)
              st_shared_v4(my_row + ((static_cast<uint32_t>(j + 4 * hseg) ^ sw) << 4), w[4 * j],
                           w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
          }
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&tmOut, stage_box, n0 + c0, row0);
          tma_store_commit();
        }
        if (p.bn_fused)
          bn_sums_from_box(p, stage_box, row0, n0 + c0 + lane, lane, bn_mu[sl], bn_rs[sl], bn_be[sl],
                           bn_da[sl], bn_dax[sl]);
      }
      tc_fence_before();
      // release the accumulator stage on the leader's barrier (remote arrive from the peer CTA)
      mbar_arrive_cluster(mapa_shared(smem_u32(&ctl->tempty[as]), 0));
      as ^= 1;
      if (as == 0) aphase ^= 1u;
    }
    if (p.bn_fused) {
#pragma unroll
      for (int sl = 0; sl < 4; ++sl) {
        const int ch = bn_n0 + (half + 2 * sl) * 32 + lane;
        atomicAdd(p.bn_sum_da + ch, bn_da[sl]);
        atomicAdd(p.bn_sum_dax + ch, bn_dax[sl]);
      }
    }
    if (lane == 0) tma_store_wait_all();
  }

  __syncwarp();         // re-converge the role warps: the cluster barrier below is .aligned
  tc_fence_before();
  cluster_sync_all();   // neither CTA may leave while the other can still signal its barriers
  if (warp == 1) tmem2_dealloc_512(tmem_base);
}

// --------------------------------------------------------------------------------------------
struct WGemmParams {
  int Ma;       // rows of D per tap (reduced operand A has Ma channels), multiple of 64
  int N;        // columns of the B operand (multiple of 64)
  int n_valid;  // columns of D actually written (<= N)
  long long P;  // reduction length (pixels / batch rows)
  int b_terms;  // 1 or 2 (dy = hi + lo)
  int block_n;  // N of one MMA: 64, 128 or 256
  int concat;   // 1: the MMA's 256 columns are [hi(128) | lo(128)] of a 128-wide layer; the two
                //    halves of the accumulator are added in the epilogue
  int cols_per_tile;  // output columns one work item covers (128 when concat, else block_n)
  int a_im2col;
  int OH, OW, pad, kw_taps, taps;   // OH, OW: the conv's OUTPUT map (rows of dy); pad: low padding along W
  int pad_h, stride;
  int tpi;      // M tiles per work item: they share the B (dy) tile, one TMEM accumulator each
  int mt;       // M tiles along the tap axis: taps, or ceil(units / 2) when flat64
  int flat64;   // Ma is an odd multiple of 64: D's rows [tap][Ma] are cut into 64-row units
                //   u = tap * ci64s + c64, and an M tile is the units (2j, 2j+1) -- its two 64-channel
                //   A boxes may come from different taps; rows past the last unit are not written
  int units, ci64s;
  int kpix;     // reduction rows per pipeline stage (32, 48 or 64)
  int dbg;      // ZKB_MEASURE builds only: 1 = no MMA issue, 2 = no TMA loads
  int ci_tiles, n_tiles, splits;
  int chunks_per_split;  // kpix-row chunks of P
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

// 256-column MMAs run the tensor pipe ~1.7x more efficiently than 128-column ones (operand
// fetch from shared memory is amortised over twice the columns), so a work item always issues
// N = 256 where the layer allows: 256 output channels at once, or, for 128-wide layers, the hi
// and lo halves of dy side by side.
__global__ void __launch_bounds__(192, 1)
wgemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB0,
                  const __grid_constant__ CUtensorMap tmB1, const WGemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_1024(smem_raw);
  SharedCtlW* ctl =
      reinterpret_cast<SharedCtlW*>(smem + static_cast<size_t>(p.num_stages) * p.stage_bytes);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // work item -> (split, tap group, ci_tile, n_tile)
  int wi = blockIdx.x;
  const int tn = wi % p.n_tiles;
  wi /= p.n_tiles;
  const int tc = wi % p.ci_tiles;
  wi /= p.ci_tiles;
  const int groups = (p.mt + p.tpi - 1) / p.tpi;
  const int group = wi % groups;     // all tap groups of a split are neighbours in launch order:
  const int split = wi / groups;     // they read the same pixels at about the same time (L2 reuse)
  const int tap0 = group * p.tpi;    // first M tile (= tap unless flat64) of this item
  const int tcount = (p.mt - tap0) < p.tpi ? (p.mt - tap0) : p.tpi;  // M tiles of this item

  const long long total_chunks = (p.P + p.kpix - 1) / p.kpix;
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
  const int box_bytes = p.kpix * 128;       // one [64 channels x kpix rows] box
  const int a_tile_bytes = 2 * box_bytes;   // 128 channels of A
  const int a_bytes = p.tpi * a_tile_bytes;
  const int boxes_per_term = p.concat ? 2 : p.block_n / 64;
  const int nB = p.b_terms * boxes_per_term;          // B boxes per stage
  const int b_term_bytes = boxes_per_term * box_bytes;
  const int mma_terms = p.concat ? 1 : p.b_terms;      // concat: hi and lo ride in one MMA

  if (n_chunks > 0) {
    if (warp == 0) {
      // Producer warp: lane l owns TMA load l of a stage so that they are issued in parallel
      // (measured r1: one elected lane issuing the 10-12 loads of a stage in turn is 10-20 % slower).
      // l < nB -> B box (term = l / boxes_per_term); l >= nB -> A box (M tile = (l-nB)/2, half = (l-nB)%2).
      const int n_loads = nB + 2 * tcount;
      const uint32_t tx_bytes = static_cast<uint32_t>(n_loads * box_bytes);
      const bool is_b = lane < nB;
      const int bt = is_b ? lane / boxes_per_term : 0, bj = is_b ? lane % boxes_per_term : 0;
      const int at = is_b ? 0 : (lane - nB) >> 1, ah = is_b ? 0 : (lane - nB) & 1;
      int tap = tap0 + at;
      int a_c = tc * 128 + ah * 64;
      if (p.flat64) {   // half ah of M tile j is unit 2j + ah (the half past the last unit re-reads it)
        int u = 2 * (tap0 + at) + ah;
        if (u >= p.units) u = p.units - 1;
        tap = u / p.ci64s;
        a_c = (u - tap * p.ci64s) * 64;
      }
      const int kh = tap / p.kw_taps;
      const int kw = tap - kh * p.kw_taps;
      const uint32_t my_off = is_b ? static_cast<uint32_t>(a_bytes + (bt * boxes_per_term + bj) * box_bytes)
                                   : static_cast<uint32_t>(at * a_tile_bytes + ah * box_bytes);
      const CUtensorMap* my_map = is_b ? (bt == 0 ? &tmB0 : &tmB1) : &tmA;
      const int my_c = is_b ? tn * p.cols_per_tile + bj * 64 : a_c;
      const unsigned hw = static_cast<unsigned>(p.OH * p.OW);
      int stage = 0;
      uint32_t phase = 0;
      for (int ch = 0; ch < n_chunks; ++ch) {
        const unsigned pix0 = static_cast<unsigned>((chunk_lo + ch) * p.kpix);
        const uint32_t full = smem_u32(&ctl->full[stage]);
        if (lane == 0) {
          mbar_wait(smem_u32(&ctl->empty[stage]), phase ^ 1u);
          if (ZKB_DBG(p, 2)) mbar_arrive(full);
          else mbar_expect_tx(full, tx_bytes);
        }
        __syncwarp();
        if (lane < n_loads && !ZKB_DBG(p, 2)) {
          const uint32_t dst = smem_u32(smem + static_cast<size_t>(stage) * p.stage_bytes) + my_off;
          if (is_b || !p.a_im2col) {
            tma_load_2d(dst, my_map, full, my_c, static_cast<int>(pix0));
          } else {
            const unsigned img = pix0 / hw;
            const unsigned rem = pix0 - img * hw;
            const unsigned oh = rem / static_cast<unsigned>(p.OW);
            const unsigned ow = rem - oh * static_cast<unsigned>(p.OW);
            tma_load_im2col_4d(dst, my_map, full, my_c, static_cast<int>(ow) * p.stride - p.pad,
                               static_cast<int>(oh) * p.stride - p.pad_h, static_cast<int>(img),
                               static_cast<uint16_t>(kw), static_cast<uint16_t>(kh));
          }
        }
        if (++stage == p.num_stages) {
          stage = 0;
          phase ^= 1u;
        }
      }
    } else if (warp == 1) {
      // The whole warp runs this loop so that every value is warp-uniform (descriptors live in
      // uniform registers and each tcgen05.mma is a single instruction instead of a per-thread
      // operand loop); one elected lane issues the MMAs and the commits.
      {
        const uint32_t idesc = umma_idesc_bf16(128, p.block_n, 1, 1);
        const int ksteps = p.kpix / 16;
        int stage = 0;
        uint32_t phase = 0;
        for (int ch = 0; ch < n_chunks; ++ch) {
          mbar_wait(smem_u32(&ctl->full[stage]), phase);
          tc_fence_after();
          const uint32_t sA = smem_u32(smem + static_cast<size_t>(stage) * p.stage_bytes);
          const uint32_t sB = sA + a_bytes;
          // Descriptors differ only in their 14-bit start-address field: one add per MMA instead of
          // rebuilding them (the single issuing thread must stay ahead of a 128-cycle MMA).
          const uint64_t a_desc0 = umma_smem_desc_sw128(sA, box_bytes, 1024);
          const uint64_t b_desc0 = umma_smem_desc_sw128(sB, box_bytes, 1024);
          if (elect_one()) {
          for (int t = 0; t < tcount; ++t) {
            const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(t * p.block_n);
            const uint64_t a_desc = a_desc0 + static_cast<uint64_t>((t * a_tile_bytes) >> 4);
            for (int bt = 0; bt < mma_terms; ++bt) {
              const uint64_t b_desc = b_desc0 + static_cast<uint64_t>((bt * b_term_bytes) >> 4);
              // 16 reduction rows per MMA = 2 groups of 8 rows (1024 B each): 2048 B per k-step
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                if (ZKB_DBG(p, 4) && bt == 1 && 2 * k >= ksteps) continue;   // see kgemm_body
                if (k < ksteps && !ZKB_DBG(p, 1))
                  umma_f16(d_tmem, a_desc + static_cast<uint64_t>(k * 128), b_desc + static_cast<uint64_t>(k * 128),
                           idesc, (ch | bt | k) != 0 ? 1u : 0u);
              }
            }
          }
          umma_commit(smem_u32(&ctl->empty[stage]));
          }
          __syncwarp();
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        if (elect_one()) umma_commit(smem_u32(&ctl->tfull));
        __syncwarp();
      }
    } else {
      const int q = warp & 3;
      mbar_wait(smem_u32(&ctl->tfull), 0);
      tc_fence_after();
      const int row_in_tap = tc * 128 + q * 32 + lane;  // channel of A
      for (int t = 0; t < tcount; ++t) {
        const long long orow = p.flat64 ? static_cast<long long>(tap0 + t) * 128 + q * 32 + lane
                                        : static_cast<long long>(tap0 + t) * p.Ma + row_in_tap;
        const bool row_ok = p.flat64 ? orow < static_cast<long long>(p.units) * 64 : row_in_tap < p.Ma;
        const uint32_t tbase =
            tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(t * p.block_n);
        for (int c = 0; c < p.cols_per_tile / 32; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(tbase + static_cast<uint32_t>(c * 32), r);
          tmem_ld_wait();
          if (p.concat) {   // add the lo half of the accumulator
            uint32_t r2[32];
            tmem_ld_32x32(tbase + static_cast<uint32_t>(128 + c * 32), r2);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j)
              r[j] = __float_as_uint(__uint_as_float(r[j]) + __uint_as_float(r2[j]));
          }
          const int col0 = tn * p.cols_per_tile + c * 32;
          if (row_ok) {
            float* o = p.out + orow * p.ldo + col0;
            if (col0 + 32 <= p.n_valid && (p.ldo & 3) == 0 && (reinterpret_cast<uintptr_t>(p.out) & 15) == 0) {
              // whole 32-column chunk, 16-byte aligned rows: vector reductions (8 instead of 32)
#pragma unroll
              for (int j = 0; j < 32; j += 4)
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(o + j),
                             "f"(__uint_as_float(r[j])), "f"(__uint_as_float(r[j + 1])),
                             "f"(__uint_as_float(r[j + 2])), "f"(__uint_as_float(r[j + 3]))
                             : "memory");
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (col0 + j < p.n_valid) atomicAdd(o + j, __uint_as_float(r[j]));
            }
          }
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

// Tile-shape overrides for A/B measurements (scripts/bench_gemm.py).  Read from the environment
// once, and only in a -DZKB_MEASURE build: the product library has no environment switches.
struct GemmTuning {
  int dbg = 0;          // ZKB_GEMM_DBG
  int pairs = -1;       // ZKB_KGEMM_PAIRS   (-1: default rule)
  int swap = -1;        // ZKB_KGEMM_SWAP
  int wgrad_tpi = 0;    // ZKB_WGRAD_TPI     (0: default rule)
  int wgrad_kpix = 0;   // ZKB_WGRAD_KPIX
  int wgrad_waves = 0;  // ZKB_WGRAD_WAVES
};

static const GemmTuning& gemm_tuning() {
  static const GemmTuning t = [] {
    GemmTuning g;
#ifdef ZKB_MEASURE
    if (const char* e = getenv("ZKB_GEMM_DBG")) g.dbg = atoi(e);
    if (const char* e = getenv("ZKB_KGEMM_PAIRS")) g.pairs = atoi(e) != 0;
    if (const char* e = getenv("ZKB_KGEMM_SWAP")) g.swap = atoi(e) != 0;
    if (const char* e = getenv("ZKB_WGRAD_TPI")) g.wgrad_tpi = atoi(e);
    if (const char* e = getenv("ZKB_WGRAD_KPIX")) g.wgrad_kpix = atoi(e) == 32 ? 32 : (atoi(e) == 48 ? 48 : 64);
    if (const char* e = getenv("ZKB_WGRAD_WAVES")) g.wgrad_waves = atoi(e) > 0 ? atoi(e) : 0;
#endif
    return g;
  }();
  return t;
}

// 'same' padding of a stride-s conv as TensorFlow defines it: out = ceil(in / s), total padding
// max((out - 1) s + k - in, 0), the smaller half in front.  s == 1: (k - 1) / 2 on both sides.
struct SameGeom {
  int out, lo, hi;
};
static SameGeom same_geometry(int in, int ksize, int stride) {
  SameGeom g;
  g.out = (in + stride - 1) / stride;
  int total = (g.out - 1) * stride + ksize - in;
  if (total < 0) total = 0;
  g.lo = total / 2;
  g.hi = total - g.lo;
  return g;
}

static int pick_block_n(int N) {
  if (N >= 256) return 256;
  if (N > 128) return 256;
  if (N > 64) return 128;
  if (N > 32) return 64;
  return 32;
}

// A_hi/A_lo: bf16.  im2col != 0: NHWC activation [NB,H,W,C], K = ksize*ksize*C, M = NB*H*W.
// Bmat: bf16 [N][K] row-major.  out: f32 or int16 [M][ldo].
struct FirstLayerFuse {   // see KGemmParams::l0_fused
  const float *x, *wq, *mean, *rstd, *beta;
  double* acc;
  int H, W;
};

struct BnBackwardFuse {   // see KGemmParams::bn_fused
  const int16_t* x;
  const float *mean, *rstd, *beta;
  double *sum_da, *sum_dax;
};

struct StatsFuse {   // see KGemmParams::st_fused
  double *sum, *sumsq;
};

struct PoolFuse {    // see KGemmParams::out_kind == 3
  uint8_t* arg;
};

// Row pairs per tile of the fused-pool forward: the most that fit 256 MMA columns with a pixel count
// the MMA accepts (a multiple of 16) and whose last 32-column accumulator read stays inside the
// 256-column stage.  0: this map cannot be tiled that way (odd sizes, rows wider than 128 pixels).
static int pool_row_pairs(int H, int W) {
  if (H <= 0 || W <= 0 || (H & 1) || (W & 1) || 2 * W > 256) return 0;
  for (int rp = 256 / (2 * W); rp >= 1; --rp) {
    const int px = rp * 2 * W;
    if (px % 16 == 0 && ((px + 31) / 32) * 32 <= 256) return rp;
  }
  return 0;
}

extern "C" int zkb_bn_stats(const void* x, int x_is_i16, long long rows, int C, double* sum,
                            double* sumsq, void* stream);
extern "C" int zkb_bn_bwd_reduce(const float* g, const void* x, int x_is_i16, long long rows, int C,
                                 const float* mean, const float* rstd, const float* beta, int ste,
                                 double* sum_da, double* sum_dax, void* stream);

static int launch_kgemm(const void* A_hi, const void* A_lo, const void* Bmat, void* out,
                        int out_kind, long long ldo, long long M, int N, int K, int im2col, int NB,
                        int H, int W, int C, int ksize, int pad, cudaStream_t stream,
                        int elem_bytes = 2, const FirstLayerFuse* l0 = nullptr,
                        const BnBackwardFuse* bn = nullptr, const StatsFuse* st = nullptr,
                        const int* sign_thr = nullptr, int stride = 1, const PoolFuse* pool = nullptr) {
  const int kelems = 128 / elem_bytes;  // K elements per 128-byte smem row
  const CUtensorMapDataType dt =
      elem_bytes == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  ZKB_REQUIRE(elem_bytes == 2 || (elem_bytes == 1 && !A_lo && (out_kind == 1 || out_kind == 2)),
              "kgemm: int8 operands are for the single-term integer forward only");
  ZKB_REQUIRE(A_hi && Bmat && (out || l0), "kgemm: null pointer");
  ZKB_REQUIRE(M > 0 && N > 0 && K > 0, "kgemm: empty problem M=%lld N=%d K=%d", M, N, K);
  ZKB_REQUIRE(K % kelems == 0, "kgemm: K=%d must be a multiple of %d", K, kelems);
  ZKB_REQUIRE(M < (1ll << 31) - 256, "kgemm: M=%lld too large", M);
  ZKB_REQUIRE(out_kind >= 0 && out_kind <= 2, "kgemm: out_kind must be 0 (f32), 1 (i16) or 2 (i8 sign)");
  ZKB_REQUIRE(!pool || (out_kind == 1 && elem_bytes == 1 && im2col && stride == 1 && st && pool->arg && N % 128 == 0 &&
                        ldo == N && pool_row_pairs(H, W) > 0),
              "kgemm: the fused max-pool needs the int8 conv forward with statistics, Cout %% 128 == 0 and an even map "
              "whose row pairs tile 256 columns (H=%d W=%d)", H, W);
  ZKB_REQUIRE(out_kind != 2 || (sign_thr && elem_bytes == 1 && (reinterpret_cast<uintptr_t>(sign_thr) & 15) == 0),
              "kgemm: the sign output needs int8 operands and 16-byte aligned thresholds");
  ZKB_REQUIRE(ldo >= N, "kgemm: ldo=%lld < N=%d", ldo, N);
  ZKB_REQUIRE((ldo * (out_kind == 0 ? 4 : (out_kind == 1 ? 2 : 1))) % 16 == 0 &&
                  (reinterpret_cast<uintptr_t>(out) & 15) == 0,
              "kgemm: output rows must be 16-byte aligned");
  ZKB_REQUIRE(!l0 || (out_kind == 0 && elem_bytes == 2 && N == 128 && im2col),
              "kgemm: the fused first-layer backward needs a 128-channel f32 dgrad");
  KGemmParams p{};
  p.M = static_cast<int>(M);
  p.N = N;
  p.num_kb = K / kelems;
  p.kelems = kelems;
  const GemmTuning& tune = gemm_tuning();
  p.dbg = tune.dbg;
  p.is_i8 = elem_bytes == 1;
  p.a_terms = A_lo ? 2 : 1;
  p.block_n = pick_block_n(N);
  p.stage_bytes = p.a_terms * kATileBytes + p.block_n * 128;
  // narrow f32-output layers (dgrad into 128 input channels): swap operand roles so that the
  // MMAs are 256 wide (see KGemmParams::swap)
  p.swap = (out_kind == 0 && elem_bytes == 2 && N % 128 == 0 && N < 256) ? 1 : 0;
  // ... and the int8 forward of a 128-channel layer with int16 output (no fused statistics there)
  if (out_kind == 1 && elem_bytes == 1 && N == 128 && ldo == N && !st) p.swap = 1;
  if (tune.swap == 0) p.swap = 0;
  else if (tune.swap == 1 && out_kind == 0 && elem_bytes == 2 && N % 128 == 0) p.swap = 1;
  if (l0) {
    p.swap = 1;
    p.l0_fused = 1;
    p.l0_x = l0->x; p.l0_wq = l0->wq; p.l0_mean = l0->mean; p.l0_rstd = l0->rstd; p.l0_beta = l0->beta;
    p.l0_acc = l0->acc; p.l0_H = l0->H; p.l0_W = l0->W;
  }
  if (pool) p.swap = 1;
  if (p.swap) {
    // 256 pixels per tile as two 128-pixel boxes (192 pixels as three 64-pixel boxes, which fits
    // three two-term stages instead of two, measured no faster in r1 and was removed)
    p.swap_boxes = 2;
    p.swap_box_px = 128;
    if (pool) {      // one box of pool_rp whole row pairs
      p.pool_rp = pool_row_pairs(H, W);
      p.pool_w = W;
      p.pool_rows = static_cast<long long>(NB) * (H / 2);
      p.pool_arg = pool->arg;
      p.swap_boxes = 1;
      p.swap_box_px = p.pool_rp * 2 * W;
    }
    p.block_n = p.swap_boxes * p.swap_box_px;
    p.stage_bytes = kATileBytes + p.a_terms * p.block_n * 128;
  }
  // CTA pairs (cta_group::2) for layers at least 256 columns wide
  // (measured r1: +3..5 % for the two-term dgrad, -8 % for the int8 forward whose stages are
  // smaller -- so pairs are the default for two-term problems only)
  bool use_pairs = !p.swap && N % 256 == 0 && p.a_terms == 2;
  if (tune.pairs >= 0) use_pairs = tune.pairs != 0 && !p.swap && N % 256 == 0;
  if (use_pairs) {
    p.block_n = 256;
    p.stage_bytes = p.a_terms * kATileBytes + kATileBytes;   // A tile(s) + this CTA's half of B
  }
  if (bn) {
    ZKB_REQUIRE(out_kind == 0 && elem_bytes == 2 && !l0 && ldo == N,
                "kgemm: fused BatchNorm-backward sums need a dense f32 dgrad output");
    if (p.swap || use_pairs) {
      p.bn_fused = 1;
      p.bn_x = bn->x; p.bn_mean = bn->mean; p.bn_rstd = bn->rstd; p.bn_beta = bn->beta;
      p.bn_sum_da = bn->sum_da; p.bn_sum_dax = bn->sum_dax;
    }
  }
  if (out_kind == 2) {
    ZKB_REQUIRE(N % p.block_n == 0, "kgemm: sign output needs N=%d to be a multiple of the %d-column tile",
                N, p.block_n);
    ZKB_REQUIRE(!p.swap && !use_pairs, "kgemm: sign output is built for the one-CTA int8 forward");
    p.sign_thr = sign_thr;
  }
  if (st) {
    ZKB_REQUIRE(out_kind == 1 && ldo == N && st->sum && st->sumsq,
                "kgemm: fused statistics need a dense int16 output and both accumulators");
    if (pool) {          // statistics of the POOLED map, accumulated by the pooling epilogue
      p.st_sum = st->sum;
      p.st_sumsq = st->sumsq;
    } else if (!p.swap && !use_pairs) {
      p.st_fused = 1;
      p.st_sum = st->sum;
      p.st_sumsq = st->sumsq;
    }
  }
  // epilogue staging: eight 4 KB TMA-store boxes, or the four [pixels][32 channels] int16 slabs of
  // the pooling epilogue
  const int epi_bytes = pool ? 4 * p.block_n * 64 : kEpiStagingBytes;
  p.num_stages = (kSmemBudget - epi_bytes) / p.stage_bytes;
  if (p.num_stages > kMaxStages) p.num_stages = kMaxStages;
  ZKB_REQUIRE(p.num_stages >= 2, "kgemm: tile does not fit in shared memory");
  p.tiles_m = static_cast<int>((M + 127) / 128);
  p.tiles_n = (N + p.block_n - 1) / p.block_n;
  if (p.swap) {
    p.tiles_m = static_cast<int>((M + p.block_n - 1) / p.block_n);
    p.tiles_n = N / 128;
  }
  if (pool) out_kind = 3;
  if (use_pairs) {
    p.tiles_m = static_cast<int>((M + 255) / 256);
    p.tiles_n = N / 256;
  }
  p.a_im2col = im2col;
  p.out_kind = out_kind;
  p.out = out;
  p.ldo = ldo;
  CUtensorMap tmA0, tmA1, tmB;
  int rc;
  if (im2col) {
    ZKB_REQUIRE(C % kelems == 0, "kgemm: channels C=%d must be a multiple of %d", C, kelems);
    ZKB_REQUIRE(K == ksize * ksize * C, "kgemm: K=%d != ksize^2*C", K);
    ZKB_REQUIRE(ksize == 2 * pad + 1, "kgemm: only 'same' windows (ksize == 2*pad+1)");
    ZKB_REQUIRE(stride == 1 || stride == 2, "kgemm: stride must be 1 or 2");
    const SameGeom gh = same_geometry(H, ksize, stride), gw = same_geometry(W, ksize, stride);
    ZKB_REQUIRE(M == static_cast<long long>(NB) * gh.out * gw.out, "kgemm: M != NB*OH*OW");
    p.OH = gh.out;
    p.OW = gw.out;
    p.pad = gw.lo;
    p.pad_h = gh.lo;
    p.stride = stride;
    p.kw_taps = ksize;
    p.cblocks = C / kelems;
    const int abox = p.swap ? p.swap_box_px : 128;   // pixel rows per A box
    rc = zkb_tmap_im2col_nhwc_ex(&tmA0, dt, elem_bytes, A_hi, NB, H, W, C, ksize, gh.lo, gh.hi, gw.lo, gw.hi,
                                 stride, kelems, abox);
    if (rc) return rc;
    tmA1 = tmA0;
    if (A_lo) {
      rc = zkb_tmap_im2col_nhwc_ex(&tmA1, dt, elem_bytes, A_lo, NB, H, W, C, ksize, gh.lo, gh.hi, gw.lo, gw.hi,
                                   stride, kelems, abox);
      if (rc) return rc;
    }
  } else {
    const int abox = p.swap ? p.swap_box_px : 128;
    rc = zkb_tmap_tiled_2d(&tmA0, dt, elem_bytes, A_hi, K, M, static_cast<uint64_t>(K) * elem_bytes,
                           kelems, abox);
    if (rc) return rc;
    tmA1 = tmA0;
    if (A_lo) {
      rc = zkb_tmap_tiled_2d(&tmA1, dt, elem_bytes, A_lo, K, M, static_cast<uint64_t>(K) * elem_bytes,
                             kelems, abox);
      if (rc) return rc;
    }
  }
  rc = zkb_tmap_tiled_2d(&tmB, dt, elem_bytes, Bmat, K, N, static_cast<uint64_t>(K) * elem_bytes, kelems,
                         (p.swap || use_pairs) ? 128 : p.block_n);
  if (rc) return rc;

  // output through TMA stores: boxes of 128 bytes x 32 rows
  CUtensorMap tmOut{};
  if (l0 || pool)
    rc = ZKB_OK;            // nothing goes through TMA stores: the epilogue reduces / pools D itself
  else if (out_kind == 0)
    rc = zkb_tmap_tiled_2d(&tmOut, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, out, N, M,
                           static_cast<uint64_t>(ldo) * 4, 32, 32);
  else if (out_kind == 1 && p.swap)
    rc = zkb_tmap_tiled_2d_plain(&tmOut, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, out, N, M,
                                 static_cast<uint64_t>(ldo) * 2, 32, 32);
  else if (out_kind == 1)
    rc = zkb_tmap_tiled_2d(&tmOut, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, out, N, M,
                           static_cast<uint64_t>(ldo) * 2, 64, 32);
  else
    rc = zkb_tmap_tiled_2d(&tmOut, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, out, N, M,
                           static_cast<uint64_t>(ldo), 128, 32);
  if (rc) return rc;

  // layout: [stages][1 KB: barriers + TMEM slot][8 x 4 KB staging boxes] (+1 KB alignment slack)
  const size_t smem = static_cast<size_t>(p.num_stages) * p.stage_bytes + 1024 + epi_bytes + 1024;
  ZKB_CUDA(cudaFuncSetAttribute(kgemm_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                227 * 1024));
  ZKB_CUDA(cudaFuncSetAttribute(kgemm_first_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                227 * 1024));
  const int tiles = p.tiles_m * p.tiles_n;
  const int sms = zkb_sm_count();
  if (use_pairs) {
    ZKB_CUDA(cudaFuncSetAttribute(kgemm2_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  227 * 1024));
    int clusters = tiles < sms / 2 ? tiles : sms / 2;
    // (fused sums: every cluster keeps one tn, see the epilogue)
    if (p.bn_fused && clusters > p.tiles_n) clusters -= clusters % p.tiles_n;
    kgemm2_bf16_kernel<<<clusters * 2, 320, smem, stream>>>(tmA0, tmA1, tmB, tmOut, p);
    ZKB_LAUNCH_CHECK("kgemm2_bf16_kernel");
    if (st) return zkb_bn_stats(out, 1, M, N, st->sum, st->sumsq, stream);
    return ZKB_OK;
  }
  int grid = tiles < sms ? tiles : sms;
  if ((p.bn_fused || p.st_fused || pool) && grid > p.tiles_n) grid -= grid % p.tiles_n;
  if (p.l0_fused)
    kgemm_first_bf16_kernel<<<grid, 320, smem, stream>>>(tmA0, tmA1, tmB, tmOut, p);
  else
    kgemm_bf16_kernel<<<grid, 320, smem, stream>>>(tmA0, tmA1, tmB, tmOut, p);
  ZKB_LAUNCH_CHECK("kgemm_bf16_kernel");
  if (st && !p.st_fused && !pool) return zkb_bn_stats(out, 1, M, N, st->sum, st->sumsq, stream);
  if (bn && !p.bn_fused)   // tile shapes without the fused epilogue: the stand-alone reduction
    return zkb_bn_bwd_reduce(static_cast<const float*>(out), bn->x, 1, M, N, bn->mean, bn->rstd,
                             bn->beta, 1, bn->sum_da, bn->sum_dax, stream);
  return ZKB_OK;
}

// A: bf16 [P][Ma] (plain) or NHWC activation [NB,H,W,Ma] viewed through tap (im2col).
// B_hi/B_lo: bf16 [P][N].  out f32 [(taps*Ma)][ldo], accumulated into (caller zeroes it).
static int launch_wgemm(const void* A, const void* B_hi, const void* B_lo, float* out,
                        long long ldo, int Ma, int N, int n_valid, long long P, int im2col, int NB,
                        int H, int W, int ksize, int pad, cudaStream_t stream, int stride = 1) {
  ZKB_REQUIRE(A && B_hi && out, "wgemm: null pointer");
  ZKB_REQUIRE(Ma > 0 && N > 0 && P > 0, "wgemm: empty problem");
  ZKB_REQUIRE(Ma % 64 == 0, "wgemm: Ma=%d must be a multiple of 64", Ma);
  ZKB_REQUIRE(N % 64 == 0, "wgemm: N=%d must be a multiple of 64", N);
  ZKB_REQUIRE(P < (1ll << 31) - 256, "wgemm: P too large");
  ZKB_REQUIRE(n_valid > 0 && n_valid <= N && ldo >= n_valid, "wgemm: bad n_valid / ldo");
  WGemmParams p{};
  p.Ma = Ma;
  p.N = N;
  p.n_valid = n_valid;
  p.P = P;
  p.b_terms = B_lo ? 2 : 1;
  p.a_im2col = im2col;
  p.taps = im2col ? ksize * ksize : 1;
  p.kw_taps = im2col ? ksize : 1;
  // MMA width: 256 columns whenever the layer allows it (see the kernel's comment)
  if (N % 256 == 0) {
    p.block_n = 256; p.concat = 0; p.cols_per_tile = 256;
  } else if (N == 128 && p.b_terms == 2) {
    p.block_n = 256; p.concat = 1; p.cols_per_tile = 128;
  } else if (N % 128 == 0) {
    p.block_n = 128; p.concat = 0; p.cols_per_tile = 128;
  } else {
    p.block_n = 64; p.concat = 0; p.cols_per_tile = 64;
  }
  // 64, 192, ... channels: M tiles of two 64-row units (see WGemmParams::flat64)
  p.flat64 = Ma % 128 != 0;
  p.ci64s = Ma / 64;
  p.units = p.taps * p.ci64s;
  p.mt = p.flat64 ? (p.units + 1) / 2 : p.taps;
  // M tiles per work item: as many accumulators as fit the 512 TMEM columns, at most a filter row
  p.tpi = 1;
  if (im2col) {
    p.tpi = 512 / p.block_n;
    if (p.tpi > ksize) p.tpi = ksize;
  }
  if (p.flat64) p.tpi = 512 / p.block_n < 4 ? 512 / p.block_n : 4;
  if (p.tpi > p.mt) p.tpi = p.mt;
  p.kpix = 64;
  const GemmTuning& tune = gemm_tuning();
  p.dbg = tune.dbg;
  if (im2col && tune.wgrad_tpi >= 1 && tune.wgrad_tpi * p.block_n <= 512) p.tpi = tune.wgrad_tpi;
  if (tune.wgrad_kpix) p.kpix = tune.wgrad_kpix;
  // H, W: the A activation's map; rows of dy index the conv's output map (equal for stride 1)
  const SameGeom wgh = same_geometry(H > 0 ? H : 1, im2col ? ksize : 1, stride);
  const SameGeom wgw = same_geometry(W > 0 ? W : 1, im2col ? ksize : 1, stride);
  p.pad = wgw.lo;
  p.pad_h = wgh.lo;
  p.stride = stride;
  p.OH = wgh.out;
  p.OW = wgw.out;
  p.ci_tiles = p.flat64 ? 1 : Ma / 128;
  p.n_tiles = N / p.cols_per_tile;
  const int boxes_b = p.b_terms * (p.concat ? 2 : p.block_n / 64);
  // Three pipeline stages hide the TMA latency, two do not (measured r1: -5..12 % with three):
  // when 64 reduction rows per stage leave room for only two stages, take 48.
  if (!tune.wgrad_kpix && kSmemBudget / ((p.tpi * 2 + boxes_b) * p.kpix * 128) < 3) p.kpix = 48;
  const int box_bytes = p.kpix * 128;
  p.stage_bytes = (p.tpi * 2 + boxes_b) * box_bytes;
  p.num_stages = kSmemBudget / p.stage_bytes;
  if (p.num_stages > kMaxStages) p.num_stages = kMaxStages;
  ZKB_REQUIRE(p.num_stages >= 2, "wgemm: stage does not fit in shared memory");
  p.out = out;
  p.ldo = ldo;
  const long long chunks = (P + p.kpix - 1) / p.kpix;
  const int sms = zkb_sm_count();
  // Split the reduction over pixels so that the grid fills FULL waves (one CTA per SM): a few
  // CTAs spilling into one more wave cost a whole extra work item of time (measured 10-13 %), and
  // one long wave beats two short ones (fewer pipeline fills and atomic epilogues; measured in the
  // step, r1) unless the output tiles divide the SMs badly.
  const int out_tiles = ((p.mt + p.tpi - 1) / p.tpi) * p.ci_tiles * p.n_tiles;
  int waves = tune.wgrad_waves;
  if (waves == 0) {
    double best = 0.0;
    for (int w = 1; w <= 3; ++w) {
      const long long g = (static_cast<long long>(w) * sms / out_tiles) * out_tiles;
      const double fill = static_cast<double>(g) / (static_cast<double>(w) * sms);
      if (fill > best + 0.05) { best = fill; waves = w; }   // more waves only for a clearly better fill
    }
    if (waves == 0) waves = 1;
  }
  long long splits = (static_cast<long long>(waves) * sms) / out_tiles;
  if (splits > chunks) splits = chunks;
  if (splits < 1) splits = 1;
  p.chunks_per_split = static_cast<int>((chunks + splits - 1) / splits);
  p.splits = static_cast<int>((chunks + p.chunks_per_split - 1) / p.chunks_per_split);

  CUtensorMap tmA, tmB0, tmB1;
  int rc;
  if (im2col) {
    ZKB_REQUIRE(stride == 1 || stride == 2, "wgemm: stride must be 1 or 2");
    ZKB_REQUIRE(P == static_cast<long long>(NB) * p.OH * p.OW, "wgemm: P != NB*OH*OW");
    ZKB_REQUIRE(ksize == 2 * pad + 1, "wgemm: only 'same' windows (ksize == 2*pad+1)");
    rc = zkb_tmap_im2col_nhwc_ex(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, A, NB, H, W, Ma, ksize,
                                 wgh.lo, wgh.hi, wgw.lo, wgw.hi, stride, 64, p.kpix);
  } else {
    rc = zkb_tmap_tiled_2d(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, A, Ma, P,
                           static_cast<uint64_t>(Ma) * 2, 64, p.kpix);
  }
  if (rc) return rc;
  rc = zkb_tmap_tiled_2d(&tmB0, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, B_hi, N, P,
                         static_cast<uint64_t>(N) * 2, 64, p.kpix);
  if (rc) return rc;
  tmB1 = tmB0;
  if (B_lo) {
    rc = zkb_tmap_tiled_2d(&tmB1, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, B_lo, N, P,
                           static_cast<uint64_t>(N) * 2, 64, p.kpix);
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

ZKB_API int zkb_bconv2d_fwd_i8(const void* act_i8, const void* wq_oc_k_i8, void* y_i16, int NB, int H,
                               int W, int Cin, int Cout, int ksize, void* stream) {
  ZKB_REQUIRE(ksize % 2 == 1, "bconv2d_fwd_i8: ksize must be odd");
  return launch_kgemm(act_i8, nullptr, wq_oc_k_i8, y_i16, 1, Cout,
                      static_cast<long long>(NB) * H * W, Cout, ksize * ksize * Cin, 1, NB, H, W,
                      Cin, ksize, ksize / 2, static_cast<cudaStream_t>(stream), 1);
}

ZKB_API int zkb_bconv2d_fwd_i8_pool_ok(int H, int W, int Cout) {
  return (pool_row_pairs(H, W) > 0 && Cout > 0 && Cout % 128 == 0) ? 1 : 0;
}

ZKB_API int zkb_bconv2d_fwd_i8_pool(const void* act_i8, const void* wq_oc_k_i8, void* p_i16, uint8_t* arg_u8,
                                    int NB, int H, int W, int Cin, int Cout, int ksize, double* sum,
                                    double* sumsq, void* stream) {
  ZKB_REQUIRE(ksize % 2 == 1, "bconv2d_fwd_i8_pool: ksize must be odd");
  ZKB_REQUIRE(p_i16 && arg_u8 && sum && sumsq, "bconv2d_fwd_i8_pool: null output");
  StatsFuse st{sum, sumsq};
  PoolFuse pool{arg_u8};
  return launch_kgemm(act_i8, nullptr, wq_oc_k_i8, p_i16, 1, Cout, static_cast<long long>(NB) * H * W, Cout,
                      ksize * ksize * Cin, 1, NB, H, W, Cin, ksize, ksize / 2, static_cast<cudaStream_t>(stream),
                      1, nullptr, nullptr, &st, nullptr, 1, &pool);
}

ZKB_API int zkb_bconv2d_fwd_i8_s2(const void* act_i8, const void* wq_oc_k_i8, void* y_i16, int NB, int H,
                                  int W, int Cin, int Cout, int ksize, void* stream) {
  ZKB_REQUIRE(ksize % 2 == 1, "bconv2d_fwd_i8_s2: ksize must be odd");
  const long long OH = (H + 1) / 2, OW = (W + 1) / 2;
  return launch_kgemm(act_i8, nullptr, wq_oc_k_i8, y_i16, 1, Cout, NB * OH * OW, Cout, ksize * ksize * Cin, 1,
                      NB, H, W, Cin, ksize, ksize / 2, static_cast<cudaStream_t>(stream), 1, nullptr, nullptr,
                      nullptr, nullptr, 2);
}

ZKB_API int zkb_bconv2d_fwd_s2(const void* act_pm1, const void* wq_oc_k, void* y_i16, int NB, int H, int W,
                               int Cin, int Cout, int ksize, void* stream) {
  ZKB_REQUIRE(ksize % 2 == 1, "bconv2d_fwd_s2: ksize must be odd");
  const long long OH = (H + 1) / 2, OW = (W + 1) / 2;
  return launch_kgemm(act_pm1, nullptr, wq_oc_k, y_i16, 1, Cout, NB * OH * OW, Cout, ksize * ksize * Cin, 1,
                      NB, H, W, Cin, ksize, ksize / 2, static_cast<cudaStream_t>(stream), 2, nullptr, nullptr,
                      nullptr, nullptr, 2);
}

ZKB_API int zkb_bconv2d_wgrad_s2(const void* act_pm1, const void* dy_hi, const void* dy_lo, float* dw_hwio,
                                 int NB, int H, int W, int Cin, int Cout, int ksize, void* stream) {
  ZKB_REQUIRE(ksize % 2 == 1, "bconv2d_wgrad_s2: ksize must be odd");
  const long long OH = (H + 1) / 2, OW = (W + 1) / 2;
  return launch_wgemm(act_pm1, dy_hi, dy_lo, dw_hwio, Cout, Cin, Cout, Cout, NB * OH * OW, 1, NB, H, W, ksize,
                      ksize / 2, static_cast<cudaStream_t>(stream), 2);
}

ZKB_API int zkb_bconv2d_fwd_i8_stats(const void* act_i8, const void* wq_oc_k_i8, void* y_i16, int NB,
                                     int H, int W, int Cin, int Cout, int ksize, double* sum,
                                     double* sumsq, void* stream) {
  ZKB_REQUIRE(ksize % 2 == 1, "bconv2d_fwd_i8_stats: ksize must be odd");
  StatsFuse st{sum, sumsq};
  return launch_kgemm(act_i8, nullptr, wq_oc_k_i8, y_i16, 1, Cout,
                      static_cast<long long>(NB) * H * W, Cout, ksize * ksize * Cin, 1, NB, H, W,
                      Cin, ksize, ksize / 2, static_cast<cudaStream_t>(stream), 1, nullptr, nullptr, &st);
}

ZKB_API int zkb_bdense_fwd_i8_stats(const void* act_i8, const void* wq_n_k_i8, void* y_i16, long long ldy,
                                    int rows, int K, int N, double* sum, double* sumsq, void* stream) {
  StatsFuse st{sum, sumsq};
  return launch_kgemm(act_i8, nullptr, wq_n_k_i8, y_i16, 1, ldy, rows, N, K, 0, 0, 0, 0, 0, 0, 0,
                      static_cast<cudaStream_t>(stream), 1, nullptr, nullptr, &st);
}

ZKB_API int zkb_bconv2d_fwd_i8_sign(const void* act_i8, const void* wq_oc_k_i8, const int* thr,
                                    void* s_out_i8, int NB, int H, int W, int Cin, int Cout,
                                    int ksize, void* stream) {
  ZKB_REQUIRE(ksize % 2 == 1, "bconv2d_fwd_i8_sign: ksize must be odd");
  return launch_kgemm(act_i8, nullptr, wq_oc_k_i8, s_out_i8, 2, Cout,
                      static_cast<long long>(NB) * H * W, Cout, ksize * ksize * Cin, 1, NB, H, W,
                      Cin, ksize, ksize / 2, static_cast<cudaStream_t>(stream), 1, nullptr, nullptr,
                      nullptr, thr);
}

ZKB_API int zkb_bdense_fwd_i8_sign(const void* act_i8, const void* wq_n_k_i8, const int* thr,
                                   void* s_out_i8, long long lds, int rows, int K, int N, void* stream) {
  return launch_kgemm(act_i8, nullptr, wq_n_k_i8, s_out_i8, 2, lds, rows, N, K, 0, 0, 0, 0, 0, 0, 0,
                      static_cast<cudaStream_t>(stream), 1, nullptr, nullptr, nullptr, thr);
}

ZKB_API int zkb_bdense_fwd_i8(const void* act_i8, const void* wq_n_k_i8, void* y_i16, long long ldy,
                              int rows, int K, int N, void* stream) {
  return launch_kgemm(act_i8, nullptr, wq_n_k_i8, y_i16, 1, ldy, rows, N, K, 0, 0, 0, 0, 0, 0, 0,
                      static_cast<cudaStream_t>(stream), 1);
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

ZKB_API int zkb_bconv2d_dgrad_bn(const void* dy_hi, const void* dy_lo, const void* wq_ic_k, void* dx_f32,
                                 int NB, int H, int W, int Cin, int Cout, int ksize,
                                 const void* x_i16, const float* mean, const float* rstd,
                                 const float* beta, double* sum_da, double* sum_dax, void* stream) {
  ZKB_REQUIRE(ksize % 2 == 1, "bconv2d_dgrad_bn: ksize must be odd");
  ZKB_REQUIRE(x_i16 && mean && rstd && beta && sum_da && sum_dax, "bconv2d_dgrad_bn: null pointer");
  BnBackwardFuse bn{static_cast<const int16_t*>(x_i16), mean, rstd, beta, sum_da, sum_dax};
  return launch_kgemm(dy_hi, dy_lo, wq_ic_k, dx_f32, 0, Cin, static_cast<long long>(NB) * H * W, Cin,
                      ksize * ksize * Cout, 1, NB, H, W, Cout, ksize, ksize / 2,
                      static_cast<cudaStream_t>(stream), 2, nullptr, &bn);
}

ZKB_API int zkb_bconv2d_dgrad_first_bwd(const void* dy_hi, const void* dy_lo, const void* wq_ic_k,
                                        int NB, int H, int W, int Cin, int Cout, int ksize,
                                        const float* x0, int H0, int W0, const float* wq0_k_co,
                                        const float* mean0, const float* rstd0, const float* beta0,
                                        double* acc, void* stream) {
  ZKB_REQUIRE(ksize == 3 && Cin == 128, "bconv2d_dgrad_first_bwd: built for 3x3 kernels into 128 channels");
  ZKB_REQUIRE(H0 == H + 2 && W0 == W + 2,
              "bconv2d_dgrad_first_bwd: first layer must be a 'valid' 3x3 conv (image %dx%d, map %dx%d)",
              H0, W0, H, W);
  ZKB_REQUIRE(x0 && wq0_k_co && mean0 && rstd0 && beta0 && acc, "bconv2d_dgrad_first_bwd: null pointer");
  FirstLayerFuse l0{x0, wq0_k_co, mean0, rstd0, beta0, acc, H0, W0};
  return launch_kgemm(dy_hi, dy_lo, wq_ic_k, nullptr, 0, Cin, static_cast<long long>(NB) * H * W, Cin,
                      ksize * ksize * Cout, 1, NB, H, W, Cout, ksize, ksize / 2,
                      static_cast<cudaStream_t>(stream), 2, &l0);
}

ZKB_API int zkb_conv_first_bwd_finalize(const double* acc, const double* gram, const float* wq0_k_co,
                                        const float* mean0, const float* rstd0, double count,
                                        double dbeta_scale, float* dbeta0, float* dw0_k_co, int Cout,
                                        void* stream) {
  ZKB_REQUIRE(acc && gram && wq0_k_co && mean0 && rstd0 && dbeta0 && dw0_k_co,
              "conv_first_bwd_finalize: null pointer");
  ZKB_REQUIRE(Cout == 128 && count > 0, "conv_first_bwd_finalize: Cout must be 128, count > 0");
  conv_first_fused_finalize_kernel<<<1, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      acc, gram, wq0_k_co, mean0, rstd0, count, dbeta_scale, dbeta0, dw0_k_co);
  ZKB_LAUNCH_CHECK("conv_first_fused_finalize_kernel");
  return ZKB_OK;
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
