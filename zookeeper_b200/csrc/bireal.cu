// bireal.cu -- the operators a Bi-Real-Net style binary residual unit adds to the BinaryNet path
// (BASELINE.json configs[3], [4]): the approx_sign and
// magnitude_aware_sign quantizers, BatchNorm with a learnt scale whose input is the INTEGER binary
// conv output times a per-channel weight magnitude, the float shortcut add, global average pooling,
// a float classifier and the 3x3 / stride-2 max-pool of the stem.  The binary contractions
// themselves are the tcgen05 GEMMs of gemm_tc.cu; everything here is HBM-bound elementwise or
// small-reduction work: 16-byte vectors along C, grid-stride over a few CTAs per SM.
//
// None of these operators is in the reference tree (larq-zoo is a third-party package the reference
// resolves model names through, examples/larq_experiment.py:52-57,59-103 is the hook); semantics
// follow the published larq / larq-zoo definitions restated in oracle/bireal_oracle.py.
#include <cuda_bf16.h>

#include "zkb_host.h"

namespace zkb {

__device__ __forceinline__ uint16_t br_bf16_rn(float f) { return __bfloat16_as_ushort(__float2bfloat16_rn(f)); }
__device__ __forceinline__ float br_bf16_f32(uint16_t h) { return __uint_as_float(static_cast<uint32_t>(h) << 16); }

// magnitude_aware_sign: alpha[c] = mean over (kh, kw, ci) of |W[kh,kw,ci,c]|  (HWIO, K rows of Cout)
__global__ void __launch_bounds__(256)
weight_abs_mean_kernel(const float* __restrict__ w, long long K, int Cout, float* __restrict__ alpha) {
  // block = 32 channels x 8 row lanes; rows strided over the lanes, lanes combined through smem
  __shared__ double part[8][33];
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  const int rl = threadIdx.x >> 5;
  double acc = 0.0;
  if (c < Cout)
    for (long long k = rl; k < K; k += 8) acc += fabsf(w[k * Cout + c]);
  part[rl][threadIdx.x & 31] = acc;
  __syncthreads();
  if (rl == 0 && c < Cout) {
    double t = 0.0;
    for (int r = 0; r < 8; ++r) t += part[r][threadIdx.x];
    alpha[c] = static_cast<float>(t / static_cast<double>(K));
  }
}

// BatchNorm(scale=True) over z = alpha * y, y the integer conv output whose sums are given:
//   mean_z = alpha mean_y, var_z = alpha^2 var_y, rstd_z = rsqrt(var_z + eps)
//   forward:  out = gamma (z - mean_z) rstd_z + beta = (y - mean_y) * gain + beta,  gain = alpha gamma rstd_z
//   backward: xhat = (y - mean_y) * rs_bwd,  rs_bwd = alpha rstd_z;  dy = gain (d - mean d - xhat mean(d xhat))
__global__ void bn_gain_finalize_kernel(const double* __restrict__ sum, const double* __restrict__ sumsq,
                                        double count, const float* __restrict__ alpha,
                                        const float* __restrict__ gamma, float eps, float momentum,
                                        int bessel, int C, float* __restrict__ mean_y,
                                        float* __restrict__ rs_bwd, float* __restrict__ gain,
                                        float* __restrict__ moving_mean, float* __restrict__ moving_var) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const double m = sum[c] / count;
  double var = sumsq[c] / count - m * m;
  if (var < 0.0) var = 0.0;
  const double a = alpha[c];
  const float var_z = static_cast<float>(a * a * var);
  const float rstd_z = rsqrtf(var_z + eps);
  mean_y[c] = static_cast<float>(m);
  rs_bwd[c] = static_cast<float>(a) * rstd_z;
  gain[c] = static_cast<float>(a) * gamma[c] * rstd_z;
  if (moving_mean) {
    const float mz = static_cast<float>(a * m);
    const float uv = (bessel && count > 1.0) ? static_cast<float>(a * a * var * (count / (count - 1.0))) : var_z;
    moving_mean[c] = moving_mean[c] * momentum + mz * (1.0f - momentum);
    moving_var[c] = moving_var[c] * momentum + uv * (1.0f - momentum);
  }
}

// out = (y - mean_y) * gain + beta (+ shortcut): 8 channels (one 16-byte int16 vector) per thread
__global__ void __launch_bounds__(256)
bn_gain_add_kernel(const int16_t* __restrict__ y, long long rows, int C, const float* __restrict__ mean_y,
                   const float* __restrict__ gain, const float* __restrict__ beta,
                   const float* __restrict__ shortcut, float* __restrict__ out) {
  const int tpr = C / 8;
  const long long total = rows * tpr;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cl = static_cast<int>(i % tpr);
    const long long off = (i / tpr) * C + cl * 8;
    const uint4 t = *reinterpret_cast<const uint4*>(y + off);
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
    float o[8];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      o[2 * e] = static_cast<float>(static_cast<int16_t>(w[e] & 0xFFFFu));
      o[2 * e + 1] = static_cast<float>(static_cast<int16_t>(w[e] >> 16));
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = cl * 8 + j;
      o[j] = (o[j] - mean_y[c]) * gain[c] + beta[c];
    }
    if (shortcut) {
      const float4 s0 = *reinterpret_cast<const float4*>(shortcut + off);
      const float4 s1 = *reinterpret_cast<const float4*>(shortcut + off + 4);
      o[0] += s0.x; o[1] += s0.y; o[2] += s0.z; o[3] += s0.w;
      o[4] += s1.x; o[5] += s1.y; o[6] += s1.z; o[7] += s1.w;
    }
    *reinterpret_cast<float4*>(out + off) = make_float4(o[0], o[1], o[2], o[3]);
    *reinterpret_cast<float4*>(out + off + 4) = make_float4(o[4], o[5], o[6], o[7]);
  }
}

// dy = gain * (d - mean_d - xhat * mean_dxhat), xhat = (y - mean_y) * rs_bwd, as the bf16 hi/lo pair
// the backward GEMMs read.  4 channels per thread (d is f32).
__global__ void __launch_bounds__(256)
bn_gain_bwd_apply_kernel(const float* __restrict__ d, const int16_t* __restrict__ y, long long rows, int C,
                         const float* __restrict__ mean_y, const float* __restrict__ rs_bwd,
                         const float* __restrict__ gain, const float* __restrict__ mean_d,
                         const float* __restrict__ mean_dx, uint16_t* __restrict__ out_hi,
                         uint16_t* __restrict__ out_lo) {
  const int tpr = C / 4;
  const long long total = rows * tpr;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cl = static_cast<int>(i % tpr);
    const long long off = (i / tpr) * C + cl * 4;
    const float4 g4 = *reinterpret_cast<const float4*>(d + off);
    const uint2 y2 = *reinterpret_cast<const uint2*>(y + off);
    const float gg[4] = {g4.x, g4.y, g4.z, g4.w};
    const float yv[4] = {static_cast<float>(static_cast<int16_t>(y2.x & 0xFFFFu)),
                         static_cast<float>(static_cast<int16_t>(y2.x >> 16)),
                         static_cast<float>(static_cast<int16_t>(y2.y & 0xFFFFu)),
                         static_cast<float>(static_cast<int16_t>(y2.y >> 16))};
    uint16_t hi[4], lo[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = cl * 4 + j;
      const float xh = (yv[j] - mean_y[c]) * rs_bwd[c];
      const float dx = gain[c] * (gg[j] - mean_d[c] - xh * mean_dx[c]);
      hi[j] = br_bf16_rn(dx);
      lo[j] = br_bf16_rn(dx - br_bf16_f32(hi[j]));
    }
    *reinterpret_cast<uint2*>(out_hi + off) = make_uint2(hi[0] | (uint32_t(hi[1]) << 16), hi[2] | (uint32_t(hi[3]) << 16));
    *reinterpret_cast<uint2*>(out_lo + off) = make_uint2(lo[0] | (uint32_t(lo[1]) << 16), lo[2] | (uint32_t(lo[3]) << 16));
  }
}

// The same pair for a FLOAT input map (BatchNorm after the 1x1 float conv of a down-sampling
// shortcut; alpha = 1 there): out = (x - mean) * gain + beta (+ shortcut), and the backward
// dx = gain * (d - mean_d - xhat * mean_dx) written as f32.  4 channels per thread.
__global__ void __launch_bounds__(256)
bn_gain_add_f32_kernel(const float* __restrict__ x, long long rows, int C, const float* __restrict__ mean,
                       const float* __restrict__ gain, const float* __restrict__ beta,
                       const float* __restrict__ shortcut, float* __restrict__ out) {
  const int tpr = C / 4;
  const long long total = rows * tpr;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c0 = static_cast<int>(i % tpr) * 4;
    const float4 v = reinterpret_cast<const float4*>(x)[i];
    float o[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) o[j] = (o[j] - mean[c0 + j]) * gain[c0 + j] + beta[c0 + j];
    if (shortcut) {
      const float4 s4 = reinterpret_cast<const float4*>(shortcut)[i];
      o[0] += s4.x; o[1] += s4.y; o[2] += s4.z; o[3] += s4.w;
    }
    reinterpret_cast<float4*>(out)[i] = make_float4(o[0], o[1], o[2], o[3]);
  }
}

__global__ void __launch_bounds__(256)
bn_gain_bwd_apply_f32_kernel(const float* __restrict__ d, const float* __restrict__ x, long long rows, int C,
                             const float* __restrict__ mean, const float* __restrict__ rs_bwd,
                             const float* __restrict__ gain, const float* __restrict__ mean_d,
                             const float* __restrict__ mean_dx, float* __restrict__ out) {
  const int tpr = C / 4;
  const long long total = rows * tpr;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c0 = static_cast<int>(i % tpr) * 4;
    const float4 g4 = reinterpret_cast<const float4*>(d)[i];
    const float4 x4 = reinterpret_cast<const float4*>(x)[i];
    const float gg[4] = {g4.x, g4.y, g4.z, g4.w}, xv[4] = {x4.x, x4.y, x4.z, x4.w};
    float o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = c0 + j;
      const float xh = (xv[j] - mean[c]) * rs_bwd[c];
      o[j] = gain[c] * (gg[j] - mean_d[c] - xh * mean_dx[c]);
    }
    reinterpret_cast<float4*>(out)[i] = make_float4(o[0], o[1], o[2], o[3]);
  }
}

// AveragePooling2D(2, 2) of the down-sampling shortcut, float NHWC ('valid': a trailing odd row /
// column is dropped), and its backward (every input of a window gets a quarter of the gradient).
__global__ void __launch_bounds__(256)
avgpool2_fwd_kernel(const float* __restrict__ x, int B, int H, int W, int C, float* __restrict__ out) {
  const int PH = H / 2, PW = W / 2, tpr = C / 4;
  const long long total = static_cast<long long>(B) * PH * PW * tpr;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cl = static_cast<int>(i % tpr);
    long long r = i / tpr;
    const int pw = static_cast<int>(r % PW);
    r /= PW;
    const int ph = static_cast<int>(r % PH);
    const long long b = r / PH;
    const float* p0 = x + ((b * H + 2 * ph) * W + 2 * pw) * C + cl * 4;
    const float4 a = *reinterpret_cast<const float4*>(p0);
    const float4 b4 = *reinterpret_cast<const float4*>(p0 + C);
    const float4 c4 = *reinterpret_cast<const float4*>(p0 + static_cast<long long>(W) * C);
    const float4 d4 = *reinterpret_cast<const float4*>(p0 + static_cast<long long>(W) * C + C);
    reinterpret_cast<float4*>(out)[i] = make_float4((a.x + b4.x + c4.x + d4.x) * 0.25f, (a.y + b4.y + c4.y + d4.y) * 0.25f,
                                                    (a.z + b4.z + c4.z + d4.z) * 0.25f, (a.w + b4.w + c4.w + d4.w) * 0.25f);
  }
}

__global__ void __launch_bounds__(256)
avgpool2_bwd_kernel(const float* __restrict__ dout, int B, int H, int W, int C, float* __restrict__ dx) {
  const int PH = H / 2, PW = W / 2, tpr = C / 4;
  const long long total = static_cast<long long>(B) * H * W * tpr;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cl = static_cast<int>(i % tpr);
    long long r = i / tpr;
    const int iw = static_cast<int>(r % W);
    r /= W;
    const int ih = static_cast<int>(r % H);
    const long long b = r / H;
    float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
    if ((ih >> 1) < PH && (iw >> 1) < PW) {
      const float4 g = *reinterpret_cast<const float4*>(dout + (((b * PH + (ih >> 1)) * PW + (iw >> 1)) * tpr + cl) * 4);
      o = make_float4(g.x * 0.25f, g.y * 0.25f, g.z * 0.25f, g.w * 0.25f);
    }
    reinterpret_cast<float4*>(dx)[i] = o;
  }
}

// approx_sign backward plus the float shortcut: out = d_short + ds * (2 - 2|h|) * [|h| <= 1]
__global__ void __launch_bounds__(256)
approx_sign_bwd_add_kernel(const float* __restrict__ ds, const float* __restrict__ h,
                           const float* __restrict__ d_short, float* __restrict__ out, long long n4) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float4 g = reinterpret_cast<const float4*>(ds)[i];
    const float4 x = reinterpret_cast<const float4*>(h)[i];
    float4 o = d_short ? reinterpret_cast<const float4*>(d_short)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    const float gx[4] = {g.x, g.y, g.z, g.w}, xx[4] = {x.x, x.y, x.z, x.w};
    float* op = &o.x;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float a = fabsf(xx[j]);
      op[j] += (a <= 1.0f) ? gx[j] * (2.0f - 2.0f * a) : 0.0f;
    }
    reinterpret_cast<float4*>(out)[i] = o;
  }
}

// global average pool [B,HW,C] -> [B,C]; thread = (b, 4 channels)
__global__ void __launch_bounds__(256)
global_avgpool_fwd_kernel(const float* __restrict__ h, int B, int HW, int C, float* __restrict__ out) {
  const int tpr = C / 4;
  const long long total = static_cast<long long>(B) * tpr;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cl = static_cast<int>(i % tpr);
    const long long b = i / tpr;
    const float* p = h + b * HW * C + cl * 4;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    for (int r = 0; r < HW; ++r) {
      const float4 v = *reinterpret_cast<const float4*>(p + static_cast<long long>(r) * C);
      a0 += v.x; a1 += v.y; a2 += v.z; a3 += v.w;
    }
    const float inv = 1.0f / static_cast<float>(HW);
    *reinterpret_cast<float4*>(out + b * C + cl * 4) = make_float4(a0 * inv, a1 * inv, a2 * inv, a3 * inv);
  }
}

__global__ void __launch_bounds__(256)
global_avgpool_bwd_kernel(const float* __restrict__ dout, int B, int HW, int C, float* __restrict__ dh) {
  const int tpr = C / 4;
  const long long total = static_cast<long long>(B) * HW * tpr;
  const float inv = 1.0f / static_cast<float>(HW);
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cl = static_cast<int>(i % tpr);
    const long long b = i / (static_cast<long long>(HW) * tpr);
    const float4 v = *reinterpret_cast<const float4*>(dout + b * C + cl * 4);
    reinterpret_cast<float4*>(dh)[i] = make_float4(v.x * inv, v.y * inv, v.z * inv, v.w * inv);
  }
}

// Float GEMMs of the zoo models: the classifier (tf.keras.layers.Dense after the pooling) and the
// 1x1 shortcut convolution of a down-sampling unit over the [pixels, C] view of the map -- the latter
// at 200 704 x 64 x 128 per direction for Bi-Real-Net-18 at batch 256, about 1.8 % of the model's
// multiply-adds but float x float, so not work for the +-1 GEMM engines.  One register-tiled FP32
// kernel serves all three directions:
//     C[i][j] (+)= sum_r A(i,r) * B(r,j) (+ bias[j])
//     A(i,r) = AT ? A[r*lda + i] : A[i*lda + r]        B(r,j) = BT ? B[j*ldb + r] : B[r*ldb + j]
// forward  y  = x w + b      : A = x,  B = w           (I, J, R) = (M, N, K)
// dgrad    dx = dy w^T       : A = dy, B = w    (BT)   (I, J, R) = (M, K, N)
// wgrad    dw = x^T dy, db   : A = x (AT), B = dy      (I, J, R) = (K, N, M), the reduction split over
//                              blockIdx.y and added atomically into a zeroed dw / db
// Tile 128 (i) x 64 (j) x 16 (r), 256 threads, 8 x 4 outputs per thread; scalar global loads, so no
// alignment or divisibility rule on any dimension (the 10- and 1000-class heads are ragged).
constexpr int kSgI = 128, kSgJ = 64, kSgR = 16;

template <bool AT, bool BT, bool SPLIT>
__global__ void __launch_bounds__(256)
sgemm_f32_kernel(const float* __restrict__ A, long long lda, const float* __restrict__ Bm, long long ldb,
                 float* __restrict__ C, long long ldc, const float* __restrict__ bias, float* __restrict__ colsum,
                 int I, int J, int R, int r_chunk, int j_tiles) {
  __shared__ __align__(16) float sA[kSgR][kSgI + 4];
  __shared__ __align__(16) float sB[kSgR][kSgJ + 4];
  const int ti = blockIdx.x / j_tiles, tj = blockIdx.x - ti * j_tiles;
  const int i0 = ti * kSgI, j0 = tj * kSgJ;
  const int r_lo = SPLIT ? static_cast<int>(blockIdx.y) * r_chunk : 0;
  const int r_hi = SPLIT ? (r_lo + r_chunk < R ? r_lo + r_chunk : R) : R;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[8][4];
#pragma unroll
  for (int a = 0; a < 8; ++a) acc[a][0] = acc[a][1] = acc[a][2] = acc[a][3] = 0.f;
  float cs[4] = {0.f, 0.f, 0.f, 0.f};
  const bool do_colsum = SPLIT && colsum != nullptr && ti == 0 && ty == 0;

  for (int r0 = r_lo; r0 < r_hi; r0 += kSgR) {
    __syncthreads();
    // A tile: kSgR x kSgI
    for (int e = threadIdx.x; e < kSgR * kSgI; e += 256) {
      int rr, ii;
      if (AT) { rr = e / kSgI; ii = e - rr * kSgI; }     // memory runs along i
      else    { ii = e / kSgR; rr = e - ii * kSgR; }     // memory runs along r
      const int r = r0 + rr, i = i0 + ii;
      float v = 0.f;
      if (r < r_hi && i < I) v = AT ? __ldg(A + static_cast<long long>(r) * lda + i) : __ldg(A + static_cast<long long>(i) * lda + r);
      sA[rr][ii] = v;
    }
    for (int e = threadIdx.x; e < kSgR * kSgJ; e += 256) {
      int rr, jj;
      if (BT) { jj = e / kSgR; rr = e - jj * kSgR; }
      else    { rr = e / kSgJ; jj = e - rr * kSgJ; }
      const int r = r0 + rr, j = j0 + jj;
      float v = 0.f;
      if (r < r_hi && j < J) v = BT ? __ldg(Bm + static_cast<long long>(j) * ldb + r) : __ldg(Bm + static_cast<long long>(r) * ldb + j);
      sB[rr][jj] = v;
    }
    __syncthreads();
#pragma unroll
    for (int rr = 0; rr < kSgR; ++rr) {
      const float4 a0 = *reinterpret_cast<const float4*>(&sA[rr][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&sA[rr][64 + ty * 4]);
      const float4 b4 = *reinterpret_cast<const float4*>(&sB[rr][tx * 4]);
      const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float bv[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[a][c] = fmaf(av[a], bv[c], acc[a][c]);
      if (do_colsum) {
#pragma unroll
        for (int c = 0; c < 4; ++c) cs[c] += bv[c];
      }
    }
  }
#pragma unroll
  for (int a = 0; a < 8; ++a) {
    const int i = i0 + (a < 4 ? ty * 4 + a : 64 + ty * 4 + (a - 4));
    if (i >= I) continue;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int j = j0 + tx * 4 + c;
      if (j >= J) continue;
      float* o = C + static_cast<long long>(i) * ldc + j;
      if (SPLIT) atomicAdd(o, acc[a][c]);
      else *o = acc[a][c] + (bias ? __ldg(bias + j) : 0.f);
    }
  }
  if (do_colsum) {
#pragma unroll
    for (int c = 0; c < 4; ++c)
      if (j0 + tx * 4 + c < J) atomicAdd(colsum + j0 + tx * 4 + c, cs[c]);
  }
}

// 3x3 max-pool, stride 2, 'same' padding (tf.keras.layers.MaxPool2D(3, 2, padding="same") of the
// Bi-Real-Net stem), float NHWC.  out = ceil(in / 2); the pad (total max((out-1)*2 + 3 - in, 0),
// the smaller half first) only shifts the window: padded positions never win.  First maximum in
// row-major window order wins (strict >), its slot 0..8 goes to arg for the backward scatter.
__global__ void __launch_bounds__(256)
maxpool3s2_fwd_kernel(const float* __restrict__ x, int B, int H, int W, int C, int OH, int OW, int pad_h,
                      int pad_w, float* __restrict__ out, uint8_t* __restrict__ arg) {
  const int tpr = C / 4;
  const long long total = static_cast<long long>(B) * OH * OW * tpr;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cl = static_cast<int>(i % tpr);
    long long r = i / tpr;
    const int ow = static_cast<int>(r % OW);
    r /= OW;
    const int oh = static_cast<int>(r % OH);
    const long long b = r / OH;
    float best[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
    uint8_t am[4] = {0, 0, 0, 0};
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
      const int ih = oh * 2 + kh - pad_h;
      if (ih < 0 || ih >= H) continue;
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const int iw = ow * 2 + kw - pad_w;
        if (iw < 0 || iw >= W) continue;
        const float4 v = *reinterpret_cast<const float4*>(x + ((b * H + ih) * W + iw) * C + cl * 4);
        const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (vv[j] > best[j]) { best[j] = vv[j]; am[j] = static_cast<uint8_t>(kh * 3 + kw); }
      }
    }
    *reinterpret_cast<float4*>(out + i * 4) = make_float4(best[0], best[1], best[2], best[3]);
    if (arg) *reinterpret_cast<uint32_t*>(arg + i * 4) = am[0] | (am[1] << 8) | (am[2] << 16) | (am[3] << 24);
  }
}

// Backward of the overlapping pool: dx[q] = sum of dout over the (at most four) windows that cover
// q and whose winning slot is q.  Gather form (no atomics): thread = (input pixel, 4 channels).
__global__ void __launch_bounds__(256)
maxpool3s2_bwd_kernel(const float* __restrict__ dout, const uint8_t* __restrict__ arg, int B, int H, int W,
                      int C, int OH, int OW, int pad_h, int pad_w, float* __restrict__ dx) {
  const int tpr = C / 4;
  const long long total = static_cast<long long>(B) * H * W * tpr;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cl = static_cast<int>(i % tpr);
    long long r = i / tpr;
    const int iw = static_cast<int>(r % W);
    r /= W;
    const int ih = static_cast<int>(r % H);
    const long long b = r / H;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    // windows (oh, ow) with oh*2 + kh - pad_h == ih for kh in 0..2
    for (int kh = 0; kh < 3; ++kh) {
      const int t = ih + pad_h - kh;
      if (t < 0 || (t & 1)) continue;
      const int oh = t >> 1;
      if (oh >= OH) continue;
      for (int kw = 0; kw < 3; ++kw) {
        const int u = iw + pad_w - kw;
        if (u < 0 || (u & 1)) continue;
        const int ow = u >> 1;
        if (ow >= OW) continue;
        const long long o = (((b * OH + oh) * OW + ow) * tpr + cl) * 4;
        const uint32_t a4 = *reinterpret_cast<const uint32_t*>(arg + o);
        const float4 g = *reinterpret_cast<const float4*>(dout + o);
        const float gg[4] = {g.x, g.y, g.z, g.w};
        const uint32_t slot = static_cast<uint32_t>(kh * 3 + kw);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (((a4 >> (8 * j)) & 0xFFu) == slot) acc[j] += gg[j];
      }
    }
    reinterpret_cast<float4*>(dx)[i] = make_float4(acc[0], acc[1], acc[2], acc[3]);
  }
}

// Zero insertion for the input gradient of a stride-2 'same' conv: the output-resolution gradient
// (hi/lo pair) is written to position (2 oh + 1 - pad_h, 2 ow + 1 - pad_w) of an input-resolution map
// and zeros everywhere else; the ordinary stride-1 dgrad GEMM over that map is then the transposed
// strided conv (ds[q] = sum_taps dy[(q - tap + pad) / 2] w[tap] wherever the division is exact).
// Thread = (input pixel, 8 channels): every element of the map is written exactly once.
__global__ void __launch_bounds__(256)
dilate2_hi_lo_kernel(const uint16_t* __restrict__ hi, const uint16_t* __restrict__ lo, int B, int H, int W,
                     int C, int OH, int OW, int off_h, int off_w, uint16_t* __restrict__ out_hi,
                     uint16_t* __restrict__ out_lo) {
  const int tpr = C / 8;
  const long long total = static_cast<long long>(B) * H * W * tpr;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cl = static_cast<int>(i % tpr);
    long long r = i / tpr;
    const int iw = static_cast<int>(r % W);
    r /= W;
    const int ih = static_cast<int>(r % H);
    const long long b = r / H;
    uint4 vh = make_uint4(0, 0, 0, 0), vl = make_uint4(0, 0, 0, 0);
    const int th = ih - off_h, tw = iw - off_w;
    if (th >= 0 && tw >= 0 && !(th & 1) && !(tw & 1) && (th >> 1) < OH && (tw >> 1) < OW) {
      const long long o = (((b * OH + (th >> 1)) * OW + (tw >> 1)) * tpr + cl) * 8;
      vh = *reinterpret_cast<const uint4*>(hi + o);
      vl = *reinterpret_cast<const uint4*>(lo + o);
    }
    *reinterpret_cast<uint4*>(out_hi + i * 8) = vh;
    *reinterpret_cast<uint4*>(out_lo + i * 8) = vl;
  }
}

}  // namespace zkb

// ============================================================================================
// C ABI (documented in include/zkb200.h)
// ============================================================================================
using namespace zkb;

#define ZKB_A16(p) ((reinterpret_cast<uintptr_t>(p) & 15) == 0)

static int br_grid(long long items, int waves = 8) {
  const long long cap = static_cast<long long>(zkb_sm_count()) * waves;
  long long g = (items + 255) / 256;
  if (g > cap) g = cap;
  return g < 1 ? 1 : static_cast<int>(g);
}

ZKB_API int zkb_weight_abs_mean(const float* w_k_co, long long K, int Cout, float* alpha, void* stream) {
  ZKB_REQUIRE(w_k_co && alpha && K > 0 && Cout > 0, "weight_abs_mean: bad arguments");
  weight_abs_mean_kernel<<<(Cout + 31) / 32, 256, 0, static_cast<cudaStream_t>(stream)>>>(w_k_co, K, Cout, alpha);
  ZKB_LAUNCH_CHECK("weight_abs_mean_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_bn_gain_finalize(const double* sum, const double* sumsq, double count, const float* alpha,
                                 const float* gamma, float eps, float momentum, int bessel, int C,
                                 float* mean_y, float* rs_bwd, float* gain, float* moving_mean,
                                 float* moving_var, void* stream) {
  ZKB_REQUIRE(sum && sumsq && alpha && gamma && mean_y && rs_bwd && gain, "bn_gain_finalize: null pointer");
  ZKB_REQUIRE(count > 0 && C > 0, "bn_gain_finalize: bad count / C");
  ZKB_REQUIRE((moving_mean == nullptr) == (moving_var == nullptr), "bn_gain_finalize: moving statistics go together");
  bn_gain_finalize_kernel<<<(C + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      sum, sumsq, count, alpha, gamma, eps, momentum, bessel, C, mean_y, rs_bwd, gain, moving_mean, moving_var);
  ZKB_LAUNCH_CHECK("bn_gain_finalize_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_bn_gain_add(const void* y_i16, long long rows, int C, const float* mean_y, const float* gain,
                            const float* beta, const float* shortcut, float* out, void* stream) {
  ZKB_REQUIRE(y_i16 && mean_y && gain && beta && out, "bn_gain_add: null pointer");
  ZKB_REQUIRE(rows >= 0 && C > 0 && C % 8 == 0, "bn_gain_add: C=%d must be a multiple of 8", C);
  ZKB_REQUIRE(ZKB_A16(y_i16) && ZKB_A16(out) && ZKB_A16(shortcut), "bn_gain_add: tensors must be 16-byte aligned");
  if (rows == 0) return ZKB_OK;
  bn_gain_add_kernel<<<br_grid(rows * (C / 8), 16), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const int16_t*>(y_i16), rows, C, mean_y, gain, beta, shortcut, out);
  ZKB_LAUNCH_CHECK("bn_gain_add_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_bn_gain_bwd_apply(const float* d, const void* y_i16, long long rows, int C,
                                  const float* mean_y, const float* rs_bwd, const float* gain,
                                  const float* mean_d, const float* mean_dx, void* out_hi, void* out_lo,
                                  void* stream) {
  ZKB_REQUIRE(d && y_i16 && mean_y && rs_bwd && gain && mean_d && mean_dx && out_hi && out_lo,
              "bn_gain_bwd_apply: null pointer");
  ZKB_REQUIRE(rows >= 0 && C > 0 && C % 4 == 0, "bn_gain_bwd_apply: C=%d must be a multiple of 4", C);
  ZKB_REQUIRE(ZKB_A16(d) && (reinterpret_cast<uintptr_t>(y_i16) & 7) == 0, "bn_gain_bwd_apply: misaligned tensors");
  if (rows == 0) return ZKB_OK;
  bn_gain_bwd_apply_kernel<<<br_grid(rows * (C / 4), 16), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      d, static_cast<const int16_t*>(y_i16), rows, C, mean_y, rs_bwd, gain, mean_d, mean_dx,
      static_cast<uint16_t*>(out_hi), static_cast<uint16_t*>(out_lo));
  ZKB_LAUNCH_CHECK("bn_gain_bwd_apply_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_approx_sign_bwd_add(const float* ds, const float* h, const float* d_short, float* out,
                                    long long n, void* stream) {
  ZKB_REQUIRE(ds && h && out && n >= 0 && n % 4 == 0, "approx_sign_bwd_add: bad arguments (n %% 4 == 0)");
  ZKB_REQUIRE(ZKB_A16(ds) && ZKB_A16(h) && ZKB_A16(out) && ZKB_A16(d_short), "approx_sign_bwd_add: misaligned tensors");
  if (n == 0) return ZKB_OK;
  approx_sign_bwd_add_kernel<<<br_grid(n / 4, 16), 256, 0, static_cast<cudaStream_t>(stream)>>>(ds, h, d_short, out, n / 4);
  ZKB_LAUNCH_CHECK("approx_sign_bwd_add_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_global_avgpool_fwd(const float* h, int B, int HW, int C, float* out, void* stream) {
  ZKB_REQUIRE(h && out && B >= 0 && HW > 0 && C > 0 && C % 4 == 0, "global_avgpool_fwd: bad arguments (C %% 4 == 0)");
  ZKB_REQUIRE(ZKB_A16(h) && ZKB_A16(out), "global_avgpool_fwd: misaligned tensors");
  if (B == 0) return ZKB_OK;
  global_avgpool_fwd_kernel<<<br_grid(static_cast<long long>(B) * (C / 4)), 256, 0, static_cast<cudaStream_t>(stream)>>>(h, B, HW, C, out);
  ZKB_LAUNCH_CHECK("global_avgpool_fwd_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_global_avgpool_bwd(const float* dout, int B, int HW, int C, float* dh, void* stream) {
  ZKB_REQUIRE(dout && dh && B >= 0 && HW > 0 && C > 0 && C % 4 == 0, "global_avgpool_bwd: bad arguments (C %% 4 == 0)");
  ZKB_REQUIRE(ZKB_A16(dout) && ZKB_A16(dh), "global_avgpool_bwd: misaligned tensors");
  if (B == 0) return ZKB_OK;
  global_avgpool_bwd_kernel<<<br_grid(static_cast<long long>(B) * HW * (C / 4), 16), 256, 0, static_cast<cudaStream_t>(stream)>>>(dout, B, HW, C, dh);
  ZKB_LAUNCH_CHECK("global_avgpool_bwd_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_dense_f32_fwd(const float* x, const float* w_k_n, const float* bias, float* y, int M, int K,
                              int N, int ldy, void* stream) {
  ZKB_REQUIRE(x && w_k_n && y && M >= 0 && K > 0 && N > 0 && ldy >= N, "dense_f32_fwd: bad arguments");
  if (M == 0) return ZKB_OK;
  const int it = (M + kSgI - 1) / kSgI, jt = (N + kSgJ - 1) / kSgJ;
  sgemm_f32_kernel<false, false, false><<<dim3(static_cast<unsigned>(it) * jt, 1), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      x, K, w_k_n, N, y, ldy, bias, nullptr, M, N, K, 0, jt);
  ZKB_LAUNCH_CHECK("sgemm_f32_kernel (forward)");
  return ZKB_OK;
}

ZKB_API int zkb_dense_f32_bwd(const float* x, const float* w_k_n, const float* dy, int lddy, float* dx,
                              float* dw_k_n, float* dbias, int M, int K, int N, void* stream) {
  ZKB_REQUIRE(x && w_k_n && dy && dw_k_n && M > 0 && K > 0 && N > 0 && lddy >= N, "dense_f32_bwd: bad arguments");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dx) {
    const int it = (M + kSgI - 1) / kSgI, jt = (K + kSgJ - 1) / kSgJ;
    sgemm_f32_kernel<false, true, false><<<dim3(static_cast<unsigned>(it) * jt, 1), 256, 0, st>>>(
        dy, lddy, w_k_n, N, dx, K, nullptr, nullptr, M, K, N, 0, jt);
    ZKB_LAUNCH_CHECK("sgemm_f32_kernel (dgrad)");
  }
  // dw = x^T dy over M rows: split the rows so that the grid fills the GPU a few times over; the
  // partial tiles are added into a zeroed dw (and dbias)
  ZKB_CUDA(cudaMemsetAsync(dw_k_n, 0, static_cast<size_t>(K) * N * sizeof(float), st));
  if (dbias) ZKB_CUDA(cudaMemsetAsync(dbias, 0, static_cast<size_t>(N) * sizeof(float), st));
  const int it = (K + kSgI - 1) / kSgI, jt = (N + kSgJ - 1) / kSgJ;
  int splits = (4 * zkb_sm_count() + it * jt - 1) / (it * jt);
  const int max_splits = (M + 4 * kSgR - 1) / (4 * kSgR);      // at least 64 rows per block
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  int r_chunk = (M + splits - 1) / splits;
  r_chunk = (r_chunk + kSgR - 1) / kSgR * kSgR;
  splits = (M + r_chunk - 1) / r_chunk;
  sgemm_f32_kernel<true, false, true><<<dim3(static_cast<unsigned>(it) * jt, static_cast<unsigned>(splits)), 256, 0, st>>>(
      x, K, dy, lddy, dw_k_n, N, nullptr, dbias, K, N, M, r_chunk, jt);
  ZKB_LAUNCH_CHECK("sgemm_f32_kernel (wgrad)");
  return ZKB_OK;
}

static void pool3s2_geometry(int in, int& out, int& pad) {
  out = (in + 1) / 2;
  int total = (out - 1) * 2 + 3 - in;
  if (total < 0) total = 0;
  pad = total / 2;
}

ZKB_API int zkb_maxpool3s2_f32_fwd(const float* x, int B, int H, int W, int C, float* out, uint8_t* arg,
                                   void* stream) {
  ZKB_REQUIRE(x && out && B >= 0 && H > 0 && W > 0 && C > 0 && C % 4 == 0, "maxpool3s2_fwd: bad arguments (C %% 4 == 0)");
  ZKB_REQUIRE(ZKB_A16(x) && ZKB_A16(out) && (reinterpret_cast<uintptr_t>(arg) & 3) == 0, "maxpool3s2_fwd: misaligned tensors");
  if (B == 0) return ZKB_OK;
  int OH, OW, ph, pw;
  pool3s2_geometry(H, OH, ph);
  pool3s2_geometry(W, OW, pw);
  maxpool3s2_fwd_kernel<<<br_grid(static_cast<long long>(B) * OH * OW * (C / 4), 16), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      x, B, H, W, C, OH, OW, ph, pw, out, arg);
  ZKB_LAUNCH_CHECK("maxpool3s2_fwd_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_maxpool3s2_f32_bwd(const float* dout, const uint8_t* arg, int B, int H, int W, int C, float* dx,
                                   void* stream) {
  ZKB_REQUIRE(dout && arg && dx && B >= 0 && H > 0 && W > 0 && C > 0 && C % 4 == 0, "maxpool3s2_bwd: bad arguments (C %% 4 == 0)");
  ZKB_REQUIRE(ZKB_A16(dout) && ZKB_A16(dx) && (reinterpret_cast<uintptr_t>(arg) & 3) == 0, "maxpool3s2_bwd: misaligned tensors");
  if (B == 0) return ZKB_OK;
  int OH, OW, ph, pw;
  pool3s2_geometry(H, OH, ph);
  pool3s2_geometry(W, OW, pw);
  maxpool3s2_bwd_kernel<<<br_grid(static_cast<long long>(B) * H * W * (C / 4), 16), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      dout, arg, B, H, W, C, OH, OW, ph, pw, dx);
  ZKB_LAUNCH_CHECK("maxpool3s2_bwd_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_dilate2_hi_lo(const void* dy_hi, const void* dy_lo, int B, int H, int W, int C, int ksize,
                              void* out_hi, void* out_lo, void* stream) {
  ZKB_REQUIRE(dy_hi && dy_lo && out_hi && out_lo, "dilate2_hi_lo: null pointer");
  ZKB_REQUIRE(B >= 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0 && ksize % 2 == 1, "dilate2_hi_lo: bad shape (C %% 8 == 0, odd ksize)");
  ZKB_REQUIRE(ZKB_A16(dy_hi) && ZKB_A16(dy_lo) && ZKB_A16(out_hi) && ZKB_A16(out_lo), "dilate2_hi_lo: misaligned tensors");
  if (B == 0) return ZKB_OK;
  const int OH = (H + 1) / 2, OW = (W + 1) / 2;
  int th = (OH - 1) * 2 + ksize - H, tw = (OW - 1) * 2 + ksize - W;
  if (th < 0) th = 0;
  if (tw < 0) tw = 0;
  // forward tap kh reads input row 2 oh + kh - pad_lo; the stride-1 dgrad over the dilated map D
  // reads D[q + pad1 - kh] with pad1 = ksize / 2, so dy[oh] must sit at row 2 oh + pad1 - pad_lo
  const int off_h = ksize / 2 - th / 2, off_w = ksize / 2 - tw / 2;
  dilate2_hi_lo_kernel<<<br_grid(static_cast<long long>(B) * H * W * (C / 8), 16), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint16_t*>(dy_hi), static_cast<const uint16_t*>(dy_lo), B, H, W, C, OH, OW, off_h, off_w,
      static_cast<uint16_t*>(out_hi), static_cast<uint16_t*>(out_lo));
  ZKB_LAUNCH_CHECK("dilate2_hi_lo_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_bn_gain_add_f32(const float* x, long long rows, int C, const float* mean, const float* gain,
                                const float* beta, const float* shortcut, float* out, void* stream) {
  ZKB_REQUIRE(x && mean && gain && beta && out, "bn_gain_add_f32: null pointer");
  ZKB_REQUIRE(rows >= 0 && C > 0 && C % 4 == 0, "bn_gain_add_f32: C=%d must be a multiple of 4", C);
  ZKB_REQUIRE(ZKB_A16(x) && ZKB_A16(out) && ZKB_A16(shortcut), "bn_gain_add_f32: tensors must be 16-byte aligned");
  if (rows == 0) return ZKB_OK;
  bn_gain_add_f32_kernel<<<br_grid(rows * (C / 4), 16), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, rows, C, mean, gain, beta, shortcut, out);
  ZKB_LAUNCH_CHECK("bn_gain_add_f32_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_bn_gain_bwd_apply_f32(const float* d, const float* x, long long rows, int C, const float* mean,
                                      const float* rs_bwd, const float* gain, const float* mean_d,
                                      const float* mean_dx, float* out, void* stream) {
  ZKB_REQUIRE(d && x && mean && rs_bwd && gain && mean_d && mean_dx && out, "bn_gain_bwd_apply_f32: null pointer");
  ZKB_REQUIRE(rows >= 0 && C > 0 && C % 4 == 0, "bn_gain_bwd_apply_f32: C=%d must be a multiple of 4", C);
  ZKB_REQUIRE(ZKB_A16(d) && ZKB_A16(x) && ZKB_A16(out), "bn_gain_bwd_apply_f32: misaligned tensors");
  if (rows == 0) return ZKB_OK;
  bn_gain_bwd_apply_f32_kernel<<<br_grid(rows * (C / 4), 16), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      d, x, rows, C, mean, rs_bwd, gain, mean_d, mean_dx, out);
  ZKB_LAUNCH_CHECK("bn_gain_bwd_apply_f32_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_avgpool2_f32_fwd(const float* x, int B, int H, int W, int C, float* out, void* stream) {
  ZKB_REQUIRE(x && out && B >= 0 && H >= 2 && W >= 2 && C > 0 && C % 4 == 0, "avgpool2_fwd: bad arguments (C %% 4 == 0)");
  ZKB_REQUIRE(ZKB_A16(x) && ZKB_A16(out), "avgpool2_fwd: misaligned tensors");
  if (B == 0) return ZKB_OK;
  avgpool2_fwd_kernel<<<br_grid(static_cast<long long>(B) * (H / 2) * (W / 2) * (C / 4), 16), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, B, H, W, C, out);
  ZKB_LAUNCH_CHECK("avgpool2_fwd_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_avgpool2_f32_bwd(const float* dout, int B, int H, int W, int C, float* dx, void* stream) {
  ZKB_REQUIRE(dout && dx && B >= 0 && H >= 2 && W >= 2 && C > 0 && C % 4 == 0, "avgpool2_bwd: bad arguments (C %% 4 == 0)");
  ZKB_REQUIRE(ZKB_A16(dout) && ZKB_A16(dx), "avgpool2_bwd: misaligned tensors");
  if (B == 0) return ZKB_OK;
  avgpool2_bwd_kernel<<<br_grid(static_cast<long long>(B) * H * W * (C / 4), 16), 256, 0, static_cast<cudaStream_t>(stream)>>>(dout, B, H, W, C, dx);
  ZKB_LAUNCH_CHECK("avgpool2_bwd_kernel");
  return ZKB_OK;
}
