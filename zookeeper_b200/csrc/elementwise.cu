// elementwise.cu -- HBM-bound kernels of the BinaryNet train step: fused preprocessing,
// first-layer (K=9) float conv, BatchNorm statistics / apply+sign / backward, 2x2 max-pool with
// first-max tie rule, softmax cross-entropy, weight sign/layout preparation, Adam + latent clip.
// All tensors are NHWC with channels innermost; every thread moves 16-byte vectors along C.
#include <cuda_bf16.h>

#include <mutex>
#include <stdlib.h>

#include "zkb_host.h"

namespace zkb {

__device__ __forceinline__ uint16_t pm1_bf16(bool nonneg) { return nonneg ? 0x3F80u : 0xBF80u; }

__device__ __forceinline__ uint16_t f32_to_bf16_rn(float f) {
  return __bfloat16_as_ushort(__float2bfloat16_rn(f));
}
__device__ __forceinline__ float bf16_to_f32(uint16_t h) {
  return __uint_as_float(static_cast<uint32_t>(h) << 16);
}

// Counter-based per-sample randomness (splitmix64 finaliser); oracle/binarynet_oracle.py
// augmentation_params() derives the same offsets from (seed, step, sample).
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// ------------------------------------------------------------------------------------------
// Preprocessing: gather + centre zero-pad + random crop + random flip + u8 -> f32 (x/127.5 - 1).
// A block works on ONE sample (blockIdx.y), so the sample's index, crop offsets and flip bit are
// worked out once per block; a thread produces 4 consecutive output floats of one row (one 16-byte
// store, fully coalesced across the warp) with a single division for its row number.  The 256
// possible results of float(px) / 127.5f - 1.0f (a true IEEE division, as the reference's
// `cast(image, float32) / (255.0 / 2.0) - 1.0` computes) are tabulated once per block in shared
// memory, so the conversion of an element is one shared-memory load -- bit-exact by construction.
// CT: channel count known at compile time (1, 3) or 0 for the generic path.
// ------------------------------------------------------------------------------------------
struct PreParams {
  const uint8_t* images;    // [n_pool, IH, IW, C]
  const long long* labels;  // [n_pool] or null
  const int* index;         // [B] sample ids, or null (identity)
  const int8_t* aug;        // [B,3] (ox, oy, flip) or null -> derive from (seed, step, sample)
  float* out;               // [B, OH, OW, C]
  int* labels_out;          // [B] or null
  float* onehot;            // [B, num_classes] or null
  int B, IH, IW, C, OH, OW, pad_size, training, binarize, num_classes;
  unsigned long long seed, step;
  const unsigned long long* step_dev;  // when non-null the step counter is read from device memory
                                       // (lets a captured CUDA graph advance the augmentation RNG)
};

__device__ __forceinline__ void sample_aug(const PreParams& p, int b, int& ox, int& oy, int& flip) {
  if (!p.training) {
    ox = oy = flip = 0;
    return;
  }
  if (p.aug) {
    ox = p.aug[3 * b];
    oy = p.aug[3 * b + 1];
    flip = p.aug[3 * b + 2];
    return;
  }
  const uint64_t step = p.step_dev ? *p.step_dev : p.step;
  const uint64_t r = mix64(mix64(p.seed ^ (step * 0xD1342543DE82EF95ull)) + static_cast<uint64_t>(b));
  const int rx = p.pad_size - p.OW + 1;
  const int ry = p.pad_size - p.OH + 1;
  ox = static_cast<int>((r & 0xFFFFu) % static_cast<uint32_t>(rx));
  oy = static_cast<int>(((r >> 16) & 0xFFFFu) % static_cast<uint32_t>(ry));
  flip = static_cast<int>((r >> 32) & 1u);
}

constexpr int kPreGroupsPerThread = 4;

// CT > 0 (and OW % 4 == 0): a work item is 4 consecutive output pixels of one row = CT float4
// stores; a thread takes kPreGroupsPerThread items 256 apart (coalesced across the block), so the
// table and the sample geometry are set up once per 4096 * CT outputs and the index arithmetic is
// per pixel, not per element.  CT == 0: generic path, one float4 of arbitrary (pixel, channel) mix.
template <int CT>
__global__ void __launch_bounds__(256)
preprocess_kernel(const PreParams p) {
  __shared__ float s_lut[256];
  __shared__ int s_geo[4];     // source sample, x shift, y shift, flip
  const int b = blockIdx.y;
  {
    float f = static_cast<float>(threadIdx.x) / 127.5f - 1.0f;
    if (p.binarize) f = f >= 0.0f ? 1.0f : -1.0f;
    s_lut[threadIdx.x] = f;
  }
  if (threadIdx.x == 0) {
    int ox, oy, flip;
    sample_aug(p, b, ox, oy, flip);
    // Training: pad to pad_size (centred), crop OHxOW at (oy, ox).  Eval: centre crop/pad to OHxOW.
    const int padded_h = p.training ? p.pad_size : p.OH;
    const int padded_w = p.training ? p.pad_size : p.OW;
    // offset of the source image inside the padded canvas (negative when cropping)
    const int off_h = (padded_h - p.IH) >= 0 ? (padded_h - p.IH) / 2 : -((p.IH - padded_h) / 2);
    const int off_w = (padded_w - p.IW) >= 0 ? (padded_w - p.IW) / 2 : -((p.IW - padded_w) / 2);
    const int src = p.index ? p.index[b] : b;
    s_geo[0] = src;
    s_geo[1] = ox - off_w;      // input column = output column (after un-flipping) + this
    s_geo[2] = oy - off_h;      // input row = output row + this
    s_geo[3] = flip;
    const long long lab = p.labels ? p.labels[src] : 0;
    if (blockIdx.x == 0) {
      if (p.labels_out) p.labels_out[b] = static_cast<int>(lab);
      if (p.onehot)
        for (int k = 0; k < p.num_classes; ++k)
          p.onehot[static_cast<long long>(b) * p.num_classes + k] = (k == lab) ? 1.0f : 0.0f;
    }
  }
  __syncthreads();
  const int dx = s_geo[1], flip = s_geo[3];
  const float pad_value = s_lut[0];                       // outside the source image: pixel value 0
  if constexpr (CT > 0) {
    const unsigned gpr = static_cast<unsigned>(p.OW) >> 2;           // 4-pixel groups per row
    const unsigned groups = static_cast<unsigned>(p.OH) * gpr;
    const uint8_t* img = p.images + static_cast<long long>(s_geo[0]) * p.IH * p.IW * CT;
    float* out_img = p.out + static_cast<long long>(b) * groups * (4 * CT);
#pragma unroll
    for (int k = 0; k < kPreGroupsPerThread; ++k) {
      const unsigned g = (blockIdx.x * kPreGroupsPerThread + k) * 256u + threadIdx.x;
      if (g >= groups) break;
      const unsigned oh = g / gpr;
      const unsigned ow0 = (g - oh * gpr) << 2;
      const int ih = static_cast<int>(oh) + s_geo[2];
      float o[4 * CT];
#pragma unroll
      for (int j = 0; j < 4 * CT; ++j) o[j] = pad_value;
      if (ih >= 0 && ih < p.IH) {
        const uint8_t* row = img + static_cast<long long>(ih) * p.IW * CT;
        // the four pixels are 4*CT CONTIGUOUS source bytes (descending pixel order when flipped),
        // starting at the lowest input column iw_first
        const int iw_first = (flip ? p.OW - 4 - static_cast<int>(ow0) : static_cast<int>(ow0)) + dx;
        const bool tail_ok = (ih + 1 < p.IH) || ((p.IW - iw_first - 4) * CT >= 3);
        if (iw_first >= 0 && iw_first + 4 <= p.IW && tail_ok) {
          // Interior group: fetch the bytes as aligned 32-bit words (at most 4 loads instead of
          // 4*CT byte loads 4*CT bytes apart across the warp) and realign them with funnel shifts.
          // Only words that hold a needed byte are read; a word may reach up to 3 bytes past the
          // group, which tail_ok keeps inside this image.
          const uint8_t* a0 = row + iw_first * CT;
          const unsigned m = static_cast<unsigned>(reinterpret_cast<uintptr_t>(a0) & 3u);
          const uint32_t* wp = reinterpret_cast<const uint32_t*>(a0 - m);
          const unsigned nw = (m + 4u * CT + 3u) >> 2;
          uint32_t w[CT + 1];
#pragma unroll
          for (int k = 0; k < CT + 1; ++k) w[k] = (static_cast<unsigned>(k) < nw) ? __ldg(wp + k) : 0u;
          uint32_t r[CT];
#pragma unroll
          for (int k = 0; k < CT; ++k) r[k] = __funnelshift_r(w[k], w[k + 1], 8u * m);
#pragma unroll
          for (int qs = 0; qs < 4; ++qs) {        // qs: pixel in ascending source order
            const int q = flip ? 3 - qs : qs;     // its place in the output group
#pragma unroll
            for (int c = 0; c < CT; ++c) {
              const int j = qs * CT + c;
              o[q * CT + c] = s_lut[(r[j >> 2] >> (8 * (j & 3))) & 0xFFu];
            }
          }
        } else {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int ow = static_cast<int>(ow0) + q;
            const int iw = (flip ? p.OW - 1 - ow : ow) + dx;
            if (iw >= 0 && iw < p.IW) {
#pragma unroll
              for (int c = 0; c < CT; ++c) o[q * CT + c] = s_lut[__ldg(row + iw * CT + c)];
            }
          }
        }
      }
      float4* dst = reinterpret_cast<float4*>(out_img + static_cast<size_t>(g) * (4 * CT));
#pragma unroll
      for (int v = 0; v < CT; ++v) dst[v] = make_float4(o[4 * v], o[4 * v + 1], o[4 * v + 2], o[4 * v + 3]);
    }
  } else {
    const int C = p.C;
    const unsigned row_elems = static_cast<unsigned>(p.OW * C);
    const unsigned vec_per_row = row_elems >> 2;
    const unsigned vec_per_img = static_cast<unsigned>(p.OH) * vec_per_row;
    const unsigned v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= vec_per_img) return;
    const unsigned oh = v / vec_per_row;
    const unsigned re = (v - oh * vec_per_row) << 2;      // first element of this vector in its row
    const int ih = static_cast<int>(oh) + s_geo[2];
    float o[4] = {pad_value, pad_value, pad_value, pad_value};
    if (ih >= 0 && ih < p.IH) {
      const uint8_t* row = p.images + (static_cast<long long>(s_geo[0]) * p.IH + ih) * p.IW * C;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const unsigned e = re + j;
        unsigned ow = e / static_cast<unsigned>(C);
        const unsigned c = e - ow * C;
        if (flip) ow = p.OW - 1 - ow;
        const int iw = static_cast<int>(ow) + dx;
        if (iw >= 0 && iw < p.IW) o[j] = s_lut[__ldg(row + iw * C + c)];
      }
    }
    float* dst = p.out + (static_cast<long long>(b) * vec_per_img + v) * 4;
    *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
  }
}

// ------------------------------------------------------------------------------------------
// First layer: float image (*) sign(W), 'valid', Cin small (K = ks*ks*Cin <= 32).  Output-write bound.
// ------------------------------------------------------------------------------------------
template <int KS, int CIN>
__global__ void __launch_bounds__(256)
conv_first_fwd_kernel(const float* __restrict__ x, const float* __restrict__ wq,  // [K][Cout]
                      float* __restrict__ y, int B, int H, int W, int Cout) {
  extern __shared__ float s_w[];  // [K][Cout]
  constexpr int K = KS * KS * CIN;
  for (int i = threadIdx.x; i < K * Cout; i += blockDim.x) s_w[i] = wq[i];
  __syncthreads();
  const int OH = H - KS + 1, OW = W - KS + 1;
  const int cvec = Cout / 4;
  const long long total = static_cast<long long>(B) * OH * OW * cvec;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cv = static_cast<int>(i % cvec);
    const long long pix = i / cvec;
    const int ow = static_cast<int>(pix % OW);
    const long long t = pix / OW;
    const int oh = static_cast<int>(t % OH);
    const int b = static_cast<int>(t / OH);
    float xv[K];
#pragma unroll
    for (int kh = 0; kh < KS; ++kh)
#pragma unroll
      for (int kw = 0; kw < KS; ++kw)
#pragma unroll
        for (int ci = 0; ci < CIN; ++ci)
          xv[(kh * KS + kw) * CIN + ci] =
              __ldg(x + ((static_cast<long long>(b) * H + oh + kh) * W + ow + kw) * CIN + ci);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const float4 w = *reinterpret_cast<const float4*>(s_w + k * Cout + cv * 4);
      acc.x = fmaf(xv[k], w.x, acc.x);
      acc.y = fmaf(xv[k], w.y, acc.y);
      acc.z = fmaf(xv[k], w.z, acc.z);
      acc.w = fmaf(xv[k], w.w, acc.w);
    }
    *reinterpret_cast<float4*>(y + pix * Cout + cv * 4) = acc;
  }
}

// dW[k][co] = sum_pix x[pix + tap(k)] * dy[pix][co].  Block = Cout threads x row-groups.
template <int KS, int CIN>
__global__ void __launch_bounds__(512)
conv_first_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                        float* __restrict__ dw, int B, int H, int W, int Cout) {
  constexpr int K = KS * KS * CIN;
  const int OH = H - KS + 1, OW = W - KS + 1;
  const long long pixels = static_cast<long long>(B) * OH * OW;
  const int co = threadIdx.x % Cout;
  const int rg = threadIdx.x / Cout;
  const int nrg = blockDim.x / Cout;
  float acc[K];
#pragma unroll
  for (int k = 0; k < K; ++k) acc[k] = 0.f;
  for (long long pix = static_cast<long long>(blockIdx.x) * nrg + rg; pix < pixels;
       pix += static_cast<long long>(gridDim.x) * nrg) {
    const int ow = static_cast<int>(pix % OW);
    const long long t = pix / OW;
    const int oh = static_cast<int>(t % OH);
    const int b = static_cast<int>(t / OH);
    const float g = dy[pix * Cout + co];
#pragma unroll
    for (int kh = 0; kh < KS; ++kh)
#pragma unroll
      for (int kw = 0; kw < KS; ++kw)
#pragma unroll
        for (int ci = 0; ci < CIN; ++ci)
          acc[(kh * KS + kw) * CIN + ci] = fmaf(
              __ldg(x + ((static_cast<long long>(b) * H + oh + kh) * W + ow + kw) * CIN + ci), g,
              acc[(kh * KS + kw) * CIN + ci]);
  }
#pragma unroll
  for (int k = 0; k < K; ++k) atomicAdd(dw + k * Cout + co, acc[k]);
}

// ------------------------------------------------------------------------------------------
// BatchNorm statistics: per-channel sum and sum of squares over rows (double accumulators in
// global memory; exact for the integer-valued conv outputs).
// ------------------------------------------------------------------------------------------
template <typename T>
struct Vec16;
template <>
struct Vec16<float> {
  static constexpr int N = 4;
  float v[4];
  __device__ __forceinline__ void load(const float* p) {
    const float4 t = *reinterpret_cast<const float4*>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  __device__ __forceinline__ float get(int i) const { return v[i]; }
};
template <>
struct Vec16<int16_t> {
  static constexpr int N = 8;
  int16_t v[8];
  __device__ __forceinline__ void load(const int16_t* p) {
    const uint4 t = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = static_cast<int16_t>(w[i] & 0xFFFFu);
      v[2 * i + 1] = static_cast<int16_t>(w[i] >> 16);
    }
  }
  __device__ __forceinline__ float get(int i) const { return static_cast<float>(v[i]); }
};

// Reduce per-thread partials (VEC channels each) across the row-lanes of a block and add the
// block totals to the global double accumulators acc0[C], acc1[C].
template <int VEC>
__device__ __forceinline__ void block_channel_reduce(const double (&a0)[VEC], const double (&a1)[VEC],
                                                     int tpr, int C, double* acc0, double* acc1) {
  // 4 KB of static shared memory, one vector slot at a time: small enough for these blocks to be
  // co-resident with a GEMM CTA (which leaves ~30 KB of the SM's shared memory free) when the
  // engine overlaps them with a wgrad kernel.
  __shared__ double s_part[2][256];
  const int lanes = blockDim.x / tpr;
  const int cl = threadIdx.x % tpr;
#pragma unroll
  for (int j = 0; j < VEC; ++j) {
    s_part[0][threadIdx.x] = a0[j];
    s_part[1][threadIdx.x] = a1[j];
    __syncthreads();
    if (threadIdx.x < tpr) {   // thread cl sums its channel (cl*VEC + j) over the row lanes
      double t0 = 0.0, t1 = 0.0;
      for (int r = 0; r < lanes; ++r) {
        t0 += s_part[0][r * tpr + cl];
        t1 += s_part[1][r * tpr + cl];
      }
      atomicAdd(acc0 + cl * VEC + j, t0);
      atomicAdd(acc1 + cl * VEC + j, t1);
    }
    __syncthreads();
  }
}

// Per-thread partial sums that stay exact for the integer maps: int16 inputs accumulate in
// 32-bit integers (|x| <= 4608: 128 rows of x fit int32, 128 rows of x^2 fit uint32) and are
// flushed to double every 128 rows; float inputs accumulate in float and flush every 32 rows.
template <typename T, int VEC>
struct StatAcc;
template <int VEC>
struct StatAcc<int16_t, VEC> {
  static constexpr int kFlush = 128;
  int s0[VEC];
  unsigned s1[VEC];
  __device__ __forceinline__ void clear() {
#pragma unroll
    for (int j = 0; j < VEC; ++j) { s0[j] = 0; s1[j] = 0u; }
  }
  __device__ __forceinline__ void add(const Vec16<int16_t>& v) {
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      const int t = v.v[j];
      s0[j] += t;
      s1[j] += static_cast<unsigned>(t * t);
    }
  }
  __device__ __forceinline__ void flush(double (&a0)[VEC], double (&a1)[VEC]) {
#pragma unroll
    for (int j = 0; j < VEC; ++j) { a0[j] += s0[j]; a1[j] += s1[j]; }
    clear();
  }
};
template <int VEC>
struct StatAcc<float, VEC> {
  static constexpr int kFlush = 32;
  float s0[VEC], s1[VEC];
  __device__ __forceinline__ void clear() {
#pragma unroll
    for (int j = 0; j < VEC; ++j) { s0[j] = 0.f; s1[j] = 0.f; }
  }
  __device__ __forceinline__ void add(const Vec16<float>& v) {
#pragma unroll
    for (int j = 0; j < VEC; ++j) { s0[j] += v.v[j]; s1[j] = fmaf(v.v[j], v.v[j], s1[j]); }
  }
  __device__ __forceinline__ void flush(double (&a0)[VEC], double (&a1)[VEC]) {
#pragma unroll
    for (int j = 0; j < VEC; ++j) { a0[j] += s0[j]; a1[j] += s1[j]; }
    clear();
  }
};

template <typename T>
__global__ void __launch_bounds__(256)
bn_stats_kernel(const T* __restrict__ x, long long rows, int C, double* __restrict__ sum,
                double* __restrict__ sumsq, long long ld = 0) {
  // ld: row pitch in elements when this launch covers a column slice of a wider map (0: pitch = C)
  constexpr int VEC = Vec16<T>::N;
  const long long pitch = ld ? ld : C;
  const int tpr = C / VEC;
  const int lanes = blockDim.x / tpr;
  const int cl = threadIdx.x % tpr;
  const int rl = threadIdx.x / tpr;
  double a0[VEC], a1[VEC];
#pragma unroll
  for (int j = 0; j < VEC; ++j) a0[j] = a1[j] = 0.0;
  if (rl < lanes) {
    StatAcc<T, VEC> acc;
    acc.clear();
    int n = 0;
    const long long step = static_cast<long long>(gridDim.x) * lanes;
    long long r = static_cast<long long>(blockIdx.x) * lanes + rl;
    const T* px = x + cl * VEC;
    // four independent 16-byte loads in flight per thread
    for (; r + 3 * step < rows; r += 4 * step) {
      Vec16<T> v0, v1, v2, v3;
      v0.load(px + r * pitch);
      v1.load(px + (r + step) * pitch);
      v2.load(px + (r + 2 * step) * pitch);
      v3.load(px + (r + 3 * step) * pitch);
      acc.add(v0); acc.add(v1); acc.add(v2); acc.add(v3);
      n += 4;
      if (n >= StatAcc<T, VEC>::kFlush - 3) { acc.flush(a0, a1); n = 0; }
    }
    for (; r < rows; r += step) {
      Vec16<T> v;
      v.load(px + r * pitch);
      acc.add(v);
      if (++n >= StatAcc<T, VEC>::kFlush - 3) { acc.flush(a0, a1); n = 0; }
    }
    acc.flush(a0, a1);
  }
  block_channel_reduce<VEC>(a0, a1, tpr, C, sum, sumsq);
}

// Per channel: mean, rstd (for normalisation, biased variance), moving-stat update.
__global__ void bn_finalize_kernel(const double* __restrict__ sum, const double* __restrict__ sumsq,
                                   double count, float eps, float momentum, int bessel, int C,
                                   float* __restrict__ mean_out, float* __restrict__ rstd_out,
                                   float* __restrict__ moving_mean, float* __restrict__ moving_var) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const double m = sum[c] / count;
  double var = sumsq[c] / count - m * m;
  if (var < 0.0) var = 0.0;
  const float meanf = static_cast<float>(m);
  const float varf = static_cast<float>(var);
  mean_out[c] = meanf;
  rstd_out[c] = rsqrtf(varf + eps);
  if (moving_mean) {
    const float uv = (bessel && count > 1.0) ? static_cast<float>(var * (count / (count - 1.0))) : varf;
    moving_mean[c] = moving_mean[c] * momentum + meanf * (1.0f - momentum);
    moving_var[c] = moving_var[c] * momentum + uv * (1.0f - momentum);
  }
}

// a = (x - mean) * rstd + beta ; s = sign(a) with sign(0) = +1, written as bf16 +-1.
template <typename T>
__global__ void __launch_bounds__(256)
bn_apply_sign_kernel(const T* __restrict__ x, long long rows, int C, const float* __restrict__ mean,
                     const float* __restrict__ rstd, const float* __restrict__ beta,
                     uint16_t* __restrict__ s_out, float* __restrict__ a_out,
                     int8_t* __restrict__ s8_out = nullptr, long long ld = 0) {
  constexpr int VEC = Vec16<T>::N;
  const long long pitch = ld ? ld : C;     // row pitch of x and of every output (see bn_stats_kernel)
  const int tpr = C / VEC;
  const long long total = rows * tpr;
  // blockDim and the grid stride are multiples of tpr, so a thread keeps one channel group:
  // its per-channel parameters live in registers for the whole loop.
  const int cl = threadIdx.x % tpr;
  float mu[VEC], rs[VEC], be[VEC];
#pragma unroll
  for (int j = 0; j < VEC; ++j) {
    mu[j] = mean[cl * VEC + j]; rs[j] = rstd[cl * VEC + j];
    be[j] = beta[cl * VEC + j];
  }
  const long long gstride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i0 = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i0 < total;
       i0 += 2 * gstride) {
   // two independent rows per iteration: both loads are issued before either is consumed
   Vec16<T> vv[2];
   const bool has2 = i0 + gstride < total;
   vv[0].load(x + (i0 / tpr) * pitch + cl * VEC);
   if (has2) vv[1].load(x + ((i0 + gstride) / tpr) * pitch + cl * VEC);
#pragma unroll
   for (int u = 0; u < 2; ++u) {
    if (u == 1 && !has2) break;
    const long long r = (i0 + u * gstride) / tpr;
    const Vec16<T>& v = vv[u];
    uint16_t s[VEC];
    float a[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      a[j] = (v.get(j) - mu[j]) * rs[j] + be[j];
      s[j] = pm1_bf16(a[j] >= 0.0f);
    }
    if (s_out) {
      uint16_t* o = s_out + r * pitch + cl * VEC;
      if (VEC == 8) {
        *reinterpret_cast<uint4*>(o) =
            make_uint4(s[0] | (uint32_t(s[1]) << 16), s[2] | (uint32_t(s[3]) << 16),
                       s[4 % VEC] | (uint32_t(s[5 % VEC]) << 16), s[6 % VEC] | (uint32_t(s[7 % VEC]) << 16));
      } else {
        *reinterpret_cast<uint2*>(o) =
            make_uint2(s[0] | (uint32_t(s[1]) << 16), s[2] | (uint32_t(s[3]) << 16));
      }
    }
    if (s8_out) {
      // the same signs as int8 (+1 / -1): operand of the kind::i8 forward GEMM
      uint32_t w8[VEC / 4];
#pragma unroll
      for (int q = 0; q < VEC / 4; ++q) {
        w8[q] = 0;
#pragma unroll
        for (int e = 0; e < 4; ++e)
          w8[q] |= ((s[4 * q + e] & 0x8000u) ? 0xFFu : 0x01u) << (8 * e);
      }
      if (VEC == 8)
        *reinterpret_cast<uint2*>(s8_out + r * pitch + cl * VEC) = make_uint2(w8[0], w8[(VEC / 4) - 1]);
      else
        *reinterpret_cast<uint32_t*>(s8_out + r * pitch + cl * VEC) = w8[0];
    }
    if (a_out) {
      float* o = a_out + r * pitch + cl * VEC;
#pragma unroll
      for (int j = 0; j < VEC; j += 4)
        *reinterpret_cast<float4*>(o + j) = make_float4(a[j], a[j + 1], a[j + 2], a[j + 3]);
    }
   }
  }
}

// Inference: sign((y - mean) * rstd + beta) for an integer y is a step function of y (rstd > 0),
// so BatchNorm + sign fold into one integer threshold per channel: the smallest y whose affine
// value -- the SAME float expression bn_apply_sign_kernel evaluates -- is >= 0.  Found by bisection
// over the int16 range; INT_MAX / INT_MIN when the sign never / always is +1.
__global__ void bn_sign_threshold_kernel(const float* __restrict__ mean, const float* __restrict__ rstd,
                                         const float* __restrict__ beta, int C, int* __restrict__ thr) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float mu = mean[c], rs = rstd[c], be = beta[c];
  auto plus = [&](int y) { return (static_cast<float>(y) - mu) * rs + be >= 0.0f; };
  int lo = -32768, hi = 32767;      // invariant: !plus(lo - 1) (virtually), plus(hi + 1) (virtually)
  if (!plus(hi)) { thr[c] = 0x7FFFFFFF; return; }
  if (plus(lo)) { thr[c] = static_cast<int>(0x80000000u); return; }
  while (hi - lo > 1) {             // !plus(lo), plus(hi)
    const int mid = lo + (hi - lo) / 2;
    if (plus(mid)) hi = mid; else lo = mid;
  }
  thr[c] = hi;
}

// Inference scale from the moving variance: rstd = rsqrt(var + eps) (as bn_finalize_kernel does
// for the batch variance).
__global__ void bn_eval_rstd_kernel(const float* __restrict__ var, float eps, int C,
                                    float* __restrict__ rstd) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < C) rstd[c] = rsqrtf(var[c] + eps);
}

// 2x2/2 max-pool of int8 signs: +1 if any of the four is +1 (= sign(BN(max y)) for rstd > 0).
__global__ void __launch_bounds__(256)
maxpool2_sign_i8_kernel(const int8_t* __restrict__ s, int B, int H, int W, int C,
                        int8_t* __restrict__ out) {
  const int PH = H / 2, PW = W / 2, tpr = C / 16;
  const long long total = static_cast<long long>(B) * PH * PW * tpr;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cl = static_cast<int>(i % tpr);
    long long r = i / tpr;
    const int pw = static_cast<int>(r % PW);
    r /= PW;
    const int ph = static_cast<int>(r % PH);
    const long long b = r / PH;
    const int8_t* base = s + ((b * H + 2 * ph) * W + 2 * pw) * C + cl * 16;
    const uint4 a = *reinterpret_cast<const uint4*>(base);
    const uint4 bq = *reinterpret_cast<const uint4*>(base + C);
    const uint4 cq = *reinterpret_cast<const uint4*>(base + static_cast<long long>(W) * C);
    const uint4 d = *reinterpret_cast<const uint4*>(base + static_cast<long long>(W) * C + C);
    // bytes are 0x01 (+1) or 0xFF (-1): the bitwise AND of the four is 0x01 as soon as one of
    // them is +1 and 0xFF only if all are -1, i.e. it is their maximum
    uint4 m;
    m.x = a.x & bq.x & cq.x & d.x;
    m.y = a.y & bq.y & cq.y & d.y;
    m.z = a.z & bq.z & cq.z & d.z;
    m.w = a.w & bq.w & cq.w & d.w;
    *reinterpret_cast<uint4*>(out + ((b * PH + ph) * PW + pw) * C + cl * 16) = m;
  }
}

// ------------------------------------------------------------------------------------------
// 2x2/2 'valid' max-pool on integer conv outputs, first maximum in row-major window order wins
// (strict > replacement), fused with the BatchNorm statistics of the pooled map.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void pool_one(const int16_t* __restrict__ y, int H, int W, int C, int PH,
                                         int PW, long long r, int cl, int16_t* __restrict__ p_out,
                                         uint8_t* __restrict__ arg_out, Vec16<int16_t>& best) {
  constexpr int VEC = 8;
  const int pw = static_cast<int>(r % PW);
  const long long t = r / PW;
  const int ph = static_cast<int>(t % PH);
  const long long b = t / PH;
  const int16_t* base = y + ((b * H + 2 * ph) * W + 2 * pw) * C + cl * VEC;
  Vec16<int16_t> v01, v10, v11;
  best.load(base);
  v01.load(base + C);
  v10.load(base + static_cast<long long>(W) * C);
  v11.load(base + static_cast<long long>(W) * C + C);
  uint8_t am[VEC];
#pragma unroll
  for (int j = 0; j < VEC; ++j) {
    uint8_t k = 0;
    if (v01.v[j] > best.v[j]) { best.v[j] = v01.v[j]; k = 1; }
    if (v10.v[j] > best.v[j]) { best.v[j] = v10.v[j]; k = 2; }
    if (v11.v[j] > best.v[j]) { best.v[j] = v11.v[j]; k = 3; }
    am[j] = k;
  }
  uint32_t w[4];
#pragma unroll
  for (int e = 0; e < 4; ++e)
    w[e] = (static_cast<uint32_t>(static_cast<uint16_t>(best.v[2 * e]))) |
           (static_cast<uint32_t>(static_cast<uint16_t>(best.v[2 * e + 1])) << 16);
  *reinterpret_cast<uint4*>(p_out + r * C + cl * VEC) = make_uint4(w[0], w[1], w[2], w[3]);
  const uint32_t alo = am[0] | (am[1] << 8) | (am[2] << 16) | (am[3] << 24);
  const uint32_t ahi = am[4] | (am[5] << 8) | (am[6] << 16) | (am[7] << 24);
  *reinterpret_cast<uint2*>(arg_out + r * C + cl * VEC) = make_uint2(alo, ahi);
}

// Register budget: 3 blocks of 256 threads per SM (the earlier form kept two pooled pixels in
// flight plus float64 accumulators, 122 registers, 2 blocks per SM: ncu showed 24 % of the warp
// slots filled and 4.0 TB/s).  The statistics accumulate as 64-bit integers (exact, converted once).
__global__ void __launch_bounds__(256, 3)
maxpool2_i16_kernel(const int16_t* __restrict__ y, int B, int H, int W, int C,
                    int16_t* __restrict__ p_out, uint8_t* __restrict__ arg_out,
                    double* __restrict__ sum, double* __restrict__ sumsq) {
  constexpr int VEC = 8;
  const int tpr = C / VEC;
  const int lanes = blockDim.x / tpr;
  const int cl = threadIdx.x % tpr;
  const int rl = threadIdx.x / tpr;
  const int PH = H / 2, PW = W / 2;
  const long long rows = static_cast<long long>(B) * PH * PW;
  long long s0[VEC], s1[VEC];
#pragma unroll
  for (int j = 0; j < VEC; ++j) s0[j] = s1[j] = 0;
  if (rl < lanes) {
    const long long step = static_cast<long long>(gridDim.x) * lanes;
    for (long long r = static_cast<long long>(blockIdx.x) * lanes + rl; r < rows; r += step) {
      Vec16<int16_t> m0;
      pool_one(y, H, W, C, PH, PW, r, cl, p_out, arg_out, m0);
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        const int t = m0.v[j];
        s0[j] += t;
        s1[j] += t * t;          // |t| <= 4608: t*t fits int32
      }
    }
  }
  double a0[VEC], a1[VEC];
#pragma unroll
  for (int j = 0; j < VEC; ++j) { a0[j] = static_cast<double>(s0[j]); a1[j] = static_cast<double>(s1[j]); }
  block_channel_reduce<VEC>(a0, a1, tpr, C, sum, sumsq);
}

// ------------------------------------------------------------------------------------------
// softmax + sparse categorical cross-entropy (mean over batch) and its gradient w.r.t. logits.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
softmax_ce_kernel(const float* __restrict__ logits, int ld, const int* __restrict__ labels, int B,
                  int NC, float inv_global_batch, float* __restrict__ dlogits,
                  double* __restrict__ loss_sum, int* __restrict__ correct) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  float loss = 0.f;
  int ok = 0;
  if (b < B) {
    const float* z = logits + static_cast<long long>(b) * ld;
    float mx = z[0];
    int am = 0;
    for (int k = 1; k < NC; ++k)
      if (z[k] > mx) { mx = z[k]; am = k; }
    float se = 0.f;
    for (int k = 0; k < NC; ++k) se += expf(z[k] - mx);
    const float lse = logf(se) + mx;
    const int lab = labels[b];
    loss = lse - z[lab];
    ok = (am == lab);
    if (dlogits) {
      float* d = dlogits + static_cast<long long>(b) * ld;
      for (int k = 0; k < ld; ++k) {
        float g = 0.f;
        if (k < NC) g = (expf(z[k] - lse) - (k == lab ? 1.f : 0.f)) * inv_global_batch;
        d[k] = g;
      }
    }
  }
  // block reduce
  __shared__ float s_loss[8];
  __shared__ int s_ok[8];
  for (int o = 16; o > 0; o >>= 1) {
    loss += __shfl_down_sync(0xffffffffu, loss, o);
    ok += __shfl_down_sync(0xffffffffu, ok, o);
  }
  if ((threadIdx.x & 31) == 0) { s_loss[threadIdx.x >> 5] = loss; s_ok[threadIdx.x >> 5] = ok; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double tl = 0.0;
    int tk = 0;
    for (int w = 0; w < (blockDim.x >> 5); ++w) { tl += s_loss[w]; tk += s_ok[w]; }
    atomicAdd(loss_sum, tl);
    atomicAdd(correct, tk);
  }
}

// The same for wide heads (the 1000-class zoo models): one warp per row, lanes strided over the
// classes -- coalesced reads and gradient writes instead of one thread walking a 4 KB row.  argmax
// keeps the first maximum (lowest index), as the thread-per-row form does.
__global__ void __launch_bounds__(256)
softmax_ce_wide_kernel(const float* __restrict__ logits, int ld, const int* __restrict__ labels, int B,
                       int NC, float inv_global_batch, float* __restrict__ dlogits,
                       double* __restrict__ loss_sum, int* __restrict__ correct) {
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= B) return;                       // whole warps leave together
  const float* z = logits + static_cast<long long>(b) * ld;
  float mx = -INFINITY;
  int am = 0x7FFFFFFF;
  for (int k = lane; k < NC; k += 32) {
    const float v = z[k];
    if (v > mx) { mx = v; am = k; }
  }
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, mx, o);
    const int oa = __shfl_xor_sync(0xffffffffu, am, o);
    if (om > mx || (om == mx && oa < am)) { mx = om; am = oa; }
  }
  float se = 0.f;
  for (int k = lane; k < NC; k += 32) se += expf(z[k] - mx);
  for (int o = 16; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
  const float lse = logf(se) + mx;
  const int lab = labels[b];
  if (dlogits) {
    float* d = dlogits + static_cast<long long>(b) * ld;
    for (int k = lane; k < ld; k += 32)
      d[k] = k < NC ? (expf(z[k] - lse) - (k == lab ? 1.f : 0.f)) * inv_global_batch : 0.f;
  }
  if (lane == 0) {
    atomicAdd(loss_sum, static_cast<double>(lse - z[lab]));
    if (am == lab) atomicAdd(correct, 1);
  }
}

// ------------------------------------------------------------------------------------------
// BatchNorm backward.  g is the gradient w.r.t. sign(a) (ste != 0: masked by |a| <= 1) or w.r.t.
// a itself (ste == 0, last layer).  Pass 1: per-channel sum(da), sum(da * xhat).
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void load4(const float* p, float (&o)[4]) {
  const float4 t = *reinterpret_cast<const float4*>(p);
  o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w;
}
__device__ __forceinline__ void load4(const int16_t* p, float (&o)[4]) {
  const uint2 t = *reinterpret_cast<const uint2*>(p);
  o[0] = static_cast<float>(static_cast<int16_t>(t.x & 0xFFFFu));
  o[1] = static_cast<float>(static_cast<int16_t>(t.x >> 16));
  o[2] = static_cast<float>(static_cast<int16_t>(t.y & 0xFFFFu));
  o[3] = static_cast<float>(static_cast<int16_t>(t.y >> 16));
}

template <typename T>
__global__ void __launch_bounds__(256)
bn_bwd_reduce_kernel(const float* __restrict__ g, const T* __restrict__ x, long long rows, int C,
                     const float* __restrict__ mean, const float* __restrict__ rstd,
                     const float* __restrict__ beta, int ste, double* __restrict__ sum_da,
                     double* __restrict__ sum_dax, long long ld = 0) {
  constexpr int VEC = 4;  // g is f32: 4 channels per 16 B
  const long long pitch = ld ? ld : C;     // row pitch of g and x (see bn_stats_kernel)
  const int tpr = C / VEC;
  const int lanes = blockDim.x / tpr;
  const int cl = threadIdx.x % tpr;
  const int rl = threadIdx.x / tpr;
  double a0[VEC], a1[VEC];
#pragma unroll
  for (int j = 0; j < VEC; ++j) a0[j] = a1[j] = 0.0;
  if (rl < lanes) {
    float mu[VEC], rs[VEC], be[VEC], f0[VEC], f1[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      mu[j] = mean[cl * VEC + j]; rs[j] = rstd[cl * VEC + j]; be[j] = beta[cl * VEC + j];
      f0[j] = f1[j] = 0.f;
    }
    int n = 0;
    const long long step = static_cast<long long>(gridDim.x) * lanes;
    // four rows per iteration: all eight loads are issued before the first is consumed
    for (long long r0 = static_cast<long long>(blockIdx.x) * lanes + rl; r0 < rows; r0 += 4 * step) {
      float gg[4][4], xv[4][4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long long r = r0 + u * step;
        if (r < rows) {
          load4(g + r * pitch + cl * VEC, gg[u]);
          load4(x + r * pitch + cl * VEC, xv[u]);
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) { gg[u][j] = 0.f; xv[u][j] = mu[j]; }   // contributes nothing
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
          const float xh = (xv[u][j] - mu[j]) * rs[j];
          const float a = xh + be[j];
          const float da = (ste && fabsf(a) > 1.0f) ? 0.f : gg[u][j];
          f0[j] += da;
          f1[j] = fmaf(da, xh, f1[j]);
        }
      }
      n += 4;
      if (n >= 64) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) { a0[j] += f0[j]; a1[j] += f1[j]; f0[j] = f1[j] = 0.f; }
        n = 0;
      }
    }
#pragma unroll
    for (int j = 0; j < VEC; ++j) { a0[j] += f0[j]; a1[j] += f1[j]; }
  }
  block_channel_reduce<VEC>(a0, a1, tpr, C, sum_da, sum_dax);
}

// Per channel: dbeta = sum(da) into the gradient arena; means for the apply pass.
__global__ void bn_bwd_finalize_kernel(const double* __restrict__ sum_da,
                                       const double* __restrict__ sum_dax, double count,
                                       double dbeta_scale, int C, float* __restrict__ dbeta,
                                       float* __restrict__ mean_da, float* __restrict__ mean_dax) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  if (dbeta) dbeta[c] = static_cast<float>(sum_da[c] * dbeta_scale);
  mean_da[c] = static_cast<float>(sum_da[c] / count);
  mean_dax[c] = static_cast<float>(sum_dax[c] / count);
}

// Pass 2: dx = rstd * (da - mean(da) - xhat * mean(da*xhat)), emitted as
//   out_mode 0: f32 [rows][C]
//   out_mode 1: bf16 hi/lo split [rows][C]
//   out_mode 2: bf16 hi/lo split scattered through the 2x2 max-pool argmax into [B,H,W,C]
//               (x, g and arg are pooled maps [B,H/2,W/2,C]; the three losers get zeros).
template <typename T>
__global__ void __launch_bounds__(256)
bn_bwd_apply_kernel(const float* __restrict__ g, const T* __restrict__ x, long long rows, int C,
                    const float* __restrict__ mean, const float* __restrict__ rstd,
                    const float* __restrict__ beta, const float* __restrict__ mean_da,
                    const float* __restrict__ mean_dax, int ste, int out_mode,
                    float* __restrict__ out_f32, uint16_t* __restrict__ out_hi,
                    uint16_t* __restrict__ out_lo, const uint8_t* __restrict__ arg, int H, int W,
                    long long ld = 0) {
  constexpr int VEC = 4;
  const long long pitch = ld ? ld : C;     // row pitch of g, x and the outputs (modes 0, 1; see bn_stats_kernel)
  const int tpr = C / VEC;
  const long long total = rows * tpr;
  const int PH = H / 2, PW = W / 2;
  const int cl = threadIdx.x % tpr;   // fixed per thread (blockDim and grid stride are multiples of tpr)
  float mu[VEC], rs[VEC], be[VEC], mda[VEC], mdx[VEC];
#pragma unroll
  for (int j = 0; j < VEC; ++j) {
    const int c = cl * VEC + j;
    mu[j] = mean[c]; rs[j] = rstd[c]; be[j] = beta[c];
    mda[j] = mean_da[c]; mdx[j] = mean_dax[c];
  }
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / tpr;
    const float4 gv = *reinterpret_cast<const float4*>(g + r * pitch + cl * VEC);
    const float gg[4] = {gv.x, gv.y, gv.z, gv.w};
    float xv[4];
    load4(x + r * pitch + cl * VEC, xv);
    float dx[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      const float xh = (xv[j] - mu[j]) * rs[j];
      const float a = xh + be[j];
      const float da = (ste && fabsf(a) > 1.0f) ? 0.f : gg[j];
      dx[j] = rs[j] * (da - mda[j] - xh * mdx[j]);
    }
    if (out_mode == 0) {
      *reinterpret_cast<float4*>(out_f32 + r * pitch + cl * VEC) = make_float4(dx[0], dx[1], dx[2], dx[3]);
      continue;
    }
    uint16_t hi[VEC], lo[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      hi[j] = f32_to_bf16_rn(dx[j]);
      lo[j] = f32_to_bf16_rn(dx[j] - bf16_to_f32(hi[j]));
    }
    if (out_mode == 1) {
      *reinterpret_cast<uint2*>(out_hi + r * pitch + cl * VEC) =
          make_uint2(hi[0] | (uint32_t(hi[1]) << 16), hi[2] | (uint32_t(hi[3]) << 16));
      *reinterpret_cast<uint2*>(out_lo + r * pitch + cl * VEC) =
          make_uint2(lo[0] | (uint32_t(lo[1]) << 16), lo[2] | (uint32_t(lo[3]) << 16));
    } else {
      const int pw = static_cast<int>(r % PW);
      const long long t = r / PW;
      const int ph = static_cast<int>(t % PH);
      const long long b = t / PH;
      const uint32_t a4 = *reinterpret_cast<const uint32_t*>(arg + r * C + cl * VEC);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const long long pos =
            ((b * H + 2 * ph + (k >> 1)) * W + 2 * pw + (k & 1)) * static_cast<long long>(C) + cl * VEC;
        uint16_t h[VEC], l[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
          const bool win = ((a4 >> (8 * j)) & 0xFFu) == static_cast<uint32_t>(k);
          h[j] = win ? hi[j] : 0;
          l[j] = win ? lo[j] : 0;
        }
        *reinterpret_cast<uint2*>(out_hi + pos) =
            make_uint2(h[0] | (uint32_t(h[1]) << 16), h[2] | (uint32_t(h[3]) << 16));
        *reinterpret_cast<uint2*>(out_lo + pos) =
            make_uint2(l[0] | (uint32_t(l[1]) << 16), l[2] | (uint32_t(l[3]) << 16));
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// Weight preparation: sign(W) in the layouts the GEMM engines read.
// ------------------------------------------------------------------------------------------
// conv: W f32 [ks,ks,Cin,Cout] (HWIO) -> fwd [Cout][tap][Cin], dgrad [Cin][tap_flipped][Cout] (bf16)
// One 32x32 (ci, co) tile per block and tap, transposed through shared memory so that the read of
// W and all three writes are coalesced.
__global__ void __launch_bounds__(256)
prep_conv_weights_kernel(const float* __restrict__ w, int ks, int Cin, int Cout,
                         uint16_t* __restrict__ w_fwd, uint16_t* __restrict__ w_dgrad,
                         int8_t* __restrict__ w_fwd8) {
  __shared__ uint8_t neg[32][33];   // 1 = negative weight
  const int taps = ks * ks;
  const int tap = blockIdx.z;
  const int ci0 = blockIdx.y * 32, co0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int kh = tap / ks, kw = tap % ks;
  const int ftap = (ks - 1 - kh) * ks + (ks - 1 - kw);
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int ci = ci0 + ty + 8 * r, co = co0 + tx;
    if (ci < Cin && co < Cout) {
      const bool nonneg = w[(static_cast<long long>(tap) * Cin + ci) * Cout + co] >= 0.0f;
      neg[ty + 8 * r][tx] = nonneg ? 0 : 1;
      w_dgrad[(static_cast<long long>(ci) * taps + ftap) * Cout + co] = pm1_bf16(nonneg);
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int co = co0 + ty + 8 * r, ci = ci0 + tx;
    if (ci < Cin && co < Cout) {
      const bool nonneg = neg[tx][ty + 8 * r] == 0;
      const long long o = (static_cast<long long>(co) * taps + tap) * Cin + ci;
      w_fwd[o] = pm1_bf16(nonneg);
      if (w_fwd8) w_fwd8[o] = nonneg ? 1 : -1;
    }
  }
}

// dense: W f32 [K][N] -> fwd [Npad][K] (rows >= N zero), dgrad [K][Npad] (cols >= N zero); same
// tile transpose.
__global__ void __launch_bounds__(256)
prep_dense_weights_kernel(const float* __restrict__ w, int K, int N, int Npad,
                          uint16_t* __restrict__ w_fwd, uint16_t* __restrict__ w_dgrad,
                          int8_t* __restrict__ w_fwd8) {
  __shared__ uint8_t code[32][33];  // 0 = +1, 1 = -1, 2 = padding
  const int k0 = blockIdx.y * 32, n0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int k = k0 + ty + 8 * r, n = n0 + tx;
    if (k < K && n < Npad) {
      uint8_t c = 2;
      if (n < N) c = w[static_cast<long long>(k) * N + n] >= 0.0f ? 0 : 1;
      code[ty + 8 * r][tx] = c;
      w_dgrad[static_cast<long long>(k) * Npad + n] = c == 2 ? 0 : pm1_bf16(c == 0);
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int n = n0 + ty + 8 * r, k = k0 + tx;
    if (k < K && n < Npad) {
      const uint8_t c = code[tx][ty + 8 * r];
      const long long o = static_cast<long long>(n) * K + k;
      w_fwd[o] = c == 2 ? 0 : pm1_bf16(c == 0);
      if (w_fwd8) w_fwd8[o] = c == 2 ? 0 : (c == 0 ? 1 : -1);
    }
  }
}

// generic elementwise sign to f32 (first-layer kernel, consumed by the SIMT conv)
__global__ void __launch_bounds__(256)
sign_f32_kernel(const float* __restrict__ w, float* __restrict__ out, long long n) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    out[i] = w[i] >= 0.0f ? 1.0f : -1.0f;
}

// ------------------------------------------------------------------------------------------
// Adam (Keras "epsilon-hat" form, bias correction folded into lr_t on the host) followed by the
// latent-weight clip to [-1, 1] for kernels (Larq weight_clip); the STE mask of the kernel
// quantizer (|w| <= 1) is applied to the incoming gradient.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
adam_clip_kernel(float* __restrict__ w, const float* __restrict__ g, float* __restrict__ m,
                 float* __restrict__ v, long long n, float lr_t_host, float b1, float b2, float eps,
                 float gscale, int is_kernel, const float* __restrict__ lr_t_dev) {
  const float lr_t = lr_t_dev ? *lr_t_dev : lr_t_host;  // device copy: for captured CUDA graphs
  const long long nv = n / 4;
  const long long tid = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = tid; i < nv; i += stride) {
    float4 w4 = reinterpret_cast<float4*>(w)[i];
    const float4 g4 = reinterpret_cast<const float4*>(g)[i];
    float4 m4 = reinterpret_cast<float4*>(m)[i];
    float4 v4 = reinterpret_cast<float4*>(v)[i];
    float* wp = &w4.x; const float* gp = &g4.x; float* mp = &m4.x; float* vp = &v4.x;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float gg = gp[j] * gscale;
      if (is_kernel && fabsf(wp[j]) > 1.0f) gg = 0.f;
      // same operation order as TF's ApplyAdam functor (fp32): m += (g-m)(1-b1), v += (g^2-v)(1-b2)
      mp[j] += (gg - mp[j]) * (1.0f - b1);
      vp[j] += (gg * gg - vp[j]) * (1.0f - b2);
      float nw = wp[j] - (mp[j] * lr_t) / (sqrtf(vp[j]) + eps);
      if (is_kernel) nw = fminf(fmaxf(nw, -1.0f), 1.0f);
      wp[j] = nw;
    }
    reinterpret_cast<float4*>(w)[i] = w4;
    reinterpret_cast<float4*>(m)[i] = m4;
    reinterpret_cast<float4*>(v)[i] = v4;
  }
  for (long long i = nv * 4 + tid; i < n; i += stride) {
    float gg = g[i] * gscale;
    if (is_kernel && fabsf(w[i]) > 1.0f) gg = 0.f;
    m[i] += (gg - m[i]) * (1.0f - b1);
    v[i] += (gg * gg - v[i]) * (1.0f - b2);
    float nw = w[i] - (m[i] * lr_t) / (sqrtf(v[i]) + eps);
    if (is_kernel) nw = fminf(fmaxf(nw, -1.0f), 1.0f);
    w[i] = nw;
  }
}

}  // namespace zkb

// ============================================================================================
// C ABI (documented in include/zkb200.h)
// ============================================================================================
using namespace zkb;

static int reduce_waves() {
#ifdef ZKB_MEASURE
  static int w = -1;     // measurement builds only: ZKB_REDUCE_WAVES, read once
  if (w < 0) {
    const char* e = getenv("ZKB_REDUCE_WAVES");
    w = (e && atoi(e) > 0) ? atoi(e) : 8;
  }
  return w;
#else
  return 8;
#endif
}

// Grid of a per-channel reduction kernel: `lanes` row lanes per 256-thread block.  Every block ends
// with one float64 atomic per channel and sum, so a block should own at least ~32 rows per lane
// (ncu, r2: the dense layers' 4096-row reductions ran 512 blocks of 8 rows and spent their time in
// 1 M contended atomics), and at most reduce_waves() blocks per SM.
static int reduce_grid(long long rows, int lanes);

static int grid_for(long long work_items, int block, int max_waves = 8) {
  const long long cap = static_cast<long long>(zkb_sm_count()) * max_waves;
  long long g = (work_items + block - 1) / block;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return static_cast<int>(g);
}

static int reduce_grid(long long rows, int lanes) {
  long long g = (rows + static_cast<long long>(lanes) * 32 - 1) / (static_cast<long long>(lanes) * 32);
  const long long cap = static_cast<long long>(zkb_sm_count()) * reduce_waves();
  if (g > cap) g = cap;
  return g < 1 ? 1 : static_cast<int>(g);
}

#define ZKB_ALIGNED16(p) ((reinterpret_cast<uintptr_t>(p) & 15) == 0)

ZKB_API int zkb_preprocess_u8(const uint8_t* images, const long long* labels, const int* index,
                              const int8_t* aug, float* out, int* labels_out, float* onehot, int B,
                              int IH, int IW, int C, int OH, int OW, int pad_size, int training,
                              int binarize, int num_classes, unsigned long long seed,
                              unsigned long long step, const unsigned long long* step_dev,
                              void* stream) {
  ZKB_REQUIRE(images && out, "preprocess: null pointer");
  ZKB_REQUIRE(B >= 0 && IH > 0 && IW > 0 && C > 0 && OH > 0 && OW > 0, "preprocess: bad shape");
  if (B == 0) return ZKB_OK;
  ZKB_REQUIRE((OW * C) % 4 == 0, "preprocess: OW*C=%d must be a multiple of 4", OW * C);
  ZKB_REQUIRE(ZKB_ALIGNED16(out), "preprocess: out must be 16-byte aligned");
  if (training)
    ZKB_REQUIRE(pad_size >= OH && pad_size >= OW, "preprocess: pad_size %d < crop size", pad_size);
  ZKB_REQUIRE(!onehot || (labels && num_classes > 0), "preprocess: one-hot needs labels");
  PreParams p{images, labels, index, aug, out, labels_out, onehot, B, IH, IW, C, OH, OW,
              pad_size, training, binarize, num_classes, seed, step, step_dev};
  ZKB_REQUIRE(B <= 65535, "preprocess: at most 65535 samples per call (got %d)", B);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if ((C == 1 || C == 3) && OW % 4 == 0) {
    const unsigned groups = static_cast<unsigned>(OH) * static_cast<unsigned>(OW / 4);
    const unsigned per_block = 256u * kPreGroupsPerThread;
    const dim3 grid((groups + per_block - 1) / per_block, B);
    if (C == 1) preprocess_kernel<1><<<grid, 256, 0, st>>>(p);
    else preprocess_kernel<3><<<grid, 256, 0, st>>>(p);
  } else {
    const unsigned vec_per_img = static_cast<unsigned>(OH) * static_cast<unsigned>(OW * C / 4);
    const dim3 grid((vec_per_img + 255) / 256, B);
    preprocess_kernel<0><<<grid, 256, 0, st>>>(p);
  }
  ZKB_LAUNCH_CHECK("preprocess_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_conv_first_fwd(const float* x, const float* wq_k_co, float* y, int B, int H, int W,
                               int Cin, int Cout, int ksize, void* stream) {
  ZKB_REQUIRE(x && wq_k_co && y, "conv_first_fwd: null pointer");
  ZKB_REQUIRE(ksize == 3 && (Cin == 1 || Cin == 3),
              "conv_first_fwd: only 3x3 kernels over 1 or 3 input channels are built");
  ZKB_REQUIRE(Cout % 4 == 0 && ZKB_ALIGNED16(y), "conv_first_fwd: Cout %% 4 and 16-byte output");
  ZKB_REQUIRE(H >= ksize && W >= ksize, "conv_first_fwd: image smaller than the kernel");
  if (B == 0) return ZKB_OK;
  const int K = ksize * ksize * Cin;
  const long long work = static_cast<long long>(B) * (H - ksize + 1) * (W - ksize + 1) * (Cout / 4);
  const int grid = grid_for(work, 256, 16);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t smem = static_cast<size_t>(K) * Cout * sizeof(float);
  if (Cin == 1)
    conv_first_fwd_kernel<3, 1><<<grid, 256, smem, st>>>(x, wq_k_co, y, B, H, W, Cout);
  else
    conv_first_fwd_kernel<3, 3><<<grid, 256, smem, st>>>(x, wq_k_co, y, B, H, W, Cout);
  ZKB_LAUNCH_CHECK("conv_first_fwd_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_conv_first_wgrad(const float* x, const float* dy, float* dw_k_co, int B, int H,
                                 int W, int Cin, int Cout, int ksize, void* stream) {
  ZKB_REQUIRE(x && dy && dw_k_co, "conv_first_wgrad: null pointer");
  ZKB_REQUIRE(ksize == 3 && (Cin == 1 || Cin == 3),
              "conv_first_wgrad: only 3x3 kernels over 1 or 3 input channels are built");
  ZKB_REQUIRE(Cout <= 512 && 512 % Cout == 0, "conv_first_wgrad: Cout must divide 512");
  if (B == 0) return ZKB_OK;
  const long long pixels = static_cast<long long>(B) * (H - ksize + 1) * (W - ksize + 1);
  const int nrg = 512 / Cout;
  long long g = (pixels + nrg * 64 - 1) / (nrg * 64);
  const long long cap = static_cast<long long>(zkb_sm_count()) * 4;
  if (g > cap) g = cap;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (Cin == 1)
    conv_first_wgrad_kernel<3, 1><<<static_cast<int>(g), 512, 0, st>>>(x, dy, dw_k_co, B, H, W, Cout);
  else
    conv_first_wgrad_kernel<3, 3><<<static_cast<int>(g), 512, 0, st>>>(x, dy, dw_k_co, B, H, W, Cout);
  ZKB_LAUNCH_CHECK("conv_first_wgrad_kernel");
  return ZKB_OK;
}

// wide != 0: maps wider than 256 vectors per row are allowed (the launcher walks column slices)
static int check_rows_c(const char* who, const void* x, long long rows, int C, int vec, int wide = 0) {
  ZKB_REQUIRE(x, "%s: null pointer", who);
  ZKB_REQUIRE(rows >= 0 && C > 0, "%s: bad shape", who);
  ZKB_REQUIRE(C % vec == 0 && (wide || C / vec <= 256), "%s: C=%d must be a multiple of %d%s", who, C, vec,
              wide ? "" : ", at most 256 vectors per row");
  ZKB_REQUIRE(ZKB_ALIGNED16(x), "%s: tensor must be 16-byte aligned", who);
  return ZKB_OK;
}

// Column slices of a [rows, C] map for the per-channel kernels: a thread owns one 16-byte vector of
// channels, a block at most 256 of them, so wider maps (the 4096-unit dense layers of BinaryAlexNet)
// are walked in slices of 256 vectors with the full row as pitch.
static int slice_width(int C, int vec) { return C / vec <= 256 ? C : 256 * vec; }

// Threads per block of the per-channel kernels: a thread keeps ONE 16-byte vector of channels for the
// whole launch, so the block is a whole number of rows -- 256 threads when the vectors per row divide
// 256 (every layer of BinaryNet), the largest multiple below otherwise (192 / 384 channels).
static int row_block(int vectors_per_row) { return (256 / vectors_per_row) * vectors_per_row; }

ZKB_API int zkb_bn_stats(const void* x, int x_is_i16, long long rows, int C, double* sum,
                         double* sumsq, void* stream) {
  const int vec = x_is_i16 ? 8 : 4;
  int rc = check_rows_c("bn_stats", x, rows, C, vec, 1);
  if (rc) return rc;
  ZKB_REQUIRE(sum && sumsq, "bn_stats: null accumulator");
  if (rows == 0) return ZKB_OK;
  const size_t smem = 0;
  const int sw = slice_width(C, vec);
  for (int c0 = 0; c0 < C; c0 += sw) {
    const int cc = C - c0 < sw ? C - c0 : sw;
    const int lanes = 256 / (cc / vec);
    const int grid = reduce_grid(rows, lanes);
    if (x_is_i16)
      bn_stats_kernel<int16_t><<<grid, row_block(cc / vec), smem, static_cast<cudaStream_t>(stream)>>>(
          static_cast<const int16_t*>(x) + c0, rows, cc, sum + c0, sumsq + c0, C);
    else
      bn_stats_kernel<float><<<grid, row_block(cc / vec), smem, static_cast<cudaStream_t>(stream)>>>(
          static_cast<const float*>(x) + c0, rows, cc, sum + c0, sumsq + c0, C);
    ZKB_LAUNCH_CHECK("bn_stats_kernel");
  }
  return ZKB_OK;
}

ZKB_API int zkb_bn_finalize(const double* sum, const double* sumsq, double count, float eps,
                            float momentum, int bessel, int C, float* mean, float* rstd,
                            float* moving_mean, float* moving_var, void* stream) {
  ZKB_REQUIRE(sum && sumsq && mean && rstd, "bn_finalize: null pointer");
  ZKB_REQUIRE(count > 0 && C > 0, "bn_finalize: bad count / C");
  ZKB_REQUIRE((moving_mean == nullptr) == (moving_var == nullptr),
              "bn_finalize: moving_mean and moving_var go together");
  bn_finalize_kernel<<<(C + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      sum, sumsq, count, eps, momentum, bessel, C, mean, rstd, moving_mean, moving_var);
  ZKB_LAUNCH_CHECK("bn_finalize_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_bn_sign_thresholds(const float* mean, const float* rstd, const float* beta, int C,
                                   int* thr_i32, void* stream) {
  ZKB_REQUIRE(mean && rstd && beta && thr_i32, "bn_sign_thresholds: null pointer");
  ZKB_REQUIRE(C > 0, "bn_sign_thresholds: C must be positive");
  bn_sign_threshold_kernel<<<(C + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(mean, rstd, beta, C,
                                                                                         thr_i32);
  ZKB_LAUNCH_CHECK("bn_sign_threshold_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_bn_eval_rstd(const float* moving_var, float eps, int C, float* rstd, void* stream) {
  ZKB_REQUIRE(moving_var && rstd, "bn_eval_rstd: null pointer");
  ZKB_REQUIRE(C > 0, "bn_eval_rstd: C must be positive");
  bn_eval_rstd_kernel<<<(C + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(moving_var, eps, C, rstd);
  ZKB_LAUNCH_CHECK("bn_eval_rstd_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_maxpool2_sign_i8(const void* s_i8, int B, int H, int W, int C, void* out_i8, void* stream) {
  ZKB_REQUIRE(s_i8 && out_i8, "maxpool2_sign_i8: null pointer");
  ZKB_REQUIRE(B >= 0 && H >= 2 && W >= 2 && C > 0 && C % 16 == 0, "maxpool2_sign_i8: bad shape (C %% 16 == 0)");
  ZKB_REQUIRE(ZKB_ALIGNED16(s_i8) && ZKB_ALIGNED16(out_i8), "maxpool2_sign_i8: pointers must be 16-byte aligned");
  if (B == 0) return ZKB_OK;
  const long long work = static_cast<long long>(B) * (H / 2) * (W / 2) * (C / 16);
  maxpool2_sign_i8_kernel<<<grid_for(work, 256, 16), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const int8_t*>(s_i8), B, H, W, C, static_cast<int8_t*>(out_i8));
  ZKB_LAUNCH_CHECK("maxpool2_sign_i8_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_bn_apply_sign(const void* x, int x_is_i16, long long rows, int C, const float* mean,
                              const float* rstd, const float* beta, void* s_out_bf16, float* a_out,
                              void* s_out_i8, void* stream) {
  const int vec = x_is_i16 ? 8 : 4;
  int rc = check_rows_c("bn_apply_sign", x, rows, C, vec, 1);
  if (rc) return rc;
  ZKB_REQUIRE(mean && rstd && beta, "bn_apply_sign: null statistics");
  ZKB_REQUIRE(s_out_bf16 || a_out || s_out_i8, "bn_apply_sign: no output requested");
  if (rows == 0) return ZKB_OK;
  const int sw = slice_width(C, vec);
  for (int c0 = 0; c0 < C; c0 += sw) {
    const int cc = C - c0 < sw ? C - c0 : sw;
    const int grid = grid_for(rows * (cc / vec), 256, 16);
    uint16_t* so = s_out_bf16 ? static_cast<uint16_t*>(s_out_bf16) + c0 : nullptr;
    float* ao = a_out ? a_out + c0 : nullptr;
    int8_t* s8 = s_out_i8 ? static_cast<int8_t*>(s_out_i8) + c0 : nullptr;
    if (x_is_i16)
      bn_apply_sign_kernel<int16_t><<<grid, row_block(cc / vec), 0, static_cast<cudaStream_t>(stream)>>>(
          static_cast<const int16_t*>(x) + c0, rows, cc, mean + c0, rstd + c0, beta + c0, so, ao, s8, C);
    else
      bn_apply_sign_kernel<float><<<grid, row_block(cc / vec), 0, static_cast<cudaStream_t>(stream)>>>(
          static_cast<const float*>(x) + c0, rows, cc, mean + c0, rstd + c0, beta + c0, so, ao, s8, C);
    ZKB_LAUNCH_CHECK("bn_apply_sign_kernel");
  }
  return ZKB_OK;
}

ZKB_API int zkb_maxpool2_i16(const void* y, int B, int H, int W, int C, void* p_out, uint8_t* arg_out,
                             double* sum, double* sumsq, void* stream) {
  int rc = check_rows_c("maxpool2_i16", y, static_cast<long long>(B) * H * W, C, 8);
  if (rc) return rc;
  ZKB_REQUIRE(p_out && arg_out && sum && sumsq, "maxpool2_i16: null output");
  ZKB_REQUIRE(H >= 2 && W >= 2, "maxpool2_i16: map smaller than the window");
  if (B == 0) return ZKB_OK;
  const long long rows = static_cast<long long>(B) * (H / 2) * (W / 2);
  const int lanes = 256 / (C / 8);
  // at least 8 pooled pixels per thread (every block ends with 2 float64 atomics per channel), at most 12 blocks per SM
  long long gmp = (rows + static_cast<long long>(lanes) * 8 - 1) / (static_cast<long long>(lanes) * 8);
  if (gmp > static_cast<long long>(zkb_sm_count()) * 12) gmp = static_cast<long long>(zkb_sm_count()) * 12;
  const int grid = gmp < 1 ? 1 : static_cast<int>(gmp);
  const size_t smem = 0;
  maxpool2_i16_kernel<<<grid, row_block(C / 8), smem, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const int16_t*>(y), B, H, W, C, static_cast<int16_t*>(p_out), arg_out, sum, sumsq);
  ZKB_LAUNCH_CHECK("maxpool2_i16_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_softmax_ce(const float* logits, int ld, const int* labels, int B, int num_classes,
                           float inv_global_batch, float* dlogits, double* loss_sum, int* correct,
                           void* stream) {
  ZKB_REQUIRE(logits && labels && loss_sum && correct, "softmax_ce: null pointer");
  ZKB_REQUIRE(num_classes > 0 && ld >= num_classes, "softmax_ce: ld < num_classes");
  if (B == 0) return ZKB_OK;
  if (num_classes > 64)      // a warp per row
    softmax_ce_wide_kernel<<<(B + 7) / 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        logits, ld, labels, B, num_classes, inv_global_batch, dlogits, loss_sum, correct);
  else
    softmax_ce_kernel<<<(B + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        logits, ld, labels, B, num_classes, inv_global_batch, dlogits, loss_sum, correct);
  ZKB_LAUNCH_CHECK("softmax_ce_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_bn_bwd_reduce(const float* g, const void* x, int x_is_i16, long long rows, int C,
                              const float* mean, const float* rstd, const float* beta, int ste,
                              double* sum_da, double* sum_dax, void* stream) {
  int rc = check_rows_c("bn_bwd_reduce", g, rows, C, 4, 1);
  if (rc) return rc;
  ZKB_REQUIRE(x && mean && rstd && beta && sum_da && sum_dax, "bn_bwd_reduce: null pointer");
  if (rows == 0) return ZKB_OK;
  const size_t smem = 0;
  const int sw = slice_width(C, 4);
  for (int c0 = 0; c0 < C; c0 += sw) {
    const int cc = C - c0 < sw ? C - c0 : sw;
    const int lanes = 256 / (cc / 4);
    const int grid = reduce_grid(rows, lanes);
    if (x_is_i16)
      bn_bwd_reduce_kernel<int16_t><<<grid, row_block(cc / 4), smem, static_cast<cudaStream_t>(stream)>>>(
          g + c0, static_cast<const int16_t*>(x) + c0, rows, cc, mean + c0, rstd + c0, beta + c0, ste, sum_da + c0,
          sum_dax + c0, C);
    else
      bn_bwd_reduce_kernel<float><<<grid, row_block(cc / 4), smem, static_cast<cudaStream_t>(stream)>>>(
          g + c0, static_cast<const float*>(x) + c0, rows, cc, mean + c0, rstd + c0, beta + c0, ste, sum_da + c0,
          sum_dax + c0, C);
    ZKB_LAUNCH_CHECK("bn_bwd_reduce_kernel");
  }
  return ZKB_OK;
}

ZKB_API int zkb_bn_bwd_finalize(const double* sum_da, const double* sum_dax, double count,
                                double dbeta_scale, int C, float* dbeta, float* mean_da,
                                float* mean_dax, void* stream) {
  ZKB_REQUIRE(sum_da && sum_dax && mean_da && mean_dax, "bn_bwd_finalize: null pointer");
  ZKB_REQUIRE(count > 0 && C > 0, "bn_bwd_finalize: bad count / C");
  bn_bwd_finalize_kernel<<<(C + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      sum_da, sum_dax, count, dbeta_scale, C, dbeta, mean_da, mean_dax);
  ZKB_LAUNCH_CHECK("bn_bwd_finalize_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_bn_bwd_apply(const float* g, const void* x, int x_is_i16, long long rows, int C,
                             const float* mean, const float* rstd, const float* beta,
                             const float* mean_da, const float* mean_dax, int ste, int out_mode,
                             float* out_f32, void* out_hi, void* out_lo, const uint8_t* arg, int H,
                             int W, void* stream) {
  int rc = check_rows_c("bn_bwd_apply", g, rows, C, 4, out_mode != 2);
  if (rc) return rc;
  ZKB_REQUIRE(x && mean && rstd && beta && mean_da && mean_dax, "bn_bwd_apply: null pointer");
  ZKB_REQUIRE(out_mode >= 0 && out_mode <= 2, "bn_bwd_apply: out_mode must be 0, 1 or 2");
  if (out_mode == 0) ZKB_REQUIRE(out_f32, "bn_bwd_apply: out_f32 is null");
  if (out_mode >= 1) ZKB_REQUIRE(out_hi && out_lo, "bn_bwd_apply: hi/lo outputs are null");
  if (out_mode == 2) {
    ZKB_REQUIRE(arg && H >= 2 && W >= 2, "bn_bwd_apply: pool scatter needs argmax and H, W");
    ZKB_REQUIRE(rows % (static_cast<long long>(H / 2) * (W / 2)) == 0,
                "bn_bwd_apply: rows is not a multiple of the pooled map size");
  }
  if (rows == 0) return ZKB_OK;
  const int sw = slice_width(C, 4);      // (the pool scatter, out_mode 2, is one slice: checked above)
  for (int c0 = 0; c0 < C; c0 += sw) {
    const int cc = C - c0 < sw ? C - c0 : sw;
    const int grid = grid_for(rows * (cc / 4), 256, 16);
    float* of = out_f32 ? out_f32 + c0 : nullptr;
    uint16_t* oh = out_hi ? static_cast<uint16_t*>(out_hi) + c0 : nullptr;
    uint16_t* ol = out_lo ? static_cast<uint16_t*>(out_lo) + c0 : nullptr;
    if (x_is_i16)
      bn_bwd_apply_kernel<int16_t><<<grid, row_block(cc / 4), 0, static_cast<cudaStream_t>(stream)>>>(
          g + c0, static_cast<const int16_t*>(x) + c0, rows, cc, mean + c0, rstd + c0, beta + c0, mean_da + c0,
          mean_dax + c0, ste, out_mode, of, oh, ol, arg, H, W, C);
    else
      bn_bwd_apply_kernel<float><<<grid, row_block(cc / 4), 0, static_cast<cudaStream_t>(stream)>>>(
          g + c0, static_cast<const float*>(x) + c0, rows, cc, mean + c0, rstd + c0, beta + c0, mean_da + c0,
          mean_dax + c0, ste, out_mode, of, oh, ol, arg, H, W, C);
    ZKB_LAUNCH_CHECK("bn_bwd_apply_kernel");
  }
  return ZKB_OK;
}

ZKB_API int zkb_prep_conv_weights(const float* w_hwio, int ksize, int Cin, int Cout, void* w_fwd,
                                  void* w_dgrad, void* w_fwd_i8, void* stream) {
  ZKB_REQUIRE(w_hwio && w_fwd && w_dgrad, "prep_conv_weights: null pointer");
  dim3 grid((Cout + 31) / 32, (Cin + 31) / 32, ksize * ksize);
  prep_conv_weights_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      w_hwio, ksize, Cin, Cout, static_cast<uint16_t*>(w_fwd), static_cast<uint16_t*>(w_dgrad),
      static_cast<int8_t*>(w_fwd_i8));
  ZKB_LAUNCH_CHECK("prep_conv_weights_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_prep_dense_weights(const float* w_k_n, int K, int N, int Npad, void* w_fwd,
                                   void* w_dgrad, void* w_fwd_i8, void* stream) {
  ZKB_REQUIRE(w_k_n && w_fwd && w_dgrad, "prep_dense_weights: null pointer");
  ZKB_REQUIRE(Npad >= N, "prep_dense_weights: Npad < N");
  dim3 grid((Npad + 31) / 32, (K + 31) / 32);
  prep_dense_weights_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      w_k_n, K, N, Npad, static_cast<uint16_t*>(w_fwd), static_cast<uint16_t*>(w_dgrad),
      static_cast<int8_t*>(w_fwd_i8));
  ZKB_LAUNCH_CHECK("prep_dense_weights_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_sign_f32(const float* w, float* out, long long n, void* stream) {
  ZKB_REQUIRE(w && out && n >= 0, "sign_f32: bad arguments");
  if (n == 0) return ZKB_OK;
  sign_f32_kernel<<<grid_for(n, 256, 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(w, out, n);
  ZKB_LAUNCH_CHECK("sign_f32_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_adam_clip(float* w, const float* g, float* m, float* v, long long n, float lr_t,
                          float beta1, float beta2, float eps, float grad_scale, int is_kernel,
                          const float* lr_t_dev, void* stream) {
  ZKB_REQUIRE(w && g && m && v && n >= 0, "adam_clip: bad arguments");
  ZKB_REQUIRE(ZKB_ALIGNED16(w) && ZKB_ALIGNED16(g) && ZKB_ALIGNED16(m) && ZKB_ALIGNED16(v),
              "adam_clip: arenas must be 16-byte aligned");
  if (n == 0) return ZKB_OK;
  adam_clip_kernel<<<grid_for(n / 4 + 1, 256, 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      w, g, m, v, n, lr_t, beta1, beta2, eps, grad_scale, is_kernel, lr_t_dev);
  ZKB_LAUNCH_CHECK("adam_clip_kernel");
  return ZKB_OK;
}

// ============================================================================================
// Allocation-free BatchNorm(train) + sign
// ============================================================================================
// Some hosts of this library are handed device buffers for inputs and outputs only and are not
// allowed to obtain device memory themselves: an operator callback running inside another
// runtime (the runtime owns the allocator and a stray cudaMalloc/cudaFree would stall the whole
// device behind its back), or a latency-bound loop that cannot afford the implicit device-wide
// synchronisation those calls carry.  The three-call sequence zkb_bn_stats -> zkb_bn_finalize ->
// zkb_bn_apply_sign needs per-channel accumulators and statistics that such a caller has nowhere
// to put.  zkb_bn_train_sign_static keeps them in the fixed array below instead.
//
// The array lives in every CUDA context that loads the library (48 KiB per GPU, until the
// process exits) and is shared by ALL calls on that GPU: it still holds the previous call's values
// on entry (the entry point clears what must start at zero), and two calls whose kernels overlap
// on the GPU -- different streams, different host threads -- corrupt each other.
//
// Channel limit: the statistics and normalising kernels give each thread of a 256-thread block one
// 16-byte vector of channels (8 int16 or 4 f32), so a row is at most 256 vectors wide and the vector
// count must divide 256: C <= 2048 for int16 input, C <= 1024 for f32 input, C / vector a power of
// two.  The array is sized for the larger of the two.
namespace zkb {
constexpr int kStaticBnMaxChannels = 2048;        // int16 input; f32 input: half of it
struct StaticBnScratch {
  double sum[kStaticBnMaxChannels];
  double sumsq[kStaticBnMaxChannels];
  float mean[kStaticBnMaxChannels];
  float rstd[kStaticBnMaxChannels];
};
static __device__ __align__(256) unsigned char g_static_bn_scratch[sizeof(StaticBnScratch)];
}  // namespace zkb

// Host code must not hand `g_static_bn_scratch` to a kernel directly: on the host that name is
// only a stand-in for the symbol.  The device address is looked up once per device, with that
// device current, and cached.
static int static_bn_scratch(StaticBnScratch** out) {
  static StaticBnScratch* cache[64] = {nullptr};
  static std::mutex mu;
  int dev = -1;
  ZKB_CUDA(cudaGetDevice(&dev));
  ZKB_REQUIRE(dev >= 0 && dev < 64, "bn_train_sign_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (cache[dev] == nullptr) {
    void* p = nullptr;
    ZKB_CUDA(cudaGetSymbolAddress(&p, g_static_bn_scratch));
    ZKB_REQUIRE(p != nullptr, "bn_train_sign_static: scratch symbol has no device address");
    cache[dev] = static_cast<StaticBnScratch*>(p);
  }
  *out = cache[dev];
  return ZKB_OK;
}

ZKB_API int zkb_bn_train_sign_static_max_channels(int x_is_i16) {
  return x_is_i16 ? kStaticBnMaxChannels : kStaticBnMaxChannels / 2;
}

ZKB_API int zkb_bn_train_sign_static(const void* x, int x_is_i16, long long rows, int C,
                                     const float* beta, float eps, float momentum, int bessel,
                                     float* moving_mean, float* moving_var, void* s_out_bf16,
                                     float* a_out, void* stream) {
  const int vec = x_is_i16 ? 8 : 4;
  const int cmax = zkb_bn_train_sign_static_max_channels(x_is_i16);
  ZKB_REQUIRE(C > 0 && C <= cmax,
              "bn_train_sign_static: C=%d exceeds the built-in scratch (at most %d channels for %s "
              "input); larger layers need the stats/finalize/apply entry points with caller-provided buffers",
              C, cmax, x_is_i16 ? "int16" : "f32");
  ZKB_REQUIRE(C % vec == 0 && 256 % (C / vec) == 0,
              "bn_train_sign_static: C=%d must be %d times a power of two (at most %d)", C, vec, cmax);
  int rc = check_rows_c("bn_train_sign_static", x, rows, C, vec);
  if (rc) return rc;
  ZKB_REQUIRE(rows > 0, "bn_train_sign_static: rows must be positive");
  ZKB_REQUIRE(beta && (s_out_bf16 || a_out), "bn_train_sign_static: null beta or no output");
  ZKB_REQUIRE((moving_mean == nullptr) == (moving_var == nullptr),
              "bn_train_sign_static: moving_mean and moving_var go together");
  StaticBnScratch* sc = nullptr;
  rc = static_bn_scratch(&sc);
  if (rc) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // the accumulators are summed into: wipe whatever the previous call left there
  ZKB_CUDA(cudaMemsetAsync(sc->sum, 0, sizeof(double) * C, st));
  ZKB_CUDA(cudaMemsetAsync(sc->sumsq, 0, sizeof(double) * C, st));
  const int lanes = 256 / (C / vec);
  const int grid_s = reduce_grid(rows, lanes);
  const size_t smem = 0;
  const int grid_a = grid_for(rows * (C / vec), 256, 16);
  if (x_is_i16) {
    bn_stats_kernel<int16_t><<<grid_s, 256, smem, st>>>(static_cast<const int16_t*>(x), rows, C,
                                                        sc->sum, sc->sumsq);
    ZKB_LAUNCH_CHECK("bn_stats_kernel");
    bn_finalize_kernel<<<(C + 127) / 128, 128, 0, st>>>(sc->sum, sc->sumsq, static_cast<double>(rows),
                                                        eps, momentum, bessel, C, sc->mean, sc->rstd,
                                                        moving_mean, moving_var);
    ZKB_LAUNCH_CHECK("bn_finalize_kernel");
    bn_apply_sign_kernel<int16_t><<<grid_a, 256, 0, st>>>(static_cast<const int16_t*>(x), rows, C,
                                                          sc->mean, sc->rstd, beta,
                                                          static_cast<uint16_t*>(s_out_bf16), a_out);
  } else {
    bn_stats_kernel<float><<<grid_s, 256, smem, st>>>(static_cast<const float*>(x), rows, C,
                                                      sc->sum, sc->sumsq);
    ZKB_LAUNCH_CHECK("bn_stats_kernel");
    bn_finalize_kernel<<<(C + 127) / 128, 128, 0, st>>>(sc->sum, sc->sumsq, static_cast<double>(rows),
                                                        eps, momentum, bessel, C, sc->mean, sc->rstd,
                                                        moving_mean, moving_var);
    ZKB_LAUNCH_CHECK("bn_finalize_kernel");
    bn_apply_sign_kernel<float><<<grid_a, 256, 0, st>>>(static_cast<const float*>(x), rows, C,
                                                        sc->mean, sc->rstd, beta,
                                                        static_cast<uint16_t*>(s_out_bf16), a_out);
  }
  ZKB_LAUNCH_CHECK("bn_apply_sign_kernel");
  return ZKB_OK;
}

// ============================================================================================
// First layer, fused with its BatchNorm: the K = ks*ks*Cin (9 or 27) float conv is recomputed from
// the image wherever its output is needed instead of being stored.  At batch 4096 the stored map
// would be 1.4 GB written once and read four times; the image it is recomputed from is 12.8 MB.
//   stats   : y on the fly -> per-channel sum / sumsq
//   bn_sign : y -> (y-mean)*rstd+beta -> sign as pm1 bf16
//   bwd_red : g (grad wrt sign output) + recomputed a, xhat -> sum(da), sum(da*xhat)
//   bwd_wgr : dy = rstd*(da - mean_da - xhat*mean_dax) on the fly -> dW[k][co] += x[pix+tap]*dy
// Thread = 4 output channels of one pixel; a block walks pixels with 256/(Cout/4) pixel lanes.
// Replaces the same reference lines as zkb_conv_first_fwd + zkb_bn_* (larq_experiment.py:62-70).
// ============================================================================================
namespace zkb {

template <int KS, int CIN>
struct FirstConv {
  static constexpr int K = KS * KS * CIN;
  // For K <= 9 each thread keeps its 4 channels of sign(W) in registers (36 floats); reading them
  // from shared memory instead costs 9 x 512 B of smem bandwidth per pixel and was the bottleneck.
  static constexpr bool kRegW = K <= 9;
  static constexpr int KR = kRegW ? K : 1;
  __device__ static __forceinline__ void load_w(const float* __restrict__ wq, int Cout, int c4,
                                                float4 (&wr)[KR]) {
    if constexpr (kRegW) {
#pragma unroll
      for (int k = 0; k < K; ++k) wr[k] = *reinterpret_cast<const float4*>(wq + k * Cout + c4);
    } else {
      wr[0] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  // Sliding KSxKS window along one output row: per pixel only the KS*CIN new taps are loaded and
  // the index arithmetic is done once per row instead of once per pixel.
  struct Row {
    float win[KS][KS * CIN];
    const float* base;
    int W;
    __device__ __forceinline__ void begin(const float* __restrict__ x, unsigned row, int H, int W_) {
      const unsigned OH = H - KS + 1;
      const unsigned b = row / OH;
      const unsigned oh = row - b * OH;
      W = W_;
      base = x + static_cast<size_t>(b * H + oh) * W * CIN;
#pragma unroll
      for (int kh = 0; kh < KS; ++kh)
#pragma unroll
        for (int j = 0; j < (KS - 1) * CIN; ++j) win[kh][j] = __ldg(base + kh * W * CIN + j);
    }
    // taps of output pixel `ow` of this row into xv (then slide by one pixel)
    __device__ __forceinline__ void step(int ow, float (&xv)[K]) {
#pragma unroll
      for (int kh = 0; kh < KS; ++kh)
#pragma unroll
        for (int ci = 0; ci < CIN; ++ci)
          win[kh][(KS - 1) * CIN + ci] = __ldg(base + (kh * W + ow + KS - 1) * CIN + ci);
#pragma unroll
      for (int kh = 0; kh < KS; ++kh)
#pragma unroll
        for (int j = 0; j < KS * CIN; ++j) xv[kh * KS * CIN + j] = win[kh][j];
#pragma unroll
      for (int kh = 0; kh < KS; ++kh)
#pragma unroll
        for (int j = 0; j < (KS - 1) * CIN; ++j) win[kh][j] = win[kh][j + CIN];
    }
  };
  // K > 9 (the 3-channel first layer): sign(W) comes from shared memory, 27 float4 loads per pixel
  // and thread against 108 FMAs -- shared-memory bandwidth, not arithmetic, bounded those kernels
  // (ncu, r2: 24 % of the CIFAR-shaped step).  NPX consecutive output pixels of a row (4; 2 in the
  // weight-gradient kernel, whose 108 accumulators leave no room for more) share their input window
  // and every weight vector that is fetched: 1/NPX of the shared-memory traffic per pixel, the same FMA order per pixel as `dot` (k ascending), hence
  // the same floats.
  template <int NPX>
  struct Group {
    float win[KS][(KS - 1 + NPX) * CIN];
    __device__ __forceinline__ void load(const float* __restrict__ row_base, int W, int ow0) {
      // columns ow0 .. ow0 + KS + NPX - 2 of the KS input rows; columns past the image edge are
      // clamped (they only feed output pixels past the row end, which the callers skip)
#pragma unroll
      for (int kh = 0; kh < KS; ++kh)
#pragma unroll
        for (int cc = 0; cc < KS - 1 + NPX; ++cc) {
          int col = ow0 + cc;
          col = col < W ? col : W - 1;
#pragma unroll
          for (int ci = 0; ci < CIN; ++ci) win[kh][cc * CIN + ci] = __ldg(row_base + (kh * W + col) * CIN + ci);
        }
    }
    __device__ __forceinline__ void dot4(const float* s_w, int Cout, int c4, float (&y)[NPX][4]) const {
#pragma unroll
      for (int q = 0; q < NPX; ++q) y[q][0] = y[q][1] = y[q][2] = y[q][3] = 0.f;
#pragma unroll
      for (int kh = 0; kh < KS; ++kh)
#pragma unroll
        for (int r = 0; r < KS * CIN; ++r) {
          const float4 w = *reinterpret_cast<const float4*>(s_w + (kh * KS * CIN + r) * Cout + c4);
#pragma unroll
          for (int q = 0; q < NPX; ++q) {
            const float x = win[kh][q * CIN + r];
            y[q][0] = fmaf(x, w.x, y[q][0]);
            y[q][1] = fmaf(x, w.y, y[q][1]);
            y[q][2] = fmaf(x, w.z, y[q][2]);
            y[q][3] = fmaf(x, w.w, y[q][3]);
          }
        }
    }
    // tap k = (kh, kw, ci) of pixel q of the quad
    __device__ __forceinline__ float tap(int q, int k) const { return win[k / (KS * CIN)][q * CIN + k % (KS * CIN)]; }
  };
  __device__ static __forceinline__ const float* row_base(const float* __restrict__ x, unsigned row, int H, int W) {
    const unsigned OH = H - KS + 1;
    const unsigned b = row / OH;
    const unsigned oh = row - b * OH;
    return x + static_cast<size_t>(b * H + oh) * W * CIN;
  }
  __device__ static __forceinline__ void dot(const float (&xv)[K], const float* s_w,
                                             const float4 (&wr)[KR], int Cout, int c4, float (&y)[4]) {
    y[0] = y[1] = y[2] = y[3] = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) {
      float4 w;
      if constexpr (kRegW) w = wr[k];
      else w = *reinterpret_cast<const float4*>(s_w + k * Cout + c4);
      y[0] = fmaf(xv[k], w.x, y[0]);
      y[1] = fmaf(xv[k], w.y, y[1]);
      y[2] = fmaf(xv[k], w.z, y[2]);
      y[3] = fmaf(xv[k], w.w, y[3]);
    }
  }
  // y[4] for channels c4..c4+3 of pixel pix; xv receives the K input taps.
  __device__ static __forceinline__ void at(const float* __restrict__ x, const float* s_w,
                                            const float4 (&wr)[KR], int Cout, long long pix, int H,
                                            int W, int c4, float (&xv)[K], float (&y)[4]) {
    // 32-bit index arithmetic (the host checks B*OH*OW < 2^31): 64-bit div/mod cost more than
    // the 9-tap convolution itself
    const unsigned OH = H - KS + 1, OW = W - KS + 1;
    const unsigned upix = static_cast<unsigned>(pix);
    const unsigned t = upix / OW;
    const unsigned ow = upix - t * OW;
    const unsigned b = t / OH;
    const unsigned oh = t - b * OH;
    const float* base = x + (static_cast<size_t>(b * H + oh) * W + ow) * CIN;
#pragma unroll
    for (int kh = 0; kh < KS; ++kh)
#pragma unroll
      for (int kw = 0; kw < KS; ++kw)
#pragma unroll
        for (int ci = 0; ci < CIN; ++ci)
          xv[(kh * KS + kw) * CIN + ci] = __ldg(base + (kh * W + kw) * CIN + ci);
    y[0] = y[1] = y[2] = y[3] = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) {
      float4 w;
      if constexpr (kRegW) w = wr[k];
      else w = *reinterpret_cast<const float4*>(s_w + k * Cout + c4);
      y[0] = fmaf(xv[k], w.x, y[0]);
      y[1] = fmaf(xv[k], w.y, y[1]);
      y[2] = fmaf(xv[k], w.z, y[2]);
      y[3] = fmaf(xv[k], w.w, y[3]);
    }
  }
};

// dynamic smem: [K * Cout floats of sign(W)] (also reused to combine the wgrad pixel lanes)
__device__ __forceinline__ float* first_conv_weights_smem() {
  extern __shared__ double s_dyn[];
  return reinterpret_cast<float*>(s_dyn);
}

template <int KS, int CIN>
__global__ void __launch_bounds__(256)
conv_first_bn_stats_kernel(const float* __restrict__ x, const float* __restrict__ wq, int B, int H,
                           int W, int Cout, double* __restrict__ sum, double* __restrict__ sumsq) {
  using FC = FirstConv<KS, CIN>;
  float* s_w = first_conv_weights_smem();
  for (int i = threadIdx.x; i < FC::K * Cout; i += blockDim.x) s_w[i] = wq[i];
  __syncthreads();
  const int tpr = Cout / 4, lanes = blockDim.x / tpr;
  const int cl = threadIdx.x % tpr, rl = threadIdx.x / tpr;
  float4 wr[FC::KR];
  FC::load_w(wq, Cout, cl * 4, wr);
  const long long pixels = static_cast<long long>(B) * (H - KS + 1) * (W - KS + 1);
  double a0[4] = {0, 0, 0, 0}, a1[4] = {0, 0, 0, 0};
  float f0[4] = {0, 0, 0, 0}, f1[4] = {0, 0, 0, 0};
  int n = 0;
  const unsigned OWr = W - KS + 1;
  const unsigned n_rows = static_cast<unsigned>(pixels / OWr);
  for (unsigned row = blockIdx.x * lanes + rl; row < n_rows; row += gridDim.x * lanes) {
   if constexpr (!FC::kRegW) {
    const float* rb = FC::row_base(x, row, H, W);
    for (unsigned ow0 = 0; ow0 < OWr; ow0 += 4) {
      typename FC::template Group<4> quad;
      quad.load(rb, W, ow0);
      float y4[4][4];
      quad.dot4(s_w, Cout, cl * 4, y4);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (ow0 + q >= OWr) break;
#pragma unroll
        for (int j = 0; j < 4; ++j) { f0[j] += y4[q][j]; f1[j] = fmaf(y4[q][j], y4[q][j], f1[j]); }
        if (++n == 32) {
#pragma unroll
          for (int j = 0; j < 4; ++j) { a0[j] += f0[j]; a1[j] += f1[j]; f0[j] = f1[j] = 0.f; }
          n = 0;
        }
      }
    }
    continue;
   }
   typename FC::Row walk;
   walk.begin(x, row, H, W);
   for (unsigned owi = 0; owi < OWr; ++owi) {
    float xv[FC::K], y[4];
    walk.step(owi, xv);
    FC::dot(xv, s_w, wr, Cout, cl * 4, y);
#pragma unroll
    for (int j = 0; j < 4; ++j) { f0[j] += y[j]; f1[j] = fmaf(y[j], y[j], f1[j]); }
    if (++n == 32) {
#pragma unroll
      for (int j = 0; j < 4; ++j) { a0[j] += f0[j]; a1[j] += f1[j]; f0[j] = f1[j] = 0.f; }
      n = 0;
    }
   }
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) { a0[j] += f0[j]; a1[j] += f1[j]; }
  block_channel_reduce<4>(a0, a1, tpr, Cout, sum, sumsq);
}

template <int KS, int CIN>
__device__ __forceinline__ void conv_first_bn_sign_body(const float* __restrict__ x, const float* __restrict__ wq, int B, int H,
                          int W, int Cout, const float* __restrict__ mean,
                          const float* __restrict__ rstd, const float* __restrict__ beta,
                          uint16_t* __restrict__ s_out, int8_t* __restrict__ s8_out) {
  using FC = FirstConv<KS, CIN>;
  float* s_w = first_conv_weights_smem();
  for (int i = threadIdx.x; i < FC::K * Cout; i += blockDim.x) s_w[i] = wq[i];
  __syncthreads();
  const int tpr = Cout / 4, lanes = blockDim.x / tpr;
  const int cl = threadIdx.x % tpr, rl = threadIdx.x / tpr;
  float4 wr[FC::KR];
  FC::load_w(wq, Cout, cl * 4, wr);
  const long long pixels = static_cast<long long>(B) * (H - KS + 1) * (W - KS + 1);
  float mu[4], rs[4], be[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) { mu[j] = mean[cl * 4 + j]; rs[j] = rstd[cl * 4 + j]; be[j] = beta[cl * 4 + j]; }
  const unsigned OWr = W - KS + 1;
  const unsigned n_rows = static_cast<unsigned>(pixels / OWr);
  if constexpr (KS == 3 && CIN == 1) {
    // One input channel: the 3x3 window slides by one column per pixel.  Three pixels per
    // iteration, so that the three column registers rotate by renaming instead of moves, and row
    // pointers that advance by three (tap loads with immediate offsets).
    const unsigned OH = H - 2;
    for (unsigned row = blockIdx.x * lanes + rl; row < n_rows; row += gridDim.x * lanes) {
      const unsigned b = row / OH, oh = row - b * OH;
      const float* r0 = x + static_cast<size_t>(b * H + oh) * W;
      const float* r1 = r0 + W;
      const float* r2 = r1 + W;
      const size_t o0 = static_cast<size_t>(row) * OWr * Cout + cl * 4;
      uint16_t* so = s_out ? s_out + o0 : nullptr;
      int8_t* so8 = s8_out ? s8_out + o0 : nullptr;
      float c0[3] = {__ldg(r0), __ldg(r1), __ldg(r2)};             // input column ow
      float c1[3] = {__ldg(r0 + 1), __ldg(r1 + 1), __ldg(r2 + 1)}; // input column ow + 1
      float c2[3];
      auto pixel = [&](const float (&a0)[3], const float (&a1)[3], const float (&a2)[3]) {
        const float xv[9] = {a0[0], a1[0], a2[0], a0[1], a1[1], a2[1], a0[2], a1[2], a2[2]};
        float y[4];
        FC::dot(xv, s_w, wr, Cout, cl * 4, y);
        uint32_t neg = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) neg |= (((y[j] - mu[j]) * rs[j] + be[j] >= 0.0f) ? 0u : 1u) << j;
        if (so) {
          *reinterpret_cast<uint2*>(so) =
              make_uint2(0x3F803F80u | ((neg & 1u) << 15) | ((neg & 2u) << 30),
                         0x3F803F80u | ((neg & 4u) << 13) | ((neg & 8u) << 28));
          so += Cout;
        }
        if (so8) {
          const uint32_t spread = (neg * 0x00204081u) & 0x01010101u;
          *reinterpret_cast<uint32_t*>(so8) = spread * 0xFEu + 0x01010101u;
          so8 += Cout;
        }
      };
      unsigned ow = 0;
      for (; ow + 3 <= OWr; ow += 3) {
        c2[0] = __ldg(r0 + 2); c2[1] = __ldg(r1 + 2); c2[2] = __ldg(r2 + 2);
        pixel(c0, c1, c2);
        c0[0] = __ldg(r0 + 3); c0[1] = __ldg(r1 + 3); c0[2] = __ldg(r2 + 3);
        pixel(c1, c2, c0);
        c1[0] = __ldg(r0 + 4); c1[1] = __ldg(r1 + 4); c1[2] = __ldg(r2 + 4);
        pixel(c2, c0, c1);
        r0 += 3; r1 += 3; r2 += 3;
      }
      if (ow < OWr) {          // one or two pixels left
        c2[0] = __ldg(r0 + 2); c2[1] = __ldg(r1 + 2); c2[2] = __ldg(r2 + 2);
        pixel(c0, c1, c2);
        if (ow + 1 < OWr) {
          c0[0] = __ldg(r0 + 3); c0[1] = __ldg(r1 + 3); c0[2] = __ldg(r2 + 3);
          pixel(c1, c2, c0);
        }
      }
    }
    return;
  }
  for (unsigned row = blockIdx.x * lanes + rl; row < n_rows; row += gridDim.x * lanes) {
   // output pointers of this thread's 4 channels, advanced by one pixel (Cout elements) per step
   const size_t o0 = static_cast<size_t>(row) * OWr * Cout + cl * 4;
   uint16_t* so = s_out ? s_out + o0 : nullptr;
   int8_t* so8 = s8_out ? s8_out + o0 : nullptr;
   if constexpr (!FC::kRegW) {
    const float* rb = FC::row_base(x, row, H, W);
    for (unsigned ow0 = 0; ow0 < OWr; ow0 += 4) {
      typename FC::template Group<4> quad;
      quad.load(rb, W, ow0);
      float y4[4][4];
      quad.dot4(s_w, Cout, cl * 4, y4);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (ow0 + q >= OWr) break;
        uint32_t neg = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) neg |= (((y4[q][j] - mu[j]) * rs[j] + be[j] >= 0.0f) ? 0u : 1u) << j;
        if (so) {
          *reinterpret_cast<uint2*>(so) =
              make_uint2(0x3F803F80u | ((neg & 1u) << 15) | ((neg & 2u) << 30),
                         0x3F803F80u | ((neg & 4u) << 13) | ((neg & 8u) << 28));
          so += Cout;
        }
        if (so8) {
          const uint32_t spread = (neg * 0x00204081u) & 0x01010101u;
          *reinterpret_cast<uint32_t*>(so8) = spread * 0xFEu + 0x01010101u;
          so8 += Cout;
        }
      }
    }
    continue;
   }
   typename FC::Row walk;
   walk.begin(x, row, H, W);
   for (unsigned owi = 0; owi < OWr; ++owi) {
    float xv[FC::K], y[4];
    walk.step(owi, xv);
    FC::dot(xv, s_w, wr, Cout, cl * 4, y);
    // bit j of neg: channel j's sign is -1 (a < 0; sign(0) = sign(-0.0) = +1 as everywhere)
    uint32_t neg = 0;
    float a[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      a[j] = (y[j] - mu[j]) * rs[j] + be[j];
      neg |= ((a[j] >= 0.0f) ? 0u : 1u) << j;
    }
    if (so) {    // bf16 +1 = 0x3F80, -1 = 0xBF80: the sign bits go to bits 15 and 31 of each pair
      *reinterpret_cast<uint2*>(so) =
          make_uint2(0x3F803F80u | ((neg & 1u) << 15) | ((neg & 2u) << 30),
                     0x3F803F80u | ((neg & 4u) << 13) | ((neg & 8u) << 28));
      so += Cout;
    }
    if (so8) {
      // one byte per channel, 0x01 or 0xFF: spread bit j to bit 8j (the four partial products of
      // the multiplication land on distinct bits), then 0x01 + 0xFE per negative byte
      const uint32_t spread = (neg * 0x00204081u) & 0x01010101u;
      *reinterpret_cast<uint32_t*>(so8) = spread * 0xFEu + 0x01010101u;
      so8 += Cout;
    }
   }
  }
}

template <int KS, int CIN>
__global__ void __launch_bounds__(256, 3)
conv_first_bn_sign_kernel(const float* __restrict__ x, const float* __restrict__ wq, int B, int H, int W,
                          int Cout, const float* __restrict__ mean, const float* __restrict__ rstd,
                          const float* __restrict__ beta, uint16_t* __restrict__ s_out,
                          int8_t* __restrict__ s8_out) {
  conv_first_bn_sign_body<KS, CIN>(x, wq, B, H, W, Cout, mean, rstd, beta, s_out, s8_out);
}

template <int KS, int CIN>
__global__ void __launch_bounds__(256)
conv_first_bwd_reduce_kernel(const float* __restrict__ x, const float* __restrict__ wq,
                             const float* __restrict__ g, int B, int H, int W, int Cout,
                             const float* __restrict__ mean, const float* __restrict__ rstd,
                             const float* __restrict__ beta, double* __restrict__ sum_da,
                             double* __restrict__ sum_dax) {
  using FC = FirstConv<KS, CIN>;
  float* s_w = first_conv_weights_smem();
  for (int i = threadIdx.x; i < FC::K * Cout; i += blockDim.x) s_w[i] = wq[i];
  __syncthreads();
  const int tpr = Cout / 4, lanes = blockDim.x / tpr;
  const int cl = threadIdx.x % tpr, rl = threadIdx.x / tpr;
  float4 wr[FC::KR];
  FC::load_w(wq, Cout, cl * 4, wr);
  const long long pixels = static_cast<long long>(B) * (H - KS + 1) * (W - KS + 1);
  float mu[4], rs[4], be[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) { mu[j] = mean[cl * 4 + j]; rs[j] = rstd[cl * 4 + j]; be[j] = beta[cl * 4 + j]; }
  double a0[4] = {0, 0, 0, 0}, a1[4] = {0, 0, 0, 0};
  float f0[4] = {0, 0, 0, 0}, f1[4] = {0, 0, 0, 0};
  int n = 0;
  const unsigned OWr = W - KS + 1;
  const unsigned n_rows = static_cast<unsigned>(pixels / OWr);
  for (unsigned row = blockIdx.x * lanes + rl; row < n_rows; row += gridDim.x * lanes) {
   if constexpr (!FC::kRegW) {
    const float* rb = FC::row_base(x, row, H, W);
    for (unsigned ow0 = 0; ow0 < OWr; ow0 += 4) {
      typename FC::template Group<4> quad;
      quad.load(rb, W, ow0);
      float gq[4][4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const unsigned ow = ow0 + q < OWr ? ow0 + q : OWr - 1;
        load4(g + (static_cast<long long>(row) * OWr + ow) * Cout + cl * 4, gq[q]);
      }
      float y4[4][4];
      quad.dot4(s_w, Cout, cl * 4, y4);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (ow0 + q >= OWr) break;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float xh = (y4[q][j] - mu[j]) * rs[j];
          const float da = (fabsf(xh + be[j]) > 1.0f) ? 0.f : gq[q][j];
          f0[j] += da;
          f1[j] = fmaf(da, xh, f1[j]);
        }
        if (++n == 64) {
#pragma unroll
          for (int j = 0; j < 4; ++j) { a0[j] += f0[j]; a1[j] += f1[j]; f0[j] = f1[j] = 0.f; }
          n = 0;
        }
      }
    }
    continue;
   }
   typename FC::Row walk;
   walk.begin(x, row, H, W);
   for (unsigned owi = 0; owi < OWr; ++owi) {
    const long long pix = static_cast<long long>(row) * OWr + owi;
    float gg[4];
    load4(g + pix * Cout + cl * 4, gg);
    float xv[FC::K], y[4];
    walk.step(owi, xv);
    FC::dot(xv, s_w, wr, Cout, cl * 4, y);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float xh = (y[j] - mu[j]) * rs[j];
      const float da = (fabsf(xh + be[j]) > 1.0f) ? 0.f : gg[j];
      f0[j] += da;
      f1[j] = fmaf(da, xh, f1[j]);
    }
    if (++n == 64) {
#pragma unroll
      for (int j = 0; j < 4; ++j) { a0[j] += f0[j]; a1[j] += f1[j]; f0[j] = f1[j] = 0.f; }
      n = 0;
    }
   }
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) { a0[j] += f0[j]; a1[j] += f1[j]; }
  block_channel_reduce<4>(a0, a1, tpr, Cout, sum_da, sum_dax);
}

template <int KS, int CIN>
__global__ void __launch_bounds__(256)
conv_first_bwd_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ wq,
                            const float* __restrict__ g, int B, int H, int W, int Cout,
                            const float* __restrict__ mean, const float* __restrict__ rstd,
                            const float* __restrict__ beta, const float* __restrict__ mean_da,
                            const float* __restrict__ mean_dax, float* __restrict__ dw) {
  using FC = FirstConv<KS, CIN>;
  float* s_w = first_conv_weights_smem();
  for (int i = threadIdx.x; i < FC::K * Cout; i += blockDim.x) s_w[i] = wq[i];
  __syncthreads();
  const int tpr = Cout / 4, lanes = blockDim.x / tpr;
  const int cl = threadIdx.x % tpr, rl = threadIdx.x / tpr;
  float4 wr[FC::KR];
  FC::load_w(wq, Cout, cl * 4, wr);
  const long long pixels = static_cast<long long>(B) * (H - KS + 1) * (W - KS + 1);
  float mu[4], rs[4], be[4], mda[4], mdx[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int c = cl * 4 + j;
    mu[j] = mean[c]; rs[j] = rstd[c]; be[j] = beta[c]; mda[j] = mean_da[c]; mdx[j] = mean_dax[c];
  }
  float acc[FC::K][4];
#pragma unroll
  for (int k = 0; k < FC::K; ++k) acc[k][0] = acc[k][1] = acc[k][2] = acc[k][3] = 0.f;
  const unsigned OWr = W - KS + 1;
  const unsigned n_rows = static_cast<unsigned>(pixels / OWr);
  for (unsigned row = blockIdx.x * lanes + rl; row < n_rows; row += gridDim.x * lanes) {
   if constexpr (!FC::kRegW) {
    const float* rb = FC::row_base(x, row, H, W);
    for (unsigned ow0 = 0; ow0 < OWr; ow0 += 2) {
      typename FC::template Group<2> quad;
      quad.load(rb, W, ow0);
      float gq[2][4];
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const unsigned ow = ow0 + q < OWr ? ow0 + q : OWr - 1;
        load4(g + (static_cast<long long>(row) * OWr + ow) * Cout + cl * 4, gq[q]);
      }
      float y4[2][4];
      quad.dot4(s_w, Cout, cl * 4, y4);
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const bool live = ow0 + q < OWr;      // (no branch: pixels past the row end contribute zeros)
        float dy[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float xh = (y4[q][j] - mu[j]) * rs[j];
          const float da = (fabsf(xh + be[j]) > 1.0f) ? 0.f : gq[q][j];
          dy[j] = live ? rs[j] * (da - mda[j] - xh * mdx[j]) : 0.f;
        }
#pragma unroll
        for (int k = 0; k < FC::K; ++k) {
          const float xk = quad.tap(q, k);
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[k][j] = fmaf(xk, dy[j], acc[k][j]);
        }
      }
    }
    continue;
   }
   typename FC::Row walk;
   walk.begin(x, row, H, W);
   for (unsigned owi = 0; owi < OWr; ++owi) {
    const long long pix = static_cast<long long>(row) * OWr + owi;
    float gg[4];
    load4(g + pix * Cout + cl * 4, gg);
    float xv[FC::K], y[4], dy[4];
    walk.step(owi, xv);
    FC::dot(xv, s_w, wr, Cout, cl * 4, y);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float xh = (y[j] - mu[j]) * rs[j];
      const float da = (fabsf(xh + be[j]) > 1.0f) ? 0.f : gg[j];
      dy[j] = rs[j] * (da - mda[j] - xh * mdx[j]);
    }
#pragma unroll
    for (int k = 0; k < FC::K; ++k)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[k][j] = fmaf(xv[k], dy[j], acc[k][j]);
   }
  }
  // combine the pixel lanes of the block through shared memory, one tap at a time
  __syncthreads();
  float* s_acc = reinterpret_cast<float*>(first_conv_weights_smem());  // weights no longer needed
#pragma unroll   // (acc[] must be indexed by constants to stay in registers)
  for (int k = 0; k < FC::K; ++k) {
#pragma unroll
    for (int j = 0; j < 4; ++j) s_acc[rl * Cout + cl * 4 + j] = acc[k][j];
    __syncthreads();
    for (int c = threadIdx.x; c < Cout; c += blockDim.x) {
      float t = 0.f;
      for (int r = 0; r < lanes; ++r) t += s_acc[r * Cout + c];
      atomicAdd(dw + k * Cout + c, t);
    }
    __syncthreads();
  }
}

// Gram matrix of the 3x3 input taps over all output pixels of a 'valid' one-channel conv:
// G[k][k'] = sum_p x_k(p) x_k'(p), S2[k] = sum_p x_k(p)  (gram = G[81] then S2[9], float64,
// accumulated into).  Everything about the first layer that is quadratic in the input and does not
// involve the incoming gradient follows from these 90 numbers: per-channel sum / sum of squares of
// the conv output (y_c = sum_k w_kc x_k) and sum_p x_k(p) y_c(p).
// (Per-thread float64 copies of the 54 running sums cost 232 registers and left one block per SM --
// ncu, r2: 12 % of the warp slots, 98 us for a 12.8 MB read.  The float partials are now folded
// into ONE float64 set per block in shared memory: a warp reduces each of its 54 floats by shuffles
// and lane 0 adds the result, every 32 pixels per thread and at the end.)
__device__ __forceinline__ void gram_flush(float (&g)[45], float (&s)[9], double (*red)[54], int warp, int lane) {
#pragma unroll
  for (int i = 0; i < 54; ++i) {
    float v = i < 45 ? g[i] : s[i - 45];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[warp][i] += static_cast<double>(v);
  }
#pragma unroll
  for (int i = 0; i < 45; ++i) g[i] = 0.f;
#pragma unroll
  for (int i = 0; i < 9; ++i) s[i] = 0.f;
}

__global__ void __launch_bounds__(256, 2)
conv_first_gram_kernel(const float* __restrict__ x, int B, int H, int W, double* __restrict__ gram) {
  const int OH = H - 2, OW = W - 2;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned n_rows = static_cast<unsigned>(B) * OH;
  __shared__ double red[8][54];
  for (int i = threadIdx.x; i < 8 * 54; i += blockDim.x) (&red[0][0])[i] = 0.0;
  __syncthreads();
  float g[45], s[9];
#pragma unroll
  for (int i = 0; i < 45; ++i) g[i] = 0.f;
#pragma unroll
  for (int i = 0; i < 9; ++i) s[i] = 0.f;
  int n = 0;     // warp-uniform: every lane runs the same number of iterations
  for (unsigned row = blockIdx.x * 8 + warp; row < n_rows; row += gridDim.x * 8) {
    const unsigned b = row / OH, oh = row - b * OH;
    const float* base = x + (static_cast<size_t>(b) * H + oh) * W;
    for (int ow0 = 0; ow0 < OW; ow0 += 32) {
      const int ow = ow0 + lane;
      if (ow < OW) {
        float xv[9];
#pragma unroll
        for (int kh = 0; kh < 3; ++kh)
#pragma unroll
          for (int kw = 0; kw < 3; ++kw) xv[kh * 3 + kw] = __ldg(base + kh * W + ow + kw);
        int t = 0;
#pragma unroll
        for (int k = 0; k < 9; ++k) {
          s[k] += xv[k];
#pragma unroll
          for (int k2 = k; k2 < 9; ++k2) { g[t] = fmaf(xv[k], xv[k2], g[t]); ++t; }
        }
      }
      if (++n == 32) {   // bound the float partial sums (|x| <= 1: at most 32 per lane and sum)
        gram_flush(g, s, red, warp, lane);
        n = 0;
      }
    }
  }
  gram_flush(g, s, red, warp, lane);
  __syncthreads();
  if (threadIdx.x < 54) {
    double v = 0.0;
    for (int w = 0; w < 8; ++w) v += red[w][threadIdx.x];
    const int i = threadIdx.x;
    if (i >= 45) {
      atomicAdd(gram + 81 + (i - 45), v);
    } else {
      int k = 0, rem = i;            // upper-triangle index -> (k, k2)
      while (rem >= 9 - k) { rem -= 9 - k; ++k; }
      const int k2 = k + rem;
      atomicAdd(gram + k * 9 + k2, v);
      if (k2 != k) atomicAdd(gram + k2 * 9 + k, v);
    }
  }
}

// sum_c = sum_k w_kc S2[k];  sumsq_c = sum_kk' w_kc w_k'c G[k][k']   (accumulated into sum / sumsq)
__global__ void conv_first_stats_from_gram_kernel(const double* __restrict__ gram,
                                                  const float* __restrict__ wq, int Cout,
                                                  double* __restrict__ sum, double* __restrict__ sumsq) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= Cout) return;
  double w[9], a = 0.0, q = 0.0;
  for (int k = 0; k < 9; ++k) { w[k] = wq[k * Cout + c]; a += w[k] * gram[81 + k]; }
  for (int k = 0; k < 9; ++k)
    for (int k2 = 0; k2 < 9; ++k2) q += w[k] * w[k2] * gram[k * 9 + k2];
  sum[c] += a;
  sumsq[c] += q;
}

}  // namespace zkb

static int first_conv_check(const char* who, const void* x, const void* wq, int B, int H, int W,
                            int Cin, int Cout, int ksize) {
  ZKB_REQUIRE(x && wq, "%s: null pointer", who);
  ZKB_REQUIRE(ksize == 3 && (Cin == 1 || Cin == 3),
              "%s: only 3x3 kernels over 1 or 3 input channels are built", who);
  ZKB_REQUIRE(Cout % 4 == 0 && Cout / 4 <= 256 && 256 % (Cout / 4) == 0,
              "%s: Cout=%d must be a multiple of 4 with Cout/4 dividing 256", who, Cout);
  ZKB_REQUIRE(B >= 0 && H >= ksize && W >= ksize, "%s: bad shape", who);
  ZKB_REQUIRE(static_cast<long long>(B) * H * W < (1ll << 31), "%s: B*H*W must stay below 2^31", who);
  return ZKB_OK;
}

static size_t first_conv_smem(int K, int Cout) {
  const size_t w = static_cast<size_t>(K) * Cout * sizeof(float);
  const size_t lanes_acc = static_cast<size_t>(256 / (Cout / 4)) * Cout * sizeof(float);
  return w > lanes_acc ? w : lanes_acc;
}

static int first_conv_grid(int B, int H, int W, int Cout, int waves = 6) {
  const long long pixels = static_cast<long long>(B) * (H - 2) * (W - 2);
  const int lanes = 256 / (Cout / 4);
  long long g = (pixels + lanes * 16 - 1) / (lanes * 16);
  const long long cap = static_cast<long long>(zkb_sm_count()) * waves;
  if (g > cap) g = cap;
  return g < 1 ? 1 : static_cast<int>(g);
}

#define ZKB_FIRST_DISPATCH(kernel, ...) ZKB_FIRST_DISPATCH_W(6, kernel, __VA_ARGS__)
#define ZKB_FIRST_DISPATCH_W(waves__, kernel, ...)                                        \
  do {                                                                                    \
    const size_t smem__ = first_conv_smem(9 * Cin, Cout);                                 \
    const int grid__ = first_conv_grid(B, H, W, Cout, waves__);                           \
    cudaStream_t st__ = static_cast<cudaStream_t>(stream);                                \
    if (Cin == 1) {                                                                       \
      kernel<3, 1><<<grid__, 256, smem__, st__>>>(__VA_ARGS__);                           \
    } else {                                                                              \
      ZKB_CUDA(cudaFuncSetAttribute(kernel<3, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                    static_cast<int>(smem__)));                           \
      kernel<3, 3><<<grid__, 256, smem__, st__>>>(__VA_ARGS__);                           \
    }                                                                                     \
    ZKB_LAUNCH_CHECK(#kernel);                                                            \
  } while (0)

ZKB_API int zkb_conv_first_bn_stats(const float* x, const float* wq_k_co, int B, int H, int W, int Cin,
                                    int Cout, int ksize, double* sum, double* sumsq, void* stream) {
  int rc = first_conv_check("conv_first_bn_stats", x, wq_k_co, B, H, W, Cin, Cout, ksize);
  if (rc) return rc;
  ZKB_REQUIRE(sum && sumsq, "conv_first_bn_stats: null accumulator");
  if (B == 0) return ZKB_OK;
  ZKB_FIRST_DISPATCH(conv_first_bn_stats_kernel, x, wq_k_co, B, H, W, Cout, sum, sumsq);
  return ZKB_OK;
}

ZKB_API int zkb_conv_first_gram(const float* x, int B, int H, int W, double* gram, void* stream) {
  ZKB_REQUIRE(x && gram, "conv_first_gram: null pointer");
  ZKB_REQUIRE(B >= 0 && H >= 3 && W >= 3, "conv_first_gram: bad shape");
  ZKB_REQUIRE(static_cast<long long>(B) * H * W < (1ll << 31), "conv_first_gram: B*H*W must stay below 2^31");
  if (B == 0) return ZKB_OK;
  const long long rows = static_cast<long long>(B) * (H - 2);
  long long grid = (rows + 31) / 32;          // about 4 rows per warp
  const long long cap = static_cast<long long>(zkb_sm_count()) * 8;
  if (grid > cap) grid = cap;
  conv_first_gram_kernel<<<static_cast<int>(grid), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, B, H, W, gram);
  ZKB_LAUNCH_CHECK("conv_first_gram_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_conv_first_stats_from_gram(const double* gram, const float* wq_k_co, int Cout, double* sum,
                                           double* sumsq, void* stream) {
  ZKB_REQUIRE(gram && wq_k_co && sum && sumsq, "conv_first_stats_from_gram: null pointer");
  ZKB_REQUIRE(Cout > 0, "conv_first_stats_from_gram: Cout must be positive");
  conv_first_stats_from_gram_kernel<<<(Cout + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      gram, wq_k_co, Cout, sum, sumsq);
  ZKB_LAUNCH_CHECK("conv_first_stats_from_gram_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_conv_first_bn_sign(const float* x, const float* wq_k_co, int B, int H, int W, int Cin,
                                   int Cout, int ksize, const float* mean, const float* rstd,
                                   const float* beta, void* s_out_bf16, void* s_out_i8, void* stream) {
  int rc = first_conv_check("conv_first_bn_sign", x, wq_k_co, B, H, W, Cin, Cout, ksize);
  if (rc) return rc;
  ZKB_REQUIRE(mean && rstd && beta && (s_out_bf16 || s_out_i8), "conv_first_bn_sign: null pointer");
  if (B == 0) return ZKB_OK;
  ZKB_FIRST_DISPATCH(conv_first_bn_sign_kernel, x, wq_k_co, B, H, W, Cout, mean, rstd, beta,
                     static_cast<uint16_t*>(s_out_bf16), static_cast<int8_t*>(s_out_i8));
  return ZKB_OK;
}

ZKB_API int zkb_conv_first_bwd_reduce(const float* x, const float* wq_k_co, const float* g, int B,
                                      int H, int W, int Cin, int Cout, int ksize, const float* mean,
                                      const float* rstd, const float* beta, double* sum_da,
                                      double* sum_dax, void* stream) {
  int rc = first_conv_check("conv_first_bwd_reduce", x, wq_k_co, B, H, W, Cin, Cout, ksize);
  if (rc) return rc;
  ZKB_REQUIRE(g && mean && rstd && beta && sum_da && sum_dax, "conv_first_bwd_reduce: null pointer");
  ZKB_REQUIRE(ZKB_ALIGNED16(g), "conv_first_bwd_reduce: g must be 16-byte aligned");
  if (B == 0) return ZKB_OK;
  ZKB_FIRST_DISPATCH(conv_first_bwd_reduce_kernel, x, wq_k_co, g, B, H, W, Cout, mean, rstd, beta,
                     sum_da, sum_dax);
  return ZKB_OK;
}

ZKB_API int zkb_conv_first_bwd_wgrad(const float* x, const float* wq_k_co, const float* g, int B,
                                     int H, int W, int Cin, int Cout, int ksize, const float* mean,
                                     const float* rstd, const float* beta, const float* mean_da,
                                     const float* mean_dax, float* dw_k_co, void* stream) {
  int rc = first_conv_check("conv_first_bwd_wgrad", x, wq_k_co, B, H, W, Cin, Cout, ksize);
  if (rc) return rc;
  ZKB_REQUIRE(g && mean && rstd && beta && mean_da && mean_dax && dw_k_co,
              "conv_first_bwd_wgrad: null pointer");
  ZKB_REQUIRE(ZKB_ALIGNED16(g), "conv_first_bwd_wgrad: g must be 16-byte aligned");
  if (B == 0) return ZKB_OK;
  // one resident block per SM (its ~250 registers per thread) and a per-block epilogue of K
  // shared-memory reductions + atomics: two long blocks per SM instead of six short ones
  ZKB_FIRST_DISPATCH_W(2, conv_first_bwd_wgrad_kernel, x, wq_k_co, g, B, H, W, Cout, mean, rstd, beta,
                       mean_da, mean_dax, dw_k_co);
  return ZKB_OK;
}
