// pool3.cu -- MaxPool2D(pool_size=3, strides=2) on NHWC maps, the overlapping pool of the larq-zoo
// models: padding "valid" in BinaryAlexNet, where it sits between
// a QuantConv2D and its BatchNormalization and so pools the INTEGER conv output (int16 here; the
// 11x11 stem's output is float), padding "same" behind Bi-Real-Net's stem.  Not in the reference tree
// (examples/larq_experiment.py:74,83,92 only has the 2x2 pool); the tie rule is the one SURVEY.md
// Appendix A.6 states for it: the first maximum in row-major window order takes the gradient --
// integer maps tie often, so it matters.
//
// HBM-bound: a thread owns 4 channels of one output (forward) or input (backward) pixel.  The
// backward is a gather -- every input pixel sums the gradients of the (at most four) windows it won --
// so it needs no atomics and can write the bf16 hi/lo pair the backward GEMMs read directly.
#include <cuda_bf16.h>

#include "zkb_host.h"

namespace zkb {

template <typename T>
struct Pool3Io;
template <>
struct Pool3Io<float> {
  static __device__ __forceinline__ void load(const float* p, float (&v)[4]) {
    const float4 t = *reinterpret_cast<const float4*>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  static __device__ __forceinline__ void store(float* p, const float (&v)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};
template <>
struct Pool3Io<int16_t> {
  static __device__ __forceinline__ void load(const int16_t* p, float (&v)[4]) {
    const uint2 t = *reinterpret_cast<const uint2*>(p);
    v[0] = static_cast<float>(static_cast<int16_t>(t.x & 0xFFFFu));
    v[1] = static_cast<float>(static_cast<int16_t>(t.x >> 16));
    v[2] = static_cast<float>(static_cast<int16_t>(t.y & 0xFFFFu));
    v[3] = static_cast<float>(static_cast<int16_t>(t.y >> 16));
  }
  static __device__ __forceinline__ void store(int16_t* p, const float (&v)[4]) {   // v holds integers
    const uint32_t a = static_cast<uint32_t>(static_cast<int>(v[0])) & 0xFFFFu, b = static_cast<uint32_t>(static_cast<int>(v[1])) << 16;
    const uint32_t c = static_cast<uint32_t>(static_cast<int>(v[2])) & 0xFFFFu, d = static_cast<uint32_t>(static_cast<int>(v[3])) << 16;
    *reinterpret_cast<uint2*>(p) = make_uint2(a | b, c | d);
  }
};

template <typename T>
__global__ void __launch_bounds__(256)
pool3s2_fwd_kernel(const T* __restrict__ x, int B, int H, int W, int C, int OH, int OW, int pad_h, int pad_w,
                   T* __restrict__ out, uint8_t* __restrict__ arg) {
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
    uint32_t am[4] = {0, 0, 0, 0};
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
      const int ih = oh * 2 + kh - pad_h;
      if (ih < 0 || ih >= H) continue;
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const int iw = ow * 2 + kw - pad_w;
        if (iw < 0 || iw >= W) continue;
        float v[4];
        Pool3Io<T>::load(x + ((b * H + ih) * W + iw) * C + cl * 4, v);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (v[j] > best[j]) { best[j] = v[j]; am[j] = static_cast<uint32_t>(kh * 3 + kw); }   // strict: the first maximum stays
      }
    }
    Pool3Io<T>::store(out + i * 4, best);
    *reinterpret_cast<uint32_t*>(arg + i * 4) = am[0] | (am[1] << 8) | (am[2] << 16) | (am[3] << 24);
  }
}

// dx[q] = sum of dout over the windows that cover q and whose winning slot is q.  HILO: the result is
// written as the bf16 pair (hi = rn(v), lo = rn(v - hi)) instead of f32.
template <bool HILO>
__global__ void __launch_bounds__(256)
pool3s2_bwd_kernel(const float* __restrict__ dout, const uint8_t* __restrict__ arg, int B, int H, int W, int C,
                   int OH, int OW, int pad_h, int pad_w, float* __restrict__ dx, uint16_t* __restrict__ dx_hi,
                   uint16_t* __restrict__ dx_lo) {
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
    for (int kh = 0; kh < 3; ++kh) {          // windows (oh, ow) with oh*2 + kh - pad_h == ih
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
    if (HILO) {
      uint16_t h[4], l[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const __nv_bfloat16 hb = __float2bfloat16_rn(acc[j]);
        h[j] = __bfloat16_as_ushort(hb);
        l[j] = __bfloat16_as_ushort(__float2bfloat16_rn(acc[j] - __bfloat162float(hb)));
      }
      *reinterpret_cast<uint2*>(dx_hi + i * 4) = make_uint2(h[0] | (static_cast<uint32_t>(h[1]) << 16), h[2] | (static_cast<uint32_t>(h[3]) << 16));
      *reinterpret_cast<uint2*>(dx_lo + i * 4) = make_uint2(l[0] | (static_cast<uint32_t>(l[1]) << 16), l[2] | (static_cast<uint32_t>(l[3]) << 16));
    } else {
      reinterpret_cast<float4*>(dx)[i] = make_float4(acc[0], acc[1], acc[2], acc[3]);
    }
  }
}

}  // namespace zkb

using namespace zkb;

namespace {
struct Pool3Geom {
  int OH, OW, ph, pw;
};
// "same": out = ceil(in / 2), total padding max((out-1)*2 + 3 - in, 0), the smaller half first;
// "valid": out = (in - 3) / 2 + 1, no padding (the last row / column of an even map is never read)
bool pool3_geometry(int H, int W, int same, Pool3Geom* g) {
  if (same) {
    g->OH = (H + 1) / 2;
    g->OW = (W + 1) / 2;
    int th = (g->OH - 1) * 2 + 3 - H, tw = (g->OW - 1) * 2 + 3 - W;
    g->ph = (th > 0 ? th : 0) / 2;
    g->pw = (tw > 0 ? tw : 0) / 2;
    return H > 0 && W > 0;
  }
  if (H < 3 || W < 3) return false;
  g->OH = (H - 3) / 2 + 1;
  g->OW = (W - 3) / 2 + 1;
  g->ph = g->pw = 0;
  return true;
}

int pool3_grid(long long items) {
  const long long cap = static_cast<long long>(zkb_sm_count()) * 16;
  long long g = (items + 255) / 256;
  if (g > cap) g = cap;
  return g < 1 ? 1 : static_cast<int>(g);
}
}  // namespace

ZKB_API int zkb_maxpool3s2_fwd(const void* x, int x_is_i16, int B, int H, int W, int C, int same, void* out,
                               uint8_t* arg, void* stream) {
  ZKB_REQUIRE(x && out && arg, "maxpool3s2_fwd: null pointer");
  ZKB_REQUIRE(B >= 0 && C > 0 && C % 4 == 0, "maxpool3s2_fwd: bad shape (C %% 4 == 0)");
  Pool3Geom g;
  ZKB_REQUIRE(pool3_geometry(H, W, same, &g), "maxpool3s2_fwd: map %dx%d is smaller than the window", H, W);
  ZKB_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(arg) & 3) == 0,
              "maxpool3s2_fwd: misaligned tensors");
  if (B == 0) return ZKB_OK;
  const int grid = pool3_grid(static_cast<long long>(B) * g.OH * g.OW * (C / 4));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (x_is_i16)
    pool3s2_fwd_kernel<int16_t><<<grid, 256, 0, st>>>(static_cast<const int16_t*>(x), B, H, W, C, g.OH, g.OW, g.ph, g.pw,
                                                     static_cast<int16_t*>(out), arg);
  else
    pool3s2_fwd_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(x), B, H, W, C, g.OH, g.OW, g.ph, g.pw,
                                                   static_cast<float*>(out), arg);
  ZKB_LAUNCH_CHECK("pool3s2_fwd_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_maxpool3s2_bwd(const float* dout, const uint8_t* arg, int B, int H, int W, int C, int same,
                               float* dx_f32, void* dx_hi, void* dx_lo, void* stream) {
  ZKB_REQUIRE(dout && arg, "maxpool3s2_bwd: null pointer");
  ZKB_REQUIRE((dx_f32 != nullptr) != (dx_hi != nullptr) && (dx_hi != nullptr) == (dx_lo != nullptr),
              "maxpool3s2_bwd: give dx_f32 or the dx_hi / dx_lo pair");
  ZKB_REQUIRE(B >= 0 && C > 0 && C % 4 == 0, "maxpool3s2_bwd: bad shape (C %% 4 == 0)");
  Pool3Geom g;
  ZKB_REQUIRE(pool3_geometry(H, W, same, &g), "maxpool3s2_bwd: map %dx%d is smaller than the window", H, W);
  ZKB_REQUIRE((reinterpret_cast<uintptr_t>(dout) & 15) == 0 && (reinterpret_cast<uintptr_t>(arg) & 3) == 0 &&
                  (reinterpret_cast<uintptr_t>(dx_f32) & 15) == 0 && (reinterpret_cast<uintptr_t>(dx_hi) & 7) == 0 &&
                  (reinterpret_cast<uintptr_t>(dx_lo) & 7) == 0,
              "maxpool3s2_bwd: misaligned tensors");
  if (B == 0) return ZKB_OK;
  const int grid = pool3_grid(static_cast<long long>(B) * H * W * (C / 4));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dx_f32)
    pool3s2_bwd_kernel<false><<<grid, 256, 0, st>>>(dout, arg, B, H, W, C, g.OH, g.OW, g.ph, g.pw, dx_f32, nullptr, nullptr);
  else
    pool3s2_bwd_kernel<true><<<grid, 256, 0, st>>>(dout, arg, B, H, W, C, g.OH, g.OW, g.ph, g.pw, nullptr,
                                                  static_cast<uint16_t*>(dx_hi), static_cast<uint16_t*>(dx_lo));
  ZKB_LAUNCH_CHECK("pool3s2_bwd_kernel");
  return ZKB_OK;
}
