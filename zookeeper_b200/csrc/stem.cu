// stem.cu -- the float stem convolutions of the larq-zoo models BASELINE.json configs[3] / [4] name
// (SURVEY.md section 8f.2, Appendix C): Bi-Real-Net-18 opens with Conv2D(64, 7, strides=2, "same") on
// the 3-channel image, BinaryAlexNet with an 11x11 / stride-4 QuantConv2D that has no input quantizer
// (float image x sign(W): the host passes the quantised kernel).  Neither is in the reference tree
// (the reference resolves model names through larq-zoo; examples/larq_experiment.py:59-103 is the
// hook); semantics restated in oracle/bireal_oracle.py.
//
// K = ksize^2 * 3 is 147 / 363 and the operand is the raw image, so this is not tensor-core work
// shaped for the binary GEMM engines (3 channels cannot be an im2col TMA operand: a pixel is 12 bytes).
// Both directions are register-tiled FP32 FMA kernels on shared-memory patches: the image patch of an
// 8 x 16 (8 x 8) tile of output pixels is read from HBM once per 64 output channels, zero-filled where
// the TensorFlow 'same' window leaves the image.  There is no input gradient (the operand is the data).
#include "zkb_host.h"

namespace zkb {

constexpr int kStemCin = 3;
constexpr int kStemCo = 64;     // output channels per block

template <int KS, int S, int TH, int TW>
__device__ __forceinline__ void stem_load_patch(float* sX, const float* __restrict__ x, int b, int iy0, int ix0,
                                                int H, int W, int nthreads) {
  constexpr int PH = (TH - 1) * S + KS, PW = (TW - 1) * S + KS;
  const float* img = x + static_cast<long long>(b) * H * W * kStemCin;
  for (int i = threadIdx.x; i < PH * PW * kStemCin; i += nthreads) {
    const int py = i / (PW * kStemCin);
    const int r = i - py * (PW * kStemCin);     // (px, ci): contiguous in the image row as well
    const int iy = iy0 + py, ix = ix0 + r / kStemCin;
    float v = 0.f;
    if (iy >= 0 && iy < H && ix >= 0 && ix < W) v = __ldg(img + (static_cast<long long>(iy) * W + ix0) * kStemCin + r);
    sX[i] = v;
  }
}

// y[b,oy,ox,co] = sum_{kh,kw,ci} x[b, oy S - pad_h + kh, ox S - pad_w + kw, ci] * w[kh,kw,ci,co]
// block: 256 threads = 16 channel lanes (4 channels each) x 16 pixel lanes (TW / (16 / TH) pixels of a row each)
template <int KS, int S, int TH, int TW>
__global__ void __launch_bounds__(256)
conv_stem_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, float* __restrict__ y, int H,
                     int W, int OH, int OW, int pad_h, int pad_w, int Cout, int tiles_x, int tiles_y) {
  constexpr int PW = (TW - 1) * S + KS, PH = (TH - 1) * S + KS;
  constexpr int GPR = 16 / TH;          // pixel lanes per tile row
  constexpr int PPT = TW / GPR;         // pixels per thread
  constexpr int KC = KS * kStemCin;     // (kw, ci) pairs of a filter row
  __shared__ __align__(16) float sX[PH * PW * kStemCin];
  __shared__ __align__(16) float sW[KC * kStemCo];
  int t = blockIdx.x;
  const int txi = t % tiles_x;
  t /= tiles_x;
  const int tyi = t % tiles_y;
  const int b = t / tiles_y;
  const int cg = blockIdx.y;
  const int oy0 = tyi * TH, ox0 = txi * TW;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int row = ty / GPR, col0 = (ty % GPR) * PPT;

  stem_load_patch<KS, S, TH, TW>(sX, x, b, oy0 * S - pad_h, ox0 * S - pad_w, H, W, 256);
  float acc[PPT][4];
#pragma unroll
  for (int j = 0; j < PPT; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;

  for (int kh = 0; kh < KS; ++kh) {
    __syncthreads();      // the patch is in place / the previous filter row has been consumed
    for (int i = threadIdx.x; i < KC * kStemCo; i += 256)
      sW[i] = __ldg(w + (static_cast<long long>(kh) * KC + (i >> 6)) * Cout + cg * kStemCo + (i & 63));
    __syncthreads();
    const float* xr = sX + ((row * S + kh) * PW + col0 * S) * kStemCin;
#pragma unroll
    for (int kc = 0; kc < KC; ++kc) {
      const float4 w4 = *reinterpret_cast<const float4*>(sW + kc * kStemCo + tx * 4);
#pragma unroll
      for (int j = 0; j < PPT; ++j) {
        const float xv = xr[j * S * kStemCin + kc];
        acc[j][0] = fmaf(xv, w4.x, acc[j][0]);
        acc[j][1] = fmaf(xv, w4.y, acc[j][1]);
        acc[j][2] = fmaf(xv, w4.z, acc[j][2]);
        acc[j][3] = fmaf(xv, w4.w, acc[j][3]);
      }
    }
  }
  const int oy = oy0 + row;
  if (oy < OH) {
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
      const int ox = ox0 + col0 + j;
      if (ox < OW)
        *reinterpret_cast<float4*>(y + ((static_cast<long long>(b) * OH + oy) * OW + ox) * Cout + cg * kStemCo + tx * 4) =
            make_float4(acc[j][0], acc[j][1], acc[j][2], acc[j][3]);
    }
  }
}

// dw[kh,kw,ci,co] += sum_{b,oy,ox} x[b, oy S - pad_h + kh, ox S - pad_w + kw, ci] * dy[b,oy,ox,co]
// block: SL pixel slices x KS filter rows x (64 / COT) channel lanes; a thread keeps the KS*3 x COT
// gradients of its filter row in registers over every tile the (persistent) block visits and adds
// them to dw once at the end.
template <int KS, int S, int TH, int TW, int COT, int SL>
__global__ void __launch_bounds__(KS * (kStemCo / COT) * SL)
conv_stem_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dw, int H,
                       int W, int OH, int OW, int pad_h, int pad_w, int Cout, int tiles_x, int tiles_y,
                       int n_tiles) {
  constexpr int PW = (TW - 1) * S + KS, PH = (TH - 1) * S + KS;
  constexpr int KC = KS * kStemCin;
  constexpr int TP = TH * TW;
  constexpr int LANES = kStemCo / COT;
  constexpr int NT = KS * LANES * SL;
  __shared__ __align__(16) float sX[PH * PW * kStemCin];
  __shared__ __align__(16) float sD[TP * kStemCo];
  const int cg = blockIdx.y;
  const int sl = threadIdx.x / (KS * LANES);
  const int rr = threadIdx.x - sl * (KS * LANES);
  const int kh = rr / LANES;
  const int co0 = (rr - kh * LANES) * COT;
  float acc[KC][COT];
#pragma unroll
  for (int kc = 0; kc < KC; ++kc)
#pragma unroll
    for (int c = 0; c < COT; ++c) acc[kc][c] = 0.f;

  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    int t = tile;
    const int txi = t % tiles_x;
    t /= tiles_x;
    const int tyi = t % tiles_y;
    const int b = t / tiles_y;
    const int oy0 = tyi * TH, ox0 = txi * TW;
    __syncthreads();      // the previous tile has been consumed
    stem_load_patch<KS, S, TH, TW>(sX, x, b, oy0 * S - pad_h, ox0 * S - pad_w, H, W, NT);
    for (int i = threadIdx.x; i < TP * (kStemCo / 4); i += NT) {
      const int p = i / (kStemCo / 4), c4 = i - p * (kStemCo / 4);
      const int oy = oy0 + p / TW, ox = ox0 + p % TW;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (oy < OH && ox < OW)
        v = __ldg(reinterpret_cast<const float4*>(dy + ((static_cast<long long>(b) * OH + oy) * OW + ox) * Cout +
                                                  cg * kStemCo + c4 * 4));
      *reinterpret_cast<float4*>(sD + p * kStemCo + c4 * 4) = v;
    }
    __syncthreads();
    for (int p = sl; p < TP; p += SL) {
      const int row = p / TW, col = p % TW;
      float dv[COT];
      if constexpr (COT == 4) {
        const float4 d4 = *reinterpret_cast<const float4*>(sD + p * kStemCo + co0);
        dv[0] = d4.x; dv[1] = d4.y; dv[2] = d4.z; dv[3] = d4.w;
      } else {
        static_assert(COT == 2, "COT is 2 or 4");
        const float2 d2 = *reinterpret_cast<const float2*>(sD + p * kStemCo + co0);
        dv[0] = d2.x; dv[1] = d2.y;
      }
      const float* xr = sX + ((row * S + kh) * PW + col * S) * kStemCin;
#pragma unroll
      for (int kc = 0; kc < KC; ++kc) {
        const float xv = xr[kc];
#pragma unroll
        for (int c = 0; c < COT; ++c) acc[kc][c] = fmaf(xv, dv[c], acc[kc][c]);
      }
    }
  }
  float* o = dw + static_cast<long long>(kh) * KC * Cout + cg * kStemCo + co0;
#pragma unroll
  for (int kc = 0; kc < KC; ++kc)
#pragma unroll
    for (int c = 0; c < COT; ++c) atomicAdd(o + static_cast<long long>(kc) * Cout + c, acc[kc][c]);
}

}  // namespace zkb

// ============================================================================================
// C ABI (documented in include/zkb200.h)
// ============================================================================================
using namespace zkb;

namespace {
struct StemGeom {
  int OH, OW, pad_h, pad_w;
};
// TensorFlow 'same': out = ceil(in / s), total padding max((out-1) s + k - in, 0), the smaller half first
StemGeom stem_geometry(int H, int W, int ksize, int stride) {
  StemGeom g;
  g.OH = (H + stride - 1) / stride;
  g.OW = (W + stride - 1) / stride;
  int th = (g.OH - 1) * stride + ksize - H, tw = (g.OW - 1) * stride + ksize - W;
  g.pad_h = (th > 0 ? th : 0) / 2;
  g.pad_w = (tw > 0 ? tw : 0) / 2;
  return g;
}

int stem_check(const char* who, const void* a, const void* b, const void* c, int B, int H, int W, int Cin, int Cout,
               int ksize, int stride) {
  ZKB_REQUIRE(a && b && c, "%s: null pointer", who);
  ZKB_REQUIRE(B >= 0 && H > 0 && W > 0, "%s: bad shape", who);
  ZKB_REQUIRE(Cin == kStemCin, "%s: built for 3-channel images (Cin=%d)", who, Cin);
  ZKB_REQUIRE(Cout > 0 && Cout % kStemCo == 0, "%s: Cout=%d must be a multiple of 64", who, Cout);
  ZKB_REQUIRE((ksize == 7 && stride == 2) || (ksize == 11 && stride == 4),
              "%s: built for the 7x7 / stride-2 and 11x11 / stride-4 stems (ksize=%d stride=%d)", who, ksize, stride);
  ZKB_REQUIRE((reinterpret_cast<uintptr_t>(a) & 15) == 0 && (reinterpret_cast<uintptr_t>(b) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(c) & 15) == 0,
              "%s: tensors must be 16-byte aligned", who);
  return ZKB_OK;
}
}  // namespace

ZKB_API int zkb_conv_stem_f32_fwd(const float* x, const float* w_hwio, float* y, int B, int H, int W, int Cin,
                                  int Cout, int ksize, int stride, void* stream) {
  int rc = stem_check("conv_stem_f32_fwd", x, w_hwio, y, B, H, W, Cin, Cout, ksize, stride);
  if (rc) return rc;
  if (B == 0) return ZKB_OK;
  const StemGeom g = stem_geometry(H, W, ksize, stride);
  constexpr int TH = 8, TW = 16;
  const int tiles_y = (g.OH + TH - 1) / TH, tiles_x = (g.OW + TW - 1) / TW;
  const long long tiles = static_cast<long long>(B) * tiles_y * tiles_x;
  ZKB_REQUIRE(tiles < (1ll << 31), "conv_stem_f32_fwd: too many tiles");
  const dim3 grid(static_cast<unsigned>(tiles), static_cast<unsigned>(Cout / kStemCo));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (ksize == 7)
    conv_stem_fwd_kernel<7, 2, TH, TW><<<grid, 256, 0, st>>>(x, w_hwio, y, H, W, g.OH, g.OW, g.pad_h, g.pad_w, Cout,
                                                            tiles_x, tiles_y);
  else
    conv_stem_fwd_kernel<11, 4, TH, TW><<<grid, 256, 0, st>>>(x, w_hwio, y, H, W, g.OH, g.OW, g.pad_h, g.pad_w, Cout,
                                                             tiles_x, tiles_y);
  ZKB_LAUNCH_CHECK("conv_stem_fwd_kernel");
  return ZKB_OK;
}

ZKB_API int zkb_conv_stem_f32_wgrad(const float* x, const float* dy, float* dw_hwio, int B, int H, int W, int Cin,
                                    int Cout, int ksize, int stride, void* stream) {
  int rc = stem_check("conv_stem_f32_wgrad", x, dy, dw_hwio, B, H, W, Cin, Cout, ksize, stride);
  if (rc) return rc;
  if (B == 0) return ZKB_OK;
  const StemGeom g = stem_geometry(H, W, ksize, stride);
  constexpr int TH = 8;
  const int TW = ksize == 7 ? 16 : 8;
  const int tiles_y = (g.OH + TH - 1) / TH, tiles_x = (g.OW + TW - 1) / TW;
  const long long tiles = static_cast<long long>(B) * tiles_y * tiles_x;
  ZKB_REQUIRE(tiles < (1ll << 31), "conv_stem_f32_wgrad: too many tiles");
  // persistent blocks: two per SM and channel group, each adds its registers to dw once
  long long gx = static_cast<long long>(zkb_sm_count()) * 2 / (Cout / kStemCo);
  if (gx < 1) gx = 1;
  if (gx > tiles) gx = tiles;
  const dim3 grid(static_cast<unsigned>(gx), static_cast<unsigned>(Cout / kStemCo));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (ksize == 7)
    conv_stem_wgrad_kernel<7, 2, TH, 16, 4, 2><<<grid, 7 * 16 * 2, 0, st>>>(
        x, dy, dw_hwio, H, W, g.OH, g.OW, g.pad_h, g.pad_w, Cout, tiles_x, tiles_y, static_cast<int>(tiles));
  else
    conv_stem_wgrad_kernel<11, 4, TH, 8, 2, 1><<<grid, 11 * 32 * 1, 0, st>>>(
        x, dy, dw_hwio, H, W, g.OH, g.OW, g.pad_h, g.pad_w, Cout, tiles_x, tiles_y, static_cast<int>(tiles));
  ZKB_LAUNCH_CHECK("conv_stem_wgrad_kernel");
  return ZKB_OK;
}
