// zkb_host.h -- host-side conventions shared by every translation unit of libzkb200.so.
//
// Status convention of the C ABI (include/zkb200.h): 0 = ok, negative = argument error detected
// before any launch, positive = cudaError_t / CUresult of the failing runtime call.  The message
// of the last failure on the calling thread is kept in a thread-local buffer.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#define ZKB_API extern "C" __attribute__((visibility("default")))

enum {
  ZKB_OK = 0,
  ZKB_ERR_ARG = -1,       // bad pointer / size / alignment
  ZKB_ERR_UNSUPPORTED = -2,
  ZKB_ERR_DRIVER = -3,    // driver entry point (tensor-map encoder) unavailable
};

int zkb_fail(int code, const char* fmt, ...);

#define ZKB_REQUIRE(cond, ...)                         \
  do {                                                 \
    if (!(cond)) return zkb_fail(ZKB_ERR_ARG, __VA_ARGS__); \
  } while (0)

#define ZKB_CUDA(call)                                                                  \
  do {                                                                                  \
    cudaError_t e__ = (call);                                                           \
    if (e__ != cudaSuccess)                                                             \
      return zkb_fail(static_cast<int>(e__), "%s failed: %s", #call, cudaGetErrorString(e__)); \
  } while (0)

#define ZKB_LAUNCH_CHECK(name)                                                               \
  do {                                                                                       \
    cudaError_t e__ = cudaGetLastError();                                                    \
    if (e__ != cudaSuccess)                                                                  \
      return zkb_fail(static_cast<int>(e__), "launch of %s failed: %s", name,                \
                      cudaGetErrorString(e__));                                              \
  } while (0)

// Tensor-map encoders (driver API reached through cudaGetDriverEntryPoint, so the library
// has no link-time dependency on libcuda and can be dlopen'ed on a box without a driver).
// dims/strides are innermost-first; strides in bytes for dims 1..rank-1.
int zkb_tmap_tiled_2d(CUtensorMap* out, CUtensorMapDataType dt, int elem_bytes, const void* base,
                      uint64_t dim0, uint64_t dim1, uint64_t stride1_bytes, uint32_t box0,
                      uint32_t box1);
// The same without shared-memory swizzle (box rows of any multiple of 16 bytes): store boxes that
// threads fill row by row.
int zkb_tmap_tiled_2d_plain(CUtensorMap* out, CUtensorMapDataType dt, int elem_bytes, const void* base,
                      uint64_t dim0, uint64_t dim1, uint64_t stride1_bytes, uint32_t box0,
                      uint32_t box1);
// NHWC activation [N,H,W,C]; 'same'-style window described by pad (lower) and ksize.
int zkb_tmap_im2col_nhwc(CUtensorMap* out, CUtensorMapDataType dt, int elem_bytes,
                         const void* base, int N, int H, int W, int C, int ksize, int pad,
                         uint32_t channels_per_pixel, uint32_t pixels_per_column);

// The same with a traversal stride and separate low / high zero padding per spatial axis (strided
// 'same' convolutions pad asymmetrically): base pixels run from -pad_lo in steps of `stride`.
int zkb_tmap_im2col_nhwc_ex(CUtensorMap* out, CUtensorMapDataType dt, int elem_bytes, const void* base,
                            int N, int H, int W, int C, int ksize, int pad_lo_h, int pad_hi_h,
                            int pad_lo_w, int pad_hi_w, int stride, uint32_t channels_per_pixel,
                            uint32_t pixels_per_column);

int zkb_sm_count();
