// uvio_klt_clahe_noalloc: CLAHE (cv::createCLAHE(10.0, Size(8,8))->apply, ov_core/src/track/TrackKLT.cpp:60-64) of a batch of
// device-resident images for callers that must not allocate: code under CUDA-graph capture, tight serving loops, harnesses
// that forbid cudaMalloc.  It is a variant of the tracker's own k_clahe_tiles / k_clahe_apply pair (kernels.cu), whose
// per-tile LUT scratch (64 tiles x 256 B per image) normally comes from a DevVec the engine grows with cudaMalloc; here the
// scratch is a file-scope __device__ array of a fixed, modest size that every GPU loading the library holds until exit.
//
// The array's NAME is only a host-side stand-in for the symbol in host code: the device address is taken once per device
// with cudaGetSymbolAddress (that device current) and passed to the kernels.  Nothing in this file calls cudaMalloc /
// cudaFree / an allocating library routine.  The arithmetic is the same image_body.h bodies the tracker runs, so results are
// bit-identical to the tracker's CLAHE path.
#include <cuda_runtime.h>
#include <stdio.h>

#include <mutex>

#include "../../include/uvio_klt.h"
#include "image_body.h"
#include "kernels.h"

namespace uvio {

constexpr int NA_MAX_IMAGES = UVIO_CLAHE_NOALLOC_MAX_IMAGES;
constexpr int NA_MAX_DIM = UVIO_CLAHE_NOALLOC_MAX_DIM;
constexpr int NA_LUT_BYTES = CLAHE_TILES * CLAHE_TILES * 256;   // per image
constexpr int NA_MAX_DEVICES = 64;

// The scratch: per-tile LUTs of one call.  512 KiB of device memory per GPU, shared by every call on that GPU.
static __device__ __align__(256) uint8_t g_clahe_lut_scratch[(size_t)NA_MAX_IMAGES * NA_LUT_BYTES];

struct NaBatch {   // by value in the kernel parameters: no job table in device memory to allocate or upload
  const uint8_t *src;
  uint8_t *dst;
  size_t src_pitch, src_image_stride, dst_pitch, dst_image_stride;
  int w, h;
};

__device__ __forceinline__ ImgView na_src(const NaBatch &b, int img) {
  return ImgView{b.src + (size_t)img * b.src_image_stride, b.w, b.h, (int)b.src_pitch};
}

// one block per (tile, image): clipped histogram -> LUT, written to the scratch (every byte of an image's 16 KiB is
// rewritten by every call, so nothing has to start at zero)
__global__ void __launch_bounds__(256) k_clahe_tiles_noalloc(NaBatch b, uint8_t *luts) {
  __shared__ unsigned sh[256];
  const ImgView I = na_src(b, blockIdx.y);
  const ClaheGeom q = clahe_geom(I.w, I.h, 10.0, CLAHE_TILES, CLAHE_TILES);
  sh[threadIdx.x] = 0;
  __syncthreads();
  clahe_tile_hist(GroupBlock(), sh, I, q, blockIdx.x % CLAHE_TILES, blockIdx.x / CLAHE_TILES);
  __syncthreads();
  if (threadIdx.x == 0) clahe_tile_lut(sh, q, luts + ((size_t)blockIdx.y * CLAHE_TILES * CLAHE_TILES + blockIdx.x) * 256);
}

// bilinear blend of the four neighbouring tile LUTs, one row band per block
__global__ void __launch_bounds__(256) k_clahe_apply_noalloc(NaBatch b, const uint8_t *luts) {
  const ImgView I = na_src(b, blockIdx.y);
  uint8_t *D = b.dst + (size_t)blockIdx.y * b.dst_image_stride;
  const ClaheGeom q = clahe_geom(I.w, I.h, 10.0, CLAHE_TILES, CLAHE_TILES);
  const uint8_t *L = luts + (size_t)blockIdx.y * NA_LUT_BYTES;
  int r0 = (int)((long long)I.h * blockIdx.x / gridDim.x), r1 = (int)((long long)I.h * (blockIdx.x + 1) / gridDim.x);
  for (int y = r0; y < r1; y++)
    for (int x = threadIdx.x; x < I.w; x += blockDim.x)
      D[(size_t)y * b.dst_pitch + x] = clahe_pixel(L, q, x, y, I.p[(size_t)y * I.pitch + x]);
}

// device address of the scratch, resolved once per device (with that device current)
static uint8_t *g_scratch_addr[NA_MAX_DEVICES];
static std::mutex g_scratch_mtx;

static cudaError_t scratch_for_current_device(int device, uint8_t **out) {
  std::lock_guard<std::mutex> lk(g_scratch_mtx);
  if (!g_scratch_addr[device]) {
    uint8_t *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void **)&p, g_clahe_lut_scratch);
    if (e != cudaSuccess) return e;
    if (!p) return cudaErrorInvalidSymbol;
    g_scratch_addr[device] = p;
  }
  *out = g_scratch_addr[device];
  return cudaSuccess;
}

}  // namespace uvio

using namespace uvio;

static int na_fail(char *err, size_t err_cap, int code, const char *fmt, long long a = 0, long long b = 0, long long c = 0) {
  if (err && err_cap) snprintf(err, err_cap, fmt, a, b, c);
  return code;
}

extern "C" int uvio_klt_clahe_noalloc(const uint8_t *src_dev, size_t src_pitch, size_t src_image_stride, uint8_t *dst_dev,
                                      size_t dst_pitch, size_t dst_image_stride, int width, int height, int n_images,
                                      void *cuda_stream, char *err, size_t err_cap) {
  if (err && err_cap) err[0] = 0;
  if (n_images == 0) return UVIO_OK;
  if (!src_dev || !dst_dev) return na_fail(err, err_cap, UVIO_E_ARG, "uvio_klt_clahe_noalloc: src_dev / dst_dev is NULL");
  if (n_images < 0 || n_images > NA_MAX_IMAGES)
    return na_fail(err, err_cap, UVIO_E_ARG,
                   "uvio_klt_clahe_noalloc: n_images = %lld exceeds the static scratch (at most %lld images per call)", n_images,
                   NA_MAX_IMAGES);
  if (width < 1 || height < 1 || width > NA_MAX_DIM || height > NA_MAX_DIM)
    return na_fail(err, err_cap, UVIO_E_ARG, "uvio_klt_clahe_noalloc: image size %lld x %lld is outside 1..%lld", width, height,
                   NA_MAX_DIM);
  if (src_pitch < (size_t)width || dst_pitch < (size_t)width || src_pitch > 0x7fffffff)
    return na_fail(err, err_cap, UVIO_E_ARG, "uvio_klt_clahe_noalloc: pitch smaller than the width (%lld) or too large", width);
  if (n_images > 1 && (src_image_stride < src_pitch * (size_t)height || dst_image_stride < dst_pitch * (size_t)height))
    return na_fail(err, err_cap, UVIO_E_ARG, "uvio_klt_clahe_noalloc: image stride smaller than pitch * height");
  int device = 0;
  cudaError_t e = cudaGetDevice(&device);
  if (e != cudaSuccess || device < 0 || device >= NA_MAX_DEVICES)
    return na_fail(err, err_cap, UVIO_E_CUDA, "uvio_klt_clahe_noalloc: no current CUDA device (cudaError %lld)", (long long)e);
  uint8_t *scratch = nullptr;
  e = scratch_for_current_device(device, &scratch);
  if (e != cudaSuccess)
    return na_fail(err, err_cap, UVIO_E_CUDA, "uvio_klt_clahe_noalloc: cudaGetSymbolAddress failed on device %lld (cudaError %lld)", device,
                   (long long)e);
  NaBatch b{src_dev, dst_dev, src_pitch, src_image_stride, dst_pitch, dst_image_stride, width, height};
  cudaStream_t st = (cudaStream_t)cuda_stream;
  k_clahe_tiles_noalloc<<<dim3(CLAHE_TILES * CLAHE_TILES, n_images), 256, 0, st>>>(b, scratch);
  k_clahe_apply_noalloc<<<dim3(HIST_BANDS * 2, n_images), 256, 0, st>>>(b, scratch);
  e = cudaGetLastError();
  if (e != cudaSuccess)
    return na_fail(err, err_cap, UVIO_E_CUDA, "uvio_klt_clahe_noalloc: kernel launch failed (cudaError %lld)", (long long)e);
  return UVIO_OK;
}
