// Execution-group abstraction shared by every kernel body in this directory.
//
// Kernel bodies are written once as "phase" code: loops of the form
//     for (int i = g.lane(); i < n; i += g.size()) { ... }   separated by   g.sync();
// On the device `g` is a warp (GroupWarp) or a thread block (GroupBlock).  tests/hostsim compiles the very same bodies
// with g++ and runs them with GroupSeq (one lane, sync = no-op) so the arithmetic can be checked against OpenCV on a
// machine without a GPU.  GroupSeq is test scaffolding only: the product library (libuvio_klt.so) is built by nvcc,
// contains device code only, and refuses to load without a CUDA device.
#pragma once
#include <math.h>
#include <stdint.h>

#ifdef __CUDACC__
#define UVIO_HD __host__ __device__ __forceinline__
#else
#define UVIO_HD inline
#endif

#define UVIO_RESTRICT __restrict__

namespace uvio {

struct GroupSeq {
  UVIO_HD int lane() const { return 0; }
  UVIO_HD int size() const { return 1; }
  UVIO_HD void sync() const {}
  static constexpr int SIZE = 1;                        // compile-time size, for per-lane register arrays
  typedef long long acc_t;                              // wide enough for one lane's share of an integer sum
  UVIO_HD unsigned reduce_add(unsigned v) const { return v; }
  // exclusive prefix sum of one value per lane, in lane order; total = sum over the group.  Every lane must call it.
  UVIO_HD int exclusive_sum(int v, int &total) const { total = v; return 0; }
};

#ifdef __CUDACC__
struct GroupWarp {
  __device__ __forceinline__ int lane() const { return threadIdx.x & 31; }
  __device__ __forceinline__ int size() const { return 32; }
  __device__ __forceinline__ void sync() const { __syncwarp(); }
  static constexpr int SIZE = 32;
  typedef int acc_t;
  __device__ __forceinline__ unsigned reduce_add(unsigned v) const { return __reduce_add_sync(0xffffffffu, v); }   // REDUX
};
struct GroupBlock {
  __device__ __forceinline__ int lane() const { return threadIdx.x; }
  __device__ __forceinline__ int size() const { return blockDim.x; }
  __device__ __forceinline__ void sync() const { __syncthreads(); }
  // block-wide exclusive prefix sum (warp shuffles + one shared word per warp); blockDim.x a multiple of 32, <= 1024
  __device__ __forceinline__ int exclusive_sum(int v, int &total) const {
    __shared__ int warp_sum[33];
    const int lane_id = threadIdx.x & 31, warp_id = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    int inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int up = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane_id >= d) inc += up;
    }
    __syncthreads();                       // the previous call's readers are done with warp_sum
    if (lane_id == 31) warp_sum[warp_id] = inc;
    __syncthreads();
    if (warp_id == 0) {
      int w = lane_id < nwarps ? warp_sum[lane_id] : 0, winc = w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, winc, d);
        if (lane_id >= d) winc += up;
      }
      if (lane_id < nwarps) warp_sum[lane_id] = winc - w;
      if (lane_id == 31) warp_sum[32] = winc;
    }
    __syncthreads();
    total = warp_sum[32];
    return warp_sum[warp_id] + inc - v;
  }
};
#endif

// cvRound: round half to even, float -> int.
UVIO_HD int round_f2i(float v) {
#ifdef __CUDA_ARCH__
  return __float2int_rn(v);
#else
  return (int)lrintf(v);
#endif
}
UVIO_HD int round_d2i(double v) {
#ifdef __CUDA_ARCH__
  return __double2int_rn(v);
#else
  return (int)lrint(v);
#endif
}
UVIO_HD int floor_f2i(float v) { return (int)floorf(v); }
UVIO_HD uint8_t sat_u8(int v) { return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); }
UVIO_HD int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// Non-contracted fp32 arithmetic: the OpenCV baseline build the oracle pins (SSE3, no FMA) never fuses a*b+c.
// The .cu files are also compiled with -fmad=false; these helpers make the intent explicit where it matters.
UVIO_HD float fmul(float a, float b) {
#ifdef __CUDA_ARCH__
  return __fmul_rn(a, b);
#else
  return a * b;
#endif
}
UVIO_HD float fadd(float a, float b) {
#ifdef __CUDA_ARCH__
  return __fadd_rn(a, b);
#else
  return a + b;
#endif
}
UVIO_HD float fsub(float a, float b) {
#ifdef __CUDA_ARCH__
  return __fsub_rn(a, b);
#else
  return a - b;
#endif
}
UVIO_HD double dmul(double a, double b) {
#ifdef __CUDA_ARCH__
  return __dmul_rn(a, b);
#else
  return a * b;
#endif
}
UVIO_HD double dadd(double a, double b) {
#ifdef __CUDA_ARCH__
  return __dadd_rn(a, b);
#else
  return a + b;
#endif
}
UVIO_HD double dsub(double a, double b) {
#ifdef __CUDA_ARCH__
  return __dsub_rn(a, b);
#else
  return a - b;
#endif
}

// BORDER_REFLECT_101 (gfedcb|abcdefgh|gfedcba), any distance.
UVIO_HD int reflect101(int p, int len) {
  if (len == 1) return 0;
  while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
  return p;
}

// The same without branches, for -len < p < 2*len - 1 (a single reflection) and len >= 2.
UVIO_HD int reflect101_once(int p, int len) {
  p = p < 0 ? -p : p;
  int q = 2 * (len - 1) - p;
  return p < q ? p : q;
}

// A pitched 8-bit image view (one pyramid level).
struct ImgView {
  const uint8_t *p;
  int w, h, pitch;
};

}  // namespace uvio
