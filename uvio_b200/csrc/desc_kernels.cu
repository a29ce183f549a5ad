// sm_100a kernels of the descriptor tracker (ov_core TrackDescriptor): thin __global__ wrappers around desc_body.h.
// FAST (k_fast), cornerSubPix (k_subpix) and the equalisation (k_preprocess) are the KLT path's kernels in kernels.cu.
#include "desc.h"
#include "orb_tables.h"

namespace uvio {

__constant__ int8_t c_orb_pairs[256][4] = ORB_PAIRS_INIT;

constexpr int DESC_THREADS = 256;
constexpr int DESC_TOPK_SMEM = 4096;
constexpr int ORB_WARPS = 4;
constexpr int MATCH_THREADS = 128;

__global__ void __launch_bounds__(DESC_THREADS) k_desc_topk(const DescUnit *units, DescGlobals X) {
  __shared__ uint32_t buf[DESC_TOPK_SMEM];
  const DescUnit &U = units[blockIdx.y >> 1];
  const int side = blockIdx.y & 1;
  if (side && !U.stereo) return;
  if ((int)blockIdx.x >= U.s[side].ncells) return;
  desc_topk(GroupBlock(), U, X, side, blockIdx.x, buf, DESC_TOPK_SMEM);
}

__global__ void __launch_bounds__(DESC_THREADS) k_desc_gather(const DescUnit *units, DescGlobals X) {
  __shared__ int s_base;
  const DescUnit &U = units[blockIdx.x >> 1];
  const int side = blockIdx.x & 1;
  if (side && !U.stereo) return;
  desc_gather(GroupBlock(), U, X, side, &s_base);
}

__global__ void __launch_bounds__(ORB_WARPS * 32) k_desc_orb(const DescUnit *units, DescGlobals X) {
  __shared__ OrbSmem sm[ORB_WARPS];
  __shared__ short rank[DESC_MAX_P];
  __shared__ int8_t pairs[1024];
  const DescUnit &U = units[blockIdx.y >> 1];
  const int side = blockIdx.y & 1;
  if (side && !U.stereo) return;
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) *X.n_sp = 0;   // k_subpix is done with the item list: re-armed for the next feed
  const DescSide &S = U.s[side];
  const int n = S.cnt[0];
  if ((int)(blockIdx.x * ORB_WARPS) >= n && blockIdx.x != 0) return;
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) pairs[i] = (&c_orb_pairs[0][0])[i];
  desc_orb_ranks(GroupBlock(), U, X, side, rank, blockIdx.x == 0);
  const int wid = threadIdx.x >> 5;
  for (int i = blockIdx.x * ORB_WARPS + wid; i < n; i += gridDim.x * ORB_WARPS) {
    const int k = rank[i];
    if (k < 0) continue;
    const float x = X.sp_out[2 * (S.slot0 + i)], y = X.sp_out[2 * (S.slot0 + i) + 1];
    orb_describe(GroupWarp(), sm[wid], S.img, round_f2i(x), round_f2i(y), pairs, S.kp_desc + (size_t)DESC_WORDS * k);
  }
}

__device__ __forceinline__ void rs_bind(RansacSmem &s, int P) {
  extern __shared__ float desc_dyn[];
  s.x0 = desc_dyn;
  s.y0 = desc_dyn + P;
  s.x1 = desc_dyn + 2 * P;
  s.y1 = desc_dyn + 3 * P;
}

__global__ void __launch_bounds__(MATCH_THREADS) k_desc_detect(const DescUnit *units, DescGlobals X, int P) {
  __shared__ RansacSmem rs;
  const DescUnit &U = units[blockIdx.x];
  if (U.stereo) {
    rs_bind(rs, P);
    __syncthreads();
    desc_match(GroupBlock(), rs, U, X, 0);
    desc_assign_stereo(GroupBlock(), U);
  } else {
    desc_assign_mono(GroupBlock(), U);
  }
}

__global__ void __launch_bounds__(MATCH_THREADS) k_desc_temporal(const DescUnit *units, DescGlobals X, int P) {
  __shared__ RansacSmem rs;
  const DescUnit &U = units[blockIdx.y];
  const int job = 1 + blockIdx.x;
  if (job == 2 && !U.stereo) return;
  rs_bind(rs, P);
  __syncthreads();
  desc_match(GroupBlock(), rs, U, X, job);
}

__global__ void __launch_bounds__(DESC_THREADS) k_desc_merge(const DescUnit *units, DescGlobals X, int P) {
  extern __shared__ int merge_dyn[];
  desc_merge(GroupBlock(), units[blockIdx.x], X, merge_dyn);
}

template <class K>
static cudaError_t smem_attr(K kernel, size_t bytes) {      // the attribute belongs to the (function, device) pair: set it every time it is needed
  if (bytes <= 48 * 1024) return cudaSuccess;
  return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
}

cudaError_t launch_desc_topk(cudaStream_t st, const DescUnit *units, int n_units, int max_cells, const DescGlobals &X) {
  if (n_units <= 0 || max_cells <= 0) return cudaSuccess;
  prof_mark(st, K_DV_TOPK);
  k_desc_topk<<<dim3(max_cells, 2 * n_units), DESC_THREADS, 0, st>>>(units, X);
  prof_mark(st, -1);
  return cudaGetLastError();
}
cudaError_t launch_desc_gather(cudaStream_t st, const DescUnit *units, int n_units, const DescGlobals &X) {
  if (n_units <= 0) return cudaSuccess;
  prof_mark(st, K_DV_LOGIC);
  k_desc_gather<<<2 * n_units, DESC_THREADS, 0, st>>>(units, X);
  prof_mark(st, -1);
  return cudaGetLastError();
}
cudaError_t launch_desc_orb(cudaStream_t st, const DescUnit *units, int n_units, int max_p, const DescGlobals &X) {
  if (n_units <= 0) return cudaSuccess;
  if (max_p > DESC_MAX_P) return cudaErrorInvalidValue;
  const int bx = (max_p + ORB_WARPS - 1) / ORB_WARPS;
  prof_mark(st, K_DV_LOGIC);
  k_desc_orb<<<dim3(bx > 0 ? bx : 1, 2 * n_units), ORB_WARPS * 32, 0, st>>>(units, X);
  prof_mark(st, -1);
  return cudaGetLastError();
}
cudaError_t launch_desc_detect(cudaStream_t st, const DescUnit *units, int n_units, int max_p, const DescGlobals &X) {
  if (n_units <= 0) return cudaSuccess;
  const size_t dyn = (size_t)max_p * 16;
  cudaError_t e = smem_attr(k_desc_detect, dyn);
  if (e != cudaSuccess) return e;
  prof_mark(st, K_RANSAC);
  k_desc_detect<<<n_units, MATCH_THREADS, dyn, st>>>(units, X, max_p);
  prof_mark(st, -1);
  return cudaGetLastError();
}
cudaError_t launch_desc_temporal(cudaStream_t st, const DescUnit *units, int n_units, int max_p, const DescGlobals &X) {
  if (n_units <= 0) return cudaSuccess;
  const size_t dyn = (size_t)max_p * 16;
  cudaError_t e = smem_attr(k_desc_temporal, dyn);
  if (e != cudaSuccess) return e;
  prof_mark(st, K_RANSAC);
  k_desc_temporal<<<dim3(2, n_units), MATCH_THREADS, dyn, st>>>(units, X, max_p);
  prof_mark(st, -1);
  return cudaGetLastError();
}
cudaError_t launch_desc_merge(cudaStream_t st, const DescUnit *units, int n_units, int max_p, const DescGlobals &X) {
  if (n_units <= 0) return cudaSuccess;
  const size_t dyn = (size_t)max_p * 8;
  cudaError_t e = smem_attr(k_desc_merge, dyn);
  if (e != cudaSuccess) return e;
  prof_mark(st, K_DV_LOGIC);
  k_desc_merge<<<n_units, DESC_THREADS, dyn, st>>>(units, X, max_p);
  prof_mark(st, -1);
  return cudaGetLastError();
}

}  // namespace uvio
