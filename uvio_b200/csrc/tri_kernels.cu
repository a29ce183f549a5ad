// sm_100a kernel of the batched feature triangulation: one thread per feature, fp64, 64-thread blocks so
// that a few hundred features spread over many SMs.  Each thread walks its feature's measurement list (a CSR row range of the
// feature store's query result, read in place) a handful of times; the clone poses (<= a few dozen) are staged in shared memory.
#include "tri.h"

namespace uvio {

constexpr int TRI_THREADS = 64;
constexpr int TRI_MAX_SMEM_CLONES = 256;

__global__ void __launch_bounds__(TRI_THREADS) k_triangulate(TriFeatures F, const TriClone *clones, int nc, TriOptions opt, TriResult *out) {
  __shared__ TriClone sc[TRI_MAX_SMEM_CLONES];
  const bool staged = nc <= TRI_MAX_SMEM_CLONES;
  if (staged) {
    for (int i = threadIdx.x; i < nc * (int)(sizeof(TriClone) / 8); i += blockDim.x) ((double *)sc)[i] = ((const double *)clones)[i];
    __syncthreads();
  }
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= F.n_features) return;
  TriResult r;
  tri_feature(F, f, staged ? sc : clones, nc, opt, r);
  out[f] = r;
}

cudaError_t launch_triangulate(cudaStream_t st, TriFeatures F, const TriClone *clones_dev, int n_clones, TriOptions opt, TriResult *out_dev) {
  if (F.n_features <= 0) return cudaSuccess;
  k_triangulate<<<(F.n_features + TRI_THREADS - 1) / TRI_THREADS, TRI_THREADS, 0, st>>>(F, clones_dev, n_clones, opt, out_dev);
  return cudaGetLastError();
}

}  // namespace uvio
