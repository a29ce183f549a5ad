// Launch wrapper of the batched triangulation kernel (tri_kernels.cu; CPU stand-in in tests/hostsim/featdb_sim.cpp).
#pragma once
#include <cuda_runtime.h>

#include "tri_body.h"

namespace uvio {
cudaError_t launch_triangulate(cudaStream_t st, TriFeatures F, const TriClone *clones_dev, int n_clones, TriOptions opt, TriResult *out_dev);
}
