// Host-callable launch wrappers of the descriptor tracker's kernels (desc_kernels.cu; bodies in desc_body.h).
#pragma once
#include <cuda_runtime.h>

#include "desc_body.h"
#include "kernels.h"

namespace uvio {

// per-cell top-K + mask test: one block per (cell, side); sides = 2 * n_units entries, side 1 of a mono unit is skipped
cudaError_t launch_desc_topk(cudaStream_t st, const DescUnit *units_dev, int n_units, int max_cells, const DescGlobals &X);
// cells joined in order -> sub-pixel items: one block per side.  Also zeroes nothing: X.n_sp is re-armed by launch_desc_orb.
cudaError_t launch_desc_gather(cudaStream_t st, const DescUnit *units_dev, int n_units, const DescGlobals &X);
// ORB's border filter and the descriptors of the refined corners: one warp per key point
cudaError_t launch_desc_orb(cudaStream_t st, const DescUnit *units_dev, int n_units, int max_p, const DescGlobals &X);
// stereo: L->R robust match + id assignment; mono: occupancy reject + id assignment.  One block per unit.
cudaError_t launch_desc_detect(cudaStream_t st, const DescUnit *units_dev, int n_units, int max_p, const DescGlobals &X);
// temporal robust matches (last -> new) of both cameras: one block per (camera, unit)
cudaError_t launch_desc_temporal(cudaStream_t st, const DescUnit *units_dev, int n_units, int max_p, const DescGlobals &X);
// id hand-over, observation list, last <- new.  One block per unit.
cudaError_t launch_desc_merge(cudaStream_t st, const DescUnit *units_dev, int n_units, int max_p, const DescGlobals &X);

}  // namespace uvio
