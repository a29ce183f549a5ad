// Launch wrappers of the feature-track store kernels (featdb_kernels.cu; CPU stand-ins in tests/hostsim/featdb_sim.cpp).
// One thread block per store; `stores` is an array of n descriptors in device memory, every argument array has n entries.
#pragma once
#include <cuda_runtime.h>

#include "featdb_body.h"

namespace uvio {

struct FdbUpdateArgs { const FdbObsIn *obs; int n; double timestamp; };
struct FdbQueryArgs { int kind; double timestamp; unsigned long long key; int remove, skip_deleted; };
struct FdbCleanupArgs { int mode; double timestamp; };
struct FdbMarkArgs { const unsigned long long *keys; int n; int what; unsigned long long new_id; };

cudaError_t launch_fdb_init(cudaStream_t st, const FdbStore *stores, int n);
cudaError_t launch_fdb_update(cudaStream_t st, const FdbStore *stores, const FdbUpdateArgs *args, int n);
cudaError_t launch_fdb_query(cudaStream_t st, const FdbStore *stores, const FdbQueryArgs *args, int n);
cudaError_t launch_fdb_cleanup(cudaStream_t st, const FdbStore *stores, const FdbCleanupArgs *args, int n);
cudaError_t launch_fdb_mark(cudaStream_t st, const FdbStore *stores, const FdbMarkArgs *args, int n);

}  // namespace uvio
