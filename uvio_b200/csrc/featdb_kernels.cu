// sm_100a kernels of the device-side feature-track store.  One 1024-thread block per store: a store
// holds the sliding window of one tracker (~10^4 observations, ~10^3 features), so a block sweeps every array in a few
// coalesced passes and the block barrier is the only synchronisation; many stores (one per tracker of a batch) = many blocks.
// These are HBM / L2 byte-moving passes: per query 20 B per stored observation read + 28 B per returned observation written.
#include "featdb.h"

namespace uvio {

constexpr int FDB_THREADS = 1024;

__global__ void __launch_bounds__(FDB_THREADS) k_fdb_init(const FdbStore *stores) {
  const FdbStore &S = stores[blockIdx.x];
  for (int s = threadIdx.x; s < S.cap_hash; s += blockDim.x) { S.h_key[s] = FDB_EMPTY; S.h_val[s] = -1; }
  if (threadIdx.x < 16) S.cnt[threadIdx.x] = 0;
  if (threadIdx.x == 0) S.batch_start[0] = 0;
  __syncthreads();
  fdb_publish(GroupBlock(), S);
}
__global__ void __launch_bounds__(FDB_THREADS) k_fdb_update(const FdbStore *stores, const FdbUpdateArgs *args) {
  const FdbUpdateArgs a = args[blockIdx.x];
  fdb_update(GroupBlock(), stores[blockIdx.x], a.obs, a.n, a.timestamp);
  fdb_publish(GroupBlock(), stores[blockIdx.x]);
}
__global__ void __launch_bounds__(FDB_THREADS) k_fdb_query(const FdbStore *stores, const FdbQueryArgs *args) {
  const FdbQueryArgs a = args[blockIdx.x];
  fdb_query(GroupBlock(), stores[blockIdx.x], a.kind, a.timestamp, a.key, a.remove, a.skip_deleted);
  fdb_publish(GroupBlock(), stores[blockIdx.x]);
}
__global__ void __launch_bounds__(FDB_THREADS) k_fdb_cleanup(const FdbStore *stores, const FdbCleanupArgs *args) {
  const FdbCleanupArgs a = args[blockIdx.x];
  fdb_cleanup(GroupBlock(), stores[blockIdx.x], a.mode, a.timestamp);
  fdb_publish(GroupBlock(), stores[blockIdx.x]);
}
__global__ void __launch_bounds__(FDB_THREADS) k_fdb_mark(const FdbStore *stores, const FdbMarkArgs *args) {
  const FdbMarkArgs a = args[blockIdx.x];
  fdb_mark(GroupBlock(), stores[blockIdx.x], a.keys, a.n, a.what, a.new_id);
  fdb_publish(GroupBlock(), stores[blockIdx.x]);
}

cudaError_t launch_fdb_init(cudaStream_t st, const FdbStore *stores, int n) {
  if (n <= 0) return cudaSuccess;
  k_fdb_init<<<n, FDB_THREADS, 0, st>>>(stores);
  return cudaGetLastError();
}
cudaError_t launch_fdb_update(cudaStream_t st, const FdbStore *stores, const FdbUpdateArgs *args, int n) {
  if (n <= 0) return cudaSuccess;
  k_fdb_update<<<n, FDB_THREADS, 0, st>>>(stores, args);
  return cudaGetLastError();
}
cudaError_t launch_fdb_query(cudaStream_t st, const FdbStore *stores, const FdbQueryArgs *args, int n) {
  if (n <= 0) return cudaSuccess;
  k_fdb_query<<<n, FDB_THREADS, 0, st>>>(stores, args);
  return cudaGetLastError();
}
cudaError_t launch_fdb_cleanup(cudaStream_t st, const FdbStore *stores, const FdbCleanupArgs *args, int n) {
  if (n <= 0) return cudaSuccess;
  k_fdb_cleanup<<<n, FDB_THREADS, 0, st>>>(stores, args);
  return cudaGetLastError();
}
cudaError_t launch_fdb_mark(cudaStream_t st, const FdbStore *stores, const FdbMarkArgs *args, int n) {
  if (n <= 0) return cudaSuccess;
  k_fdb_mark<<<n, FDB_THREADS, 0, st>>>(stores, args);
  return cudaGetLastError();
}

}  // namespace uvio
