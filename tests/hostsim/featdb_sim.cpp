// TEST SCAFFOLDING ONLY — see tests/hostsim/cuda_runtime.h.  CPU stand-ins for the launch wrappers of featdb_kernels.cu: the
// same featdb_body.h code with one lane (GroupSeq), one "block" per store.
#include "../../uvio_b200/csrc/featdb.h"

namespace uvio {

cudaError_t launch_fdb_init(cudaStream_t, const FdbStore *stores, int n) {
  for (int k = 0; k < n; k++) {
    const FdbStore &S = stores[k];
    for (int s = 0; s < S.cap_hash; s++) { S.h_key[s] = FDB_EMPTY; S.h_val[s] = -1; }
    for (int i = 0; i < 16; i++) S.cnt[i] = 0;
    S.batch_start[0] = 0;
    fdb_publish(GroupSeq(), S);
  }
  return 0;
}
cudaError_t launch_fdb_update(cudaStream_t, const FdbStore *stores, const FdbUpdateArgs *a, int n) {
  for (int k = 0; k < n; k++) { fdb_update(GroupSeq(), stores[k], a[k].obs, a[k].n, a[k].timestamp); fdb_publish(GroupSeq(), stores[k]); }
  return 0;
}
cudaError_t launch_fdb_query(cudaStream_t, const FdbStore *stores, const FdbQueryArgs *a, int n) {
  for (int k = 0; k < n; k++) { fdb_query(GroupSeq(), stores[k], a[k].kind, a[k].timestamp, a[k].key, a[k].remove, a[k].skip_deleted); fdb_publish(GroupSeq(), stores[k]); }
  return 0;
}
cudaError_t launch_fdb_cleanup(cudaStream_t, const FdbStore *stores, const FdbCleanupArgs *a, int n) {
  for (int k = 0; k < n; k++) { fdb_cleanup(GroupSeq(), stores[k], a[k].mode, a[k].timestamp); fdb_publish(GroupSeq(), stores[k]); }
  return 0;
}
cudaError_t launch_fdb_mark(cudaStream_t, const FdbStore *stores, const FdbMarkArgs *a, int n) {
  for (int k = 0; k < n; k++) { fdb_mark(GroupSeq(), stores[k], a[k].keys, a[k].n, a[k].what, a[k].new_id); fdb_publish(GroupSeq(), stores[k]); }
  return 0;
}

}  // namespace uvio

#include "../../uvio_b200/csrc/tri.h"
namespace uvio {
cudaError_t launch_triangulate(cudaStream_t, TriFeatures F, const TriClone *clones, int n_clones, TriOptions opt, TriResult *out) {
  for (int f = 0; f < F.n_features; f++) tri_feature(F, f, clones, n_clones, opt, out[f]);
  return 0;
}
}  // namespace uvio
