// C ABI of the device-side feature-track store (include/uvio_featdb.h) over the kernels of featdb_kernels.cu.
// Host side = memory management and argument marshalling only; every decision about features and measurements is taken by
// the kernels (featdb_body.h).  Calls are synchronous: when one returns, its effect is in HBM and its counts are on the host.
#include "../../include/uvio_featdb.h"

#include <string.h>

#include <algorithm>
#include <mutex>
#include <new>
#include <functional>
#include <set>
#include <utility>
#include <vector>

#include "featdb.h"

using namespace uvio;

struct uvio_featdb {
  int device = 0;
  cudaStream_t stream = nullptr;
  std::mutex mtx;
  FdbStore S;                    // host copy of the descriptor (device pointers)
  FdbStore *S_dev = nullptr;
  void *slab = nullptr;
  int *cnt_h = nullptr;          // mapped: the kernels' copy of cnt[16] (cnt[8..10] = the last query's counts)
  FdbObsIn *in_h = nullptr, *in_d = nullptr;   // staging of one update batch
  unsigned long long *keys_h = nullptr, *keys_d = nullptr;
  void *args_h = nullptr, *args_d = nullptr;   // mapped: one argument struct
  int cap_in = 0;
  std::vector<UvioObs> scratch;
};

namespace {

template <class T>
T *carve(char *&p, size_t n) {
  T *r = (T *)p;
  p += (n * sizeof(T) + 255) / 256 * 256;
  return r;
}
template <class T>
size_t need(size_t n) { return (n * sizeof(T) + 255) / 256 * 256; }

int pow2_at_least(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

#define FDB_TRY(expr)                            \
  do {                                           \
    if ((expr) != cudaSuccess) return UVIO_E_CUDA; \
  } while (0)

int sync_status(uvio_featdb *db) {
  if (cudaStreamSynchronize(db->stream) != cudaSuccess) return UVIO_E_CUDA;
  return db->cnt_h[3] == 0 ? UVIO_OK : UVIO_E_SIZE;
}

int run_update(uvio_featdb *db, double timestamp, const UvioObs *obs, int n) {
  if (n <= 0) return UVIO_OK;
  if (n > db->cap_in) return UVIO_E_SIZE;
  for (int i = 0; i < n; i++) db->in_h[i] = FdbObsIn{(unsigned long long)obs[i].id, (unsigned long long)obs[i].cam_id, obs[i].u, obs[i].v, obs[i].un, obs[i].vn};
  FDB_TRY(cudaMemcpyAsync(db->in_d, db->in_h, sizeof(FdbObsIn) * (size_t)n, cudaMemcpyHostToDevice, db->stream));
  FdbUpdateArgs *a = (FdbUpdateArgs *)db->args_h;
  *a = FdbUpdateArgs{db->in_d, n, timestamp};
  FDB_TRY(launch_fdb_update(db->stream, db->S_dev, (const FdbUpdateArgs *)db->args_d, 1));
  return sync_status(db);
}

}  // namespace

extern "C" {

int uvio_featdb_create(int cuda_device, int max_observations, int max_features, uvio_featdb **out) {
  if (!out || max_observations < 16 || max_features < 16 || max_observations > (1 << 26) || max_features > (1 << 24)) return UVIO_E_ARG;
  *out = nullptr;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || cuda_device < 0 || cuda_device >= count) return UVIO_E_CUDA;   // no CPU fallback
  if (cudaSetDevice(cuda_device) != cudaSuccess) return UVIO_E_CUDA;
  uvio_featdb *db = new (std::nothrow) uvio_featdb;
  if (!db) return UVIO_E_ARG;
  db->device = cuda_device;
  FdbStore &S = db->S;
  memset(&S, 0, sizeof(S));
  S.cap_obs = max_observations;
  S.cap_feat = max_features;
  S.cap_hash = pow2_at_least(2 * max_features);
  S.cap_batch = std::max(64, max_observations / 8);
  const size_t nq = (size_t)S.cap_feat * FDB_MAXCAM + 2;
  const size_t nc = std::max(std::max((size_t)S.cap_obs, nq), (size_t)S.cap_batch) + 2;
  size_t bytes = need<unsigned long long>(S.cap_obs) + need<double>(S.cap_obs) + need<FdbMeas>(S.cap_obs) + 2 * need<int>(S.cap_obs) +
                 need<int>(S.cap_batch + 2) + 2 * need<unsigned long long>(S.cap_feat) + 2 * need<uint8_t>(S.cap_feat) +
                 need<unsigned long long>(S.cap_hash) + need<int>(S.cap_hash) + 4 * need<int>(nc) + 2 * need<uint8_t>(S.cap_feat) +
                 need<int>(S.cap_feat + 2) + need<int>(2048) + 2 * need<int>(nc) + need<unsigned long long>(nc) + need<double>(nc) +
                 need<FdbMeas>(nc) + 2 * need<int>(nc) + 2 * need<unsigned long long>(S.cap_feat) + need<uint8_t>(S.cap_feat) +
                 need<int>(S.cap_feat + 2) + 2 * need<int>(nq) + need<double>(S.cap_obs) + 2 * need<float>(2 * (size_t)S.cap_obs) + 4096;
  auto fail = [&](int code) {
    uvio_featdb_destroy(db);
    return code;
  };
  if (cudaMalloc(&db->slab, bytes) != cudaSuccess) return fail(UVIO_E_CUDA);
  if (cudaMemset(db->slab, 0, bytes) != cudaSuccess) return fail(UVIO_E_CUDA);
  char *p = (char *)db->slab;
  S.o_key = carve<unsigned long long>(p, S.cap_obs);
  S.o_t = carve<double>(p, S.cap_obs);
  S.o_m = carve<FdbMeas>(p, S.cap_obs);
  S.o_cam = carve<int>(p, S.cap_obs);
  S.o_feat = carve<int>(p, S.cap_obs);
  S.batch_start = carve<int>(p, S.cap_batch + 2);
  S.f_key = carve<unsigned long long>(p, S.cap_feat);
  S.f_featid = carve<unsigned long long>(p, S.cap_feat);
  S.f_del = carve<uint8_t>(p, S.cap_feat);
  S.f_alive = carve<uint8_t>(p, S.cap_feat);
  S.h_key = carve<unsigned long long>(p, S.cap_hash);
  S.h_val = carve<int>(p, S.cap_hash);
  S.q_cnt = carve<int>(p, nc);
  S.q_first = carve<int>(p, nc);
  S.q_last = carve<int>(p, nc);
  S.q_cursor = carve<int>(p, nc);
  S.q_hit = carve<uint8_t>(p, S.cap_feat);
  S.q_sel = carve<uint8_t>(p, S.cap_feat);
  S.q_featpos = carve<int>(p, S.cap_feat + 2);
  S.q_part = carve<int>(p, 2048);
  S.c_keep = carve<int>(p, nc);
  S.c_pos = carve<int>(p, nc);
  S.c_u64 = carve<unsigned long long>(p, nc);
  S.c_f64 = carve<double>(p, nc);
  S.c_m = carve<FdbMeas>(p, nc);
  S.c_i32a = carve<int>(p, nc);
  S.c_i32b = carve<int>(p, nc);
  S.r_featid = carve<unsigned long long>(p, S.cap_feat);
  S.r_key = carve<unsigned long long>(p, S.cap_feat);
  S.r_del = carve<uint8_t>(p, S.cap_feat);
  S.r_row0 = carve<int>(p, S.cap_feat + 2);
  S.r_row_cam = carve<int>(p, nq);
  S.r_off = carve<int>(p, nq);
  S.r_t = carve<double>(p, S.cap_obs);
  S.r_uv = carve<float>(p, 2 * (size_t)S.cap_obs);
  S.r_uvn = carve<float>(p, 2 * (size_t)S.cap_obs);
  S.cnt = carve<int>(p, 16);
  S.r_cnt = S.cnt + 8;
  int *cnt_d = nullptr;
  if (cudaHostAlloc((void **)&db->cnt_h, 16 * sizeof(int), cudaHostAllocMapped) != cudaSuccess ||
      cudaHostGetDevicePointer((void **)&cnt_d, db->cnt_h, 0) != cudaSuccess)
    return fail(UVIO_E_CUDA);
  S.cnt_pub = cnt_d;
  if ((size_t)(p - (char *)db->slab) > bytes) return fail(UVIO_E_ARG);
  db->cap_in = S.cap_obs;
  if (cudaHostAlloc((void **)&db->in_h, sizeof(FdbObsIn) * (size_t)db->cap_in, cudaHostAllocDefault) != cudaSuccess ||
      cudaMalloc((void **)&db->in_d, sizeof(FdbObsIn) * (size_t)db->cap_in) != cudaSuccess ||
      cudaHostAlloc((void **)&db->keys_h, sizeof(unsigned long long) * (size_t)S.cap_feat, cudaHostAllocDefault) != cudaSuccess ||
      cudaMalloc((void **)&db->keys_d, sizeof(unsigned long long) * (size_t)S.cap_feat) != cudaSuccess ||
      cudaHostAlloc(&db->args_h, 256, cudaHostAllocMapped) != cudaSuccess || cudaHostGetDevicePointer(&db->args_d, db->args_h, 0) != cudaSuccess ||
      cudaMalloc((void **)&db->S_dev, sizeof(FdbStore)) != cudaSuccess ||
      cudaStreamCreateWithFlags(&db->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaMemcpy(db->S_dev, &S, sizeof(FdbStore), cudaMemcpyHostToDevice) != cudaSuccess ||
      launch_fdb_init(db->stream, db->S_dev, 1) != cudaSuccess || cudaStreamSynchronize(db->stream) != cudaSuccess)
    return fail(UVIO_E_CUDA);
  *out = db;
  return UVIO_OK;
}

void uvio_featdb_destroy(uvio_featdb *db) {
  if (!db) return;
  cudaSetDevice(db->device);
  if (db->stream) {
    cudaStreamSynchronize(db->stream);
    cudaStreamDestroy(db->stream);
  }
  if (db->slab) cudaFree(db->slab);
  if (db->cnt_h) cudaFreeHost(db->cnt_h);
  if (db->in_h) cudaFreeHost(db->in_h);
  if (db->in_d) cudaFree(db->in_d);
  if (db->keys_h) cudaFreeHost(db->keys_h);
  if (db->keys_d) cudaFree(db->keys_d);
  if (db->args_h) cudaFreeHost(db->args_h);
  if (db->S_dev) cudaFree(db->S_dev);
  delete db;
}

int uvio_featdb_update(uvio_featdb *db, double timestamp, const UvioObs *obs, int n) {
  if (!db || n < 0 || (n > 0 && !obs)) return UVIO_E_ARG;
  std::lock_guard<std::mutex> lk(db->mtx);
  if (cudaSetDevice(db->device) != cudaSuccess) return UVIO_E_CUDA;
  for (int i = 0; i < n; i++)
    if (obs[i].cam_id >= (size_t)FDB_MAXCAM || (unsigned long long)obs[i].id == FDB_EMPTY) return UVIO_E_ARG;
  // a batch may hold one observation per (id, camera): a list that repeats one (never the tracker's) is cut there
  std::set<std::pair<unsigned long long, unsigned long long>> seen;
  int b = 0;
  for (int i = 0; i < n; i++) {
    const std::pair<unsigned long long, unsigned long long> k((unsigned long long)obs[i].id, (unsigned long long)obs[i].cam_id);
    if (!seen.insert(k).second) {
      int rc = run_update(db, timestamp, obs + b, i - b);
      if (rc != UVIO_OK) return rc;
      b = i;
      seen.clear();
      seen.insert(k);
    }
  }
  return run_update(db, timestamp, obs + b, n - b);
}

int uvio_featdb_update_from_tracker(uvio_featdb *db, uvio_klt *tracker, double timestamp) {
  if (!db || !tracker) return UVIO_E_ARG;
  int n = 0;
  int rc = uvio_klt_get_new_observations(tracker, nullptr, 0, &n);
  if (rc != UVIO_OK) return rc;
  std::vector<UvioObs> tmp((size_t)std::max(n, 1));
  rc = uvio_klt_get_new_observations(tracker, tmp.data(), n, &n);
  if (rc != UVIO_OK) return rc;
  return uvio_featdb_update(db, timestamp, tmp.data(), n);
}

int uvio_featdb_query(uvio_featdb *db, int kind, double timestamp, size_t key, int remove, int skip_deleted, int *n_features, int *n_rows,
                      int *n_observations) {
  if (!db || kind < 0 || kind > 4) return UVIO_E_ARG;
  std::lock_guard<std::mutex> lk(db->mtx);
  if (cudaSetDevice(db->device) != cudaSuccess) return UVIO_E_CUDA;
  FdbQueryArgs *a = (FdbQueryArgs *)db->args_h;
  *a = FdbQueryArgs{kind, timestamp, (unsigned long long)key, remove, skip_deleted};
  FDB_TRY(launch_fdb_query(db->stream, db->S_dev, (const FdbQueryArgs *)db->args_d, 1));
  int rc = sync_status(db);
  if (n_features) *n_features = db->cnt_h[8];
  if (n_rows) *n_rows = db->cnt_h[9];
  if (n_observations) *n_observations = db->cnt_h[10];
  return rc;
}

int uvio_featdb_fetch(uvio_featdb *db, size_t *featid, size_t *key, uint8_t *to_delete, int *row0, int *row_cam, int *off, double *t, float *uv,
                      float *uvn) {
  if (!db) return UVIO_E_ARG;
  std::lock_guard<std::mutex> lk(db->mtx);
  if (cudaSetDevice(db->device) != cudaSuccess) return UVIO_E_CUDA;
  const size_t nf = (size_t)db->cnt_h[8], nr = (size_t)db->cnt_h[9], no = (size_t)db->cnt_h[10];
  const FdbStore &S = db->S;
  static_assert(sizeof(size_t) == sizeof(unsigned long long), "size_t is 64 bits on every host this ABI targets");
  if (featid && nf) FDB_TRY(cudaMemcpyAsync(featid, S.r_featid, nf * 8, cudaMemcpyDeviceToHost, db->stream));
  if (key && nf) FDB_TRY(cudaMemcpyAsync(key, S.r_key, nf * 8, cudaMemcpyDeviceToHost, db->stream));
  if (to_delete && nf) FDB_TRY(cudaMemcpyAsync(to_delete, S.r_del, nf, cudaMemcpyDeviceToHost, db->stream));
  if (row0) FDB_TRY(cudaMemcpyAsync(row0, S.r_row0, (nf + 1) * 4, cudaMemcpyDeviceToHost, db->stream));
  if (row_cam && nr) FDB_TRY(cudaMemcpyAsync(row_cam, S.r_row_cam, nr * 4, cudaMemcpyDeviceToHost, db->stream));
  if (off) FDB_TRY(cudaMemcpyAsync(off, S.r_off, (nr + 1) * 4, cudaMemcpyDeviceToHost, db->stream));
  if (t && no) FDB_TRY(cudaMemcpyAsync(t, S.r_t, no * 8, cudaMemcpyDeviceToHost, db->stream));
  if (uv && no) FDB_TRY(cudaMemcpyAsync(uv, S.r_uv, no * 8, cudaMemcpyDeviceToHost, db->stream));
  if (uvn && no) FDB_TRY(cudaMemcpyAsync(uvn, S.r_uvn, no * 8, cudaMemcpyDeviceToHost, db->stream));
  FDB_TRY(cudaStreamSynchronize(db->stream));
  return UVIO_OK;
}

int uvio_featdb_result_device(uvio_featdb *db, const void *ptrs[9]) {
  if (!db || !ptrs) return UVIO_E_ARG;
  const FdbStore &S = db->S;
  const void *p[9] = {S.r_featid, S.r_key, S.r_del, S.r_row0, S.r_row_cam, S.r_off, S.r_t, S.r_uv, S.r_uvn};
  memcpy(ptrs, p, sizeof(p));
  return UVIO_OK;
}

static int mark(uvio_featdb *db, const size_t *keys, int n, int what, size_t new_id) {
  if (!db || n < 0 || (n > 0 && !keys)) return UVIO_E_ARG;
  if (n == 0) return UVIO_OK;
  std::lock_guard<std::mutex> lk(db->mtx);
  if (cudaSetDevice(db->device) != cudaSuccess) return UVIO_E_CUDA;
  for (int b = 0; b < n; b += db->S.cap_feat) {
    const int m = std::min(db->S.cap_feat, n - b);
    for (int i = 0; i < m; i++) db->keys_h[i] = (unsigned long long)keys[b + i];
    FDB_TRY(cudaMemcpyAsync(db->keys_d, db->keys_h, 8 * (size_t)m, cudaMemcpyHostToDevice, db->stream));
    FdbMarkArgs *a = (FdbMarkArgs *)db->args_h;
    *a = FdbMarkArgs{db->keys_d, m, what, (unsigned long long)new_id};
    FDB_TRY(launch_fdb_mark(db->stream, db->S_dev, (const FdbMarkArgs *)db->args_d, 1));
    FDB_TRY(cudaStreamSynchronize(db->stream));
  }
  return UVIO_OK;
}

int uvio_featdb_mark_to_delete(uvio_featdb *db, const size_t *keys, int n) { return mark(db, keys, n, 0, 0); }
int uvio_featdb_change_feat_id(uvio_featdb *db, size_t id_old, size_t id_new) { return mark(db, &id_old, 1, 1, id_new); }

static int cleanup(uvio_featdb *db, int mode, double timestamp) {
  if (!db) return UVIO_E_ARG;
  std::lock_guard<std::mutex> lk(db->mtx);
  if (cudaSetDevice(db->device) != cudaSuccess) return UVIO_E_CUDA;
  FdbCleanupArgs *a = (FdbCleanupArgs *)db->args_h;
  *a = FdbCleanupArgs{mode, timestamp};
  FDB_TRY(launch_fdb_cleanup(db->stream, db->S_dev, (const FdbCleanupArgs *)db->args_d, 1));
  return sync_status(db);
}
int uvio_featdb_cleanup(uvio_featdb *db) { return cleanup(db, 0, 0.0); }
int uvio_featdb_cleanup_measurements(uvio_featdb *db, double timestamp) { return cleanup(db, 1, timestamp); }
int uvio_featdb_cleanup_measurements_exact(uvio_featdb *db, double timestamp) { return cleanup(db, 2, timestamp); }

}  // extern "C"

// ---- batched entry points: the kernels take one block per store, so n stores cost one launch and one synchronisation ----------
namespace {

struct BatchCtx {                       // per device: the launch tables of one batched call (descriptors, arguments, staged input)
  std::mutex mtx;
  cudaStream_t st = nullptr;
  char *host = nullptr, *dev = nullptr;
  size_t cap = 0;
};
BatchCtx g_batch[64];

int batch_reserve(BatchCtx &B, size_t bytes) {
  if (!B.st && cudaStreamCreateWithFlags(&B.st, cudaStreamNonBlocking) != cudaSuccess) return UVIO_E_CUDA;
  if (bytes <= B.cap) return UVIO_OK;
  if (B.host) cudaFreeHost(B.host);
  if (B.dev) cudaFree(B.dev);
  B.host = B.dev = nullptr;
  B.cap = 0;
  const size_t want = bytes * 2;
  if (cudaHostAlloc((void **)&B.host, want, cudaHostAllocDefault) != cudaSuccess || cudaMalloc((void **)&B.dev, want) != cudaSuccess) return UVIO_E_CUDA;
  B.cap = want;
  return UVIO_OK;
}

// locks every store (address order), checks they share a device, returns the device's batch context
int batch_begin(uvio_featdb *const *dbs, int n, std::vector<std::unique_lock<std::mutex>> &locks, BatchCtx **out) {
  if (!dbs || n <= 0) return UVIO_E_ARG;
  for (int i = 0; i < n; i++) {
    if (!dbs[i] || dbs[i]->device != dbs[0]->device) return UVIO_E_ARG;
    for (int k = 0; k < i; k++)
      if (dbs[k] == dbs[i]) return UVIO_E_ARG;
  }
  std::vector<uvio_featdb *> order(dbs, dbs + n);
  std::sort(order.begin(), order.end(), [](uvio_featdb *a, uvio_featdb *b) { return std::less<uvio_featdb *>()(a, b); });
  for (uvio_featdb *d : order) locks.emplace_back(d->mtx);
  const int dev = dbs[0]->device;
  if (dev < 0 || dev >= 64 || cudaSetDevice(dev) != cudaSuccess) return UVIO_E_CUDA;
  *out = &g_batch[dev];
  return UVIO_OK;
}

template <class Args>
int batch_run(BatchCtx &B, uvio_featdb *const *dbs, int n, const std::vector<Args> &args, size_t extra_bytes, const void *extra,
              cudaError_t (*launch)(cudaStream_t, const FdbStore *, const Args *, int), int *status, size_t extra_off) {
  // table layout: [FdbStore x n][Args x n][extra]
  const size_t o_args = (sizeof(FdbStore) * (size_t)n + 255) / 256 * 256;
  FdbStore *hs = (FdbStore *)B.host;
  for (int i = 0; i < n; i++) hs[i] = dbs[i]->S;
  memcpy(B.host + o_args, args.data(), sizeof(Args) * (size_t)n);
  (void)extra;
  const size_t total = extra_off + extra_bytes;
  FDB_TRY(cudaMemcpyAsync(B.dev, B.host, total, cudaMemcpyHostToDevice, B.st));
  FDB_TRY(launch(B.st, (const FdbStore *)B.dev, (const Args *)(B.dev + o_args), n));
  FDB_TRY(cudaStreamSynchronize(B.st));
  int worst = UVIO_OK;
  for (int i = 0; i < n; i++) {
    const int rc = dbs[i]->cnt_h[3] == 0 ? UVIO_OK : UVIO_E_SIZE;
    if (status) status[i] = rc;
    worst = std::max(worst, rc);
  }
  return worst;
}

size_t batch_extra_off(int n, size_t args_size) {
  const size_t o_args = (sizeof(FdbStore) * (size_t)n + 255) / 256 * 256;
  return (o_args + args_size * (size_t)n + 255) / 256 * 256;
}

}  // namespace

extern "C" {

int uvio_featdb_update_batch(uvio_featdb *const *dbs, int n, const double *timestamps, const UvioObs *obs, const int *counts, int *status) {
  if (!timestamps || !counts) return UVIO_E_ARG;
  std::vector<std::unique_lock<std::mutex>> locks;
  BatchCtx *Bp = nullptr;
  int rc = batch_begin(dbs, n, locks, &Bp);
  if (rc != UVIO_OK) return rc;
  BatchCtx &B = *Bp;
  std::lock_guard<std::mutex> lb(B.mtx);
  size_t total = 0;
  for (int i = 0; i < n; i++) {
    if (counts[i] < 0 || counts[i] > dbs[i]->cap_in) return UVIO_E_SIZE;
    total += (size_t)counts[i];
  }
  if (total && !obs) return UVIO_E_ARG;
  const size_t off = batch_extra_off(n, sizeof(FdbUpdateArgs));
  rc = batch_reserve(B, off + sizeof(FdbObsIn) * total + 256);
  if (rc != UVIO_OK) return rc;
  FdbObsIn *in = (FdbObsIn *)(B.host + off);
  std::vector<FdbUpdateArgs> args(n);
  // a batch may hold one observation per (id, camera) and store (the tracker's lists never repeat one): checked with a small
  // open-addressing table per store; a list that does is refused here -- feed it through uvio_featdb_update, which cuts it
  std::vector<unsigned long long> tab;
  size_t at = 0;
  for (int i = 0; i < n; i++) {
    const int m = counts[i];
    size_t cap = 16;
    while (cap < 4 * (size_t)m) cap <<= 1;
    tab.assign(cap, ~0ull);
    for (int k = 0; k < m; k++) {
      const UvioObs &o = obs[at + k];
      if (o.cam_id >= (size_t)FDB_MAXCAM || (unsigned long long)o.id == FDB_EMPTY) return UVIO_E_ARG;
      const unsigned long long key = ((unsigned long long)o.id << 2) | (unsigned long long)o.cam_id;
      size_t h = (size_t)((key * 0x9E3779B97F4A7C15ull) >> 20) & (cap - 1);
      while (tab[h] != ~0ull) {
        if (tab[h] == key) return UVIO_E_ARG;
        h = (h + 1) & (cap - 1);
      }
      tab[h] = key;
      in[at + k] = FdbObsIn{(unsigned long long)o.id, (unsigned long long)o.cam_id, o.u, o.v, o.un, o.vn};
    }
    args[i] = FdbUpdateArgs{(const FdbObsIn *)(B.dev + off) + at, m, timestamps[i]};
    at += (size_t)m;
  }
  return batch_run<FdbUpdateArgs>(B, dbs, n, args, sizeof(FdbObsIn) * total, nullptr, launch_fdb_update, status, off);
}

int uvio_featdb_update_from_trackers(uvio_featdb *const *dbs, uvio_klt *const *trackers, int n, const double *timestamps, int *status) {
  if (!dbs || !trackers || n <= 0) return UVIO_E_ARG;
  std::vector<int> counts(n, 0);
  std::vector<UvioObs> all;
  std::vector<UvioObs> tmp;
  for (int i = 0; i < n; i++) {
    int m = 0;
    int rc = uvio_klt_get_new_observations(trackers[i], nullptr, 0, &m);
    if (rc != UVIO_OK && rc != UVIO_E_SIZE) return rc;
    tmp.resize((size_t)std::max(m, 1));
    rc = uvio_klt_get_new_observations(trackers[i], tmp.data(), m, &m);
    if (rc != UVIO_OK) return rc;
    counts[i] = m;
    all.insert(all.end(), tmp.begin(), tmp.begin() + m);
  }
  return uvio_featdb_update_batch(dbs, n, timestamps, all.data(), counts.data(), status);
}

int uvio_featdb_update_from_desc_trackers(uvio_featdb *const *dbs, uvio_desc *const *trackers, int n, const double *timestamps, int *status) {
  if (!dbs || !trackers || n <= 0) return UVIO_E_ARG;
  std::vector<int> counts(n, 0);
  std::vector<UvioObs> all, tmp;
  for (int i = 0; i < n; i++) {
    int m = 0;
    int rc = uvio_desc_get_new_observations(trackers[i], nullptr, 0, &m);
    if (rc != UVIO_OK && rc != UVIO_E_SIZE) return rc;
    tmp.resize((size_t)std::max(m, 1));
    rc = uvio_desc_get_new_observations(trackers[i], tmp.data(), m, &m);
    if (rc != UVIO_OK) return rc;
    counts[i] = m;
    all.insert(all.end(), tmp.begin(), tmp.begin() + m);
  }
  return uvio_featdb_update_batch(dbs, n, timestamps, all.data(), counts.data(), status);
}

int uvio_featdb_query_batch(uvio_featdb *const *dbs, int n, int kind, const double *timestamps, const size_t *keys, int remove, int skip_deleted,
                            int *n_features, int *n_rows, int *n_observations, int *status) {
  if (kind < 0 || kind > 4 || !timestamps || (kind == UVIO_FDB_GET_FEATURE && !keys)) return UVIO_E_ARG;
  std::vector<std::unique_lock<std::mutex>> locks;
  BatchCtx *Bp = nullptr;
  int rc = batch_begin(dbs, n, locks, &Bp);
  if (rc != UVIO_OK) return rc;
  BatchCtx &B = *Bp;
  std::lock_guard<std::mutex> lb(B.mtx);
  const size_t off = batch_extra_off(n, sizeof(FdbQueryArgs));
  rc = batch_reserve(B, off + 256);
  if (rc != UVIO_OK) return rc;
  std::vector<FdbQueryArgs> args(n);
  for (int i = 0; i < n; i++) args[i] = FdbQueryArgs{kind, timestamps[i], keys ? (unsigned long long)keys[i] : 0ull, remove, skip_deleted};
  rc = batch_run<FdbQueryArgs>(B, dbs, n, args, 0, nullptr, launch_fdb_query, status, off);
  for (int i = 0; i < n; i++) {
    if (n_features) n_features[i] = dbs[i]->cnt_h[8];
    if (n_rows) n_rows[i] = dbs[i]->cnt_h[9];
    if (n_observations) n_observations[i] = dbs[i]->cnt_h[10];
  }
  return rc;
}

int uvio_featdb_cleanup_batch(uvio_featdb *const *dbs, int n, int mode, const double *timestamps, int *status) {
  if (mode < 0 || mode > 2 || (mode != 0 && !timestamps)) return UVIO_E_ARG;
  std::vector<std::unique_lock<std::mutex>> locks;
  BatchCtx *Bp = nullptr;
  int rc = batch_begin(dbs, n, locks, &Bp);
  if (rc != UVIO_OK) return rc;
  BatchCtx &B = *Bp;
  std::lock_guard<std::mutex> lb(B.mtx);
  const size_t off = batch_extra_off(n, sizeof(FdbCleanupArgs));
  rc = batch_reserve(B, off + 256);
  if (rc != UVIO_OK) return rc;
  std::vector<FdbCleanupArgs> args(n);
  for (int i = 0; i < n; i++) args[i] = FdbCleanupArgs{mode, timestamps ? timestamps[i] : 0.0};
  return batch_run<FdbCleanupArgs>(B, dbs, n, args, 0, nullptr, launch_fdb_cleanup, status, off);
}

int uvio_featdb_size(uvio_featdb *db, int *n_features, int *n_observations) {
  if (!db) return UVIO_E_ARG;
  std::lock_guard<std::mutex> lk(db->mtx);
  if (n_features) *n_features = db->cnt_h[4];
  if (n_observations) *n_observations = db->cnt_h[0];
  return UVIO_OK;
}

}  // extern "C"

// ---- batched triangulation (include/uvio_tri.h) --------------------------------------------------------------------------
#include "../../include/uvio_tri.h"
#include "tri.h"

static_assert(sizeof(UvioClonePose) == sizeof(TriClone) && sizeof(UvioTriOptions) == sizeof(TriOptions) && sizeof(UvioTriResult) == sizeof(TriResult),
              "the ABI structs and the kernel's structs are the same bytes");

namespace {
struct TriScratch {           // per call: clone table + results in device memory
  TriClone *clones = nullptr;
  TriResult *out = nullptr;
  ~TriScratch() {
    if (clones) cudaFree(clones);
    if (out) cudaFree(out);
  }
};
int run_triangulate(cudaStream_t st, TriFeatures F, const UvioClonePose *clones, int n_clones, const UvioTriOptions *opt, UvioTriResult *out) {
  if (F.n_features == 0) return UVIO_OK;
  TriScratch s;
  FDB_TRY(cudaMalloc((void **)&s.clones, sizeof(TriClone) * (size_t)std::max(n_clones, 1)));
  FDB_TRY(cudaMalloc((void **)&s.out, sizeof(TriResult) * (size_t)F.n_features));
  if (n_clones) FDB_TRY(cudaMemcpyAsync(s.clones, clones, sizeof(TriClone) * (size_t)n_clones, cudaMemcpyHostToDevice, st));
  TriOptions o;
  memcpy(&o, opt, sizeof(o));
  FDB_TRY(launch_triangulate(st, F, s.clones, n_clones, o, s.out));
  FDB_TRY(cudaMemcpyAsync(out, s.out, sizeof(TriResult) * (size_t)F.n_features, cudaMemcpyDeviceToHost, st));
  FDB_TRY(cudaStreamSynchronize(st));
  for (int f = 0; f < F.n_features; f++)
    if (out[f].ok_triangulation < 0 || out[f].ok_refine < 0) return UVIO_E_ARG;   // a measurement without a clone pose (.at() throws in the reference)
  return UVIO_OK;
}
}  // namespace

extern "C" {

void uvio_tri_default_options(UvioTriOptions *o) {   // FeatureInitializerOptions.h:35-69
  if (!o) return;
  o->triangulate_1d = 0;
  o->refine_features = 1;
  o->max_runs = 5;
  o->init_lamda = 1e-3;
  o->max_lamda = 1e10;
  o->min_dx = 1e-6;
  o->min_dcost = 1e-6;
  o->lam_mult = 10;
  o->min_dist = 0.10;
  o->max_dist = 60;
  o->max_baseline = 40;
  o->max_cond_number = 10000;
}

int uvio_tri_triangulate_featdb(uvio_featdb *db, const UvioClonePose *clones, int n_clones, const UvioTriOptions *opt, UvioTriResult *out) {
  if (!db || n_clones < 0 || (n_clones > 0 && !clones) || !opt || !out) return UVIO_E_ARG;
  std::lock_guard<std::mutex> lk(db->mtx);
  if (cudaSetDevice(db->device) != cudaSuccess) return UVIO_E_CUDA;
  const FdbStore &S = db->S;
  TriFeatures F{db->cnt_h[8], S.r_row0, S.r_row_cam, S.r_off, S.r_t, S.r_uvn};
  return run_triangulate(db->stream, F, clones, n_clones, opt, out);
}

int uvio_tri_triangulate(int cuda_device, int n_features, const int *row0, const int *row_cam, const int *off, const double *t, const float *uvn,
                         const UvioClonePose *clones, int n_clones, const UvioTriOptions *opt, UvioTriResult *out) {
  if (n_features < 0 || n_clones < 0 || (n_clones > 0 && !clones) || !opt) return UVIO_E_ARG;
  if (n_features == 0) return UVIO_OK;
  if (!row0 || !row_cam || !off || !t || !uvn || !out) return UVIO_E_ARG;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || cuda_device < 0 || cuda_device >= count) return UVIO_E_CUDA;   // no CPU fallback
  if (cudaSetDevice(cuda_device) != cudaSuccess) return UVIO_E_CUDA;
  const int nr = row0[n_features], no = off[nr];
  if (nr < 0 || no < 0) return UVIO_E_ARG;
  char *slab = nullptr;
  const size_t b0 = need<int>(n_features + 1), b1 = need<int>(nr + 1), b2 = need<int>(nr + 1), b3 = need<double>(no + 1), b4 = need<float>(2 * (size_t)no + 2);
  FDB_TRY(cudaMalloc((void **)&slab, b0 + b1 + b2 + b3 + b4));
  struct Free { char *p; ~Free() { cudaFree(p); } } guard{slab};
  char *p = slab;
  int *d_row0 = carve<int>(p, n_features + 1), *d_cam = carve<int>(p, nr + 1), *d_off = carve<int>(p, nr + 1);
  double *d_t = carve<double>(p, no + 1);
  float *d_uvn = carve<float>(p, 2 * (size_t)no + 2);
  cudaStream_t st = nullptr;   // the default stream: this entry point is a synchronous convenience
  FDB_TRY(cudaMemcpyAsync(d_row0, row0, 4 * (size_t)(n_features + 1), cudaMemcpyHostToDevice, st));
  if (nr) FDB_TRY(cudaMemcpyAsync(d_cam, row_cam, 4 * (size_t)nr, cudaMemcpyHostToDevice, st));
  FDB_TRY(cudaMemcpyAsync(d_off, off, 4 * (size_t)(nr + 1), cudaMemcpyHostToDevice, st));
  if (no) FDB_TRY(cudaMemcpyAsync(d_t, t, 8 * (size_t)no, cudaMemcpyHostToDevice, st));
  if (no) FDB_TRY(cudaMemcpyAsync(d_uvn, uvn, 8 * (size_t)no, cudaMemcpyHostToDevice, st));
  TriFeatures F{n_features, d_row0, d_cam, d_off, d_t, d_uvn};
  return run_triangulate(st, F, clones, n_clones, opt, out);
}

}  // extern "C"
