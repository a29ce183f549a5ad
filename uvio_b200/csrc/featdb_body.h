// Device-side feature-track store: the bodies of its three kernels (update, query, cleanup), written against the execution
// group abstraction of simt.h -- one thread block per store on the device (GroupBlock), one lane in tests/hostsim (GroupSeq).
//
// What it restates (ov_core/src/feat/FeatureDatabase.cpp, ov_core/src/feat/Feature.cpp):
//   update_feature                 FeatureDatabase.cpp:59-85     fdb_update
//   get_feature                    :36-57                        fdb_query kind 3
//   features_not_containing_newer  :87-126                       fdb_query kind 0
//   features_containing_older      :128-167                      fdb_query kind 1
//   features_containing            :169-209                      fdb_query kind 2
//   cleanup                        :211-224                      fdb_cleanup mode 0
//   cleanup_measurements           :226-243 (Feature.cpp:84-106) fdb_cleanup mode 1
//   cleanup_measurements_exact     :245-263 (Feature.cpp:60-82)  fdb_cleanup mode 2
//
// Layout in HBM (per store): the observations as a structure-of-arrays LOG in arrival order, cut into BATCHES (one per
// update call; a batch holds at most one observation per (feature, camera), which is what lets a block place the
// observations of a feature in time order without atomics: batches are walked in order, lanes take the observations of a
// batch); the features as arrays in INSERTION order (key, featid, to_delete, alive); an open-addressing hash key -> feature
// index that is rebuilt whenever features are dropped, so it never holds tombstones.
#pragma once
#include "simt.h"

namespace uvio {

constexpr int FDB_MAXCAM = 4;
constexpr unsigned long long FDB_EMPTY = ~0ull;      // hash slot free (no feature may use this key)
constexpr int FDB_PROVISIONAL = 0x40000000;          // h_val >= this while an update is electing the creator of a new feature

struct FdbObsIn {            // one update_feature call (== UvioObs)
  unsigned long long id, cam;
  float u, v, un, vn;
};
struct FdbMeas { float u, v, un, vn; };

struct FdbStore {
  int cap_obs, cap_feat, cap_hash, cap_batch;        // cap_hash is a power of two >= 2 * cap_feat
  // log, arrival order
  unsigned long long *o_key;
  double *o_t;
  FdbMeas *o_m;
  int *o_cam;
  int *o_feat;                                       // feature index of the observation
  int *batch_start;                                  // [cap_batch + 1]
  // features, insertion order
  unsigned long long *f_key, *f_featid;
  uint8_t *f_del, *f_alive;
  // hash
  unsigned long long *h_key;
  int *h_val;
  // counters: [0] observations, [1] features (alive or not, until compacted), [2] batches, [3] status of the last kernel
  // (0 ok, 1 would overflow: nothing changed), [4] alive features, [5] scratch flag, [8..10] = r_cnt.  Device memory; every
  // kernel ends by copying the 16 words to cnt_pub (mapped host memory), which is all the host reads after a call.
  int *cnt;
  int *cnt_pub;
  // per-query scratch, [cap_feat * FDB_MAXCAM] unless noted
  int *q_cnt, *q_first, *q_last, *q_cursor;
  uint8_t *q_hit, *q_sel;                            // [cap_feat]
  int *q_featpos;                                    // [cap_feat + 1] exclusive scan of q_sel
  int *q_part;                                       // [2048] partial sums of the block scans
  // compaction scratch
  int *c_keep, *c_pos;                               // [max(cap_obs, cap_feat) + 1]
  unsigned long long *c_u64;
  double *c_f64;
  FdbMeas *c_m;
  int *c_i32a, *c_i32b;
  // result of the last query
  int *r_cnt;                                        // = cnt + 8: [0] features, [1] rows, [2] observations
  unsigned long long *r_featid, *r_key;
  uint8_t *r_del;
  int *r_row0;                                       // [features + 1]
  int *r_row_cam;                                    // [rows]
  int *r_off;                                        // [rows + 1]
  double *r_t;
  float *r_uv, *r_uvn;                               // [2 * observations]
};

// ---- atomics (sequential stand-ins on the host: GroupSeq runs one lane) -------------------------------------------------
#ifdef __CUDA_ARCH__
__device__ __forceinline__ unsigned long long fdb_cas64(unsigned long long *p, unsigned long long cmp, unsigned long long v) { return atomicCAS(p, cmp, v); }
__device__ __forceinline__ void fdb_min(int *p, int v) { atomicMin(p, v); }
__device__ __forceinline__ void fdb_max(int *p, int v) { atomicMax(p, v); }
__device__ __forceinline__ void fdb_add(int *p, int v) { atomicAdd(p, v); }
__device__ __forceinline__ int fdb_add_ret(int *p, int v) { return atomicAdd(p, v); }
#else
inline unsigned long long fdb_cas64(unsigned long long *p, unsigned long long cmp, unsigned long long v) { unsigned long long o = *p; if (o == cmp) *p = v; return o; }
inline void fdb_min(int *p, int v) { if (v < *p) *p = v; }
inline void fdb_max(int *p, int v) { if (v > *p) *p = v; }
inline void fdb_add(int *p, int v) { *p += v; }
inline int fdb_add_ret(int *p, int v) { int o = *p; *p += v; return o; }
#endif

// loads of words other lanes of the block update with atomics in the same kernel: read at L2, never a stale L1 line
#ifdef __CUDA_ARCH__
__device__ __forceinline__ int fdb_ldi(const int *p) { return __ldcg(p); }
__device__ __forceinline__ unsigned long long fdb_ld64(const unsigned long long *p) { return __ldcg(p); }
#else
inline int fdb_ldi(const int *p) { return *p; }
inline unsigned long long fdb_ld64(const unsigned long long *p) { return *p; }
#endif

UVIO_HD unsigned fdb_hash(unsigned long long k) {   // splitmix64 finaliser
  k ^= k >> 30; k *= 0xbf58476d1ce4e5b9ull;
  k ^= k >> 27; k *= 0x94d049bb133111ebull;
  k ^= k >> 31;
  return (unsigned)k;
}
// slot holding `key`, or the free slot where the probe sequence of `key` ends
UVIO_HD int fdb_probe(const FdbStore &S, unsigned long long key) {
  unsigned s = fdb_hash(key) & (unsigned)(S.cap_hash - 1);
  for (int n = 0; n < S.cap_hash; n++) {
    unsigned long long k = fdb_ld64(&S.h_key[s]);
    if (k == key || k == FDB_EMPTY) return (int)s;
    s = (s + 1) & (unsigned)(S.cap_hash - 1);
  }
  return -1;
}
UVIO_HD int fdb_find(const FdbStore &S, unsigned long long key) {
  int s = fdb_probe(S, key);
  return (s >= 0 && fdb_ld64(&S.h_key[s]) == key) ? fdb_ldi(&S.h_val[s]) : -1;
}

// exclusive scan of in[0..n) into out[0..n], out[n] = total; in == out is allowed.  Every lane takes a contiguous chunk.
template <class G>
UVIO_HD void fdb_scan(G g, const int *in, int *out, int n, int *part) {
  (void)part;
  const int lanes = g.size(), chunk = (n + lanes - 1) / lanes;
  const int b = g.lane() * chunk, e = b + chunk < n ? b + chunk : n;
  int s = 0;
  for (int i = b; i < e; i++) s += in[i];
  int total;
  int run = g.exclusive_sum(s, total);     // lane order == chunk order
  if (g.lane() == 0) out[n] = total;
  for (int i = b; i < e; i++) { int v = in[i]; out[i] = run; run += v; }
  g.sync();
}

template <class G>
UVIO_HD void fdb_rebuild_hash(G g, const FdbStore &S, int n_feat) {
  for (int s = g.lane(); s < S.cap_hash; s += g.size()) { S.h_key[s] = FDB_EMPTY; S.h_val[s] = -1; }
  g.sync();
  for (int f = g.lane(); f < n_feat; f += g.size()) {
    if (!S.f_alive[f]) continue;
    const unsigned long long key = S.f_key[f];
    unsigned s = fdb_hash(key) & (unsigned)(S.cap_hash - 1);
    while (true) {                                       // keys are unique: claim the first free slot of the probe sequence
      if (fdb_cas64(&S.h_key[s], FDB_EMPTY, key) == FDB_EMPTY) { S.h_val[s] = f; break; }
      s = (s + 1) & (unsigned)(S.cap_hash - 1);
    }
  }
  g.sync();
}

// ---- update_feature x n (one batch) ------------------------------------------------------------------------------------
// The caller guarantees at most one observation per (id, cam) in the batch and n <= g.size() * anything (lanes stride).
template <class G>
UVIO_HD void fdb_update(G g, const FdbStore &S, const FdbObsIn *in, int n, double timestamp) {
  const int n_obs = S.cnt[0], n_feat = S.cnt[1], n_batch = S.cnt[2];
  g.sync();
  // 1. look every id up; ids that are not there yet elect their first observation (batch order) as the creator
  if (g.lane() == 0) S.cnt[5] = 0;
  g.sync();
  for (int i = g.lane(); i < n; i += g.size()) {
    const unsigned long long key = in[i].id;
    unsigned s = fdb_hash(key) & (unsigned)(S.cap_hash - 1);
    int probes = 0;
    for (; probes < S.cap_hash; probes++) {
      unsigned long long k = fdb_ld64(&S.h_key[s]);
      if (k == FDB_EMPTY) {
        const unsigned long long old = fdb_cas64(&S.h_key[s], FDB_EMPTY, key);
        k = old == FDB_EMPTY ? key : old;
      }
      if (k == key) break;
      s = (s + 1) & (unsigned)(S.cap_hash - 1);
    }
    if (probes == S.cap_hash) fdb_max(&S.cnt[5], 1);      // the table is full: more new features than the store can take
    S.c_i32a[i] = (int)s;                                // slot of this observation's key
    S.c_keep[i] = 0;
  }
  g.sync();
  if (fdb_ldi(&S.cnt[5])) {                              // undo the claimed slots and report: the store is unchanged
    g.sync();
    fdb_rebuild_hash(g, S, n_feat);
    if (g.lane() == 0) S.cnt[3] = 1;
    g.sync();
    return;
  }
  for (int i = g.lane(); i < n; i += g.size()) {
    int *v = &S.h_val[S.c_i32a[i]];
    if (fdb_ldi(v) < 0) fdb_max(v, FDB_PROVISIONAL + n);   // a slot claimed just now (-1): open the election (all contenders write the same value)
  }
  g.sync();
  for (int i = g.lane(); i < n; i += g.size()) {
    int *v = &S.h_val[S.c_i32a[i]];
    if (fdb_ldi(v) >= FDB_PROVISIONAL) fdb_min(v, FDB_PROVISIONAL + i);
  }
  g.sync();
  for (int i = g.lane(); i < n; i += g.size()) S.c_keep[i] = fdb_ldi(&S.h_val[S.c_i32a[i]]) == FDB_PROVISIONAL + i;
  g.sync();
  fdb_scan(g, S.c_keep, S.c_pos, n, S.q_part);           // creators in batch order -> new feature indices
  const int n_new = S.c_pos[n];
  const bool fits = n_obs + n <= S.cap_obs && n_feat + n_new <= S.cap_feat && n_batch + 1 <= S.cap_batch;
  if (!fits) {                                           // undo the claimed slots and report: the store is unchanged
    g.sync();
    fdb_rebuild_hash(g, S, n_feat);
    if (g.lane() == 0) S.cnt[3] = 1;
    g.sync();
    return;
  }
  for (int i = g.lane(); i < n; i += g.size()) {
    if (!S.c_keep[i]) continue;
    const int f = n_feat + S.c_pos[i];
    S.f_key[f] = in[i].id;                               // FeatureDatabase.cpp:76-84: a new Feature with featid = id
    S.f_featid[f] = in[i].id;
    S.f_del[f] = 0;
    S.f_alive[f] = 1;
    S.h_val[S.c_i32a[i]] = f;
  }
  g.sync();
  for (int i = g.lane(); i < n; i += g.size()) {         // :66-73 / :79-82: append (u, v), (u_n, v_n), timestamp
    const int at = n_obs + i;
    S.o_key[at] = in[i].id;
    S.o_t[at] = timestamp;
    S.o_m[at] = FdbMeas{in[i].u, in[i].v, in[i].un, in[i].vn};
    S.o_cam[at] = (int)in[i].cam;
    S.o_feat[at] = fdb_ldi(&S.h_val[S.c_i32a[i]]);
  }
  g.sync();
  if (g.lane() == 0) {
    S.cnt[0] = n_obs + n;
    S.cnt[1] = n_feat + n_new;
    S.cnt[4] += n_new;
    if (n > 0) {
      S.batch_start[n_batch] = n_obs;
      S.batch_start[n_batch + 1] = n_obs + n;
      S.cnt[2] = n_batch + 1;
    }
    S.cnt[3] = 0;
  }
  g.sync();
}

// ---- compaction: drops dead observations / features (stable), remaps indices, rebuilds the hash ---------------------------
// c_i32b[i] != 0 marks observation i as dead; features with f_alive == 0 go, and so do their observations.  drop_empty: features
// left without a measurement go too (cleanup_measurements, FeatureDatabase.cpp:237-241).
template <class G>
UVIO_HD void fdb_compact(G g, const FdbStore &S, bool drop_empty) {   // c_i32b[i] != 0 marks observation i as dead
  const int n_obs = S.cnt[0], n_feat = S.cnt[1], n_batch = S.cnt[2];
  g.sync();
  // measurements per feature among the survivors
  for (int f = g.lane(); f < n_feat; f += g.size()) S.q_cnt[f] = 0;
  g.sync();
  for (int i = g.lane(); i < n_obs; i += g.size()) {
    const bool keep = !S.c_i32b[i] && S.f_alive[S.o_feat[i]];
    S.c_keep[i] = keep;
    if (keep) fdb_add(&S.q_cnt[S.o_feat[i]], 1);
  }
  g.sync();
  if (drop_empty) {
    for (int f = g.lane(); f < n_feat; f += g.size())
      if (S.f_alive[f] && fdb_ldi(&S.q_cnt[f]) < 1) S.f_alive[f] = 0;
    g.sync();
  }
  // observations: stable compaction through the scratch arrays
  fdb_scan(g, S.c_keep, S.c_pos, n_obs, S.q_part);
  const int m_obs = S.c_pos[n_obs];
  for (int i = g.lane(); i < n_obs; i += g.size()) {
    if (!S.c_keep[i]) continue;
    const int d = S.c_pos[i];
    S.c_u64[d] = S.o_key[i];
    S.c_f64[d] = S.o_t[i];
    S.c_m[d] = S.o_m[i];
    S.c_i32a[d] = S.o_cam[i] | (S.o_feat[i] << 3);
  }
  g.sync();
  // batches: new start = survivors before the old start; empty batches disappear
  for (int b = g.lane(); b <= n_batch; b += g.size()) S.q_cursor[b] = S.c_pos[S.batch_start[b]];
  g.sync();
  for (int i = g.lane(); i < m_obs; i += g.size()) {
    S.o_key[i] = S.c_u64[i];
    S.o_t[i] = S.c_f64[i];
    S.o_m[i] = S.c_m[i];
    S.o_cam[i] = S.c_i32a[i] & 7;
    S.o_feat[i] = S.c_i32a[i] >> 3;
  }
  g.sync();
  if (g.lane() == 0) {                                   // few batches (one per frame in the window): one lane
    int nb = 0;
    for (int b = 0; b < n_batch; b++)
      if (S.q_cursor[b + 1] > S.q_cursor[b]) {
        S.batch_start[nb] = S.q_cursor[b];
        S.batch_start[nb + 1] = S.q_cursor[b + 1];
        nb++;
      }
    if (nb == 0) S.batch_start[0] = 0;
    S.cnt[2] = nb;
    S.cnt[0] = m_obs;
  }
  g.sync();
  // features: stable compaction, then the observations' feature indices follow
  for (int f = g.lane(); f < n_feat; f += g.size()) S.c_keep[f] = S.f_alive[f];
  g.sync();
  fdb_scan(g, S.c_keep, S.c_pos, n_feat, S.q_part);
  const int m_feat = S.c_pos[n_feat];
  for (int f = g.lane(); f < n_feat; f += g.size()) {
    if (!S.c_keep[f]) continue;
    const int d = S.c_pos[f];
    S.c_u64[d] = S.f_key[f];
    ((unsigned long long *)S.c_f64)[d] = S.f_featid[f];
    S.c_i32a[d] = S.f_del[f];
  }
  g.sync();
  for (int i = g.lane(); i < m_obs; i += g.size()) S.o_feat[i] = S.c_pos[S.o_feat[i]];
  g.sync();
  for (int f = g.lane(); f < m_feat; f += g.size()) {
    S.f_key[f] = S.c_u64[f];
    S.f_featid[f] = ((unsigned long long *)S.c_f64)[f];
    S.f_del[f] = (uint8_t)S.c_i32a[f];
    S.f_alive[f] = 1;
  }
  g.sync();
  if (g.lane() == 0) {
    S.cnt[1] = m_feat;
    S.cnt[4] = m_feat;
  }
  g.sync();
  fdb_rebuild_hash(g, S, m_feat);
}

// ---- cleanup / cleanup_measurements / cleanup_measurements_exact -----------------------------------------------------------
template <class G>
UVIO_HD void fdb_cleanup(G g, const FdbStore &S, int mode, double timestamp) {
  const int n_obs = S.cnt[0], n_feat = S.cnt[1];
  g.sync();
  if (mode == 0) {                                        // FeatureDatabase.cpp:214-222
    for (int f = g.lane(); f < n_feat; f += g.size())
      if (S.f_alive[f] && S.f_del[f]) S.f_alive[f] = 0;
  }
  for (int i = g.lane(); i < n_obs; i += g.size()) {
    const double t = S.o_t[i];
    S.c_i32b[i] = mode == 1 ? (t <= timestamp) : mode == 2 ? (t == timestamp) : 0;   // Feature.cpp:95 (<=), :71 (== via std::find)
  }
  g.sync();
  fdb_compact(g, S, mode != 0);
  if (g.lane() == 0) S.cnt[3] = 0;
  g.sync();
}

// ---- the queries ------------------------------------------------------------------------------------------------------------
template <class G>
UVIO_HD void fdb_query(G g, const FdbStore &S, int kind, double timestamp, unsigned long long key, int remove, int skip_deleted) {
  const int n_obs = S.cnt[0], n_feat = S.cnt[1], n_batch = S.cnt[2];
  g.sync();
  for (int k = g.lane(); k < n_feat * FDB_MAXCAM; k += g.size()) {
    S.q_cnt[k] = 0;
    S.q_first[k] = 0x7fffffff;
    S.q_last[k] = -1;
    S.q_cursor[k] = 0;
  }
  for (int f = g.lane(); f < n_feat; f += g.size()) S.q_hit[f] = 0;
  g.sync();
  // per (feature, camera): count, first and last measurement (by arrival: the vectors' front() and back()); exact-time hits
  for (int i = g.lane(); i < n_obs; i += g.size()) {
    const int f = S.o_feat[i], k = f * FDB_MAXCAM + S.o_cam[i];
    fdb_add(&S.q_cnt[k], 1);
    fdb_min(&S.q_first[k], i);
    fdb_max(&S.q_last[k], i);
    if (S.o_t[i] == timestamp) S.q_hit[f] = 1;            // :189 std::find(..., timestamp)
  }
  g.sync();
  const int wanted = kind == 3 ? fdb_find(S, key) : -1;
  for (int f = g.lane(); f < n_feat; f += g.size()) {
    bool sel = S.f_alive[f] != 0;
    if (sel && skip_deleted && kind != 3 && kind != 4 && S.f_del[f]) sel = false;   // :97-100, :138-141, :179-182
    if (sel) {
      if (kind == 0) {                                    // :104: any camera whose newest measurement is >= timestamp
        bool newer = false;
        for (int c = 0; c < FDB_MAXCAM; c++) newer |= fdb_ldi(&S.q_cnt[f * FDB_MAXCAM + c]) > 0 && S.o_t[fdb_ldi(&S.q_last[f * FDB_MAXCAM + c])] >= timestamp;
        sel = !newer;
      } else if (kind == 1) {                             // :145: any camera whose oldest measurement is < timestamp
        bool older = false;
        for (int c = 0; c < FDB_MAXCAM; c++) older |= fdb_ldi(&S.q_cnt[f * FDB_MAXCAM + c]) > 0 && S.o_t[fdb_ldi(&S.q_first[f * FDB_MAXCAM + c])] < timestamp;
        sel = older;
      } else if (kind == 2) {
        sel = S.q_hit[f] != 0;
      } else if (kind == 3) {
        sel = f == wanted;
      }
    }
    S.q_sel[f] = sel;
    S.c_keep[f] = sel;
  }
  g.sync();
  fdb_scan(g, S.c_keep, S.q_featpos, n_feat, S.q_part);
  const int n_sel = S.q_featpos[n_feat];
  // rows: one per camera with measurements
  for (int f = g.lane(); f < n_feat; f += g.size()) {
    int rows = 0;
    if (S.q_sel[f])
      for (int c = 0; c < FDB_MAXCAM; c++) rows += fdb_ldi(&S.q_cnt[f * FDB_MAXCAM + c]) > 0;
    S.c_keep[f] = rows;
  }
  g.sync();
  fdb_scan(g, S.c_keep, S.c_pos, n_feat, S.q_part);       // c_pos[f] = first row of feature f
  const int n_rows = S.c_pos[n_feat];
  for (int f = g.lane(); f < n_feat; f += g.size()) {
    if (!S.q_sel[f]) continue;
    const int k = S.q_featpos[f];
    S.r_featid[k] = S.f_featid[f];
    S.r_key[k] = S.f_key[f];
    S.r_del[k] = S.f_del[f];
    S.r_row0[k] = S.c_pos[f];
    int r = S.c_pos[f];
    for (int c = 0; c < FDB_MAXCAM; c++)
      if (fdb_ldi(&S.q_cnt[f * FDB_MAXCAM + c]) > 0) {
        S.r_row_cam[r] = c;
        S.c_i32a[r] = fdb_ldi(&S.q_cnt[f * FDB_MAXCAM + c]);   // observations of row r
        S.q_first[f * FDB_MAXCAM + c] = r;                // from here on: the row of (f, c)
        r++;
      }
  }
  if (g.lane() == 0) S.r_row0[n_sel] = n_rows;
  g.sync();
  fdb_scan(g, S.c_i32a, S.r_off, n_rows, S.q_part);
  const int n_out = S.r_off[n_rows];
  // observations: batches in order, the lanes take the observations of a batch (at most one per (feature, camera))
  for (int b = 0; b < n_batch; b++) {
    for (int i = S.batch_start[b] + g.lane(); i < S.batch_start[b + 1]; i += g.size()) {
      const int f = S.o_feat[i];
      if (!S.q_sel[f]) continue;
      const int k = f * FDB_MAXCAM + S.o_cam[i];
      const int at = S.r_off[S.q_first[k]] + S.q_cursor[k];
      S.q_cursor[k] += 1;
      S.r_t[at] = S.o_t[i];
      S.r_uv[2 * at] = S.o_m[i].u;
      S.r_uv[2 * at + 1] = S.o_m[i].v;
      S.r_uvn[2 * at] = S.o_m[i].un;
      S.r_uvn[2 * at + 1] = S.o_m[i].vn;
    }
    g.sync();
  }
  if (g.lane() == 0) {
    S.r_cnt[0] = n_sel;
    S.r_cnt[1] = n_rows;
    S.r_cnt[2] = n_out;
    S.cnt[3] = 0;
  }
  g.sync();
  if (remove && n_sel > 0) {                               // :118-119 etc.: erase the returned features from the map
    for (int f = g.lane(); f < n_feat; f += g.size())
      if (S.q_sel[f]) S.f_alive[f] = 0;
    for (int i = g.lane(); i < n_obs; i += g.size()) S.c_i32b[i] = 0;
    g.sync();
    fdb_compact(g, S, false);
  }
}

template <class G>
UVIO_HD void fdb_publish(G g, const FdbStore &S) {
  g.sync();
  for (int k = g.lane(); k < 16; k += g.size()) S.cnt_pub[k] = S.cnt[k];
}

// feat->to_delete = true for the given keys; change_feat_id's database half
template <class G>
UVIO_HD void fdb_mark(G g, const FdbStore &S, const unsigned long long *keys, int n, int what, unsigned long long new_id) {
  for (int i = g.lane(); i < n; i += g.size()) {
    const int f = fdb_find(S, keys[i]);
    if (f < 0) continue;
    if (what == 0) S.f_del[f] = 1;
    else S.f_featid[f] = new_id;                          // TrackBase.cpp:236 (the erase / insert around it act on a copy)
  }
  g.sync();
}

}  // namespace uvio
