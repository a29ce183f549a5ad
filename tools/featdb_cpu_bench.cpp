// CPU arm of tools/rows_bench.py (row n1): the per-frame FeatureDatabase calls the estimator makes (VioManager.cpp:369, :579-591)
// on the dependency-free C++ mirror of ov_core::FeatureDatabase (tests/cpp/ov_core_mirror.h: the reference's
// unordered_map<size_t, shared_ptr<Feature>> + mutex, statement by statement), one store, one thread.
//   featdb_cpu_bench <obs.bin> <n_frames> <repeat>     obs.bin: per frame  int32 n, then n x {u64 id, u64 cam, f32 u, v, un, vn}
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "ov_core_mirror.h"

struct Rec { unsigned long long id, cam; float u, v, un, vn; };

int main(int argc, char **argv) {
  if (argc < 4) return 2;
  FILE *f = fopen(argv[1], "rb");
  const int n_frames = atoi(argv[2]), repeat = atoi(argv[3]);
  std::vector<std::vector<Rec>> frames(n_frames);
  for (int k = 0; k < n_frames; k++) {
    int n = 0;
    if (fread(&n, 4, 1, f) != 1) return 3;
    frames[k].resize(n);
    if (n && fread(frames[k].data(), sizeof(Rec), n, f) != (size_t)n) return 3;
  }
  fclose(f);
  double total = 0;
  long calls = 0, returned = 0;
  for (int rep = 0; rep < repeat; rep++) {
    ov_core::FeatureDatabase db;
    auto t0 = std::chrono::steady_clock::now();
    for (int k = 0; k < n_frames; k++) {
      const double t = 0.05 * k;
      for (const Rec &r : frames[k]) db.update_feature((size_t)r.id, t, (size_t)r.cam, r.u, r.v, r.un, r.vn);
      calls += (long)frames[k].size();
      auto lost = db.features_not_containing_newer(t, false, true);   // VioManager.cpp:369
      returned += (long)lost.size();
      for (size_t i = 0; i < lost.size(); i += 2) lost[i]->to_delete = true;   // the updaters mark what they used
      db.cleanup();                                                    // :579
      if (k > 5) db.cleanup_measurements(t - 0.25);                    // :589
    }
    total += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  }
  printf("{\"frames\": %d, \"seconds\": %.6f, \"frames_per_sec\": %.1f, \"update_feature_calls\": %ld, \"features_returned\": %ld}\n", n_frames * repeat, total,
         n_frames * repeat / total, calls, returned);
  return 0;
}
