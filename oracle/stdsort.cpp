// ORACLE — TEST INFRASTRUCTURE ONLY.
// Permutation produced by libstdc++ std::sort with the reference's comparator
// (Grider_FAST::compare_response, ov_core/src/track/Grider_FAST.h:57, called at Grider_GRID.h:128):
//     static bool compare_response(cv::KeyPoint first, cv::KeyPoint second) { return first.response > second.response; }
// std::sort is not stable, and the tie order decides which corners make the per-cell top-K, so the oracle asks the real
// std::sort instead of guessing.
#include <algorithm>
#include <cstdint>
#include <vector>

namespace {
struct Kp {  // stands in for cv::KeyPoint: only .response takes part in the comparison
  float x, y, size, angle, response;
  int octave, class_id;
  int src_index;
};
bool compare_response(Kp first, Kp second) { return first.response > second.response; }
}  // namespace

extern "C" void oracle_std_sort_perm(const float *response, int n, int32_t *perm_out) {
  std::vector<Kp> v(n);
  for (int i = 0; i < n; i++) {
    v[i] = Kp{0.f, 0.f, 7.f, -1.f, response[i], 0, -1, i};
  }
  std::sort(v.begin(), v.end(), compare_response);
  for (int i = 0; i < n; i++) perm_out[i] = v[i].src_index;
}
