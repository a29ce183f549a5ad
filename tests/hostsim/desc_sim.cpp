// TEST SCAFFOLDING ONLY: CPU stand-ins of the descriptor tracker's launch wrappers (uvio_b200/csrc/desc.h) running the same
// kernel bodies (desc_body.h) with GroupSeq, so that desc_engine.cpp can be exercised by `pytest -m "not gpu"`.
#include <vector>

#include "../../uvio_b200/csrc/desc.h"
#include "../../uvio_b200/csrc/orb_tables.h"

namespace uvio {

static const int8_t g_pairs[256][4] = ORB_PAIRS_INIT;

cudaError_t launch_desc_topk(cudaStream_t, const DescUnit *units, int n_units, int, const DescGlobals &X) {
  std::vector<uint32_t> buf(64);   // smaller than most cells: the in-place sort path is exercised too
  for (int u = 0; u < n_units; u++)
    for (int side = 0; side < (units[u].stereo ? 2 : 1); side++)
      for (int r = 0; r < units[u].s[side].ncells; r++) desc_topk(GroupSeq(), units[u], X, side, r, buf.data(), (int)buf.size());
  return 0;
}
cudaError_t launch_desc_gather(cudaStream_t, const DescUnit *units, int n_units, const DescGlobals &X) {
  int base = 0;
  for (int u = 0; u < n_units; u++)
    for (int side = 0; side < (units[u].stereo ? 2 : 1); side++) desc_gather(GroupSeq(), units[u], X, side, &base);
  return 0;
}
cudaError_t launch_desc_orb(cudaStream_t, const DescUnit *units, int n_units, int max_p, const DescGlobals &X) {
  std::vector<short> rank(max_p + 1);
  OrbSmem sm;
  *X.n_sp = 0;
  for (int u = 0; u < n_units; u++)
    for (int side = 0; side < (units[u].stereo ? 2 : 1); side++) {
      const DescSide &S = units[u].s[side];
      const int n = S.cnt[0];
      desc_orb_ranks(GroupSeq(), units[u], X, side, rank.data(), true);
      for (int i = 0; i < n; i++) {
        if (rank[i] < 0) continue;
        const float x = X.sp_out[2 * (S.slot0 + i)], y = X.sp_out[2 * (S.slot0 + i) + 1];
        orb_describe(GroupSeq(), sm, S.img, round_f2i(x), round_f2i(y), &g_pairs[0][0], S.kp_desc + (size_t)DESC_WORDS * rank[i]);
      }
    }
  return 0;
}
static void bind(RansacSmem &rs, std::vector<float> &dyn, int P) {
  dyn.assign(4 * (size_t)P + 4, 0.f);
  rs.x0 = dyn.data();
  rs.y0 = dyn.data() + P;
  rs.x1 = dyn.data() + 2 * P;
  rs.y1 = dyn.data() + 3 * P;
}
cudaError_t launch_desc_detect(cudaStream_t, const DescUnit *units, int n_units, int max_p, const DescGlobals &X) {
  static RansacSmem rs;
  std::vector<float> dyn;
  bind(rs, dyn, max_p);
  for (int u = 0; u < n_units; u++) {
    if (units[u].stereo) {
      desc_match(GroupSeq(), rs, units[u], X, 0);
      desc_assign_stereo(GroupSeq(), units[u]);
    } else {
      desc_assign_mono(GroupSeq(), units[u]);
    }
  }
  return 0;
}
cudaError_t launch_desc_temporal(cudaStream_t, const DescUnit *units, int n_units, int max_p, const DescGlobals &X) {
  static RansacSmem rs;
  std::vector<float> dyn;
  bind(rs, dyn, max_p);
  for (int u = 0; u < n_units; u++)
    for (int job = 1; job <= (units[u].stereo ? 2 : 1); job++) desc_match(GroupSeq(), rs, units[u], X, job);
  return 0;
}
cudaError_t launch_desc_merge(cudaStream_t, const DescUnit *units, int n_units, int max_p, const DescGlobals &X) {
  std::vector<int> inv(2 * (size_t)max_p + 2);
  for (int u = 0; u < n_units; u++) desc_merge(GroupSeq(), units[u], X, inv.data());
  return 0;
}

}  // namespace uvio
