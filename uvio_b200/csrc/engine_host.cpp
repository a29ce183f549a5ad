// The round-1 pipeline (UVIO_KLT_PIPELINE=host): three GPU rounds per feed with the reference's order-dependent bookkeeping on host
// threads between them (hostlogic.h), detections issued ahead, the optional background part.
#include "engine_internal.h"

namespace uvio {

static const uint8_t *mask_ptr(CamState &c, int slot) { return c.mask_nonzero[slot] ? c.mask[slot].data() : nullptr; }

// emit FAST jobs for the cells of one side
static void emit_fast_jobs(Lane *ctx, Side &s, int threshold, int &njobs, size_t &nwords) {
  HOSTPROF(emit_fast_jobs);
  const int cols = s.cam->w, rows = s.cam->h;
  s.job_of_cell.assign(s.cells.size(), -1);
  for (size_t r = 0; r < s.cells.size(); r++) {
    int x = s.cells[r].first * s.geom.size_x, y = s.cells[r].second * s.geom.size_y;
    if (x + s.geom.size_x > cols || y + s.geom.size_y > rows) continue;   // Grider_GRID.h:117
    FastJob j;
    j.img = s.cam->pyr[s.det_slot];
    j.x0 = x; j.y0 = y; j.w = s.geom.size_x; j.h = s.geom.size_y;
    j.out_off = (int)nwords;
    j.out_cap = ((j.w + 1) / 2) * ((j.h + 1) / 2);
    j.threshold = threshold;
    ctx->fast_jobs.h[njobs] = j;
    s.job_of_cell[r] = njobs++;
    nwords += (size_t)j.out_cap;
  }
}

// Phase A: occupancy passes and the list of cells to run FAST on.
// ahead: run at the end of a feed for the NEXT one (the frame that is current now will be "last" then).
static void phase_a(Unit &u, bool ahead = false) {
  HOSTPROF(phase_a);
  const Params &P = u.rig->prm;
  Side &L = u.L;
  const int thr = std::min(20, (int)(0.5 * P.num_features));
  {
    CamState &c = *L.cam;
    L.det_slot = (u.first || ahead) ? c.cur : c.last();
    L.det_mask = mask_ptr(c, L.det_slot);
    { HOSTPROF(a_occ_init); occ_init(L.occ, c.w, c.h, P.min_px_dist, P.grid_x, P.grid_y); }
    HOSTPROF(a_left_rest);
    if (!u.first) occ_feed(L.occ, c.pts_last.data(), c.ids_last.data(), c.pts_last.size(), L.det_mask, nullptr);
    const int need = P.num_features - (int)L.occ.pts.size();
    // mono returns when need < thr (TrackKLT.cpp:470); stereo detects when need > thr (:609)
    L.detect = u.stereo ? (need > thr) : !(need < thr);
    L.cells.clear();
    if (L.detect) {
      valid_cells(L.occ, L.det_mask, P.num_features, L.cells);
      L.geom = grider_geom(c.w, c.h, P.num_features, P.grid_x, P.grid_y);
    }
  }
  if (u.stereo) {
    Side &R = u.R;
    CamState &c = *R.cam;
    R.det_slot = (u.first || ahead) ? c.cur : c.last();
    R.det_mask = mask_ptr(c, R.det_slot);
    occ_init(R.occ, c.w, c.h, P.min_px_dist, P.grid_x, P.grid_y);
    // The right pass (:722-780) runs after the left detection appended new stereo points to pts1, but it is a plain
    // in-order loop: its first |pts_last| steps only see the old points, and the new ids are fresh, so they cannot
    // change the stereo-id exemption of an old point. Run those steps now; phase C continues the same loop.
    IdTable &stereo_ids = u.tab_a;
    { HOSTPROF(a_idtable); stereo_ids.fill(L.occ.ids.data(), L.occ.ids.size()); }
    HOSTPROF(a_right_feed);
    if (!u.first) occ_feed(R.occ, c.pts_last.data(), c.ids_last.data(), c.pts_last.size(), R.det_mask, &stereo_ids);
    // Cells: new points only raise counts and lower the need, so the cells valid now are a superset of the cells the
    // reference will extract from; FAST + top-K + sub-pixel run on the superset, phase C keeps the right subset.
    const int need = P.num_features - (int)R.occ.pts.size();
    R.detect = need > thr;
    R.cells.clear();
    if (R.detect) {
      valid_cells(R.occ, R.det_mask, P.num_features, R.cells);
      R.geom = grider_geom(c.w, c.h, P.num_features, P.grid_x, P.grid_y);
    }
  }
}

// Phase B: per-cell top-K (reference sort order), left mask test, sub-pixel / stereo-LK items.
static void phase_b(Lane *ctx, Unit &u, int &nsp, int &nlk, std::vector<uint32_t> &tmp, std::vector<uint32_t> &grid) {
  HOSTPROF(phase_b);
  for (int side = 0; side < (u.stereo ? 2 : 1); side++) {
    Side &s = side ? u.R : u.L;
    if (s.top.size() < s.cells.size()) s.top.resize(s.cells.size());   // entries beyond cells.size() are stale scratch
    for (size_t r = 0; r < s.cells.size(); r++) s.top[r].clear();
    if (!s.detect) continue;
    HOSTPROF(b_topk);
    for (size_t r = 0; r < s.cells.size(); r++) {
      int j = s.job_of_cell[r];
      if (j < 0) continue;
      const FastJob &job = ctx->fast_jobs.h[j];
#ifdef UVIO_HOSTPROF
      { static long long cells = 0, corners = 0; cells++; corners += ctx->fast_cnt.h[j]; if ((cells & 255) == 0) fprintf(stderr, "[hostprof] cells %lld corners %lld\n", cells, corners); }
#endif
      cell_top_k(ctx->fast_out.h + job.out_off, ctx->fast_cnt.h[j], job.w, job.h, job.x0, job.y0, s.geom.per_cell, tmp, grid, s.top[r]);
    }
  }
  // left: mask test against mask_updated = mask | painted boxes (Grider_GRID.h:142-148), then sub-pixel
  Side &L = u.L;
  u.selL.clear();
  u.sp0L = nsp;
  if (L.detect) {
    HOSTPROF(b_masktest);
    const int cols = L.cam->w, rows = L.cam->h;
    for (size_t r = 0; r < L.cells.size(); r++)
      for (auto &k : L.top[r]) {
        int x = (int)k.x, y = (int)k.y;
        if (x < 0 || x > cols || y < 0 || y > rows) continue;
        size_t at = (size_t)y * cols + x;
        if ((L.det_mask && L.det_mask[at] > 127) || occ_boxed(L.occ, at)) continue;
        u.selL.push_back(k);
      }
  }
  for (auto &k : u.selL) ctx->sp_items.h[nsp++] = SubpixItem{k.x, k.y, L.cam->pyr[L.det_slot], 0};
  u.lk0 = nlk;
  if (u.stereo) {
    // left -> right LK of every selected left corner, chained on the sub-pixel output (item i <-> sub-pixel item i)
    for (size_t i = 0; i < u.selL.size(); i++)
      ctx->lk_items.h[nlk++] = LkItem{0, 0, 0, 0, L.cam->pyr[L.det_slot], u.R.cam->pyr[u.R.det_slot]};
  }
}

static void phase_b_right(Lane *ctx, Unit &u, int &nsp) {
  HOSTPROF(phase_b_right);
  Side &R = u.R;
  R.sp_of_cell.assign(R.cells.size(), -1);
  if (!R.detect) return;
  for (size_t r = 0; r < R.cells.size(); r++) {
    R.sp_of_cell[r] = nsp;
    for (auto &k : R.top[r]) ctx->sp_items.h[nsp++] = SubpixItem{k.x, k.y, R.cam->pyr[R.det_slot], 0};
  }
}

// Phase C: finish the detection with the refined corners (min-distance reject, ids, right side), then either store the
// first-frame result or emit the temporal LK / RANSAC work.
static int phase_c_detect(Lane *ctx, Unit &u) {
  HOSTPROF(phase_c_detect);
  const Params &P = u.rig->prm;
  Side &L = u.L;
  const int thr = std::min(20, (int)(0.5 * P.num_features));
  std::vector<Pt> &refined = u.refined;
  refined.resize(u.selL.size());
  for (size_t i = 0; i < u.selL.size(); i++) refined[i] = Pt{ctx->sp_out.h[2 * (u.sp0L + i)], ctx->sp_out.h[2 * (u.sp0L + i) + 1]};
  std::vector<int> &keep = u.keep;
  close_reject(L.occ, refined.data(), refined.size(), keep);       // TrackKLT.cpp:497-512 / :635-650
  std::vector<Pt> &new_r = u.new_r;
  std::vector<size_t> &new_r_ids = u.new_r_ids;
  new_r.clear();
  new_r_ids.clear();
  if (!u.stereo) {
    for (int i : keep) {                                            // :519-527
      L.occ.pts.push_back(refined[i]);
      L.occ.ids.push_back(++u.rig->currid);
    }
  } else {
    Side &R = u.R;
    const int cols0 = L.cam->w, rows0 = L.cam->h, cols1 = R.cam->w, rows1 = R.cam->h;
    for (int i : keep) {                                            // :675-708
      const Pt p0 = refined[i];
      const LkOut o = ctx->lk_out.h[u.lk0 + i];
      const bool oob_left = (int)p0.x < 0 || (int)p0.x >= cols0 || (int)p0.y < 0 || (int)p0.y >= rows0;
      const bool oob_right = (int)o.x < 0 || (int)o.x >= cols1 || (int)o.y < 0 || (int)o.y >= rows1;
      if (!oob_left && !oob_right && ctx->lk_status.h[u.lk0 + i] == 1) {
        size_t id = ++u.rig->currid;
        L.occ.pts.push_back(p0);
        L.occ.ids.push_back(id);
        new_r.push_back(Pt{o.x, o.y});
        new_r_ids.push_back(id);
      } else if (!oob_left) {
        L.occ.pts.push_back(p0);
        L.occ.ids.push_back(++u.rig->currid);
      }
    }
    // right: continue the occupancy loop over the appended stereo points (:722-780), then the right top-up (:784-826)
    IdTable &stereo_ids = u.tab_a;
    stereo_ids.fill(L.occ.ids.data(), L.occ.ids.size());
    occ_feed(R.occ, new_r.data(), new_r_ids.data(), new_r.size(), R.det_mask, &stereo_ids);
    const int need = P.num_features - (int)R.occ.pts.size();
    if (need > thr) {
      std::vector<std::pair<int, int>> cells;
      valid_cells(R.occ, R.det_mask, P.num_features, cells);
      std::vector<Pt> &ext = u.ext;
      ext.clear();
      const uint8_t *mask0 = (cols0 == cols1 && rows0 == rows1) ? L.det_mask : nullptr;   // :721 clones mask0 (sic)
      for (auto &cell : cells) {
        size_t r = 0;
        while (r < R.cells.size() && R.cells[r] != cell) r++;
        if (r == R.cells.size() || !R.detect) return E_ARG;   // cannot happen: R.cells is a superset (phase_a)
        for (size_t k = 0; k < R.top[r].size(); k++) {
          const KeyPointLite &kp = R.top[r][k];
          int x = (int)kp.x, y = (int)kp.y;
          if (x < 0 || x > cols1 || y < 0 || y > rows1) continue;
          size_t at = (size_t)y * cols1 + x;
          if ((mask0 && mask0[at] > 127) || occ_boxed(R.occ, at)) continue;
          int sp = R.sp_of_cell[r] + (int)k;
          ext.push_back(Pt{ctx->sp_out.h[2 * sp], ctx->sp_out.h[2 * sp + 1]});
        }
      }
      close_reject(R.occ, ext.data(), ext.size(), keep);            // :811-825
      for (int i : keep) {
        R.occ.pts.push_back(ext[i]);
        R.occ.ids.push_back(++u.rig->currid);
      }
    }
    R.pts.swap(R.occ.pts);
    R.ids.swap(R.occ.ids);
  }
  L.pts.swap(L.occ.pts);
  L.ids.swap(L.occ.ids);
  return OK;
}

static void store_last(Side &s, std::vector<Pt> &pts, std::vector<size_t> &ids) {
  s.cam->pts_last.swap(pts);
  s.cam->ids_last.swap(ids);
  s.cam->have_last = true;
}

static void emit_temporal(Lane *ctx, Unit &u, int &nlk, int &nrs) {
  HOSTPROF(emit_temporal);
  for (int side = 0; side < (u.stereo ? 2 : 1); side++) {
    Side &s = side ? u.R : u.L;
    CamState &c = *s.cam;
    s.lk_off = nlk;
    s.lk_n = 0;
    s.lk_run = s.pts.size() >= 10;                                  // TrackKLT.cpp:848-852
    if (!s.lk_run) continue;
    s.lk_n = (int)s.pts.size();
    for (auto &p : s.pts) ctx->lk_items.h[nlk++] = LkItem{p.x, p.y, p.x, p.y, c.pyr[c.last()], c.pyr[c.cur]};
    RansacJob j;
    j.off = s.lk_off;
    j.n = s.lk_n;
    j.cam0 = j.cam1 = c.cam_model_idx;
    j.thr = 2.0 / c.max_focal;                                      // :870-873 (id0 == id1 for the temporal match)
    j.run = 1;
    ctx->rs_jobs.h[nrs++] = j;
  }
}

// Phase D: merge / filter the matches and produce the feature-database observations (TrackKLT.cpp:143-189, :285-381).
static void phase_d(Lane *ctx, Unit &u) {
  HOSTPROF(phase_d);
  Rig &rig = *u.rig;
  Side &L = u.L, &R = u.R;
  auto new_pt = [&](Side &s, size_t i) { const LkOut o = ctx->lk_out.h[s.lk_off + i]; return Pt{o.x, o.y}; };
  auto ok = [&](Side &s, size_t i) { return s.lk_run && ctx->rs_mask.h[s.lk_off + i]; };
  auto add_obs = [&](Side &s, size_t i, size_t id, Pt p) {
    rig.obs.push_back(Obs{id, s.cam->cam_id, p.x, p.y, ctx->p1n.h[2 * (s.lk_off + i)], ctx->p1n.h[2 * (s.lk_off + i) + 1]});
  };
  std::vector<Pt> &good_l = u.good_l, &good_r = u.good_r;   // scratch kept in the unit; swapped into pts_last at the end
  std::vector<size_t> &ids_l = u.ids_l, &ids_r = u.ids_r;
  good_l.clear(); good_r.clear(); ids_l.clear(); ids_r.clear();
  good_l.reserve(L.pts.size()); ids_l.reserve(L.pts.size());
  good_r.reserve(R.pts.size()); ids_r.reserve(R.pts.size());
  rig.obs.reserve(L.pts.size() + R.pts.size());
  if (!u.stereo) {
    CamState &c = *L.cam;
    if (L.pts.empty()) {                                            // :143-152 (mask_ll empty)
      store_last(L, good_l, ids_l);
      return;
    }
    const uint8_t *mask = mask_ptr(c, c.cur);                       // :166 tests the current message mask
    for (size_t i = 0; i < L.pts.size(); i++) {                     // :159-173
      if (!L.lk_run) break;
      Pt p = new_pt(L, i);
      if (p.x < 0 || p.y < 0 || (int)p.x >= c.w || (int)p.y >= c.h) continue;
      if (mask && mask[(size_t)(int)p.y * c.w + (int)p.x] > 127) continue;
      if (ok(L, i)) {
        good_l.push_back(p);
        ids_l.push_back(L.ids[i]);
        add_obs(L, i, L.ids[i], p);
      }
    }
    store_last(L, good_l, ids_l);
    return;
  }
  if (L.pts.empty() && R.pts.empty()) {                             // :285-299
    store_last(L, good_l, ids_l);
    store_last(R, good_r, ids_r);
    return;
  }
  const int cols_l = L.cam->w, rows_l = L.cam->h, cols_r = R.cam->w, rows_r = R.cam->h;
  IdTable &first_right = u.tab_a;                                   // id -> first index in ids_right_old (:314-320)
  first_right.fill(R.ids.data(), R.ids.size());
  std::vector<size_t> &obs_l_idx = u.obs_l_idx, &obs_r_idx = u.obs_r_idx;
  obs_l_idx.clear(); obs_r_idx.clear();
  obs_l_idx.reserve(L.pts.size()); obs_r_idx.reserve(R.pts.size());
  if (L.lk_run)
    for (size_t i = 0; i < L.pts.size(); i++) {                     // :306-338
      Pt p = new_pt(L, i);
      if (p.x < 0 || p.y < 0 || (int)p.x > cols_l || (int)p.y > rows_l) continue;   // '>' on the left (:308-309)
      const int ir = first_right.find(L.ids[i]);
      const bool found_right = ir >= 0;
      if (ok(L, i) && found_right && ok(R, (size_t)ir)) {
        Pt q = new_pt(R, (size_t)ir);
        if (q.x < 0 || q.y < 0 || (int)q.x >= cols_r || (int)q.y >= rows_r) continue;
        good_l.push_back(p);
        ids_l.push_back(L.ids[i]);
        obs_l_idx.push_back(i);
        good_r.push_back(q);
        ids_r.push_back(R.ids[ir]);
        obs_r_idx.push_back((size_t)ir);
      } else if (ok(L, i)) {
        good_l.push_back(p);
        ids_l.push_back(L.ids[i]);
        obs_l_idx.push_back(i);
      }
    }
  if (R.lk_run) {
    IdTable &added = u.tab_b;
    added.reset(ids_r.size() + R.pts.size());
    for (size_t k = 0; k < ids_r.size(); k++) added.insert(ids_r[k], (int)k);
    for (size_t i = 0; i < R.pts.size(); i++) {                     // :341-354
      Pt q = new_pt(R, i);
      if (q.x < 0 || q.y < 0 || (int)q.x >= cols_r || (int)q.y >= rows_r) continue;
      if (ok(R, i) && added.find(R.ids[i]) < 0) {
        good_r.push_back(q);
        ids_r.push_back(R.ids[i]);
        obs_r_idx.push_back(i);
        added.insert(R.ids[i], (int)i);
      }
    }
  }
  for (size_t k = 0; k < good_l.size(); k++) add_obs(L, obs_l_idx[k], ids_l[k], good_l[k]);   // :357-366 left, then right
  for (size_t k = 0; k < good_r.size(); k++) add_obs(R, obs_r_idx[k], ids_r[k], good_r[k]);
  store_last(L, good_l, ids_l);
  store_last(R, good_r, ids_r);
}

cudaError_t Pass::wait_new_frames() {
  cudaError_t e = cudaStreamWaitEvent(ctx->stream, ctx->up_done, 0);
  for (int i = 0; i < Context::PF_RING && e == cudaSuccess; i++)
    if ((wait_evts >> i) & 1u) e = cudaStreamWaitEvent(ctx->stream, ctx->dev->pf_evt[i], 0);
  return e;
}

int Pass::issue_detection(bool ahead, bool &any_first, int &max_cw, int &max_ch) {
  Context *dev = ctx->dev;
  size_t ncells = 0, nwords = 0;
  max_cw = max_ch = 1;
  any_first = false;
  for (auto &u : units) {
    if (u.failed) continue;
    phase_a(u, ahead);
    any_first |= u.first;
    for (int side = 0; side < (u.stereo ? 2 : 1); side++) {
      Side &s = side ? u.R : u.L;
      if (!s.detect) continue;
      ncells += s.cells.size();
      nwords += s.cells.size() * (size_t)(((s.geom.size_x + 1) / 2) * ((s.geom.size_y + 1) / 2));
      max_cw = std::max(max_cw, s.geom.size_x);
      max_ch = std::max(max_ch, s.geom.size_y);
    }
  }
  CU_TRY(ctx->fast_jobs.reserve(ncells));
  CU_TRY(ctx->fast_cnt.reserve(ncells));
  CU_TRY(ctx->fast_out.reserve(nwords));
  CU_TRY(ctx->fast_out_dev.reserve(nwords));
  CU_TRY(ctx->fast_cnt_dev.reserve(ncells, true));
  CU_TRY(ctx->fast_ticket.reserve(ncells, true));
  njobs = 0;
  nwords = 0;
  for (auto &u : units) {
    if (u.failed) continue;
    for (int side = 0; side < (u.stereo ? 2 : 1); side++) {
      Side &s = side ? u.R : u.L;
      if (s.detect) emit_fast_jobs(ctx, s, u.rig->prm.fast_threshold, njobs, nwords);
    }
  }
  CU_TRY(ctx->fast_jobs.upload(ctx->stream, (size_t)njobs));
  if (!any_first) {
    HOSTPROF(s0_fast_launch);
    CU_TRY(launch_fast(ctx->stream, ctx->fast_jobs.d, njobs, max_cw, max_ch, dev->pyr_dev.d, ctx->fast_out_dev.d, ctx->fast_cnt_dev.d,
                       ctx->fast_ticket.d, ctx->fast_out.d, ctx->fast_cnt.d));
  }
  return OK;
}

bool ahead_enabled() {   // UVIO_KLT_AHEAD=0: every feed issues its own detection (A/B measurements)
  static const bool v = [] {
    const char *e = getenv("UVIO_KLT_AHEAD");
    return !(e && e[0] == '0');
  }();
  return v;
}

// A background detection that will not be used: the id counters it advanced go back to where they were.  Every tracker in
// the list is unfed since (a feed of any of them through another lane cancels first: feed_batch).
void cancel_ahead(Lane &ln) {
  Lane::Ahead &a = ln.ahead;
  if (a.valid && a.level == 2) {
    for (size_t i = a.unit.size(); i-- > 0;) {   // in reverse: a tracker with several units ends at its earliest value
      a.unit[i]->rig->currid = a.currid_before[i];
      a.unit[i]->rig->bg_lane = -1;
    }
  }
  a.valid = false;
  a.level = 0;
}

// The detection issued when the previous pass of this lane ended serves this pass if the units are the same and nothing
// it read has changed: same trackers (by serial: an address may be reused), no set_num_features / change_feat_id since,
// no feed of any of them through another lane since, same stereo / mono shape, and no unit starting over (a tracker that
// lost every point detects on the NEW frame).
bool Pass::ahead_usable() {
  const Lane::Ahead &a = ctx->ahead;
  if (!a.valid || a.unit.size() != units.size()) return false;
  for (size_t i = 0; i < units.v.size(); i++) {
    const Unit *u = units.v[i];
    if (u != a.unit[i] || u->rig->serial != a.serial[i] || u->rig->epoch != a.epoch[i] || u->rig->feeds != a.feeds[i] + 1 ||
        (char)u->stereo != a.stereo[i] || u->first || u->failed || u->L.cam != a.cam_l[i] || (u->stereo && u->R.cam != a.cam_r[i]))
      return false;
  }
  return true;
}

bool background_enabled() {   // UVIO_KLT_BACKGROUND=1 turns it on (off until it has been through the GPU suite)
  static const bool v = [] {
    const char *e = getenv("UVIO_KLT_BACKGROUND");
    return e && e[0] == '1';
  }();
  return v;
}

// End of a pass: issue the next pass's detection (see Lane::Ahead).  All or nothing per lane.
void Pass::issue_ahead() {
  Lane::Ahead &a = ctx->ahead;
  a.valid = false;
  a.level = 0;
  // (when FAST itself is being timed, every feed issues its own detection inside the timed window)
  if (!ahead_enabled() || (ctx->prof_on && ((ctx->prof.mask >> K_FAST) & 1u)) || rc != OK || units.size() == 0) return;
  for (auto &u : units) {
    if (u.failed || u.req->status != OK) return;
    const bool empty = u.stereo ? (u.L.cam->pts_last.empty() && u.R.cam->pts_last.empty()) : u.L.cam->pts_last.empty();
    if (empty) return;                              // its next feed starts over on the new frame (TrackKLT.cpp:110, :221)
  }
  a.serial.clear(); a.epoch.clear(); a.feeds.clear(); a.unit.clear(); a.stereo.clear(); a.cam_l.clear(); a.cam_r.clear();
  for (auto &u : units) {
    u.first = false;
    a.cam_l.push_back(u.L.cam);
    a.cam_r.push_back(u.stereo ? u.R.cam : nullptr);
    a.serial.push_back(u.rig->serial);
    a.epoch.push_back(u.rig->epoch);
    a.feeds.push_back(u.rig->feeds);
    a.unit.push_back(&u);
    a.stereo.push_back((char)u.stereo);
  }
  bool any_first = false;
  int cw, ch;
  if (issue_detection(true, any_first, cw, ch) != OK || any_first) return;
  a.njobs = njobs;
  a.valid = true;
  a.level = 1;
}

// After feed_batch has returned: the rest of the NEXT feed's detection for the units of this pass -- wait for the FAST that
// issue_ahead launched, per-cell top-K, sub-pixel + stereo LK (round 2), then ids and the right-image pass -- so that the
// next feed starts at its temporal LK.  Nothing here touches what the caller may read meanwhile (observations, pts_last);
// the trackers' id counters are advanced and put back by cancel_ahead if the result is not used.
void Pass::background() {
  Lane::Ahead &a = ctx->ahead;
  if (!a.valid || a.level != 1 || !background_enabled()) return;
  if (ctx->prof_on && ((ctx->prof.mask >> K_SUBPIX) & 1u)) return;   // sub-pixel timed: keep it inside the timed window
  cudaSetDevice(ctx->dev->device);
  g_prof = ctx->prof_on ? &ctx->prof : nullptr;
  auto give_up = [&] {
    a.valid = false;
    a.level = 0;
    g_prof = nullptr;
  };
  if (ctx->sync() != OK) return give_up();
  njobs = a.njobs;
  tmark = now_us();
  if (stage1() != OK || ctx->sync() != OK) return give_up();
  a.currid_before.clear();
  for (auto &u : units) a.currid_before.push_back(u.rig->currid);
  bool ok = true;
  for (auto &u : units) ok = ok && phase_c_detect(ctx, u) == OK;
  a.level = 2;            // from here on cancel_ahead puts the id counters back
  a.nsp = nsp;
  a.nlk = nlk;
  for (auto &u : units) u.rig->bg_lane = ctx->index;
  if (!ok) cancel_ahead(*ctx);
  g_prof = nullptr;
}

int Pass::stage0() {
  int e = stage0_book();
  if (e != OK) return e;
  return stage0_issue();
}

// host bookkeeping of the new messages: frame slots, masks, pre-processing jobs, units (shared by the host and the device path)
int Pass::stage0_book() {
  Context *dev = ctx->dev;
  t0 = tmark = now_us();
  ctx->prof_begin();
  g_prof = ctx->prof_on ? &ctx->prof : nullptr;
  // ---- host bookkeeping of the new messages: frame slots, masks, pre-processing jobs, units ---------------------------------
  int n_mask_jobs = 0;
  size_t mask_in_bytes = 0, mask_out_bytes = 0;
  std::vector<PreItem> &items = ctx->pre_items;
  items.clear();
  wait_evts = 0;
  std::vector<HostImage> &zero_masks = ctx->zero_masks;   // masks of this pass already found to be all zero
  zero_masks.clear();
  unsigned long long need_seq = 0;   // last announced batch this pass consumes: its event must have been recorded
  for (int i = 0; i < n; i++) {
    FeedRequest &rq = *reqs[i];
    Rig &rig = *rq.rig;
    const Params &P = rig.prm;
    rig.feeds++;
    for (int k = 0; k < rq.n; k++) {
      CamState &c = *rig.find_cam(rq.sensor_ids[k]);
      // frame number n_fed + 1 of this camera: announced earlier (submit_batch; already on its way or done) or uploaded now
      const bool announced = c.n_sub > c.n_fed;
      if (!announced) c.n_sub++;
      c.n_fed++;
      c.cur = (int)(c.n_fed % PYR_SLOTS);
      if (announced) {
        wait_evts |= 1u << c.sub_evt[c.cur];
        need_seq = std::max(need_seq, c.sub_seq[c.cur]);
      }
      else items.push_back(PreItem{&c, rq.images[k], c.cur, P.histogram_method, P.downsample_cameras != 0, (int)(c.n_sub & 1u)});
      // the same mask memory given to several cameras of this call (a rig's shared mask, a batch's common one) is scanned once
      const HostImage &mk = rq.masks[k];
      bool known_zero = false;
      for (const HostImage &z : zero_masks) known_zero |= z.p == mk.p && z.stride == mk.stride && z.w == mk.w && z.h == mk.h;
      if (known_zero) c.mask_nonzero[c.cur] = false;
      else if (!P.downsample_cameras) c.mask_nonzero[c.cur] = copy_mask(mk, c.mask[c.cur], c.w, c.h);
      else c.mask_nonzero[c.cur] = mask_any_set(mk, c.in_w, c.in_h);
      if (!c.mask_nonzero[c.cur] && mk.p && !known_zero && zero_masks.size() < 16) zero_masks.push_back(mk);
      if (P.downsample_cameras && c.mask_nonzero[c.cur]) {   // halved on the device below
        n_mask_jobs++;
        mask_in_bytes += (size_t)((c.in_w + 15) / 16 * 16) * c.in_h;
        mask_out_bytes += (size_t)((c.w + 3) / 4 * 4) * c.h;
      }
    }
    // units (TrackKLT.cpp:80-93)
    auto make_side = [&](Side &s, int k) { s.cam = rig.find_cam(rq.sensor_ids[k]); s.msg = k; };
    if (rq.n == 2 && P.use_stereo) {
      Unit &u = *rig.unit_at(0);
      u.req = &rq; u.stereo = true;
      make_side(u.L, 0);
      make_side(u.R, 1);
      u.first = u.L.cam->no_points() && u.R.cam->no_points();                                         // :221
      if (!u.first && (!u.L.cam->have_last || !u.R.cam->have_last)) { fail(rq, E_BADMSG); continue; }
      units.push_back(&u);
    } else {
      for (int k = 0; k < rq.n; k++) {
        Unit &u = *rig.unit_at((size_t)k);
        u.req = &rq; u.stereo = false;
        make_side(u.L, k);
        u.first = u.L.cam->no_points();                                                                // :110
        units.push_back(&u);
      }
    }
  }
  if (need_seq && dev->wait_issued(need_seq) != OK) { ctx->cuda_err = dev->cuda_err; return E_CUDA; }
  if (n_mask_jobs > 0) {
    // downsample_cameras with masks that mask something: cv::pyrDown of the mask (VioManager.cpp:275-277).  The host
    // bookkeeping below reads the halved mask (first frames detect on the frame being fed), so this small job runs on the
    // idle main stream and is waited for; the frames themselves go up on the upload stream meanwhile.
    CU_TRY(ctx->mask_jobs.reserve((size_t)n_mask_jobs));
    CU_TRY(ctx->mask_in.reserve(mask_in_bytes));
    CU_TRY(ctx->mask_out.reserve(mask_out_bytes));
    size_t in_off = 0, out_off = 0;
    int nj = 0, max_w = 1, max_h = 1;
    for (int i = 0; i < n; i++) {
      FeedRequest &rq = *reqs[i];
      if (!rq.rig->prm.downsample_cameras) continue;
      for (int k = 0; k < rq.n; k++) {
        CamState &c = *rq.rig->find_cam(rq.sensor_ids[k]);
        if (!c.mask_nonzero[c.cur]) continue;
        const int ip = (c.in_w + 15) / 16 * 16, op = (c.w + 3) / 4 * 4;
        CU_TRY(cudaMemcpy2DAsync(ctx->mask_in.d + in_off, (size_t)ip, rq.masks[k].p, rq.masks[k].stride, (size_t)c.in_w, (size_t)c.in_h,
                                 cudaMemcpyDefault, ctx->stream));
        ctx->mask_jobs.h[nj++] = DownJob{ImgView{ctx->mask_in.d + in_off, c.in_w, c.in_h, ip}, ImgView{ctx->mask_out.d + out_off, c.w, c.h, op}};
        in_off += (size_t)ip * c.in_h;
        out_off += (size_t)op * c.h;
        max_w = std::max(max_w, c.w);
        max_h = std::max(max_h, c.h);
        ctx->stats[2] += (unsigned long long)c.in_w * c.in_h;
        ctx->stats[3] += (unsigned long long)c.w * c.h;
      }
    }
    CU_TRY(ctx->mask_jobs.upload(ctx->stream, (size_t)nj));
    CU_TRY(launch_downsample(ctx->stream, ctx->mask_jobs.d, nj, max_w, max_h));
    ctx->stats[1] += 1;
    if (ctx->sync() != OK) return E_CUDA;
    out_off = 0;
    for (int i = 0; i < n; i++) {
      FeedRequest &rq = *reqs[i];
      if (!rq.rig->prm.downsample_cameras) continue;
      for (int k = 0; k < rq.n; k++) {
        CamState &c = *rq.rig->find_cam(rq.sensor_ids[k]);
        if (!c.mask_nonzero[c.cur]) continue;
        const int op = (c.w + 3) / 4 * 4;
        for (int y = 0; y < c.h; y++) memcpy(c.mask[c.cur].data() + (size_t)y * c.w, ctx->mask_out.h + out_off + (size_t)y * op, (size_t)c.w);
        out_off += (size_t)op * c.h;
      }
    }
  }
  if (!ctx->sp_mask_ready) {
    CU_TRY(ctx->sp_mask.reserve(SP_N));
    CU_TRY(upload_subpix_mask(ctx->sp_mask.d, ctx->stream));
    ctx->sp_mask_ready = true;
  }
  return OK;
}

int Pass::stage0_issue() {
  Context *dev = ctx->dev;
  std::vector<PreItem> &items = ctx->pre_items;
  // ---- round 1a: FAST.  Detection runs on the previous frame's pyramid, which is already in HBM, so it is issued ahead of
  // the upload (its job list would otherwise queue behind the images on the copy engine) -- or was issued when the previous
  // feed of these trackers ended.  First frames detect on the frame that is being uploaded and wait for it.
  int max_cw = 1, max_ch = 1;
  bool any_first = false;
  if (ahead_usable()) {
    njobs = ctx->ahead.njobs;
    if (ctx->ahead.level == 2) {
      detected = true;
      nsp = ctx->ahead.nsp;
      nlk = ctx->ahead.nlk;
      for (auto &u : units) u.rig->bg_lane = -1;
    }
    ctx->ahead.valid = false;
    ctx->ahead.level = 0;
  } else {
    cancel_ahead(*ctx);
    int e = issue_detection(false, any_first, max_cw, max_ch);
    if (e != OK) return e;
  }
  // ---- round 1b: upload + equalise + pyramid on the upload stream (frames announced by submit_batch are on their way) -------
  npre = (int)items.size();
  {
    const PreBuffers B{&ctx->pre_jobs, &ctx->down_jobs, &ctx->clahe_lut};
    int e = issue_preprocess(dev, ctx->up_stream, B, items, ctx->stats);
    if (e != OK) { ctx->cuda_err = dev->cuda_err; return e; }
    CU_TRY(cudaEventRecord(ctx->up_done, ctx->up_stream));
  }
  if (any_first) {
    CU_TRY(wait_new_frames());
    CU_TRY(launch_fast(ctx->stream, ctx->fast_jobs.d, njobs, max_cw, max_ch, dev->pyr_dev.d, ctx->fast_out_dev.d, ctx->fast_cnt_dev.d,
                       ctx->fast_ticket.d, ctx->fast_out.d, ctx->fast_cnt.d));
  }
  seg(0);
  return OK;
}

int Pass::stage1() {
  Context *dev = ctx->dev;
  ctx->stats[1] += njobs > 0;
  for (int j = 0; j < njobs; j++) ctx->stats[5] += (unsigned long long)ctx->fast_jobs.h[j].w * ctx->fast_jobs.h[j].h;
  for (int j = 0; j < njobs; j++) ctx->stats[3] += 4ull * ctx->fast_cnt.h[j] + 4ull;
  t1 = now_us();
  // ---- round 2: sub-pixel + stereo LK --------------------------------------------------------------------------------
  size_t sp_bound = 0;
  for (auto &u : units)
    for (int side = 0; side < (u.stereo ? 2 : 1); side++) {
      Side &s = side ? u.R : u.L;
      if (s.detect) sp_bound += s.cells.size() * (size_t)s.geom.per_cell;
    }
  CU_TRY(ctx->sp_items.reserve(sp_bound));
  CU_TRY(ctx->sp_out.reserve(2 * sp_bound));
  CU_TRY(ctx->sp_out_dev.reserve(2 * sp_bound));
  CU_TRY(ctx->lk_items.reserve(sp_bound));
  CU_TRY(ctx->lk_out.reserve(sp_bound));
  CU_TRY(ctx->lk_status.reserve(sp_bound));
  CU_TRY(ctx->lk_out_dev.reserve(sp_bound));
  CU_TRY(ctx->lk_status_dev.reserve(sp_bound));
  nsp = 0;
  nlk = 0;
  {
    // sub-pixel items: all left selections first (so that stereo-LK item i chains on sub-pixel item i), then the right
    std::vector<uint32_t> tmp;
    std::vector<uint32_t> grid;
    for (auto &u : units) phase_b(ctx, u, nsp, nlk, tmp, grid);
    // a unit's lk0 must equal its sp0L for the chaining: true when only stereo units select points; mono units add
    // sub-pixel items without LK items, so stereo units are given explicit coordinates below instead
    for (auto &u : units)
      if (u.stereo) phase_b_right(ctx, u, nsp);
  }
  HOSTPROF(s1_launches);
  CU_TRY(ctx->sp_items.upload(ctx->stream, (size_t)nsp));
  CU_TRY(launch_subpix(ctx->stream, ctx->sp_items.d, nsp, dev->pyr_dev.d, ctx->sp_mask.d, ctx->sp_out.d, ctx->sp_out_dev.d));
  // stereo LK items read their start point from the sub-pixel output: build the index map
  bool chained = true;
  for (auto &u : units) chained &= (!u.stereo || u.lk0 == u.sp0L);
  if (nlk > 0) {
    if (!chained) {
      // mixed batch: wait for the refined corners and fill the LK items on the host
      if (ctx->sync() != OK) return E_CUDA;
      for (auto &u : units)
        if (u.stereo)
          for (size_t i = 0; i < u.selL.size(); i++) {
            LkItem &it = ctx->lk_items.h[u.lk0 + i];
            it.px = it.nx = ctx->sp_out.h[2 * (u.sp0L + i)];
            it.py = it.ny = ctx->sp_out.h[2 * (u.sp0L + i) + 1];
          }
    }
    CU_TRY(ctx->lk_items.upload(ctx->stream, (size_t)nlk));
    // chained and with its items in mapped memory, the launch may overlap the sub-pixel kernel (it waits for it on the device)
    CU_TRY(launch_lk(ctx->stream, ctx->lk_items.d, nlk, dev->pyr_dev.d, chained ? ctx->sp_out_dev.d : nullptr, 30, 0.01,
                     ctx->lk_out.d, ctx->lk_status.d, ctx->lk_out_dev.d, ctx->lk_status_dev.d, chained && ctx->lk_items.mapped && pdl_on()));
  }
  seg(2);
  return OK;
}

int Pass::stage2() {
  Context *dev = ctx->dev;
  ctx->stats[1] += (nsp > 0) + (nlk > 0);
  ctx->stats[3] += 8ull * nsp + 9ull * nlk;
  // ---- host: finish detection; round 3: temporal LK + RANSAC -----------------------------------------------------------
  size_t lk_bound = 0;
  for (auto &u : units) {
    int e = detected ? OK : phase_c_detect(ctx, u);
    if (e != OK) { fail(*u.req, e); u.failed = true; continue; }
    if (u.first) {                                                   // TrackKLT.cpp:110-123 / :221-240: store, no DB update
      store_last(u.L, u.L.pts, u.L.ids);
      if (u.stereo) store_last(u.R, u.R.pts, u.R.ids);
      continue;
    }
    lk_bound += u.L.pts.size() + (u.stereo ? u.R.pts.size() : 0);
  }
  t2 = now_us();
  CU_TRY(ctx->lk_items.reserve(lk_bound));
  CU_TRY(ctx->lk_out.reserve(lk_bound));
  CU_TRY(ctx->lk_status.reserve(lk_bound));
  CU_TRY(ctx->lk_out_dev.reserve(lk_bound));
  CU_TRY(ctx->lk_status_dev.reserve(lk_bound));
  CU_TRY(ctx->rs_jobs.reserve(2 * units.size()));
  CU_TRY(ctx->p0n_dev.reserve(2 * lk_bound));
  CU_TRY(ctx->p1n_dev.reserve(2 * lk_bound));
  CU_TRY(ctx->p1n.reserve(2 * lk_bound));
  CU_TRY(ctx->rs_mask.reserve(lk_bound));
  CU_TRY(ctx->rs_mask_dev.reserve(lk_bound));
  nlk = 0;
  nrs = 0;
  int max_n = 1;
  for (auto &u : units) {
    if (u.first || u.failed) continue;
    emit_temporal(ctx, u, nlk, nrs);
    max_n = std::max(max_n, std::max(u.L.lk_n, u.R.lk_n));
  }
  HOSTPROF(s2_launches);
  CU_TRY(wait_new_frames());   // the temporal LK reads the new pyramids
  CU_TRY(ctx->lk_items.upload(ctx->stream, (size_t)nlk));
  CU_TRY(ctx->rs_jobs.upload(ctx->stream, (size_t)nrs));
  CU_TRY(launch_lk(ctx->stream, ctx->lk_items.d, nlk, dev->pyr_dev.d, nullptr, 30, 0.01, ctx->lk_out.d, ctx->lk_status.d,
                   ctx->lk_out_dev.d, ctx->lk_status_dev.d));
  CU_TRY(launch_ransac(ctx->stream, ctx->rs_jobs.d, nrs, max_n, ctx->lk_items.d, ctx->lk_out_dev.d, ctx->lk_status_dev.d,
                       dev->cam_dev.d, ctx->p0n_dev.d, ctx->p1n_dev.d, ctx->p1n.d, ctx->rs_mask.d, ctx->rs_mask_dev.d,
                       nlk > 0 && ctx->rs_jobs.mapped && ctx->lk_items.mapped && pdl_on()));
  seg(4);
  return OK;
}

int Pass::stage3() {
  ctx->stats[1] += (nlk > 0) + (nrs > 0);
  ctx->stats[3] += 9ull * nlk + (nrs > 0 ? 9ull * nlk : 0ull);
  ctx->stats[0] += (unsigned long long)n;
  t3 = now_us();
  for (auto &u : units)
    if (!u.first && !u.failed) phase_d(ctx, u);
  for (auto &u : units) {   // a device-side copy of this state, if any, is stale now
    u.rig->dv_dirty = true;
    u.rig->dv_mirror_stale = false;
    for (auto &c : u.rig->cams) c.dv_n_last = -1;
  }
  const double t4 = now_us();
  issue_ahead();
  seg(6);
  ctx->seg_us[7] += 1;
  ctx->prof_end();
  for (int i = 0; i < n; i++) {
    float *s = reqs[i]->rig->stage_us;
    s[0] = (float)(t1 - t0);   // upload + equalise + pyramid (+ FAST of round 1)
    s[1] = (float)(t2 - t1);   // detection (top-K, sub-pixel, id assignment)
    s[2] = (float)(t3 - t2);   // temporal KLT + RANSAC
    s[3] = 0.f;                // stereo KLT is part of the detection round here
    s[4] = (float)(t4 - t3);   // merge + observations
  }
  return OK;
}



}  // namespace uvio
