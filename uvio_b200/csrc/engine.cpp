// Per-device context and the batched frame pipeline of the GPU TrackKLT.  See engine.h for the division of labour.
//
// One frame of a batch of rigs costs three device round trips, whatever the batch size:
//   round 1  upload + equalise + pyramid of the new frames; FAST on every cell that may need new features
//            (left cells, and a superset of the right cells that is cheap to prove sufficient - see phase_a)
//   round 2  cornerSubPix of the per-cell top-K corners; left->right LK of the new left corners (chained on the device)
//   round 3  temporal LK of every tracked point, undistortion, fundamental-matrix RANSAC
// Between rounds the host runs the reference's order-dependent bookkeeping (hostlogic.h).
#include "engine.h"

#include <stdio.h>

#include <chrono>
#include <map>
#include <memory>

#include "hostlogic.h"

namespace uvio {

#define CU_TRY(expr)                                          \
  do {                                                        \
    cudaError_t e__ = (expr);                                 \
    if (e__ != cudaSuccess) { ctx->cuda_err = (int)e__; return E_CUDA; } \
  } while (0)

template <class T>
struct MappedVec {   // pinned host memory mapped into the device address space (job lists in, small results out)
  T *h = nullptr, *d = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t want) {   // contents are not preserved; call only while the stream is idle
    if (want <= cap) return cudaSuccess;
    size_t ncap = want + want / 2 + 64;
    if (h) cudaFreeHost(h);
    h = d = nullptr;
    cap = 0;
    cudaError_t e = cudaHostAlloc((void **)&h, ncap * sizeof(T), cudaHostAllocMapped);
    if (e != cudaSuccess) return e;
    e = cudaHostGetDevicePointer((void **)&d, h, 0);
    if (e != cudaSuccess) return e;
    cap = ncap;
    return cudaSuccess;
  }
  ~MappedVec() { if (h) cudaFreeHost(h); }
};

template <class T>
struct DevVec {
  T *d = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t want, bool zero = false) {
    if (want <= cap) return cudaSuccess;
    size_t ncap = want + want / 2 + 64;
    if (d) cudaFree(d);
    d = nullptr;
    cap = 0;
    cudaError_t e = cudaMalloc((void **)&d, ncap * sizeof(T));
    if (e != cudaSuccess) return e;
    if (zero) e = cudaMemset(d, 0, ncap * sizeof(T));
    cap = ncap;
    return e;
  }
  ~DevVec() { if (d) cudaFree(d); }
};

class Context {
 public:
  int device = 0;
  cudaStream_t stream = nullptr;
  int cuda_err = 0;
  std::mutex mtx;
  // tables (host copy + device copy, re-uploaded when dirty)
  std::vector<PyrDesc> pyr_tab;
  std::vector<ImgView> raw_tab;
  std::vector<CamModel> cam_tab;
  DevVec<PyrDesc> pyr_dev;
  DevVec<ImgView> raw_dev;
  DevVec<CamModel> cam_dev;
  bool tables_dirty = true;
  std::vector<void *> allocs;   // image memory of all cameras
  // round 1
  MappedVec<PreJob> pre_jobs;
  DevVec<unsigned> hist, hist_ticket;
  DevVec<uint8_t> lut, clahe_lut;
  MappedVec<FastJob> fast_jobs;
  MappedVec<int> fast_thr;
  DevVec<uint32_t> fast_out_dev;
  DevVec<unsigned> fast_cnt_dev, fast_ticket;
  MappedVec<uint32_t> fast_out;
  MappedVec<unsigned> fast_cnt;
  // round 2 / 3
  DevVec<float> sp_mask;
  bool sp_mask_ready = false;
  MappedVec<SubpixItem> sp_items;
  MappedVec<float> sp_out;
  DevVec<float> sp_out_dev;
  MappedVec<LkItem> lk_items;
  MappedVec<LkOut> lk_out;
  MappedVec<uint8_t> lk_status;
  DevVec<LkOut> lk_out_dev;
  DevVec<uint8_t> lk_status_dev;
  MappedVec<RansacJob> rs_jobs;
  DevVec<float> p0n_dev, p1n_dev;
  MappedVec<float> p1n;
  MappedVec<uint8_t> rs_mask;
  DevVec<uint8_t> rs_mask_dev;

  int sync() {
    cudaError_t e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) { cuda_err = (int)e; return E_CUDA; }
    return OK;
  }
  int upload_tables() {
    Context *ctx = this;
    if (!tables_dirty) return OK;
    CU_TRY(cudaStreamSynchronize(stream));
    CU_TRY(pyr_dev.reserve(pyr_tab.size()));
    CU_TRY(raw_dev.reserve(raw_tab.size()));
    CU_TRY(cam_dev.reserve(cam_tab.size()));
    if (!pyr_tab.empty()) CU_TRY(cudaMemcpy(pyr_dev.d, pyr_tab.data(), pyr_tab.size() * sizeof(PyrDesc), cudaMemcpyHostToDevice));
    if (!raw_tab.empty()) CU_TRY(cudaMemcpy(raw_dev.d, raw_tab.data(), raw_tab.size() * sizeof(ImgView), cudaMemcpyHostToDevice));
    if (!cam_tab.empty()) CU_TRY(cudaMemcpy(cam_dev.d, cam_tab.data(), cam_tab.size() * sizeof(CamModel), cudaMemcpyHostToDevice));
    tables_dirty = false;
    return OK;
  }
  // device memory for one camera: a raw upload buffer and two pyramids
  int alloc_camera(CamState &c, int w, int h, int pyr_levels, int win) {
    Context *ctx = this;
    const int L = pyr_max_level(w, h, pyr_levels, win);
    if (L + 1 > MAX_LEVELS) return E_ARG;
    auto pitch_of = [](int ww) { return (ww + 127) / 128 * 128; };
    size_t raw_bytes = (size_t)pitch_of(w) * h, pyr_bytes = 0;
    int lw = w, lh = h;
    for (int l = 0; l <= L; l++) {
      pyr_bytes += (size_t)pitch_of(lw) * lh;
      lw = (lw + 1) / 2;
      lh = (lh + 1) / 2;
    }
    uint8_t *base = nullptr;
    CU_TRY(cudaMalloc((void **)&base, raw_bytes + 2 * pyr_bytes + 256));
    CU_TRY(cudaMemset(base, 0, raw_bytes + 2 * pyr_bytes + 256));
    allocs.push_back(base);
    c.w = w;
    c.h = h;
    c.raw = (int)raw_tab.size();
    raw_tab.push_back(ImgView{base, w, h, pitch_of(w)});
    uint8_t *p = base + raw_bytes;
    for (int s = 0; s < 2; s++) {
      PyrDesc d;
      memset(&d, 0, sizeof(d));
      d.nlv = L + 1;
      lw = w;
      lh = h;
      for (int l = 0; l <= L; l++) {
        d.lv[l] = ImgView{p, lw, lh, pitch_of(lw)};
        p += (size_t)pitch_of(lw) * lh;
        lw = (lw + 1) / 2;
        lh = (lh + 1) / 2;
      }
      c.pyr[s] = (int)pyr_tab.size();
      pyr_tab.push_back(d);
      c.mask[s].assign((size_t)w * h, 0);
      c.mask_nonzero[s] = false;
    }
    c.paint.assign((size_t)w * h, 0);
    tables_dirty = true;
    return OK;
  }
  ~Context() {
    for (void *p : allocs) cudaFree(p);
    if (stream) cudaStreamDestroy(stream);
  }
};

Context *context_for_device(int device, int *err) {
  static std::mutex m;
  static std::map<int, std::unique_ptr<Context>> table;
  std::lock_guard<std::mutex> lk(m);
  auto it = table.find(device);
  if (it != table.end()) return it->second.get();
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || device < 0 || device >= count) {   // no CUDA device: there is no other way to run this tracker
    if (err) *err = e != cudaSuccess ? (int)e : (int)cudaErrorInvalidDevice;
    return nullptr;
  }
  std::unique_ptr<Context> c(new Context);
  c->device = device;
  if ((e = cudaSetDevice(device)) != cudaSuccess || (e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess) {
    if (err) *err = (int)e;
    return nullptr;
  }
  Context *p = c.get();
  table[device] = std::move(c);
  return p;
}

Rig::Rig(Context *c, const Params &p, const CamCalib *cc, int n_cams) : ctx(c), prm(p) {
  currid = 4 * (size_t)p.num_aruco + 1;   // TrackBase.cpp:34
  std::lock_guard<std::mutex> lk(c->mtx);
  for (int i = 0; i < n_cams; i++) {
    CamState s;
    s.cam_id = cc[i].cam_id;
    s.cam_model_idx = (int)c->cam_tab.size();
    s.max_focal = cc[i].model.k[0] > cc[i].model.k[1] ? cc[i].model.k[0] : cc[i].model.k[1];
    c->cam_tab.push_back(cc[i].model);
    cam_index[s.cam_id] = (int)cams.size();
    cams.push_back(std::move(s));
  }
  c->tables_dirty = true;
}
Rig::~Rig() {}
CamState *Rig::find_cam(size_t id) {
  auto it = cam_index.find(id);
  return it == cam_index.end() ? nullptr : &cams[it->second];
}
