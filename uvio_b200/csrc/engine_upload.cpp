// Upload of the new frames (copies merged where host and device layouts allow, cut into pieces), the optional halving and the
// image-wide passes (equalisation / CLAHE, pyramid levels) of one batch of frames.
#include "engine_internal.h"

namespace uvio {

// H2D of the frames (merged into as few copies as their layout allows), the optional halving, equalisation and pyramids,
// all on stream st.  Frames already in device memory are consumed where they lie.
// copy_st / copied: when given, the copies go to a stream of their own and st waits for them through the event, so that the
// upload of one batch overlaps the kernels of the batch before it (the frames alternate between two upload buffers).
int issue_preprocess(Context *ctx, cudaStream_t st, const PreBuffers &B, const std::vector<PreItem> &items, unsigned long long *stats,
                            int group, cudaStream_t copy_st, cudaEvent_t copied) {
  const bool split = copy_st != nullptr && copy_st != st;
  if (!split) copy_st = st;   // (with a copy stream the groups still apply: each group's kernels wait for its own copies only)
  HOSTPROF(s0_memcpy);
  const int nimg = (int)items.size();
  if (nimg == 0) return OK;
  int any_clahe = 0, max_band = 0, ndown = 0, max_dw = 1, max_dh = 1, n_down_jobs = 0, max_w = 0, max_h = 0, max_nlv = 0;
  std::vector<char> on_dev((size_t)nimg, 0);
  for (auto &it : items) {
    any_clahe |= it.method == 2;
    n_down_jobs += it.downsample;
  }
  CU_TRY(B.pre->reserve((size_t)nimg));
  if (n_down_jobs) CU_TRY(B.down->reserve((size_t)n_down_jobs));
  if (any_clahe) CU_TRY(B.clahe->reserve((size_t)nimg * CLAHE_TILES * CLAHE_TILES * 256));
  // Copies whose source and destination both continue the previous one are merged (a batch of tightly packed frames is one copy).
  uint8_t *pend_dst = nullptr;
  const uint8_t *pend_src = nullptr;
  size_t pend_bytes = 0;
  // A long copy is cut into pieces: the feed that runs meanwhile needs the same PCIe link for its launches and its results, and
  // a transfer of many megabytes ahead of them stretches its chain (round 1: ~100 us per host round trip behind a 20 MB copy,
  // tools/pcie_probe.py).  Between pieces the link takes their requests.  Measured with the device-side pipeline (64 streams,
  // UVIO_KLT_COPYPROF): 1 MiB pieces upload at 42 GB/s beside the kernels, 2 MiB at 47, 3 MiB at 49, 8 MiB at 51, whole batches
  // at 54; with the chain replaying its CUDA graphs the end-to-end rate peaks at 3 - 4 MiB (67.4 k frames/s; 65.0 k at 2 MiB,
  // 66.5 k at 8 MiB; 58 k at round 1's 1 MiB).
  static const size_t piece = [] {
    const char *e = getenv("UVIO_KLT_UPLOAD_PIECE");
    long v = e ? atol(e) : (3l << 20);
    return (size_t)(v < 4096 ? 4096 : v);
  }();
  // UVIO_KLT_COPY_STREAMS=2: a second stream (and copy engine) for every other piece of an announced batch
  struct Second {
    cudaStream_t st = nullptr;
    cudaEvent_t joined = nullptr;
    Second() {
      const char *e = getenv("UVIO_KLT_COPY_STREAMS");
      if (e && atoi(e) >= 2) {
        int lo = 0, hi = 0;
        cudaDeviceGetStreamPriorityRange(&lo, &hi);
        if (cudaStreamCreateWithPriority(&st, cudaStreamNonBlocking, hi) != cudaSuccess) st = nullptr;
        if (st && cudaEventCreateWithFlags(&joined, cudaEventDisableTiming) != cudaSuccess) st = nullptr;
      }
    }
  };
  static Second second;      // (one device per process in the configurations this knob is measured on)
  cudaStream_t copy_st2 = split ? second.st : nullptr;
  unsigned n_piece = 0;
  if (copy_st2) {               // whatever the first copy stream has been told to wait for (the readers of the upload buffers) holds for the second
    CU_TRY(cudaEventRecord(second.joined, copy_st));
    CU_TRY(cudaStreamWaitEvent(copy_st2, second.joined, 0));
  }
  auto flush = [&]() -> cudaError_t {
    cudaError_t e = cudaSuccess;
    for (size_t off = 0; off < pend_bytes && e == cudaSuccess; off += piece) {
      // with a second copy stream the pieces alternate between two copy engines: the gap between two pieces of one stream is
      // filled by the other's transfer (1 MiB pieces on one stream reach 42 GB/s beside the tracker's kernels, whole batches 54)
      cudaStream_t cs = (copy_st2 && (n_piece++ & 1u)) ? copy_st2 : copy_st;
      e = cudaMemcpyAsync(pend_dst + off, pend_src + off, std::min(piece, pend_bytes - off), cudaMemcpyHostToDevice, cs);
    }
    pend_bytes = 0;
    return e;
  };
  // Pass 1: the job lists (one small copy each, ahead of the frames).
  { HOSTPROF(pre_pass1);
  for (int j = 0; j < nimg; j++) {
    const PreItem &it = items[j];
    CamState &c = *it.cam;
    const HostImage &im = it.im;
    B.pre->h[j] = PreJob{ctx->raw_tab[it.downsample ? c.raw : c.raw_in[it.par]], c.pyr[it.slot], it.method};
    max_band = std::max(max_band, ((c.h + 7) / 8 + 1) * c.w);
    max_w = std::max(max_w, c.w);
    max_h = std::max(max_h, c.h);
    max_nlv = std::max(max_nlv, ctx->pyr_tab[c.pyr[it.slot]].nlv);
    // where does the frame live?  A frame that starts where the previous one ended lies in the same allocation (a packed ring of
    // frames is asked about once per batch: the query costs about a microsecond, 128 of them delay the copies of a batch by more
    // than its pre-processing takes)
    bool on_device;
    if (j > 0 && items[j - 1].im.p + items[j - 1].im.stride * (size_t)items[j - 1].im.h == im.p) {
      on_device = on_dev[j - 1] != 0;
    } else {
      cudaPointerAttributes attr;
      on_device = cudaPointerGetAttributes(&attr, im.p) == cudaSuccess && (attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged);
      if (!on_device) cudaGetLastError();   // an unregistered host pointer is not an error
    }
    on_dev[j] = on_device;
    if (it.downsample) {
      const ImgView full = on_device ? ImgView{im.p, im.w, im.h, (int)im.stride} : ctx->raw_tab[c.raw_in[it.par]];
      B.down->h[ndown++] = DownJob{full, ctx->raw_tab[c.raw]};
      max_dw = std::max(max_dw, c.w);
      max_dh = std::max(max_dh, c.h);
    } else if (on_device) {
      B.pre->h[j].src = ImgView{im.p, im.w, im.h, (int)im.stride};
    }
  }
  }
  CU_TRY(B.pre->upload(copy_st, (size_t)nimg));
  if (ndown > 0) CU_TRY(B.down->upload(copy_st, (size_t)ndown));
  // Pass 2: group by group, the frames go up and are (halved and) pre-processed.  A feed issues one group; submit_batch cuts
  // a batch into several so that the pyramids of the first frames are built while the last ones are still on the link, and
  // so that no launch takes every SM from the rounds of the feed that is running meanwhile.
  if (group <= 0 || group > nimg) group = nimg;
  int down_at = 0;
  // UVIO_KLT_COPYPROF=1: CUDA-event time of each batch's copies on the copy stream (how fast the upload runs beside the kernels
  // of the feed in flight), averaged and printed when the process ends.  Development aid.
  struct CopyProf {
    bool on = getenv("UVIO_KLT_COPYPROF") != nullptr;
    cudaEvent_t ev[8][2] = {};
    bool used[8] = {};
    unsigned long long n = 0, bytes = 0;
    double ms = 0;
    unsigned k = 0;
    ~CopyProf() {
      if (on && n) fprintf(stderr, "[copyprof] %llu batches, %.3f ms per batch, %.2f GB/s while the tracker runs\n", n, ms / n, bytes / (ms * 1e6));
    }
  };
  static CopyProf cp;
  const bool cp_now = cp.on && split;
  unsigned cp_slot = 0;
  if (cp_now) {
    cp_slot = cp.k++ % 8;
    if (!cp.ev[cp_slot][0]) { cudaEventCreate(&cp.ev[cp_slot][0]); cudaEventCreate(&cp.ev[cp_slot][1]); }
    if (cp.used[cp_slot]) {
      float t = 0;
      if (cudaEventSynchronize(cp.ev[cp_slot][1]) == cudaSuccess && cudaEventElapsedTime(&t, cp.ev[cp_slot][0], cp.ev[cp_slot][1]) == cudaSuccess) {
        cp.ms += t;
        cp.n++;
      }
    }
    cudaEventRecord(cp.ev[cp_slot][0], copy_st);
  }
  const unsigned long long cp_bytes0 = stats[2];
  HOSTPROF(pre_pass2);
  for (int g0 = 0; g0 < nimg; g0 += group) {
    const int g1 = std::min(nimg, g0 + group);
    int gdown = 0;
    for (int j = g0; j < g1; j++) {
      const PreItem &it = items[j];
      const HostImage &im = it.im;
      const ImgView &raw = ctx->raw_tab[it.cam->raw_in[it.par]];
      gdown += it.downsample;
      if (on_dev[j]) continue;
      stats[2] += (unsigned long long)im.w * im.h;
      if ((size_t)raw.pitch == im.stride) {
        const size_t bytes = (size_t)raw.pitch * (im.h - 1) + im.w, span = (size_t)raw.pitch * im.h;
        if (pend_bytes && pend_src + pend_bytes == im.p && pend_dst + pend_bytes == raw.p) {
          pend_bytes += span;
        } else {
          CU_TRY(flush());
          pend_dst = (uint8_t *)raw.p;
          pend_src = im.p;
          pend_bytes = span;
        }
        if (span != bytes && (size_t)im.w != im.stride) {   // never read past the caller's last row
          pend_bytes -= span - bytes;
          CU_TRY(flush());
        }
      } else {
        CU_TRY(flush());
        CU_TRY(cudaMemcpy2DAsync((void *)raw.p, (size_t)raw.pitch, im.p, im.stride, (size_t)im.w, (size_t)im.h, cudaMemcpyDefault, copy_st));
      }
    }
    CU_TRY(flush());
    if (split && copy_st2) {          // the pieces that went through the second stream belong to this group too
      CU_TRY(cudaEventRecord(second.joined, copy_st2));
      CU_TRY(cudaStreamWaitEvent(copy_st, second.joined, 0));
    }
    if (split) {
      CU_TRY(cudaEventRecord(copied, copy_st));
      CU_TRY(cudaStreamWaitEvent(st, copied, 0));
    }
    if (gdown > 0) {   // downsample_cameras: halve the frames where they landed, in front of the pre-processing
      CU_TRY(launch_downsample(st, B.down->d + down_at, gdown, max_dw, max_dh));
      down_at += gdown;
      stats[1] += 1;
    }
    if (cp_now && g1 == nimg) {
      cudaEventRecord(cp.ev[cp_slot][1], copy_st);
      cp.used[cp_slot] = true;
      cp.bytes += stats[2] - cp_bytes0;
    }
    HOSTPROF(s0_launch_preprocess);
    CU_TRY(launch_preprocess(st, B.pre->d + g0, g1 - g0, ctx->pyr_dev.d, any_clahe ? B.clahe->d + (size_t)g0 * CLAHE_TILES * CLAHE_TILES * 256 : nullptr,
                             any_clahe, max_band, max_w, max_h, max_nlv));
    stats[1] += (unsigned)(1 + 2 * any_clahe + pyramid_launches(max_w, max_h, max_nlv));
  }
  return OK;
}


}  // namespace uvio
