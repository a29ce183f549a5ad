// sm_100a kernels of the TrackKLT hot path.  The arithmetic lives in the *_body.h headers (one body per stage, written
// against an execution-group abstraction); this file binds the bodies to thread blocks / warps, shared memory and the
// launch geometry.  Built with -fmad=false: the CPU tracker's OpenCV build does not fuse multiply-adds.
#include "kernels.h"

#include <cooperative_groups.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "fast_body.h"
#include "image_body.h"
#include "lk_body.h"
#include "ransac_body.h"
#include "subpix_body.h"
#include "track_body.h"

namespace cg = cooperative_groups;

namespace uvio {

// ---------------------------------------------------------------------------------------------------------------------
// Pre-processing of one uploaded frame by one thread-block cluster: equalizeHist (histogram -> LUT -> apply) and every
// level of the optical-flow pyramid in a single launch.  The eight blocks of a cluster share an image; the steps that
// need the whole image (histogram complete before the LUT, level l-1 complete before level l) are separated by cluster
// barriers instead of kernel boundaries, and the partial histograms are combined through distributed shared memory.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int PRE_CLUSTER = 8;      // blocks per image (portable maximum)
#ifndef UVIO_PRE_THREADS
#define UVIO_PRE_THREADS 512
#endif
constexpr int PRE_THREADS = UVIO_PRE_THREADS;
constexpr int PRE_NSUB = 8;         // private histograms per block (two warps share one)
constexpr int PRE_MAX_BAND_BYTES = 96 * 1024;   // largest band of rows a block stages in shared memory (2 blocks/SM stay resident)

// ---- TMA (bulk asynchronous copy) helpers: global -> shared, completion counted in bytes on an mbarrier ----------------
__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // make the initialised barrier visible to the async proxy
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_1d(void *dst_smem, const void *src_gmem, unsigned bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_addr(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE_%=;\n"
      "bra WAIT_LOOP_%=;\n"
      "WAIT_DONE_%=:\n"
      "}\n" ::"r"(smem_addr(bar)),
      "r"(parity)
      : "memory");
}

__global__ void __cluster_dims__(PRE_CLUSTER, 1, 1) __launch_bounds__(PRE_THREADS, 1024 / PRE_THREADS)
    k_preprocess(const PreJob *jobs, const PyrDesc *pyr_tab, int band_bytes, int with_pyramid) {
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned rank = cluster.block_rank();
  const PreJob job = jobs[blockIdx.y];
  ImgView I = job.src;
  const PyrDesc &P = pyr_tab[job.pyr];
  const ImgView D0 = P.lv[0];
  extern __shared__ __align__(128) uint8_t band[];   // this block's rows of the frame, staged by TMA (band_bytes of them)
  __shared__ __align__(8) uint64_t band_bar;
  __shared__ unsigned sh[PRE_NSUB * 256];
  __shared__ unsigned bh[256];      // this block's histogram (read by the whole cluster)
  __shared__ unsigned cdf[256];
  __shared__ uint8_t lut[256];
  __shared__ int s_first;
  GroupBlock g;
  int r0 = (int)((long long)I.h * rank / PRE_CLUSTER), r1 = (int)((long long)I.h * (rank + 1) / PRE_CLUSTER);
  uint8_t *dst0 = (uint8_t *)D0.p;
  // Stage the band in shared memory with bulk asynchronous copies (one per row, or one for the whole band when the rows are
  // contiguous): the frame is then read from global memory once for both the histogram and the LUT pass.
  const bool staged = job.method != 2 && (I.w & 15) == 0 && (I.pitch & 15) == 0 && ((((size_t)I.p) & 15) == 0) &&
                      (size_t)(r1 - r0) * I.w <= (size_t)band_bytes && r1 > r0;
  if (staged) {
    const int rows = r1 - r0;
    if (threadIdx.x == 0) mbar_init(&band_bar, 1);
    __syncthreads();
    if (threadIdx.x == 0) mbar_expect_tx(&band_bar, (unsigned)(rows * I.w));
    __syncthreads();
    if (I.pitch == I.w) {
      if (threadIdx.x == 0) tma_load_1d(band, I.p + (size_t)r0 * I.pitch, (unsigned)(rows * I.w), &band_bar);
    } else {
      for (int r = threadIdx.x; r < rows; r += PRE_THREADS)
        tma_load_1d(band + (size_t)r * I.w, I.p + (size_t)(r0 + r) * I.pitch, (unsigned)I.w, &band_bar);
    }
    mbar_wait(&band_bar, 0);
    I = ImgView{band, I.w, rows, I.w};   // from here on "the image" of this block is its staged band
    dst0 += (size_t)r0 * D0.pitch;
    r1 = rows;
    r0 = 0;
  }
  const int total_px = job.src.w * job.src.h;
  if (job.method == 1) {   // uniform over the cluster
    for (int k = threadIdx.x; k < PRE_NSUB * 256; k += PRE_THREADS) sh[k] = 0;
    __syncthreads();
    hist_rows(g, sh, I, r0, r1, PRE_NSUB);
    __syncthreads();
    if (threadIdx.x < 256) {
      unsigned v = 0;
#pragma unroll
      for (int k = 0; k < PRE_NSUB; k++) v += sh[k * 256 + threadIdx.x];
      bh[threadIdx.x] = v;
    }
    cluster.sync();
    // every block builds the LUT from the eight partial histograms (App. A.1)
    if (threadIdx.x < 256) {
      unsigned hv = 0;
#pragma unroll
      for (int r = 0; r < PRE_CLUSTER; r++) hv += cluster.map_shared_rank(bh, r)[threadIdx.x];
      cdf[threadIdx.x] = hv;
      if (threadIdx.x == 0) s_first = 256;
    }
    __syncthreads();
    if (threadIdx.x < 256 && cdf[threadIdx.x]) atomicMin(&s_first, (int)threadIdx.x);
    for (int d = 1; d < 256; d <<= 1) {   // inclusive scan
      unsigned v = (threadIdx.x < 256 && threadIdx.x >= d) ? cdf[threadIdx.x - d] : 0;
      __syncthreads();
      if (threadIdx.x < 256) cdf[threadIdx.x] += v;
      __syncthreads();
    }
    if (threadIdx.x < 256) {
      const int i0 = s_first, total = total_px;
      const int h0 = (int)(cdf[i0] - (i0 ? cdf[i0 - 1] : 0));
      uint8_t out;
      if (h0 == total) out = (uint8_t)i0;
      else if ((int)threadIdx.x <= i0) out = 0;
      else {
        float scale = 255.f / (float)(total - h0);
        out = sat_u8(round_f2i(fmul((float)(int)(cdf[threadIdx.x] - cdf[i0]), scale)));
      }
      lut[threadIdx.x] = out;
    }
  } else if (threadIdx.x < 256) {
    lut[threadIdx.x] = (uint8_t)threadIdx.x;
  }
  __syncthreads();
  if (job.method != 2) lut_rows(g, lut, I, dst0, D0.pitch, r0, r1);   // CLAHE wrote level 0 already
  // pyramid: one lane per 4 x PD_ROWS output strip, the whole cluster per level
  const int gtid = rank * PRE_THREADS + threadIdx.x;
  for (int l = 1; l < P.nlv && with_pyramid; l++) {
    cluster.sync();   // level l-1 is complete (also keeps every block alive while its histogram may still be read)
    const ImgView S = P.lv[l - 1], D = P.lv[l];
    const int nsx = (D.w + PD_COLS - 1) / PD_COLS, nsy = (D.h + PD_ROWS - 1) / PD_ROWS;
    for (int sidx = gtid; sidx < nsx * nsy; sidx += PRE_CLUSTER * PRE_THREADS) {
      int sy = sidx / nsx, sx = sidx - sy * nsx;
      pyrdown_strip(S, (uint8_t *)D.p, D.w, D.h, D.pitch, sx * PD_COLS, sy * PD_ROWS);
    }
  }
  cluster.sync();
}

// ---------------------------------------------------------------------------------------------------------------------
// CLAHE: per-tile clipped LUTs -> bilinear apply
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_clahe_tiles(const PreJob *jobs, uint8_t *luts) {
  const PreJob job = jobs[blockIdx.y];
  if (job.method != 2) return;
  __shared__ unsigned sh[256];
  const ImgView I = job.src;
  const ClaheGeom q = clahe_geom(I.w, I.h, 10.0, CLAHE_TILES, CLAHE_TILES);
  sh[threadIdx.x] = 0;
  __syncthreads();
  clahe_tile_hist(GroupBlock(), sh, I, q, blockIdx.x % CLAHE_TILES, blockIdx.x / CLAHE_TILES);
  __syncthreads();
  if (threadIdx.x == 0) clahe_tile_lut(sh, q, luts + ((size_t)blockIdx.y * CLAHE_TILES * CLAHE_TILES + blockIdx.x) * 256);
}

__global__ void __launch_bounds__(256) k_clahe_apply(const PreJob *jobs, const PyrDesc *pyr_tab,
                                                     const uint8_t *luts) {
  const PreJob job = jobs[blockIdx.y];
  if (job.method != 2) return;
  const ImgView I = job.src;
  const ImgView D = pyr_tab[job.pyr].lv[0];
  const ClaheGeom q = clahe_geom(I.w, I.h, 10.0, CLAHE_TILES, CLAHE_TILES);
  const uint8_t *L = luts + (size_t)blockIdx.y * CLAHE_TILES * CLAHE_TILES * 256;
  int r0 = (int)((long long)I.h * blockIdx.x / gridDim.x), r1 = (int)((long long)I.h * (blockIdx.x + 1) / gridDim.x);
  for (int y = r0; y < r1; y++)
    for (int x = threadIdx.x; x < I.w; x += blockDim.x)
      ((uint8_t *)D.p)[(size_t)y * D.pitch + x] = clahe_pixel(L, q, x, y, I.p[(size_t)y * I.pitch + x]);
}

// One pyramid level of every job in its own launch: one lane per 4 x 4 output strip, grid = (strip chunks, jobs).  With the
// levels inside k_preprocess a cluster of 8 x 512 lanes walks 5,640 / 1,410 / 360 / 96 strips of a 752x480 frame between
// cluster barriers (lanes mostly idle, 55 % of warp samples at the barriers); here every level of the whole batch is one
// barrier-free launch that fills the GPU.
__global__ void __launch_bounds__(256) k_pyrdown_level(const PreJob *jobs, const PyrDesc *pyr_tab, int level) {
  const PyrDesc &P = pyr_tab[jobs[blockIdx.y].pyr];
  if (level >= P.nlv) return;
  const ImgView S = P.lv[level - 1], D = P.lv[level];
  const int nsx = (D.w + PD_COLS - 1) / PD_COLS, nsy = (D.h + PD_ROWS - 1) / PD_ROWS;
  for (int sidx = blockIdx.x * 256 + threadIdx.x; sidx < nsx * nsy; sidx += gridDim.x * 256) {
    int sy = sidx / nsx, sx = sidx - sy * nsx;
    pyrdown_strip(S, (uint8_t *)D.p, D.w, D.h, D.pitch, sx * PD_COLS, sy * PD_ROWS);
  }
}

// Levels `first` .. nlv-1 of every job in ONE launch: one block per image walks the small levels (a 188x120 level is 1,410
// strips: three rounds of a 512-lane block) with a block barrier between them, instead of a launch of a few microseconds each.
__global__ void __launch_bounds__(512) k_pyrdown_tail(const PreJob *jobs, const PyrDesc *pyr_tab, int first) {
  const PyrDesc &P = pyr_tab[jobs[blockIdx.x].pyr];
  for (int level = first; level < P.nlv; level++) {
    const ImgView S = P.lv[level - 1], D = P.lv[level];
    const int nsx = (D.w + PD_COLS - 1) / PD_COLS, nsy = (D.h + PD_ROWS - 1) / PD_ROWS;
    for (int sidx = threadIdx.x; sidx < nsx * nsy; sidx += blockDim.x) {
      int sy = sidx / nsx, sx = sidx - sy * nsx;
      pyrdown_strip(S, (uint8_t *)D.p, D.w, D.h, D.pitch, sx * PD_COLS, sy * PD_ROWS);
    }
    __threadfence_block();
    __syncthreads();   // the level just written is the next one's source
  }
}

cudaError_t launch_preprocess(cudaStream_t st, const PreJob *jobs, int njobs, const PyrDesc *pyr_tab,
                              uint8_t *clahe_lut, int any_clahe, int max_band_bytes, int max_w, int max_h, int max_nlv) {
  if (njobs <= 0) return cudaSuccess;
  // dynamic shared memory for the staged band of rows (0 = bands too large: the kernel reads global memory directly)
  int band_bytes = max_band_bytes <= PRE_MAX_BAND_BYTES ? (max_band_bytes + 127) / 128 * 128 : 0;
  {
    static std::mutex m;
    static bool attr_set[64] = {false};   // per device: the attribute belongs to the (function, device) pair
    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lk(m);
    if (device < 0 || device >= 64 || !attr_set[device]) {
      e = cudaFuncSetAttribute(k_preprocess, cudaFuncAttributeMaxDynamicSharedMemorySize, PRE_MAX_BAND_BYTES);
      if (e != cudaSuccess) return e;
      if (device >= 0 && device < 64) attr_set[device] = true;
    }
  }
  if (any_clahe) {
    prof_mark(st, K_CLAHE_TILES);
    k_clahe_tiles<<<dim3(CLAHE_TILES * CLAHE_TILES, njobs), 256, 0, st>>>(jobs, clahe_lut);
    prof_mark(st, K_CLAHE_APPLY);
    k_clahe_apply<<<dim3(HIST_BANDS * 2, njobs), 256, 0, st>>>(jobs, pyr_tab, clahe_lut);
  }
  // max_nlv > 1 with the frame size given: the pyramid levels go in launches of their own (UVIO_KLT_PYR=fused keeps them inside)
  static const bool fused_env = [] { const char *e = getenv("UVIO_KLT_PYR"); return e && !strcmp(e, "fused"); }();
  const bool split = max_nlv > 1 && max_w > 0 && max_h > 0 && !fused_env;
  prof_mark(st, K_PREPROCESS);
  k_preprocess<<<dim3(PRE_CLUSTER, njobs), PRE_THREADS, band_bytes, st>>>(jobs, pyr_tab, band_bytes, split ? 0 : 1);
  if (split) {
    // level 1 (and level 2 of large frames) fills the GPU as a launch of its own; the rest is one block per image
    int w = max_w, h = max_h, l = 1;
    for (; l < max_nlv; l++) {
      w = (w + 1) / 2;
      h = (h + 1) / 2;
      const int strips = ((w + PD_COLS - 1) / PD_COLS) * ((h + PD_ROWS - 1) / PD_ROWS);
      if (l > 1 && strips <= 4096) break;
      prof_mark(st, K_PYRDOWN);
      k_pyrdown_level<<<dim3((strips + 255) / 256, njobs), 256, 0, st>>>(jobs, pyr_tab, l);
    }
    if (l < max_nlv) {
      prof_mark(st, K_PYRDOWN);
      k_pyrdown_tail<<<njobs, 512, 0, st>>>(jobs, pyr_tab, l);
    }
  }
  prof_mark(st, -1);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// downsample_cameras (ov_msckf/src/core/VioManager.cpp:268-278): cv::pyrDown of the message's images and masks to
// Size(cols / 2.0, rows / 2.0), in front of feed_new_camera.  One lane per 4x4 output strip, grid = (strip chunks, jobs);
// an HBM-bound byte pass (1 B read per source pixel, 1/4 B written).
// ---------------------------------------------------------------------------------------------------------------------
constexpr int DOWN_THREADS = 256;

__global__ void __launch_bounds__(DOWN_THREADS) k_downsample(const DownJob *jobs) {
  const DownJob job = jobs[blockIdx.y];
  const ImgView S = job.src, D = job.dst;
  const int nsx = (D.w + PD_COLS - 1) / PD_COLS, nsy = (D.h + PD_ROWS - 1) / PD_ROWS;
  for (int sidx = blockIdx.x * DOWN_THREADS + threadIdx.x; sidx < nsx * nsy; sidx += gridDim.x * DOWN_THREADS) {
    int sy = sidx / nsx, sx = sidx - sy * nsx;
    pyrdown_strip(S, (uint8_t *)D.p, D.w, D.h, D.pitch, sx * PD_COLS, sy * PD_ROWS);
  }
}

cudaError_t launch_downsample(cudaStream_t st, const DownJob *jobs, int njobs, int max_dst_w, int max_dst_h) {
  if (njobs <= 0) return cudaSuccess;
  const int strips = ((max_dst_w + PD_COLS - 1) / PD_COLS) * ((max_dst_h + PD_ROWS - 1) / PD_ROWS);
  const int bx = (strips + DOWN_THREADS - 1) / DOWN_THREADS;
  prof_mark(st, K_DOWNSAMPLE);
  k_downsample<<<dim3(bx > 0 ? bx : 1, njobs), DOWN_THREADS, 0, st>>>(jobs);
  prof_mark(st, -1);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// FAST on grid cells. grid = (max tiles per cell, jobs). The last tile of a job to finish compacts the job's list into
// mapped host memory, so the host sees exactly `count` words per cell and nothing else crosses PCIe.
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_fast(const FastJob *jobs, const PyrDesc *pyr_tab, uint32_t *out_dev,
                                              unsigned *count_dev, unsigned *ticket_dev, uint32_t *out_host, unsigned *count_host) {
  const FastJob job = jobs[blockIdx.y];
  const int tiles_x = (job.w + FT_W - 1) / FT_W, tiles_y = (job.h + FT_H - 1) / FT_H;
  const int ntiles = tiles_x * tiles_y;
  if ((int)blockIdx.x >= ntiles) return;
  __shared__ FastSmem s;
  __shared__ int s_last;
  const ImgView I = pyr_tab[job.img].lv[0];
  fast_tile(GroupBlock(), s, I, job, (blockIdx.x % tiles_x) * FT_W, (blockIdx.x / tiles_x) * FT_H, job.threshold, out_dev,
            &count_dev[blockIdx.y]);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(&ticket_dev[blockIdx.y], 1u) == (unsigned)ntiles - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // One thread reads the count and re-arms the counters; the others take the count from shared memory.  (Every thread
  // used to read count_dev itself while thread 0 went on to zero it: a warp scheduled late read 0 and skipped its share
  // of the copy, leaving stale words in the host list -- seen as one tracker in ~6 % of 19-tracker runs diverging.)
  __shared__ unsigned s_n;
  if (threadIdx.x == 0) {
    unsigned c = __ldcg(&count_dev[blockIdx.y]);
    s_n = c > (unsigned)job.out_cap ? (unsigned)job.out_cap : c;
    count_dev[blockIdx.y] = 0;
    ticket_dev[blockIdx.y] = 0;
  }
  __syncthreads();
  const unsigned n = s_n;
  for (unsigned i = threadIdx.x; i < n; i += blockDim.x) out_host[job.out_off + i] = __ldcg(&out_dev[job.out_off + i]);
  if (threadIdx.x == 0) count_host[blockIdx.y] = n;
}

cudaError_t launch_fast(cudaStream_t st, const FastJob *jobs, int njobs, int max_cell_w, int max_cell_h, const PyrDesc *pyr_tab,
                        uint32_t *out_dev, unsigned *count_dev, unsigned *ticket_dev,
                        uint32_t *out_host, unsigned *count_host) {
  if (njobs <= 0) return cudaSuccess;
  int tiles = ((max_cell_w + FT_W - 1) / FT_W) * ((max_cell_h + FT_H - 1) / FT_H);
  prof_mark(st, K_FAST);
  k_fast<<<dim3(tiles, njobs), 256, 0, st>>>(jobs, pyr_tab, out_dev, count_dev, ticket_dev, out_host, count_host);
  prof_mark(st, -1);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// cornerSubPix: one warp per point
// ---------------------------------------------------------------------------------------------------------------------
constexpr int WARPS_PER_BLOCK = 4;

// Programmatic dependent launch (sm_90+): a kernel launched with the attribute may be scheduled while the kernel before it
// in the stream is still running; it reads what does not depend on that kernel (its job list, in host memory: one PCIe
// round trip that would otherwise follow the launch), then waits for the predecessor's results with pdl_wait().
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_release_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <class... KArgs, class... Args>
static cudaError_t launch_after(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, bool overlap_predecessor,
                                Args... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = overlap_predecessor ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32) k_subpix(const SubpixItem *items, int n, const PyrDesc *pyr_tab,
                                                                 const float *mask, float *out_xy, float *out_xy_dev, const int *n_dev) {
  __shared__ SubpixSmem sm[WARPS_PER_BLOCK];
  __shared__ float smask[SP_N];
  pdl_release_dependents();   // the stereo LK chained on this launch may fetch its items meanwhile
  for (int i = threadIdx.x; i < SP_N; i += blockDim.x) smask[i] = mask[i];
  __syncthreads();
  const int wid = threadIdx.x >> 5;
  // n_dev: a compact item list built on the device (the device-side pipeline): its length is read here, warps stride over it
  // and every item names the slot its result goes to; otherwise item i -> slot i, one warp each
  if (n_dev) n = *n_dev;
  for (int i = blockIdx.x * WARPS_PER_BLOCK + wid; i < n; i += gridDim.x * WARPS_PER_BLOCK) {
    const SubpixItem it = items[i];
    if (it.pyr < 0) continue;
    float x = it.x, y = it.y;
    subpix_point(GroupWarp(), sm[wid], pyr_tab[it.pyr].lv[0], smask, 20, 0.001 * 0.001, x, y);
    if ((threadIdx.x & 31) == 0) {
      const int at = n_dev ? it.pad : i;
      out_xy[2 * at] = x;
      out_xy[2 * at + 1] = y;
      if (out_xy_dev) {
        out_xy_dev[2 * at] = x;
        out_xy_dev[2 * at + 1] = y;
      }
    }
    __syncwarp();
  }
}

cudaError_t launch_subpix(cudaStream_t st, const SubpixItem *items, int n, const PyrDesc *pyr_tab, const float *mask_dev,
                          float *out_xy, float *out_xy_dev, const int *n_dev) {
  if (n <= 0) return cudaSuccess;
  int blocks = (n + WARPS_PER_BLOCK - 1) / WARPS_PER_BLOCK;
  if (n_dev && blocks > 148 * 4) blocks = 148 * 4;
  prof_mark(st, K_SUBPIX);
  k_subpix<<<blocks, WARPS_PER_BLOCK * 32, 0, st>>>(items, n, pyr_tab, mask_dev, out_xy, out_xy_dev, n_dev);
  prof_mark(st, -1);
  return cudaGetLastError();
}

cudaError_t upload_subpix_mask(float *mask_dev, cudaStream_t st) {
  // the Gaussian weights are evaluated with the host's expf, as the CPU tracker does (App. A.5)
  float m[SP_N];
  for (int i = 0; i < SP_W; i++) {
    float y = (float)(i - SP_WIN) / SP_WIN;
    float vy = expf(-y * y);
    for (int j = 0; j < SP_W; j++) {
      float x = (float)(j - SP_WIN) / SP_WIN;
      m[i * SP_W + j] = (float)(vy * expf(-x * x));
    }
  }
  cudaError_t e = cudaMemcpyAsync(mask_dev, m, sizeof(m), cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) return e;
  return cudaStreamSynchronize(st);
}

// ---------------------------------------------------------------------------------------------------------------------
// pyramidal LK: one warp per feature, all levels inside the kernel
// ---------------------------------------------------------------------------------------------------------------------
// work counters for the roofline arithmetic: [0] templates staged (point-levels), [1] iterations, [2] points,
// [3] iterations whose mismatch sums ran as ordered fp32 chains (the others took the exact integer reduction)
constexpr int LK_WORK_SLOTS = 64;   // the counters are spread over slots so that the atomics of one launch do not queue up
__device__ unsigned long long g_lk_work[LK_WORK_SLOTS][4];

// one warp per block: features differ a lot in iteration count, and single-warp blocks let the SM refill warp by warp
constexpr int LK_WARPS_PER_BLOCK = 1;

#ifdef UVIO_LK_MIN_BLOCKS
#define LK_BOUNDS __launch_bounds__(LK_WARPS_PER_BLOCK * 32, UVIO_LK_MIN_BLOCKS)
#else
#define LK_BOUNDS __launch_bounds__(LK_WARPS_PER_BLOCK * 32, 28)   // 72 registers: 28 single-warp blocks per SM (measured 3 % faster than 79 / 25)
#endif
__global__ void LK_BOUNDS k_lk(const LkItem *items, int n, const PyrDesc *pyr_tab,
                                                             const float *init_from_dev, int max_count, double eps_sq,
                                                             LkOut *out_host, uint8_t *status_host, LkOut *out_dev,
                                                             uint8_t *status_dev, const int *out_idx, const int *n_dev) {
  __shared__ LkSmem sm[LK_WARPS_PER_BLOCK];
  const int wid = threadIdx.x >> 5;
  pdl_release_dependents();   // the RANSAC launch behind this one may start on what it can do without the tracks
  // n_dev: a compact item list built on the device; its length is read here and the (single-warp) blocks stride over it
  if (n_dev) n = *n_dev;
  bool waited = false;
  for (int i = blockIdx.x * LK_WARPS_PER_BLOCK + wid; i < n; i += gridDim.x * LK_WARPS_PER_BLOCK) {
    const LkItem it = items[i];
    float px = it.px, py = it.py, nx = it.nx, ny = it.ny;
    if (it.prev_pyr < 0) continue;   // an unused slot of the device-side pipeline
    if (!waited) {
      pdl_wait();               // the refined corners (init_from_dev) and the kernel before this one's use of out_dev / status_dev
      waited = true;
    }
    const int at = out_idx ? out_idx[i] : i;     // results of a compact item list land where the full list has them
    if (init_from_dev) {
      px = nx = init_from_dev[2 * at];
      py = ny = init_from_dev[2 * at + 1];
    }
    const PyrDesc &P0 = pyr_tab[it.prev_pyr], &P1 = pyr_tab[it.next_pyr];
    LkParams prm;
    prm.max_level = P0.nlv - 1;
    prm.max_count = max_count;
    prm.eps_sq = eps_sq;
    prm.min_eig = 1e-4;
    uint8_t status;
    int work[3] = {0, 0, 0};
    lk_track_point(GroupWarp(), sm[wid], P0.lv, P1.lv, prm, px, py, nx, ny, status, work);
    if ((threadIdx.x & 31) == 0) {
      unsigned long long *wk = g_lk_work[blockIdx.x % LK_WORK_SLOTS];
      atomicAdd(&wk[0], (unsigned long long)work[0]);
      atomicAdd(&wk[1], (unsigned long long)work[1]);
      atomicAdd(&wk[2], 1ull);
      if (work[2]) atomicAdd(&wk[3], (unsigned long long)work[2]);
      LkOut o{nx, ny};
      out_host[at] = o;
      status_host[at] = status;
      out_dev[at] = o;
      status_dev[at] = status;
    }
    __syncwarp();
  }
}

cudaError_t read_lk_work(unsigned long long out[4], int reset) {
  unsigned long long h[LK_WORK_SLOTS][4];
  cudaError_t e = cudaMemcpyFromSymbol(h, g_lk_work, sizeof(h));
  out[0] = out[1] = out[2] = out[3] = 0;
  for (int s = 0; s < LK_WORK_SLOTS; s++)
    for (int k = 0; k < 4; k++) out[k] += h[s][k];
  if (e == cudaSuccess && reset) {
    memset(h, 0, sizeof(h));
    e = cudaMemcpyToSymbol(g_lk_work, h, sizeof(h));
  }
  return e;
}

cudaError_t launch_lk(cudaStream_t st, const LkItem *items, int n, const PyrDesc *pyr_tab, const float *init_from_dev, int max_count,
                      double eps, LkOut *out_host, uint8_t *status_host, LkOut *out_dev, uint8_t *status_dev, bool overlap,
                      const int *out_idx, const int *n_dev) {
  if (n <= 0) return cudaSuccess;
  int blocks = (n + LK_WARPS_PER_BLOCK - 1) / LK_WARPS_PER_BLOCK;
  if (n_dev && blocks > 148 * 28) blocks = 148 * 28;   // one resident wave of single-warp blocks striding over the device-built list
  prof_mark(st, K_LK);
  cudaError_t e = launch_after(k_lk, dim3(blocks), dim3(LK_WARPS_PER_BLOCK * 32), 0, st, overlap,
                               items, n, pyr_tab, init_from_dev, max_count, eps * eps, out_host, status_host, out_dev, status_dev, out_idx, n_dev);
  prof_mark(st, -1);
  return e != cudaSuccess ? e : cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// undistort + fundamental-matrix RANSAC: one block per (rig, camera)
// ---------------------------------------------------------------------------------------------------------------------
constexpr int RANSAC_THREADS = 128;

__global__ void __launch_bounds__(RANSAC_THREADS) k_ransac(const RansacJob *jobs, const LkItem *items, const LkOut *lk,
                                                           const uint8_t *status, const CamModel *cams, float *p0n, float *p1n,
                                                           float *p1n_host, uint8_t *mask_host, uint8_t *mask_dev) {
  extern __shared__ float dyn[];
  __shared__ RansacSmem s;
  const RansacJob job = jobs[blockIdx.x];
  const int n = job.n;
  const CamModel c0 = cams[job.cam0], c1 = cams[job.cam1];
  // the old points come from the job list (host): undistorted while the LK launch ahead of this one is still tracking
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int k = job.off + i;
    float a, b;
    undistort_point(c0, items[k].px, items[k].py, a, b);
    p0n[2 * k] = a;
    p0n[2 * k + 1] = b;
  }
  pdl_wait();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int k = job.off + i;
    float a, b;
    const LkOut o = lk[k];
    undistort_point(c1, o.x, o.y, a, b);
    p1n[2 * k] = a;
    p1n[2 * k + 1] = b;
    p1n_host[2 * k] = a;
    p1n_host[2 * k + 1] = b;
  }
  __syncthreads();
  const bool run = job.run && n >= 7;
  if (run) {
    s.x0 = dyn;
    s.y0 = dyn + n;
    s.x1 = dyn + 2 * n;
    s.y1 = dyn + 3 * n;
    __syncthreads();
    ransac_job(GroupBlock(), s, p0n + 2 * job.off, p1n + 2 * job.off, n, job.thr, 0.999, 1000, mask_dev + job.off);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x)
    mask_host[job.off + i] = run ? (uint8_t)(status[job.off + i] && mask_dev[job.off + i]) : (uint8_t)0;
}

cudaError_t launch_ransac(cudaStream_t st, const RansacJob *jobs, int njobs, int max_n, const LkItem *items, const LkOut *lk_dev,
                          const uint8_t *status_dev, const CamModel *cams_dev, float *p0n_dev, float *p1n_dev, float *p1n_host,
                          uint8_t *mask_host, uint8_t *mask_dev, bool overlap) {
  if (njobs <= 0) return cudaSuccess;
  size_t dyn = (size_t)4 * max_n * sizeof(float);
  if (dyn + sizeof(RansacSmem) > 48 * 1024) {
    // the attribute belongs to the (function, device) pair and lanes of several devices launch concurrently
    static std::mutex m;
    static size_t configured[64] = {0};
    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lk(m);
    if (device < 0 || device >= 64 || dyn > configured[device]) {
      e = cudaFuncSetAttribute(k_ransac, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
      if (e != cudaSuccess) return e;
      if (device >= 0 && device < 64) configured[device] = dyn;
    }
  }
  prof_mark(st, K_RANSAC);
  cudaError_t e = launch_after(k_ransac, dim3(njobs), dim3(RANSAC_THREADS), dyn, st, overlap, jobs, items, lk_dev, status_dev, cams_dev,
                               p0n_dev, p1n_dev, p1n_host, mask_host, mask_dev);
  prof_mark(st, -1);
  return e != cudaSuccess ? e : cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// Device-side bookkeeping (track_body.h): the kernels between the pixel kernels of a feed.  One block per tracker.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int DV_THREADS = 256;
constexpr int DV_THREADS_MAX = 1024;   // per-tracker bookkeeping blocks of trackers with more than 512 points per camera (their O(N^2) membership loops and
                                       // scans are spread over four times the lanes; the kernels need at most 48 registers)
static int dv_threads(int max_p) { return max_p > 512 ? DV_THREADS_MAX : DV_THREADS; }
constexpr int DV_TOPK_SMEM = 4096;   // corner words a cell's top-K block sorts in shared memory (more: in the job's own global region)

// dynamic shared memory of the bookkeeping kernels: two id lists and one point list of P entries each (DvSmem)
__device__ __forceinline__ DvSmem dv_smem(int P) {
  extern __shared__ __align__(16) unsigned long long dv_dyn[];
  DvSmem sm;
  sm.ids0 = dv_dyn;
  sm.ids1 = dv_dyn + P;
  sm.pts = (float *)(dv_dyn + 2 * P);
  return sm;
}

// A block works on a private copy of its unit's descriptor in shared memory, with this feed's bits applied: the descriptor
// array itself stays in device memory unchanged from feed to feed.
__device__ __forceinline__ const DvUnit &dv_load_unit(DvUnit &sU, const DvUnit *units, int u, unsigned bits) {
  __syncthreads();                                         // the previous unit's readers are done
  const unsigned *src = (const unsigned *)&units[u];
  unsigned *dst = (unsigned *)&sU;
  for (int i = threadIdx.x; i < (int)(sizeof(DvUnit) / 4); i += blockDim.x) dst[i] = src[i];
  __syncthreads();
  if (threadIdx.x == 0) dv_apply_feed(sU, bits);
  __syncthreads();
  return sU;
}

template <int T>
__global__ void __launch_bounds__(T) k_dv_occupancy(const DvUnit *units, const DvHead *heads, DvGlobals X, int P, DvFeedBits bits) {
  __shared__ DvUnit sU;
  const DvHead h = heads[blockIdx.x];
  if (blockIdx.x == 0 && threadIdx.x < 4) X.n_items[threadIdx.x] = 0;   // the item lists select / assign append to (later kernels of this chain)
  for (int u = h.first; u < h.first + h.count; u++) {
    if (threadIdx.x == 0) X.feed_bits[u] = bits.v[u];                   // for the kernels after this one (they may be replayed from a graph)
    dv_occupancy(GroupBlock(), dv_load_unit(sU, units, u, bits.v[u]), X, dv_smem(P));
  }
}

// FAST over the cells the occupancy kernel listed: a fixed grid walks (active job, tile) pairs
__global__ void __launch_bounds__(256) k_dv_fast(DvGlobals X, const PyrDesc *pyr_tab, int max_tiles) {
  __shared__ FastSmem s;
  const int n_active = *X.fast_n_active;
  for (int w = blockIdx.x; w < n_active * max_tiles; w += gridDim.x) {
    const int a = w / max_tiles, tile = w - a * max_tiles;
    const int slot = X.fast_active[a];
    const FastJob job = X.fast_jobs[slot];
    const int tiles_x = (job.w + FT_W - 1) / FT_W, tiles_y = (job.h + FT_H - 1) / FT_H;
    if (tile >= tiles_x * tiles_y) continue;
    fast_tile(GroupBlock(), s, pyr_tab[job.img].lv[0], job, (tile % tiles_x) * FT_W, (tile / tiles_x) * FT_H, job.threshold, X.fast_out,
              &X.fast_cnt[slot]);
  }
}

// per-cell top-K in std::sort order: one block per job slot; grid = (2C max, units)
__global__ void __launch_bounds__(DV_THREADS) k_dv_topk(const DvUnit *units, DvGlobals X) {
  __shared__ uint32_t buf[DV_TOPK_SMEM];
  __shared__ DvUnit sU;
  {
    const DvUnit &G = units[blockIdx.y];
    const int side0 = blockIdx.x / G.C;
    if (side0 > 1 || (side0 == 1 && !G.stereo)) return;
  }
  const DvUnit &U = dv_load_unit(sU, units, blockIdx.y, X.feed_bits[blockIdx.y]);
  const int side = blockIdx.x / U.C, r = blockIdx.x - side * U.C;
  dv_topk(GroupBlock(), U, X, side, r, buf, DV_TOPK_SMEM);
}

template <int T>
__global__ void __launch_bounds__(T) k_dv_select(const DvUnit *units, const DvHead *heads, DvGlobals X, int P) {
  const DvHead h = heads[blockIdx.x];
  if (blockIdx.x == 0 && threadIdx.x == 0) *X.fast_n_active = 0;   // FAST and the top-K are done with the list: re-armed for the next feed
  __shared__ DvUnit sU;
  for (int u = h.first; u < h.first + h.count; u++) dv_select(GroupBlock(), dv_load_unit(sU, units, u, X.feed_bits[u]), X, dv_smem(P));
}
template <int T>
__global__ void __launch_bounds__(T) k_dv_assign(const DvUnit *units, const DvHead *heads, DvGlobals X, int P) {
  const DvHead h = heads[blockIdx.x];
  __shared__ DvUnit sU;
  for (int u = h.first; u < h.first + h.count; u++) dv_assign(GroupBlock(), dv_load_unit(sU, units, u, X.feed_bits[u]), X, dv_smem(P));
}
template <int T>
__global__ void __launch_bounds__(T) k_dv_merge(const DvUnit *units, const DvHead *heads, DvGlobals X, int P) {
  const DvHead h = heads[blockIdx.x];
  __shared__ DvUnit sU;
  for (int u = h.first; u < h.first + h.count; u++) dv_merge(GroupBlock(), dv_load_unit(sU, units, u, X.feed_bits[u]), X, dv_smem(P));
}

template <class K>
static cudaError_t dv_smem_attr(K kernel, size_t bytes) {   // the attribute belongs to the (function, device) pair
  if (bytes <= 48 * 1024) return cudaSuccess;
  return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
}
static size_t dv_smem_bytes(int P) { return (size_t)P * 24; }

cudaError_t launch_dv_occupancy(cudaStream_t st, const DvUnit *units, const DvHead *heads, int n_heads, const DvGlobals &X, int max_p,
                                const DvFeedBits &bits) {
  if (n_heads <= 0) return cudaSuccess;
  cudaError_t e = dv_threads(max_p) == DV_THREADS ? dv_smem_attr(k_dv_occupancy<DV_THREADS>, dv_smem_bytes(max_p)) : dv_smem_attr(k_dv_occupancy<DV_THREADS_MAX>, dv_smem_bytes(max_p));
  if (e != cudaSuccess) return e;
  prof_mark(st, K_DV_LOGIC);
  if (dv_threads(max_p) == DV_THREADS) k_dv_occupancy<DV_THREADS><<<n_heads, DV_THREADS, dv_smem_bytes(max_p), st>>>(units, heads, X, max_p, bits);
  else k_dv_occupancy<DV_THREADS_MAX><<<n_heads, DV_THREADS_MAX, dv_smem_bytes(max_p), st>>>(units, heads, X, max_p, bits);
  prof_mark(st, -1);
  return cudaGetLastError();
}
cudaError_t launch_dv_fast(cudaStream_t st, const DvGlobals &X, const PyrDesc *pyr_tab, int n_slots, int max_cell_w, int max_cell_h) {
  if (n_slots <= 0) return cudaSuccess;
  const int tiles = ((max_cell_w + FT_W - 1) / FT_W) * ((max_cell_h + FT_H - 1) / FT_H);
  const long work = (long)n_slots * tiles;
  const int grid = (int)(work < 148 * 8 ? work : 148 * 8);   // 8 blocks of 256 threads per SM
  prof_mark(st, K_FAST);
  k_dv_fast<<<grid, 256, 0, st>>>(X, pyr_tab, tiles);
  prof_mark(st, -1);
  return cudaGetLastError();
}
cudaError_t launch_dv_topk(cudaStream_t st, const DvUnit *units, int n_units, int max_c, const DvGlobals &X) {
  if (n_units <= 0) return cudaSuccess;
  prof_mark(st, K_DV_TOPK);
  k_dv_topk<<<dim3(2 * max_c, n_units), DV_THREADS, 0, st>>>(units, X);
  prof_mark(st, -1);
  return cudaGetLastError();
}
cudaError_t launch_dv_select(cudaStream_t st, const DvUnit *units, const DvHead *heads, int n_heads, const DvGlobals &X, int max_p) {
  if (n_heads <= 0) return cudaSuccess;
  cudaError_t e = dv_threads(max_p) == DV_THREADS ? dv_smem_attr(k_dv_select<DV_THREADS>, dv_smem_bytes(max_p)) : dv_smem_attr(k_dv_select<DV_THREADS_MAX>, dv_smem_bytes(max_p));
  if (e != cudaSuccess) return e;
  prof_mark(st, K_DV_LOGIC);
  if (dv_threads(max_p) == DV_THREADS) k_dv_select<DV_THREADS><<<n_heads, DV_THREADS, dv_smem_bytes(max_p), st>>>(units, heads, X, max_p);
  else k_dv_select<DV_THREADS_MAX><<<n_heads, DV_THREADS_MAX, dv_smem_bytes(max_p), st>>>(units, heads, X, max_p);
  prof_mark(st, -1);
  return cudaGetLastError();
}
cudaError_t launch_dv_assign(cudaStream_t st, const DvUnit *units, const DvHead *heads, int n_heads, const DvGlobals &X, int max_p) {
  if (n_heads <= 0) return cudaSuccess;
  cudaError_t e = dv_threads(max_p) == DV_THREADS ? dv_smem_attr(k_dv_assign<DV_THREADS>, dv_smem_bytes(max_p)) : dv_smem_attr(k_dv_assign<DV_THREADS_MAX>, dv_smem_bytes(max_p));
  if (e != cudaSuccess) return e;
  prof_mark(st, K_DV_LOGIC);
  if (dv_threads(max_p) == DV_THREADS) k_dv_assign<DV_THREADS><<<n_heads, DV_THREADS, dv_smem_bytes(max_p), st>>>(units, heads, X, max_p);
  else k_dv_assign<DV_THREADS_MAX><<<n_heads, DV_THREADS_MAX, dv_smem_bytes(max_p), st>>>(units, heads, X, max_p);
  prof_mark(st, -1);
  return cudaGetLastError();
}
cudaError_t launch_dv_merge(cudaStream_t st, const DvUnit *units, const DvHead *heads, int n_heads, const DvGlobals &X, int max_p) {
  if (n_heads <= 0) return cudaSuccess;
  cudaError_t e = dv_threads(max_p) == DV_THREADS ? dv_smem_attr(k_dv_merge<DV_THREADS>, dv_smem_bytes(max_p)) : dv_smem_attr(k_dv_merge<DV_THREADS_MAX>, dv_smem_bytes(max_p));
  if (e != cudaSuccess) return e;
  prof_mark(st, K_DV_LOGIC);
  if (dv_threads(max_p) == DV_THREADS) k_dv_merge<DV_THREADS><<<n_heads, DV_THREADS, dv_smem_bytes(max_p), st>>>(units, heads, X, max_p);
  else k_dv_merge<DV_THREADS_MAX><<<n_heads, DV_THREADS_MAX, dv_smem_bytes(max_p), st>>>(units, heads, X, max_p);
  prof_mark(st, -1);
  return cudaGetLastError();
}

}  // namespace uvio
