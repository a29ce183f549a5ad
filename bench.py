#!/usr/bin/env python
"""bench.py — stereo frames/s of the TrackKLT hot path (BASELINE.json metric) on N B200s.

  python bench.py --gpus N --steps K --warmup W            (N > 1: launched by torch.distributed.run, one rank per GPU)
  python bench.py --config {euroc_mav,tum_vi,rs_d455,hd_1000,uzhfpv_8}     one of BASELINE.json's five configurations
  python bench.py --impl reference ...                     the CPU tracker (cv2-driven oracle) on the host cores

Workload of the default line (config.workload): the configuration the metric is quoted on — euroc_mav, 752x480 stereo, 200
features per camera, 5x5 grid, FAST threshold 20, histogram equalisation, the reference's pyramid (5 requested levels).
S independent synthetic camera streams live on a GPU (streams are the unit that shards; no data-path
collective).  One "step" feeds R consecutive stereo frames to every stream: R batched uvio_klt_feed_batch calls, S * R
frames.  value = frames all ranks processed / max-over-ranks device time with the frames already resident in HBM; e2e = the
same calls on pinned HOST frames, the upload and the result read-back inside the timed region.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import mmap
import os
import subprocess
import sys
import time

# One hardware work queue per CUDA stream (the default of 8 makes streams share queues).  Must be set before the CUDA context exists.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "stereo_frames_per_sec_tracked_TrackKLT_752x480"
UNIT = "frames/s"
# BASELINE.json configs -> (tracker configuration, streams per GPU)
BENCH_CONFIGS = {"euroc_mav": ("EUROC_MAV", 64), "tum_vi": ("TUM_VI", 64), "rs_d455": ("RS_D455", 64), "hd_1000": ("HD_1000", 16),
                 "uzhfpv_8": ("UZHFPV_INDOOR", 8)}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="euroc_mav", choices=sorted(BENCH_CONFIGS))
    ap.add_argument("--streams", type=int, default=int(os.environ.get("UVIO_BENCH_STREAMS", "0")), help="streams per GPU (0: the config's default)")
    ap.add_argument("--frames-per-step", type=int, default=32, help="consecutive frames fed to every stream in one step")
    ap.add_argument("--pool", type=int, default=40, help="distinct frames per stream; a run walks them forwards and backwards")
    ap.add_argument("--distinct", type=int, default=8, help="distinct synthetic sequences per GPU (others are time-shifted copies)")
    ap.add_argument("--cpu-frames", type=int, default=120, help="frames of the bounded CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def workload_config(name="euroc_mav"):
    from uvio_b200 import configs
    return getattr(configs, BENCH_CONFIGS[name][0])


def config_dict(cfg, streams, frames_per_step, extra=None):
    hist = {0: "no equalisation", 1: "equalizeHist", 2: "CLAHE"}[cfg.histogram]
    d = {"workload": f"TrackKLT {cfg.name} {cfg.width}x{cfg.height} stereo, {cfg.num_features} feats/cam, "
                     f"{cfg.grid_x}x{cfg.grid_y} grid, fast_thresh {cfg.fast_threshold}, {hist}, pyr_levels {cfg.pyr_levels}"
                     + (" (reference-faithful; effective top level 4)" if cfg.name == "euroc_mav" else ""),
         "streams_per_gpu": streams, "frames_per_step": streams * frames_per_step, "feeds_per_step": frames_per_step,
         "l2_policy": "every feed reads frames that were last touched a whole pool ago (pool >> 126 MB L2); no explicit flush"}
    if extra:
        d.update(extra)
    return d


# ---------------------------------------------------------------------------------------------------------------------
# synthetic input: D distinct sequences per rank, stream s replays sequence s % D shifted by 5*(s // D) frames.  A run walks
# the pool forwards, then backwards (camera motion reversed), and so on: continuous motion from a bounded amount of memory.
# ---------------------------------------------------------------------------------------------------------------------
def make_frames(cfg, rank, streams, distinct, n_pool):
    from uvio_b200.synth import SynthStereoStream
    distinct = max(1, min(distinct, streams))
    shift = 5
    need = n_pool + shift * ((streams - 1) // distinct)
    seqs = []
    for d in range(distinct):
        s = SynthStereoStream(cfg.width, cfg.height, stream_id=1000 * rank + d, events=False)
        seqs.append([s.frame(k) for k in range(need)])
    return [[seqs[s % distinct][k + shift * (s // distinct)] for s in range(streams)] for k in range(n_pool)]


def pool_index(t, n_pool):
    """frame of the pool that feed number t reads: 0, 1, .., n-1, n-2, .., 1, 0, 1, .."""
    if n_pool <= 1:
        return 0
    period = 2 * n_pool - 2
    k = t % period
    return k if k < n_pool else period - k


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={gpu_index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.p.terminate()
        try:
            out, _ = self.p.communicate(timeout=5)
        except Exception:
            self.p.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


_keep_alive = []


def pinned_hugepage_buffer(nbytes):
    """Host buffer the way a camera driver that wants the full PCIe rate allocates its frame ring: a
    2 MiB-aligned anonymous mapping advised to transparent huge pages, touched, then cudaHostRegister'ed.  On this pool's
    hosts such a buffer uploads 2-4x faster than a cudaHostAlloc one of the same size (tools/h2d_probe.py)."""
    import torch
    H = 2 << 20
    size = (nbytes + H - 1) // H * H + H
    m = mmap.mmap(-1, size, flags=mmap.MAP_PRIVATE | mmap.MAP_ANONYMOUS)
    a = np.frombuffer(m, dtype=np.uint8)
    off = (-a.ctypes.data) % H
    try:
        m.madvise(mmap.MADV_HUGEPAGE)
    except Exception:
        pass
    a[:] = 0
    v = a[off:off + nbytes]
    r = torch.cuda.cudart().cudaHostRegister(v.ctypes.data, (nbytes + 4095) // 4096 * 4096, 0)
    if int(r) != 0:
        raise RuntimeError(f"cudaHostRegister failed: {r}")
    _keep_alive.append((m, a))
    return v


def bind_near_gpu(local):
    """Keep this rank's threads (and, by first touch, its frame ring) on the NUMA node the GPU's PCIe root hangs off:
    with one rank per GPU on a two-socket host, uploads that cross the socket link share it with every other rank.
    Applied only when sysfs names a node and at least four of its CPUs are in this process's affinity mask; returns what
    was done for the JSON line."""
    import torch
    try:
        pr = torch.cuda.get_device_properties(local)
        bus = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node < 0:
            return {"gpu_pci": bus, "numa_node": node, "bound": False}
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        mine = os.sched_getaffinity(0)
        use = cpus & mine
        if len(use) < 4 or use == mine:
            return {"gpu_pci": bus, "numa_node": node, "bound": False, "cpus_allowed": len(mine), "cpus_on_node": len(use)}
        os.sched_setaffinity(0, use)
        return {"gpu_pci": bus, "numa_node": node, "bound": True, "cpus": len(use), "cpus_before": len(mine)}
    except Exception as e:                                   # no sysfs entry, no permission: run unbound
        return {"bound": False, "why": f"{type(e).__name__}: {e}"[:120]}


def h2d_only_leg(hv, dev, device, feeds, dist):
    """The uploads of the end-to-end leg and nothing else: one feed's frames per iteration, sliding through the same host
    ring, every rank at once.  What it reaches is the ceiling the host's PCIe/memory fabric puts on the end-to-end figure."""
    import torch
    n_pool = hv.shape[0]
    src = [torch.from_numpy(hv[t]) for t in range(n_pool)]
    dst = dev[0]
    for t in range(4):
        dst.copy_(src[t % n_pool], non_blocking=True)
    torch.cuda.synchronize(device)
    if dist is not None:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for t in range(feeds):
        dst.copy_(src[pool_index(t, n_pool)], non_blocking=True)
    e1.record()
    torch.cuda.synchronize(device)
    sec = e0.elapsed_time(e1) / 1e3
    if dist is not None:
        tt = torch.tensor([sec], dtype=torch.float64, device=f"cuda:{device}")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        sec = float(tt.item())
        dist.barrier()
    return sec


# ---------------------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------------------
def _frame_structs(track, base_ptr, n_pool, streams, h, w, keep):
    """UvioFrame[n_pool][streams] over a [n_pool, S, 2, H, W] uint8 block starting at base_ptr (host or device memory).
    Masks are NULL (= all zeros, include/uvio_klt.h): a caller that passes real masks pays a host scan of each per feed."""
    ids = (C.c_size_t * 2)(0, 1)
    strides = (C.c_size_t * 2)(w, w)
    ws = (C.c_int * 2)(w, w)
    hs = (C.c_int * 2)(h, h)
    masks = (C.c_void_p * 2)(None, None)
    keep.extend([ids, strides, ws, hs, masks])
    img_bytes = h * w
    steps = []
    for t in range(n_pool):
        arr = (track.UvioFrame * streams)()
        for s in range(streams):
            p0 = base_ptr + ((t * streams + s) * 2) * img_bytes
            imgs = (C.c_void_p * 2)(p0, p0 + img_bytes)
            keep.append(imgs)
            f = arr[s]
            f.timestamp = 0.05 * t
            f.n_img = 2
            f.sensor_ids = C.cast(ids, C.POINTER(C.c_size_t))
            f.img = C.cast(imgs, C.POINTER(C.c_void_p))
            f.img_stride = C.cast(strides, C.POINTER(C.c_size_t))
            f.mask = C.cast(masks, C.POINTER(C.c_void_p))
            f.mask_stride = C.cast(strides, C.POINTER(C.c_size_t))
            f.width = C.cast(ws, C.POINTER(C.c_int))
            f.height = C.cast(hs, C.POINTER(C.c_int))
        steps.append(arr)
    return steps


def _run_leg(track, L, cfg, device, streams, pool, warm_feeds, timed_feeds, dist, obs_buf, counts, sample_clocks, prof_enable, announce=True,
             sync_after_submit=False):
    """warm_feeds + timed_feeds batched feeds on fresh trackers; returns (seconds max-over-ranks, n_obs, clocks, stats, profile).
    announce: feed t+1 is announced (uvio_klt_submit_batch: its upload and pre-processing start) right before feed t is
    issued, so every timed feed has exactly one upload inside the timed region -- the one of the feed after it."""
    import torch
    trackers = [track.make_tracker(cfg, device=device) for _ in range(streams)]
    hs = (C.c_void_p * streams)(*[t._h for t in trackers])
    cap = obs_buf.shape[1]
    n_obs = 0
    n_pool = len(pool)
    if announce:
        rc = L.uvio_klt_submit_batch(hs, pool[pool_index(0, n_pool)], streams)
        if rc != 0:
            raise RuntimeError(f"uvio_klt_submit_batch failed: {rc}")
    call_s = [0.0, 0.0, 0.0]          # wall-clock of the three ABI calls of a feed, summed over the timed feeds

    def one(t, timed):
        t0 = time.perf_counter()
        if announce:
            rc = L.uvio_klt_submit_batch(hs, pool[pool_index(t + 1, n_pool)], streams)
            if rc != 0:
                raise RuntimeError(f"uvio_klt_submit_batch failed: {rc}")
            if sync_after_submit:               # the announced frames' image passes run to completion with the GPU to themselves
                time.sleep(0.0005)              # (the library's uploader thread issues them: give it the time to, then wait for the device)
                torch.cuda.synchronize(device)
        t1 = time.perf_counter()
        rc = L.uvio_klt_feed_batch(hs, pool[pool_index(t, n_pool)], streams)
        if rc != 0:
            raise RuntimeError(f"uvio_klt_feed_batch failed: {rc} (cuda {L.uvio_klt_cuda_error(trackers[0]._h)})")
        t2 = time.perf_counter()
        L.uvio_klt_get_new_observations_batch(hs, streams, obs_buf.ctypes.data, cap, counts)
        t3 = time.perf_counter()
        if timed:
            call_s[0] += t1 - t0
            call_s[1] += t2 - t1
            call_s[2] += t3 - t2
        return int(sum(counts))

    for t in range(warm_feeds):
        one(t, False)
    torch.cuda.synchronize(device)
    if dist is not None:
        dist.barrier()
    track.profile(device, enable=prof_enable)
    track.host_profile(device, reset=True)
    s0 = track.device_stats(device)
    clocks = ClockSampler(device) if sample_clocks else None
    L.uvio_klt_timer_start(device)
    t_wall = time.perf_counter()
    for t in range(warm_feeds, warm_feeds + timed_feeds):
        n_obs += one(t, True)
    ms = C.c_float()
    L.uvio_klt_timer_stop(device, C.byref(ms))
    torch.cuda.synchronize(device)
    wall = time.perf_counter() - t_wall
    clk = clocks.stop() if clocks else None
    prof = track.profile(device, enable=0)
    hostp = track.host_profile(device, reset=True)
    s1 = track.device_stats(device)
    sec = ms.value / 1e3
    if dist is not None:
        tt = torch.tensor([sec], dtype=torch.float64, device=f"cuda:{device}")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        sec = float(tt.item())
        dist.barrier()
    delta = {k: s1[k] - s0[k] for k in s0}
    passes = max(1.0, hostp.pop("lane_passes"))
    delta["host_us_per_group_pass"] = {k: round(v / passes, 1) for k, v in hostp.items()}
    delta["groups"] = passes / max(1, timed_feeds)
    delta["call_us_per_feed"] = {"submit_batch": round(1e6 * call_s[0] / timed_feeds, 1), "feed_batch": round(1e6 * call_s[1] / timed_feeds, 1),
                                 "get_new_observations_batch": round(1e6 * call_s[2] / timed_feeds, 1)}
    for t in trackers:
        t.close()
    return sec, n_obs, clk, delta, prof, wall


def _latency_leg(track, L, cfg, device, pool, n_frames, pause_s=0.0):
    """One stream fed frame by frame through uvio_klt_feed (host frames): wall-clock of the synchronous call."""
    t = track.make_tracker(cfg, device=device)
    obs = np.zeros((4096,), track.OBS_DTYPE)
    n = C.c_int()
    lat = []
    for k in range(n_frames):
        if pause_s:
            time.sleep(pause_s)
        t0 = time.perf_counter()
        rc = L.uvio_klt_feed(t._h, C.byref(pool[pool_index(k, len(pool))][0]))
        L.uvio_klt_get_new_observations(t._h, obs.ctypes.data, 4096, C.byref(n))
        lat.append(1e3 * (time.perf_counter() - t0))
        if rc != 0:
            raise RuntimeError(f"uvio_klt_feed failed: {rc}")
    t.close()
    lat = sorted(lat[5:])
    return {"p50": lat[len(lat) // 2], "p99": lat[min(len(lat) - 1, int(0.99 * len(lat)))], "frames": len(lat),
            "what": f"uvio_klt_feed + uvio_klt_get_new_observations, one {cfg.name} stereo stream, pinned host frames"
                    + (f", frames {1e3 * pause_s:.0f} ms apart" if pause_s else ", back to back")}


def run_ours(args):
    import torch
    from uvio_b200 import build, shard, track
    build.build()
    L = track.lib()
    cfg = workload_config(args.config)
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the tracker has no CPU path")
    torch.cuda.set_device(local)
    numa = bind_near_gpu(local) if os.environ.get("UVIO_BENCH_NUMA", "1") != "0" else {"bound": False, "why": "UVIO_BENCH_NUMA=0"}
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist_mod.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
        dist = dist_mod
    S = args.streams or BENCH_CONFIGS[args.config][1]
    W, K, R, NP = args.warmup, args.steps, max(1, args.frames_per_step), max(2, args.pool)
    announce = os.environ.get("UVIO_BENCH_ANNOUNCE", "1") != "0"
    h, w = cfg.height, cfg.width
    frames = make_frames(cfg, rank, S, args.distinct, NP)
    hv_flat = pinned_hugepage_buffer(NP * S * 2 * h * w)
    hv = hv_flat.reshape(NP, S, 2, h, w)
    for t in range(NP):
        for s in range(S):
            hv[t, s, 0] = frames[t][s][0]
            hv[t, s, 1] = frames[t][s][1]
    del frames
    dev = torch.from_numpy(hv).to(f"cuda:{local}")
    keep = []
    dev_pool = _frame_structs(track, dev.data_ptr(), NP, S, h, w, keep)
    host_pool = _frame_structs(track, hv.ctypes.data, NP, S, h, w, keep)
    obs_buf = np.zeros((S, 4096), track.OBS_DTYPE)
    counts = (C.c_int * S)()
    # leg 1: frames resident in HBM; leg 2: end to end from pinned host memory (upload + results inside the timed region).  Both
    # run the way a caller runs the library: no event records between kernels (bracketing the k_lk launches costs 5 % and keeps
    # the chain from replaying its CUDA graphs: 72.0 k / 65.4 k frames/s without against 68.1 k / 61.8 k with, UVIO_BENCH_LKPROF=1).
    # leg 1b repeats leg 1 -- same frames, same warm-up, same number of feeds -- with two event records around every launch of the
    # dominant kernel (k_lk): the kernel's live duration for the roofline.  leg 3 brackets every kernel for the shares.
    lk_only = 1 | ((1 << track.KERNEL_NAMES.index("k_lk")) << 1)
    in_legs = lk_only if os.environ.get("UVIO_BENCH_LKPROF", "0") == "1" else 0
    wf, tf = W * R, K * R
    sec, n_obs, clk, st, prof, wall = _run_leg(track, L, cfg, local, S, dev_pool, wf, tf, dist, obs_buf, counts, True, in_legs, announce)
    sec2, n_obs2, _c2, st2, _p2, wall2 = _run_leg(track, L, cfg, local, S, host_pool, wf, tf, dist, obs_buf, counts, False, in_legs, announce)
    if in_legs:
        sec1b, st1b, prof1b = sec, st, prof
    else:
        sec1b, _n1b, _c1b, st1b, prof1b, _w1b = _run_leg(track, L, cfg, local, S, dev_pool, wf, tf, dist, obs_buf, counts, False, lk_only, announce)
    sec3, _n3, _c3, st3, prof3, _w3 = _run_leg(track, L, cfg, local, S, dev_pool, R, min(tf, 4 * R), dist, obs_buf, counts, False, 1, announce)
    # leg 4: frames resident, every announcement followed by a device synchronisation before the feed is issued: the image-wide
    # passes of the announced frames then run with the GPU to themselves, and their event brackets are the kernels' own durations
    sec4, _n4, _c4, st4, prof4, _w4 = _run_leg(track, L, cfg, local, S, dev_pool, R, min(tf, 2 * R), dist, obs_buf, counts, False, 1, True, True)
    sec_h2d = h2d_only_leg(hv, dev, local, min(tf, 8 * R), dist)
    h2d_feed_bytes = float(S * 2 * h * w)
    lat = _latency_leg(track, L, cfg, local, host_pool, 60)
    lat_paced = _latency_leg(track, L, cfg, local, host_pool, 60, pause_s=0.003)
    # the one collective of the path: run statistics all-gathered over NCCL / NVLink
    gathered = shard.gather_stats(dist, {"frames": S * tf, "observations": n_obs, "seconds": sec, "launches": st["launches"],
                                         "h2d_bytes": st2["h2d_bytes"], "d2h_bytes": st2["d2h_bytes"]}, device=f"cuda:{local}")
    total_frames = world * S * tf
    value = total_frames / sec
    e2e = total_frames / sec2
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": 1e3 * sec / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8 images, int32 FAST/LK inner products, fp32 LK normal equations, fp64 cornerSubPix/RANSAC",
        "data": "synthetic (uvio_b200.synth, seeded)",
        "config": config_dict(cfg, S, R, {"ms_per_feed": 1e3 * sec / tf, "timed_region_s": sec,
                                          "single_stream_frame_latency_ms": lat, "single_stream_frame_latency_ms_paced": lat_paced,
                                          "observations_per_frame": n_obs / max(1, S * tf),
                                          "pipeline": os.environ.get("UVIO_KLT_PIPELINE", "device"),
                                          "masks": "NULL (documented as all zeros)"}),
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": st2["h2d_bytes"] / K,
                "d2h_bytes_per_step": st2["d2h_bytes"] / K, "ms_per_step": 1e3 * sec2 / K, "ms_per_feed": 1e3 * sec2 / tf,
                "upload_only": {"GB_per_s_per_gpu": h2d_feed_bytes * min(tf, 8 * R) / sec_h2d / 1e9,
                                "frames_per_sec_ceiling": world * S * min(tf, 8 * R) / sec_h2d,
                                "what": "the same frames copied from the same host ring with nothing else running, all ranks at "
                                        "once (max over ranks): what the host's PCIe / memory fabric allows the end-to-end figure"},
                "host_numa": numa,
                "note": "uvio_klt_feed_batch + uvio_klt_get_new_observations_batch on pinned host frames (2 MiB-aligned transparent-huge-page "
                        "mapping, cudaHostRegister'ed)"
                        + ("; feed t+1 is announced with uvio_klt_submit_batch before feed t is issued, so its upload overlaps feed t "
                           "(exactly one upload of a feed's frames inside the timed region per timed feed)" if announce else "")},
        "gpu_launches": int(st["launches"]),
        "host": {"groups": st["groups"], "resident_us_per_group_pass": st["host_us_per_group_pass"],
                 "e2e_us_per_group_pass": st2["host_us_per_group_pass"],
                 "resident_call_us_per_feed": st["call_us_per_feed"], "e2e_call_us_per_feed": st2["call_us_per_feed"]},
        "stats_gather": {"backend": "nccl" if dist is not None else "none (1 rank)", "job": gathered["job"],
                         "frames_per_rank": [r["frames"] for r in gathered["per_rank"]]},
        "clocks": clk,
    }
    line["config"]["ms_per_feed_with_all_kernels_bracketed_by_events"] = 1e3 * sec3 / max(1, min(tf, 4 * R))
    line["config"]["ms_per_feed_with_k_lk_bracketed_by_events"] = 1e3 * sec1b / tf
    return line, (prof1b, prof3, st3, min(tf, 4 * R), prof4, min(tf, 2 * R)), st1b, tf, S, cfg, rank, world, dist


# ---------------------------------------------------------------------------------------------------------------------
# CPU arm: the cv2-driven restatement of the reference (oracle/) — the one place bench.py may execute oracle code
# ---------------------------------------------------------------------------------------------------------------------
def _cpu_worker(job):
    """One independent stream on one host thread; returns (frames, seconds, seconds inside the OpenCV kernels or None)."""
    stream_id, n_warm, n_frames, cv_threads, cfg_name, split = job
    import cv2
    cv2.setNumThreads(cv_threads)
    cv2.setRNGSeed(0)                                   # VioManager.cpp:66-67
    from oracle import trackklt_cv2 as T
    from uvio_b200.synth import SynthStereoStream
    cfg = workload_config(cfg_name)
    inside = [0.0]
    if split:                                           # time spent inside the OpenCV kernels; the rest is the Python driver
        def timed(fn):
            def wrapped(*a, **k):
                t0 = time.perf_counter()
                try:
                    return fn(*a, **k)
                finally:
                    inside[0] += time.perf_counter() - t0
            return wrapped
        for name in ("equalize", "build_pyramid", "fast_detect", "corner_subpix", "lk_track", "fundamental_ransac_mask"):
            setattr(T, name, timed(getattr(T, name)))
        T.Cam.undistort = timed(T.Cam.undistort)
    s = SynthStereoStream(cfg.width, cfg.height, stream_id=stream_id, events=False)
    frames = [s.frame(k) for k in range(n_warm + n_frames)]
    masks = [np.zeros((cfg.height, cfg.width), np.uint8)] * 2
    o = T.make_oracle(cfg)
    for k in range(n_warm):
        o.feed_new_camera(0.05 * k, [0, 1], list(frames[k]), masks)
    inside[0] = 0.0
    t0 = time.perf_counter()
    for k in range(n_warm, n_warm + n_frames):
        o.feed_new_camera(0.05 * k, [0, 1], list(frames[k]), masks)
    return n_frames, time.perf_counter() - t0, (inside[0] if split else None)


def cpu_throughput(n_procs, n_frames, n_warm=10, cv_threads=1, cfg_name="euroc_mav", split=False):
    """Aggregate frames/s of n_procs independent streams, one process each (SURVEY 8(d) CPU baseline plan)."""
    import multiprocessing as mp
    jobs = [(9000 + i, n_warm, n_frames, cv_threads, cfg_name, split) for i in range(n_procs)]
    with mp.get_context("spawn").Pool(n_procs) as pool:      # always a child process: the wrappers of `split` stay there
        res = pool.map(_cpu_worker, jobs)
    return sum(f / s for f, s, _i in res), max(s for _f, s, _i in res), res


def c_restatement_seconds_per_frame(cfg, n=6):
    """The no-Python cross-check: the plain-C restatement of every OpenCV routine on the path (oracle/stages.c),
    single thread, on the work of one steady-state stereo frame: equalise + pyramid of both images, FAST on 5 of the 25 cells
    per camera (a steady-state top-up), cornerSubPix of 30 corners, LK of 2 x num_features points."""
    from oracle import cstages
    from uvio_b200.synth import SynthStereoStream
    s = SynthStereoStream(cfg.width, cfg.height, stream_id=77, events=False)
    fr = [s.frame(k) for k in range(n + 1)]
    sx, sy = cfg.width // cfg.grid_x, cfg.height // cfg.grid_y
    rng = np.random.default_rng(1)
    pts = np.c_[rng.uniform(20, cfg.width - 20, cfg.num_features), rng.uniform(20, cfg.height - 20, cfg.num_features)].astype(np.float32)
    prev = None
    t_total = 0.0
    for k in range(n + 1):
        t0 = time.perf_counter()
        cur = []
        for img in fr[k]:
            if cfg.histogram == 2:
                eq = cstages.clahe(img)
            elif cfg.histogram == 1:
                eq = cstages.equalize_hist(img)
            else:
                eq = img
            cur.append(cstages.build_pyramid(eq, cfg.pyr_levels))
        for lv in cur:
            for c in range(5):
                gx, gy = c % cfg.grid_x, c // cfg.grid_x
                cstages.fast(np.ascontiguousarray(lv[0][gy * sy:(gy + 1) * sy, gx * sx:(gx + 1) * sx]), cfg.fast_threshold)
            cstages.corner_subpix(lv[0], pts[:15])
        if prev is not None:
            for a, b in zip(prev, cur):
                cstages.lk(a, b, pts, pts)
        dt = time.perf_counter() - t0
        if prev is not None:
            t_total += dt
        prev = cur
    return t_total / n


def cpu_baseline_entry(n_frames, cfg_name):
    import cv2
    cores = os.cpu_count() or 1
    cfg = workload_config(cfg_name)
    # reference default: one tracker with cv::setNumThreads(4) (VioManagerOptions.h:412)
    fps, sec, res = cpu_throughput(1, n_frames, cv_threads=4, cfg_name=cfg_name, split=True)
    inside = res[0][2]
    entry = {"value": fps, "unit": UNIT, "cores": 4, "kind": "port",
             "sample": f"{n_frames} stereo frames of one synthetic {cfg.name} stream after 10 warm-up frames, cv2 {cv2.__version__} (IPP off), "
                       f"cv::setNumThreads(4) as the reference default; host has {cores} logical cores; {sec:.1f} s of CPU work",
             "what": "oracle/trackklt_cv2.py: the reference's TrackKLT control flow restated line by line over the real OpenCV kernels",
             "python_driver_share": round(1.0 - inside / sec, 3),
             "python_driver_note": "share of the sample's wall time spent OUTSIDE the OpenCV kernels (the restated control flow in Python): "
                                   "the C++ reference is at most this much faster than the port"}
    try:
        fps_all, sec_all, _r = cpu_throughput(1, max(30, n_frames // 2), cv_threads=-1, cfg_name=cfg_name)
        entry["all_threads"] = {"value": fps_all, "unit": UNIT, "what": "the same stream with cv::setNumThreads(-1) (all cores)", "cores": cores}
    except Exception as e:      # noqa: BLE001
        entry["all_threads"] = {"error": str(e)}
    try:
        spf = c_restatement_seconds_per_frame(cfg)
        entry["c_restatement"] = {"value": 1.0 / spf, "unit": UNIT, "cores": 1,
                                  "what": "oracle/stages.c (plain C, no Python, no OpenCV, one thread) on the kernels of one steady-state stereo "
                                          "frame: equalise + pyramid x2, FAST on 5 cells per camera, cornerSubPix of 30 corners, LK of 2 x N points"}
    except Exception as e:      # noqa: BLE001
        entry["c_restatement"] = {"error": str(e)}
    return entry


def run_reference(args):
    """--impl reference: the CPU tracker on all host cores, one independent stream per process, cv2 single-threaded each."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None
    import cv2
    cfg = workload_config(args.config)
    cores = os.cpu_count() or 1
    procs = max(1, min(cores, 64))
    per_step = 4                                       # frames per stream per "step" (bounded sample of a step's frames)
    n_frames = max(8, per_step * args.steps)
    n_warm = max(3, per_step * args.warmup)
    fps, sec, _res = cpu_throughput(procs, n_frames, n_warm=n_warm, cv_threads=1, cfg_name=args.config)
    sample = (f"{procs} independent streams x {n_frames} stereo frames ({per_step} per step) after {n_warm} warm-up frames, one process "
              f"per stream, cv2 {cv2.__version__} single-threaded per process, IPP off")
    S = args.streams or BENCH_CONFIGS[args.config][1]
    return {
        "impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sec / max(1, args.steps), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8 / fp32 / fp64 (OpenCV CPU)", "data": "synthetic (uvio_b200.synth, seeded)",
        "config": config_dict(cfg, S, max(1, args.frames_per_step),
                              {"note": f"CPU arm: {procs} concurrent host processes stand in for the streams; each step is a bounded sample "
                                       f"({per_step} frames per process) of the GPU arm's step"}),
        "cpu_baseline": {"value": fps, "unit": UNIT, "cores": procs, "kind": "port", "sample": sample},
        "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


# ---------------------------------------------------------------------------------------------------------------------
# rooflines.  k_lk (the kernel with the largest share of GPU time) is bound by instruction issue, not by HBM: its entry is
# warp-instructions per second against 148 SMs x 4 schedulers x SM clock, instructions per feature taken from the committed
# ncu capture of the same kernel, duration measured live with CUDA events around every launch of the timed region.  The
# image-wide pass (k_preprocess) is the HBM-shaped kernel of the path: roofline_image_passes.
# ---------------------------------------------------------------------------------------------------------------------
def _levels(cfg):
    lv = [(cfg.width, cfg.height)]
    w, h = cfg.width, cfg.height
    for _ in range(cfg.pyr_levels):
        nw, nh = (w + 1) // 2, (h + 1) // 2
        if nw <= 15 or nh <= 15:
            break
        w, h = nw, nh
        lv.append((w, h))
    return lv


def _load_profile(name):
    for rnd in ("r2", "r1"):
        p = os.path.join(ROOT, "profiles", f"{rnd}_{name}")
        if os.path.exists(p):
            return json.load(open(p)), os.path.relpath(p, ROOT)
    return None, None


def roofline_entry(profs, stats, feeds, S, cfg, clk):
    prof_timed, prof_all, stats_all, feeds_prof, prof_alone, feeds_alone = profs   # feeds_prof: feeds of the leg that bracketed every kernel
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak_hbm = float(json.load(open(peaks_path))["hbm_gbs"])
        peak_src = "MEASURED_PEAKS.json hbm_gbs (measured)"
    else:
        peak_hbm, peak_src = 6650.0, "B200_PROFILING.md fallback"
    total_ms = sum(ms for ms, _n in prof_all.values())
    share = {k: round(v[0] / total_ms, 4) for k, v in prof_all.items() if total_ms and v[0] > 0}
    lv = _levels(cfg)
    P = cfg.width * cfg.height
    sig = sum(a * b for a, b in lv)
    # ---- dominant kernel -------------------------------------------------------------------------------------------------
    name = max(prof_all.items(), key=lambda kv: kv[1][0])[0]
    ms, launches = prof_timed.get(name, (0.0, 0))
    in_timed = launches > 0
    if not in_timed:
        ms, launches = prof_all[name]
    entry = {"kernel": name, "launches": launches, "avg_launch_ms": ms / max(1, launches),
             "kernel_time_share_of_gpu_time": share, "sum_of_kernel_durations_ms_per_feed": total_ms / max(1, feeds_prof),
             "timing": ("CUDA events around every launch of this kernel, on the launching streams, in a leg of its own that repeats the timed "
                        "region of `value` (same frames, warm-up and number of feeds; config.ms_per_feed_with_k_lk_bracketed_by_events): the "
                        "throughput legs carry no event records; the shares come from a short leg with every kernel bracketed") if in_timed else
                       "separate profiling leg (the kernel was not bracketed in the timed region)"}
    if name == "k_lk":
        cap, cap_path = _load_profile("lk_ncu.json")
        feats = stats["lk_point_levels"] / max(1, len(lv))                      # features tracked in the timed leg
        sm_mhz = (clk or {}).get("sm_mhz") or 1965.0
        peak = 148 * 4 * sm_mhz * 1e6 / 1e9                                     # G warp-instructions / s
        if cap:
            inst = cap["smsp__inst_executed_sum"] / cap["features_in_launch"] * feats
            ach = inst / (ms / 1e3) / 1e9
            entry.update({"bound": "issue", "unit": "G warp-inst/s", "achieved": ach, "peak": peak, "frac": ach / peak,
                          "peak_source": f"148 SMs x 4 schedulers x {sm_mhz:.0f} MHz (SM clock sampled during the timed region)",
                          "warp_instructions_per_feature": cap["smsp__inst_executed_sum"] / cap["features_in_launch"],
                          "instructions_source": cap_path + ": " + cap.get("capture", ""),
                          "fp32_pipe_utilisation": cap.get("fma_pipe_utilisation"), "alu_pipe_utilisation": cap.get("alu_pipe_utilisation"),
                          "ncu_issue_slot_utilisation": cap.get("issue_active"), "ncu_duration_us": cap.get("duration_us"),
                          "traffic": (cap["dram_bytes_read"] + cap["dram_bytes_write"]) / cap["features_in_launch"] * feats / max(1, launches),
                          "algorithmic_bytes_per_launch": (324 * stats["lk_point_levels"] + 256 * stats["lk_iterations"]) / max(1, launches),
                          "features_per_launch": feats / max(1, launches),
                          "alone": ({"achieved": cap["smsp__inst_executed_sum"] / (cap["duration_us_sum"] * 1e-6) / 1e9,
                                     "frac": cap["smsp__inst_executed_sum"] / (cap["duration_us_sum"] * 1e-6) / 1e9 / peak,
                                     "what": "the same kernel with the GPU to itself: warp instructions / duration summed over every k_lk launch of the "
                                             "capture (ncu serialises launches); `achieved` above divides by event-bracketed durations of launches that "
                                             "share the SMs with the other kernels of two trackers' groups"} if cap.get("duration_us_sum") else None),
                          "note": "LK is an L2 / shared-memory gather with ordered fp32 accumulation: bound by instruction issue, not by HBM "
                                  "(traffic = DRAM bytes per launch scaled from the capture; algorithmic bytes = 324 B per point-level + 256 B "
                                  "per iteration, SURVEY 8(d), for reference)"})
        else:
            entry.update({"bound": "issue", "achieved": None, "frac": None, "note": "no ncu capture of k_lk under profiles/"})
    else:
        imgs = 2 * S * feeds
        alg = {"k_preprocess": imgs * (3 * P + (sig - lv[-1][0] * lv[-1][1]) + (sig - P)), "k_fast": stats["fast_pixels"]}.get(name)
        if alg and ms > 0:
            ach = alg / (ms / 1e3) / 1e9
            entry.update({"bound": "hbm", "unit": "GB/s", "achieved": ach, "peak": peak_hbm, "peak_source": peak_src, "frac": ach / peak_hbm,
                          "traffic": None, "algorithmic_bytes_per_launch": alg / max(1, launches)})
        else:
            entry.update({"bound": "latency", "achieved": None, "frac": None,
                          "note": "the dominant kernel is latency-bound per-tracker work with no meaningful byte or flop count"})
    # ---- the image-wide passes: equalisation (k_preprocess, k_clahe_*) and the pyramid levels (k_pyrdown: launches of their own) ------
    # timed in the leg that synchronises the device after every announcement, so that the announced frames' passes run alone;
    # in the throughput legs the same launches queue behind the LK blocks of the feed in flight and their event brackets
    # measure the wait (reported as `under_load`)
    image_pass = None
    def _ms(pr, names):
        return sum(pr.get(k, (0.0, 0))[0] for k in names), sum(pr.get(k, (0.0, 0))[1] for k in names)
    eq_k, pyr_k = ("k_preprocess", "k_clahe_tiles", "k_clahe_apply"), ("k_pyrdown",)
    ems, el = _ms(prof_alone, eq_k)
    dms, dl = _ms(prof_alone, pyr_k)
    if el > 0 and ems > 0:
        imgs = 2 * S * feeds_alone
        eq_bytes = 3 * P + (3 * P if cfg.histogram == 2 else 0)                  # hist read + apply read + write (CLAHE: its two passes in front)
        pyr_bytes = (sig - lv[-1][0] * lv[-1][1]) + (sig - P)                   # every level read once to be halved, every lower level written once
        per_img = eq_bytes + pyr_bytes
        ach = imgs * per_img / ((ems + dms) / 1e3) / 1e9
        pl = prof_alone.get("k_preprocess", (0.0, 1))[1]
        ums, _ul = _ms(prof_all, eq_k + pyr_k)
        image_pass = {"bound": "hbm", "kernel": "k_preprocess + k_pyrdown" if dl else "k_preprocess", "unit": "GB/s", "achieved": ach, "peak": peak_hbm,
                      "peak_source": peak_src, "frac": ach / peak_hbm, "launches": el + dl, "avg_launch_ms": (ems + dms) / max(1, pl),
                      "images_per_launch": imgs / max(1, pl), "algorithmic_bytes_per_image": per_img, "traffic": None,
                      "parts": {"equalise": {"ms": ems, "launches": el, "GB_per_s": imgs * (eq_bytes if dl else per_img) / (ems / 1e3) / 1e9},
                                "pyramid": ({"ms": dms, "launches": dl, "GB_per_s": imgs * pyr_bytes / (dms / 1e3) / 1e9} if dl and dms > 0 else None)},
                      "under_load": ({"achieved": 2 * S * feeds_prof * per_img / (ums / 1e3) / 1e9, "what": "the same launches in the announced legs, "
                                      "queued behind the LK of the feed in flight (event brackets include the wait)"} if ums > 0 else None),
                      "timing": "CUDA events around every launch, on the launching stream, in a leg that synchronises the device between the "
                                "announcement of a feed (uvio_klt_submit_batch: its image passes) and the feed before it (resident frames)"}
        cap, cap_path = _load_profile("preprocess_traffic.json")
        if cap:
            image_pass["traffic"] = (cap["dram_bytes_read"] + cap["dram_bytes_write"]) / cap["images_in_launch"] * imgs / max(1, pl)
            image_pass["traffic_source"] = cap_path + ": " + cap.get("capture", "") + " (DRAM bytes of the k_preprocess launch only)"
    entry["image_pass"] = image_pass
    return entry


def main():
    args = parse_args()
    if args.impl == "reference":
        line = run_reference(args)
        if line is not None:
            print(json.dumps(line), flush=True)
        return
    line, prof, st, feeds, S, cfg, rank, world, dist = run_ours(args)
    if rank == 0:
        line["roofline"] = roofline_entry(prof, st, feeds, S, cfg, line.get("clocks"))
        line["roofline_image_passes"] = line["roofline"].pop("image_pass")
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_entry(args.cpu_frames, args.config)
        else:
            line["cpu_baseline"] = None
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
