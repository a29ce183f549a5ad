"""Host->device copy rate of the box for the shapes bench.py's end-to-end leg uses, by how the host buffer was allocated.
Finding this probes: a 23 MB pinned buffer uploads at ~17 GB/s while 3 MB and 256 MB ones reach ~51 GB/s."""
import ctypes, json, mmap, os
import numpy as np
import torch

rt = torch.cuda.cudart()
torch.cuda.init()

def timed(fn, reps=20):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    fn(); torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps / 1e3

def pinned_torch(n):
    return torch.empty(n, dtype=torch.uint8).pin_memory()

_keep = []
def pinned_huge(n):
    """2 MiB-aligned anonymous mapping advised to transparent huge pages, touched, then cudaHostRegister'ed."""
    H = 2 << 20
    size = (n + H - 1) // H * H + H
    m = mmap.mmap(-1, size, flags=mmap.MAP_PRIVATE | mmap.MAP_ANONYMOUS)
    a = np.frombuffer(m, dtype=np.uint8)
    base = a.ctypes.data
    off = (-base) % H
    try:
        m.madvise(mmap.MADV_HUGEPAGE)
    except Exception as e:
        print("madvise failed", e)
    a[:] = 1
    v = a[off:off + n]
    r = rt.cudaHostRegister(v.ctypes.data, (n + 4095) // 4096 * 4096, 0)
    assert int(r) == 0, r
    _keep.append((m, a))
    return torch.from_numpy(v)

out = {}
for k in ("enabled", "defrag", "shmem_enabled"):
    try:
        out["thp_" + k] = open("/sys/kernel/mm/transparent_hugepage/" + k).read().strip()
    except Exception as e:
        out["thp_" + k] = str(e)
try:
    out["anon_huge_kb_before"] = [l for l in open("/proc/meminfo") if "AnonHugePages" in l][0].split()[1]
except Exception:
    pass
frame = 2 * 752 * 480
for label, alloc in (("torch_pin", pinned_torch), ("huge_reg", pinned_huge)):
    for mb_name, n in (("2.9MB", 4 * frame), ("5.8MB", 8 * frame), ("11.6MB", 16 * frame), ("23.1MB", 32 * frame), ("46MB", 64 * frame),
                       ("128MB", 128 << 20), ("808MB", 35 * 32 * frame)):
        h = alloc(n)
        d = torch.empty(n, dtype=torch.uint8, device="cuda")
        s = timed(lambda: d.copy_(h, non_blocking=True), reps=10)
        out[f"{label}_{mb_name}"] = round(n / s / 1e9, 1)
        if mb_name == "808MB":   # the bench's pattern: 23.1 MB windows sliding through one large buffer
            step = 32 * frame
            dd = torch.empty(step, dtype=torch.uint8, device="cuda")
            i = [0]
            def slide():
                k = i[0] % 35
                dd.copy_(h[k * step:(k + 1) * step], non_blocking=True)
                i[0] += 1
            s = timed(slide, reps=35)
            out[f"{label}_808MB_sliding_23.1MB"] = round(step / s / 1e9, 1)
            # and split in 8 lane-sized copies on 8 streams, as the lanes issue them
            streams = [torch.cuda.Stream() for _ in range(8)]
            lane = step // 8
            def lanes():
                k = i[0] % 35
                evs = []
                for l, st in enumerate(streams):
                    with torch.cuda.stream(st):
                        dd[l * lane:(l + 1) * lane].copy_(h[k * step + l * lane:k * step + (l + 1) * lane], non_blocking=True)
                        e = torch.cuda.Event(); e.record(st); evs.append(e)
                for e in evs:
                    torch.cuda.current_stream().wait_event(e)
                i[0] += 1
            s = timed(lanes, reps=35)
            out[f"{label}_808MB_sliding_8lanes"] = round(step / s / 1e9, 1)
        del h, d
try:
    out["anon_huge_kb_after"] = [l for l in open("/proc/meminfo") if "AnonHugePages" in l][0].split()[1]
except Exception:
    pass
print(json.dumps(out))
