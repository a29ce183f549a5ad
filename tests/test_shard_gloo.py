"""CPU, world_size 2, gloo: the N > 1 path of bench.py — static stream assignment and the statistics all-gather."""
import os
import socket

import torch.distributed as dist
import torch.multiprocessing as mp

from uvio_b200.shard import assign_streams, gather_stats


def test_assign_streams_partitions_exactly():
    for n, world in ((64, 8), (7, 2), (3, 4)):
        parts = [assign_streams(n, world, r) for r in range(world)]
        assert sorted(s for p in parts for s in p) == list(range(n))
        assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = assign_streams(10, world, rank)
    res = gather_stats(dist, {"frames": 100.0 * len(mine), "observations": 5.0, "seconds": 1.0 + rank, "launches": 7})
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, res))


def test_gather_stats_world_size_2():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for r in (0, 1):
        job = out[r]["job"]
        assert job["frames"] == 1000.0 and job["seconds"] == 2.0 and job["frames_per_sec"] == 500.0 and job["launches"] == 14
        assert [row["frames"] for row in out[r]["per_rank"]] == [500.0, 500.0]
