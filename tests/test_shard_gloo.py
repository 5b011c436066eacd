"""Multi-GPU host logic on CPU (not gpu): view-range sharding and the counter gather, world_size 2 over gloo."""
import importlib
import os
import socket
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_view_ranges_partition_the_batch():
    shard = importlib.import_module("xash3d-fwgs_b200.shard")
    for n in (0, 1, 7, 256, 4096 * 8 + 3):
        for world in (1, 2, 3, 8):
            r = [shard.view_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[i][1] == r[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard.view_range(10, 2, 2)


def _worker(rank, world, port, q):
    import torch.distributed as dist

    sys.path.insert(0, ROOT)
    shard = importlib.import_module("xash3d-fwgs_b200.shard")
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    lo, hi = shard.view_range(257, rank, world)
    local = {"views": hi - lo, "pixels": (hi - lo) * 320 * 240, "overflow_edges": rank}
    tot = shard.gather_counters(local)
    q.put((rank, lo, hi, tot))
    dist.destroy_process_group()


def test_counter_gather_world2_gloo():
    import torch.multiprocessing as mp

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(timeout=60)
    assert res[0][1:3] == (0, 129) and res[1][1:3] == (129, 257)
    for _, _, _, tot in res:
        assert tot == {"views": 257, "pixels": 257 * 320 * 240, "overflow_edges": 1}
