"""CPU: the N>1 host-side path (track sharding + measurement aggregation) with the gloo backend, world_size 2."""
import os
import socket

import pytest

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_tracks, q):
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import __graft_entry__ as g

    umx = g.load_package()
    from importlib import import_module

    sharding = import_module("umx_cpp_b200.sharding")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = sharding.shard_tracks(n_tracks, rank, world)
    gathered = [None] * world
    dist.all_gather_object(gathered, list(mine))
    # every rank demixes len(mine) tracks of 180 s; pretend rank r needed (100 + 10 r) ms
    total_s, max_ms = sharding.aggregate(len(mine) * 180.0, 100.0 + 10.0 * rank, dist)
    # the segment plan is rank-independent host logic (no GPU needed)
    plan = umx.segment_plan(7938000)
    q.put((rank, gathered, total_s, max_ms, len(plan)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n_tracks", [5, 64])
def test_track_sharding_world2(n_tracks):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_tracks, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, gathered, total_s, max_ms, n_seg in res:
        flat = [t for part in gathered for t in part]
        assert sorted(flat) == list(range(n_tracks))          # every track exactly once
        assert abs(len(gathered[0]) - len(gathered[1])) <= 1  # balanced
        assert total_s == n_tracks * 180.0                    # whole-job audio seconds
        assert max_ms == 110.0                                # max over ranks, not a mean
        assert n_seg == 5


def test_shard_tracks_edges():
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import __graft_entry__ as g

    g.load_package()
    from importlib import import_module

    sh = import_module("umx_cpp_b200.sharding")
    assert list(sh.shard_tracks(256, 3, 8)) == list(range(96, 128))
    assert [len(sh.shard_tracks(10, r, 4)) for r in range(4)] == [3, 3, 2, 2]
    assert list(sh.shard_tracks(0, 0, 2)) == [] and list(sh.shard_tracks(1, 1, 2)) == []
    with pytest.raises(ValueError):
        sh.shard_tracks(4, 2, 2)
