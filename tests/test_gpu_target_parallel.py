"""GPU (>= 2 devices): target-parallel mode (BASELINE config 5).  One process per GPU; rank j runs its share of
the 4 stem networks, the fc3 masks are all-gathered over NCCL before the Wiener stage, and every rank writes
its own stems.  The union of the ranks' stems must equal the single-GPU result."""
import os
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, tmp, n_buf):
    import sys

    sys.path.insert(0, ROOT)
    import __graft_entry__ as g
    from oracle import synth

    umx = g.load_package()
    idfile = os.path.join(tmp, "nccl_id")
    if rank == 0:
        with open(idfile + ".tmp", "wb") as f:
            f.write(umx.nccl_unique_id())
        os.replace(idfile + ".tmp", idfile)
    t0 = time.time()
    while not os.path.exists(idfile):
        if time.time() - t0 > 120:
            raise RuntimeError("no NCCL id")
        time.sleep(0.05)
    uid = open(idfile, "rb").read()
    model = umx.load_umx_model(synth.make_ggml(hidden=1024, seed=1234), device=rank)
    tracks = [synth.make_audio(n, 60 + i) for i, n in enumerate((30000, 41000))]
    sess = umx.Session(model, len(tracks), n_buf)
    if rank == 0:  # single-GPU result first
        ref_out = sess.demix(tracks)
        np.save(os.path.join(tmp, "single.npy"), np.stack([np.pad(o, ((0, 0), (0, 0), (0, 41000 - o.shape[2]))) for o in ref_out]))
    sess.init_target_parallel(uid, rank, world)
    out = sess.demix(tracks)
    nt = 4 // world
    np.save(os.path.join(tmp, f"tp_rank{rank}.npy"),
            np.stack([np.pad(o[rank * nt:(rank + 1) * nt], ((0, 0), (0, 0), (0, 41000 - o.shape[2]))) for o in out]))


def _run(world, tmp_path):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_worker, args=(r, world, str(tmp_path), 16384)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=600)
        assert p.exitcode == 0
    single = np.load(tmp_path / "single.npy")  # (tracks, 4, 2, n)
    nt = 4 // world
    for r in range(world):
        part = np.load(tmp_path / f"tp_rank{r}.npy")  # (tracks, nt, 2, n)
        assert np.abs(part - single[:, r * nt:(r + 1) * nt]).max() < 2e-6


def test_target_parallel_matches_single_gpu(umx, tmp_path):
    n = umx.lib().umx_cuda_device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    _run(4 if n >= 4 else 2, tmp_path)
