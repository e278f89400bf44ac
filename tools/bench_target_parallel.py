"""BASELINE.json configs[4]: ONE track, the 4 stem networks split over `world` GPUs (umx_cuda_session_init_target_parallel),
fc3 masks all-gathered over NCCL before Wiener-EM.  One process per GPU; wall clock of rank 0 around demix (pinned host buffers),
best of 3 after a warm-up.  python tools/bench_target_parallel.py [world] [seconds]"""
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def worker(rank, world, tmp, seconds):
    sys.path.insert(0, ROOT)
    import numpy as np
    import torch
    import __graft_entry__ as g
    from oracle import synth

    umx = g.load_package()
    idfile = os.path.join(tmp, "nccl_id")
    if rank == 0:
        with open(idfile + ".tmp", "wb") as f:
            f.write(umx.nccl_unique_id())
        os.replace(idfile + ".tmp", idfile)
    while not os.path.exists(idfile):
        time.sleep(0.05)
    uid = open(idfile, "rb").read()
    torch.cuda.set_device(rank)
    model = umx.load_umx_model(synth.make_ggml(hidden=1024, seed=1234), device=rank)
    n = int(seconds * 44100)
    a = synth.make_audio(n, 7)
    host_in = torch.from_numpy(np.ascontiguousarray(a.T)).pin_memory()
    host_out = torch.empty((4, n, 2), dtype=torch.float32).pin_memory()
    sess = umx.Session(model, 1, umx.SEGMENT_SAMPLES)
    if world > 1:
        sess.init_target_parallel(uid, rank, world)
    ts = []
    for it in range(4):
        open(os.path.join(tmp, f"ready_{it}_{rank}"), "w").close()
        while not all(os.path.exists(os.path.join(tmp, f"ready_{it}_{r}")) for r in range(world)):
            time.sleep(0.001)
        t = time.perf_counter()
        sess.demix_into([host_in.numpy()], [host_out.numpy()])
        ts.append(time.perf_counter() - t)
    if rank == 0:
        best = min(ts[1:])
        print(f"target-parallel, world {world}, 1 x {seconds:.0f} s track: {best * 1e3:.1f} ms -> {seconds / best:.0f} audio-s/s (rank 0, its {4 // world} stems)")


if __name__ == "__main__":
    import torch.multiprocessing as mp

    world = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    seconds = float(sys.argv[2]) if len(sys.argv) > 2 else 420.0
    ctx = mp.get_context("spawn")
    with tempfile.TemporaryDirectory() as td:
        ps = [ctx.Process(target=worker, args=(r, world, td, seconds)) for r in range(world)]
        for p in ps:
            p.start()
        for p in ps:
            p.join(timeout=900)
