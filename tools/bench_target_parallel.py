"""BASELINE.json configs[4]: the 4 stem networks split over `world` GPUs (umx_cuda_session_init_target_parallel), fc3 masks
all-gathered over NCCL before Wiener-EM (on its own stream in the segment wavefront, so it overlaps the next segment's network).
One process per GPU; wall clock of rank 0 around demix (pinned host buffers), best of 3 after a warm-up; the all-gather's own
time comes from the library's stage timers and gives the achieved NVLink rate.
    python tools/bench_target_parallel.py [world] [seconds] [tracks]"""
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def worker(rank, world, tmp, seconds, tracks):
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
    host_in = [torch.from_numpy(np.ascontiguousarray(synth.make_audio(n, 7 + i).T)).pin_memory() for i in range(tracks)]
    host_out = [torch.empty((4, n, 2), dtype=torch.float32).pin_memory() for _ in range(tracks)]
    sess = umx.Session(model, tracks, umx.SEGMENT_SAMPLES)
    if world > 1:
        sess.init_target_parallel(uid, rank, world)
    sess.set_profiling(True)
    ts = []
    for it in range(4):
        open(os.path.join(tmp, f"ready_{it}_{rank}"), "w").close()
        while not all(os.path.exists(os.path.join(tmp, f"ready_{it}_{r}")) for r in range(world)):
            time.sleep(0.001)
        t = time.perf_counter()
        sess.demix_into([h.numpy() for h in host_in], [h.numpy() for h in host_out])
        ts.append(time.perf_counter() - t)
    if rank == 0:
        best = min(ts[1:])
        line = (f"target-parallel, world {world}, {tracks} x {seconds:.0f} s track(s): {best * 1e3:.1f} ms -> "
                f"{tracks * seconds / best:.0f} audio-s/s (rank 0, its {4 // world} stems)")
        if world > 1:
            segs = len(umx.segment_plan(n))
            ag = sess.stage_ms("allgather")  # busy time of the all-gather stream in the last call
            recv = (4 - 4 // world) * tracks * 2584 * 4104 * 4 * segs  # bytes this rank receives per call
            line += (f"; all-gather of the masks: {segs} x {recv / segs / 1e6:.0f} MB received per rank in {ag:.2f} ms = "
                     f"{recv / 1e9 / (ag / 1e3):.0f} GB/s per rank (NVLink 5: 900 GB/s per direction)")
        print(line, flush=True)


if __name__ == "__main__":
    import torch.multiprocessing as mp

    world = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    seconds = float(sys.argv[2]) if len(sys.argv) > 2 else 420.0
    tracks = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    ctx = mp.get_context("spawn")
    with tempfile.TemporaryDirectory() as td:
        ps = [ctx.Process(target=worker, args=(r, world, td, seconds, tracks)) for r in range(world)]
        for p in ps:
            p.start()
        for p in ps:
            p.join(timeout=900)
