"""Copy-only ceiling of bench.py's end-to-end arm: N processes (one per GPU), each streaming exactly the bench's own copy pattern
per step — `--tracks` x 180 s tracks: 8 B/sample host->device and 32 B/sample device->host, pinned memory, one cudaMemcpyAsync
per track and segment (H2D) / per stem, track and segment (D2H), both directions at once on two streams — with NO compute.
Prints one JSON line (rank 0): GB/s per rank and node-wide, and the per-step time that bounds e2e from below.

    python tools/host_copy_ceiling.py                       (1 GPU)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 tools/host_copy_ceiling.py
Options: --no-affinity (do not pin the rank's threads to its share of the host cores), --one-direction h2d|d2h."""
import argparse
import json
import os
import time

import torch
import torch.distributed as dist

ap = argparse.ArgumentParser()
ap.add_argument("--tracks", type=int, default=32)
ap.add_argument("--track-seconds", type=float, default=180.0)
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--segments", type=int, default=5)
ap.add_argument("--no-affinity", action="store_true")
ap.add_argument("--one-direction", default="", choices=["", "h2d", "d2h"])
ap.add_argument("--d2h-2d", action="store_true", help="one strided torch copy_ per track instead of one contiguous copy per stem "
                "(torch stages such copies itself: ~30x slower than the cudaMemcpy2DAsync the library issues; for the record only)")
a = ap.parse_args()
rank, local, world = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("LOCAL_RANK", 0), ("WORLD_SIZE", 1)))
if world > 1 and not a.no_affinity and hasattr(os, "sched_setaffinity"):
    cpus = sorted(os.sched_getaffinity(0))
    per = max(1, len(cpus) // world)
    os.sched_setaffinity(0, set(cpus[local * per:(local + 1) * per]) or set(cpus))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
n, T, S = int(a.track_seconds * 44100), a.tracks, a.segments
h_in = [torch.empty((n, 2), dtype=torch.float32).pin_memory() for _ in range(T)]
h_out = [torch.empty((4, n, 2), dtype=torch.float32).pin_memory() for _ in range(T)]
d_in = [torch.empty((n, 2), dtype=torch.float32, device=dev) for _ in range(T)]
d_out = [torch.zeros((4, n, 2), dtype=torch.float32, device=dev) for _ in range(T)]
s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
bounds = [n * k // S for k in range(S + 1)]


def step():
    for k in range(S):
        lo, hi = bounds[k], bounds[k + 1]
        if a.one_direction in ("", "h2d"):
            with torch.cuda.stream(s_in):
                for i in range(T):
                    d_in[i][lo:hi].copy_(h_in[i][lo:hi], non_blocking=True)
        if a.one_direction in ("", "d2h"):
            with torch.cuda.stream(s_out):
                for i in range(T):
                    if not a.d2h_2d:
                        for j in range(4):
                            h_out[i][j, lo:hi].copy_(d_out[i][j, lo:hi], non_blocking=True)
                    else:
                        h_out[i][:, lo:hi].copy_(d_out[i][:, lo:hi], non_blocking=True)  # 4 rows of (hi-lo)*8 bytes: one 2-D copy


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


step()
barrier()
t0 = time.perf_counter()
for _ in range(a.steps):
    step()
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / a.steps
t = torch.tensor([dt], device=dev)
if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
dt_max = float(t.item())
if rank == 0:
    b_in = T * n * 8 if a.one_direction in ("", "h2d") else 0
    b_out = T * n * 32 if a.one_direction in ("", "d2h") else 0
    print(json.dumps({"n_gpus": world, "tracks_per_gpu": T, "h2d_bytes_per_step": b_in, "d2h_bytes_per_step": b_out,
                      "ms_per_step_max_over_ranks": dt_max * 1e3, "gbs_per_rank": (b_in + b_out) / dt_max / 1e9,
                      "gbs_node": world * (b_in + b_out) / dt_max / 1e9,
                      "e2e_ceiling_audio_s_per_s": world * T * a.track_seconds / dt_max,
                      "affinity": (not a.no_affinity) and world > 1, "host_cpus": os.cpu_count(),
                      "direction": a.one_direction or "both", "d2h_copies": "torch strided copy_ per track" if a.d2h_2d else "contiguous, one per stem"}))
if world > 1:
    dist.destroy_process_group()
