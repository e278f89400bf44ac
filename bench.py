#!/usr/bin/env python
"""bench.py — audio-seconds demixed per second (UMX-L, 4 stems, 44.1 kHz stereo) on N B200s.

A "step" is one pass of the whole hot path (STFT -> 4 networks with streaming BiLSTM -> Wiener-EM ->
iSTFT -> segment recombine) over one batch of synthetic tracks.  Per GPU the batch is
`--tracks` (default 32) synthetic 3-minute stereo tracks, so 8 GPUs together process BASELINE.json
configs[3] (256 x 3 min, sharded, no collective); the per-GPU work is fixed => weak scaling.

  value      whole-job audio-seconds / second with the tracks already resident in HBM
             (umx_cuda_demix_device), timed with CUDA events on the launching stream, max over ranks
  e2e        the same through the host-buffer C-ABI call (umx_cuda_demix): pinned host audio in,
             4 stems out, host<->device copies inside the timed region
  roofline   dominant kernel of the step, algorithmic bytes/flops / its CUDA-event time
  cpu_baseline   the compiled, unmodified reference (oracle/_ref) on the host cores, bounded sample

`--impl reference` times only the reference's CPU implementation (rank 0).
"""
from __future__ import annotations

import argparse
import gzip
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# one hardware work queue per stream of the segment wavefront (must be set before the CUDA context is created)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

SR = 44100
SEG = 2646000
STRIDE = 1984500
FRAMES = 2584
# algorithmic work per (track, 60 s segment), SURVEY.md 8(d)
ALGO = {
    "stft": ("hbm", 148.2e6 + 30.7e6),   # wave in + spec out + mag out, + the cropped network input
    "istft": ("hbm", 423.5e6),           # 4 stem spectrograms in, 4 waveforms out
    "wiener": ("hbm", 847.2e6),          # two passes over mix + 4 masks, 4 stem spectrograms out
    "gemm": ("tensor", 453.2e9),         # fc1 + 6 W_ih + fc2 + fc3, 4 targets
    "lstm": ("tensor", 130.1e9),         # W_hh h recurrence, 4 targets x 3 layers x 2 directions
}


# dram__bytes_read.sum + dram__bytes_write.sum per launch of the stage's kernels, from ncu captures of one 32-track
# segment batch (profiles/r01_ncu_gemm_launch_metrics.csv: fc1, 3 x W_ih, fc2, fc3; profiles/r01_ncu_full_*.txt)
TRAFFIC_32 = {
    "gemm": (4.26e9 + 3 * 6.84e9 + 4.10e9 + 6.97e9) / 6,  # algorithmic: 2.33, 6.77, 4.06, 6.78 GB
    "istft": 16.4e9,                                       # algorithmic 13.6 GB (spectra in, accumulators read+written)
    "lstm": 6.77e9,                                        # input projection read (5.4 GB) + layer output planes written (1.3 GB)
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm": d["hbm_gbs"], "tensor": d["bf16_tflops_sustained"], "src": "measured (MEASURED_PEAKS.json)"}
    return {"hbm": 6650.0, "tensor": 1400.0, "src": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 8:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def synth_model_file(tmpdir: str) -> str:
    from oracle import synth

    p = os.path.join(tmpdir, "ggml-model-umxl-u8-synth.bin.gz")
    with open(p, "wb") as f:
        f.write(gzip.compress(synth.make_ggml(hidden=1024, seed=1234), 1))
    return p


REF_SAMPLE_S = 7.5      # bounded CPU sample: one 7.5 s window = 323 of the 2584 frames of a stock window
REF_FULL_FRAMES = 2584


def stock_rate(t_sample_s: float, sample_frames: int, track_seconds: float):
    """audio-s/s of the reference's STOCK path on one benched track, from the time of a bounded sample window.
    The stock CLI runs every segment as a full 60 s / 2584-frame window however short the chunk is (umx.cpp:161-163),
    and a track of `track_seconds` has ceil((n + 22050 - 4033) / 1,984,500) segments (umx.cpp:214); the reference's cost is linear
    in frames (every stage is per frame; checked against one real full window by --impl reference)."""
    n = int(track_seconds * SR)
    segs = (n + 22050 - 4033 + STRIDE - 1) // STRIDE
    t_window = t_sample_s * REF_FULL_FRAMES / sample_frames
    return track_seconds / (segs * t_window), t_window, segs


def reference_sample(model_path: str, warmup: int, steps: int, window_s: float = REF_SAMPLE_S):
    """The compiled, unmodified reference (oracle/_ref) through umx_inference on one bounded window per step.
    Returns (mean seconds per step, frames per window)."""
    from oracle import ref, synth

    rm = ref.Model(model_path)
    n_buf = int(window_s * SR)
    st = ref.LstmState(512, ref.nb_frames(n_buf))
    audio = synth.make_audio(n_buf, 7)
    for _ in range(warmup):
        ref.umx_inference(rm, st, audio, n_buf)
    t0 = time.time()
    for _ in range(steps):
        ref.umx_inference(rm, st, audio, n_buf)
    return (time.time() - t0) / steps, ref.nb_frames(n_buf)


def run_reference(args, rank):
    """--impl reference: the reference's own CPU implementation of the path, rank 0 only.  Each step is a bounded sample (one
    7.5 s window through umx_inference); `value` is what the STOCK path costs for the benched track (full 60 s windows), and one
    real full window is timed after the K steps to check the extrapolation."""
    if rank != 0:
        return
    with tempfile.TemporaryDirectory() as td:
        mp = synth_model_file(td)
        dt, frames = reference_sample(mp, args.warmup, args.steps)
        value, t_window, segs = stock_rate(dt, frames, args.track_seconds)
        full = None
        if not args.no_full_window:
            from oracle import ref, synth

            rm = ref.Model(mp)
            st = ref.LstmState(512, ref.nb_frames(SEG))
            audio = synth.make_audio(SEG, 7)
            t0 = time.time()
            ref.umx_inference(rm, st, audio, SEG)
            t_full = time.time() - t0
            full = {"seconds_per_full_window": t_full, "value_from_full_window": args.track_seconds / (segs * t_full),
                    "extrapolated_seconds_per_full_window": t_window}
    cores = os.cpu_count()
    sample = (f"compiled unmodified reference (oracle/_ref, Eigen GEMM/GEMV, -O3 -march=x86-64-v3 -ffast-math), synthetic UMX-L "
              f"u8/u16 weights: one {REF_SAMPLE_S} s window ({frames} frames) through umx_inference per step = {dt:.2f} s; value = "
              f"{args.track_seconds:.0f} s track / ({segs} stock windows of {REF_FULL_FRAMES} frames x {t_window:.1f} s); "
              f"{cores} host threads available to Eigen, single-threaded in practice (SURVEY F6)")
    print(json.dumps({
        "impl": "reference", "metric": "audio-sec demixed/sec (UMX-L, 4 stems, 44.1kHz stereo)", "value": value,
        "unit": "audio-s/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"UMX-L 4-stem full pipeline, synthetic {args.track_seconds:.0f} s stereo 44.1 kHz tracks, stock path "
                               f"({segs} full 60 s windows per track), timed on a bounded sample per step"},
        "cpu_baseline": {"value": value, "unit": "audio-s/s", "cores": cores, "kind": "reference", "sample": sample},
        "stock_window_check": full,
        "e2e": {"value": value, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    # a run that stops making progress prints where it stands and exits instead of hanging its caller
    import faulthandler

    faulthandler.dump_traceback_later(int(os.environ.get("UMX_BENCH_WATCHDOG_S", "1500")), exit=True)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--tracks", type=int, default=32, help="tracks per GPU (32 x 8 GPUs = configs[3])")
    ap.add_argument("--track-seconds", type=float, default=180.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--gemm-terms", type=int, default=0)
    ap.add_argument("--gemm-impl", type=int, default=-1, help="session option: 0 = CTA pairs in order, single-CTA tiles in the wavefront (default), 2 = pairs always, 3 = single-CTA always")
    ap.add_argument("--pipeline", type=int, default=-1, help="session option: 0 = every kernel in order on one stream, 3 (default) = segment wavefront")
    ap.add_argument("--lane-targets", type=int, default=0, help="session option: targets per stream of the wavefront (1, 2, 4)")
    ap.add_argument("--timeline", default="", help="write the stage intervals of the last timed step to this file (JSON)")
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs only: skip the host-buffer arm (the line then carries no e2e value)")
    ap.add_argument("--no-in-order", action="store_true", help="skip the in-order pass that yields roofline_in_order")
    ap.add_argument("--no-full-window", action="store_true", help="--impl reference: skip the one real 60 s window after the timed steps")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist

    import __graft_entry__ as g
    from oracle import synth  # synthetic inputs only (test infrastructure is not on the timed path)

    umx = g.load_package()
    from importlib import import_module

    sharding = import_module("umx_cpp_b200.sharding")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)

    n = int(args.track_seconds * SR)
    T = args.tracks
    model = umx.load_umx_model(synth.make_ggml(hidden=1024, seed=1234), device=local_rank)
    sess = umx.Session(model, max_tracks=T, n_buf=SEG)
    if args.gemm_terms:
        sess.set_option("gemm_terms", args.gemm_terms)
    if args.gemm_impl >= 0:
        sess.set_option("gemm_impl", args.gemm_impl)
    if args.pipeline >= 0:
        sess.set_option("pipeline", args.pipeline)
    if args.lane_targets:
        sess.set_option("lane_targets", args.lane_targets)
    pipeline = args.pipeline if args.pipeline >= 0 else 3
    stream = torch.cuda.current_stream()
    sess.set_stream(stream.cuda_stream)
    # Stage timers: CUDA event pairs the library records on the launching streams.  Recording costs no synchronisation (the
    # events are pooled and only read after the timed region), so they stay on inside the timed region.
    sess.set_profiling(True)

    # pin this rank's threads to its share of the host cores: the pinned buffers below are then first touched (and the copy
    # engines fed) from there
    if world > 1 and hasattr(os, "sched_setaffinity"):
        try:
            cpus = sorted(os.sched_getaffinity(0))
            per = max(1, len(cpus) // world)
            os.sched_setaffinity(0, set(cpus[local_rank * per:(local_rank + 1) * per]) or set(cpus))
        except OSError:
            pass

    host_in = [torch.empty((n, 2), dtype=torch.float32).pin_memory() for _ in range(T)]
    my_tracks = sharding.shard_tracks(T * world, rank, world)  # weak scaling: every rank owns T tracks of the job
    for i, trk in enumerate(my_tracks):
        host_in[i].numpy()[:] = synth.make_audio(n, 7 + trk).T
    d_in = [h.to(dev) for h in host_in]
    d_out = [torch.empty((4, n, 2), dtype=torch.float32, device=dev) for _ in range(T)]
    ns = [n] * T
    segs = len(umx.segment_plan(n))

    def step_device():
        sess.demix_device([t.data_ptr() for t in d_in], ns, [t.data_ptr() for t in d_out])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def read_stages():
        return {k: max(sess.stage_ms(k), 0.0) for k in list(ALGO) + ["finalize", "h2d", "d2h"]}

    for _ in range(args.warmup):
        step_device()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches = 0
    ev0.record(stream)
    for _ in range(args.steps):
        step_device()
        launches += sess.launches()
    ev1.record(stream)
    barrier()
    clocks = sampler.stop()
    ms = ev0.elapsed_time(ev1) / args.steps
    stage_ms = read_stages()  # the LAST timed step (the library keeps the most recent call's timers)
    if args.timeline and rank == 0:
        json.dump(sess.timeline(), open(args.timeline, "w"))
    audio_s, ms_max = sharding.aggregate(T * args.track_seconds, ms, dist if world > 1 else None, dev)
    value = audio_s / (ms_max / 1e3)

    # ---- the same step once more with every kernel in order on one stream (outside the timed region): per-stage times that
    # do not overlap, i.e. the numbers an ncu launch list of one step reproduces
    stage_ms_in_order, ms_in_order = None, None
    if not args.no_in_order:
        sess.set_option("pipeline", 0)
        step_device()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        step_device()
        e1.record(stream)
        torch.cuda.synchronize()
        ms_in_order = e0.elapsed_time(e1)
        stage_ms_in_order = read_stages()
        sess.set_option("pipeline", pipeline)

    # ---- e2e: host buffers through umx_cuda_demix, copies inside the timed region ----
    e2e_value, e2e_ms_max, k_e2e, checksum, copy_ms = None, None, 0, None, {}
    e2e_sync_value, e2e_sync_ms = None, None
    if not args.no_e2e:
        # two sets of output buffers: consecutive calls in flight must not share them
        host_out = [[torch.empty((4, n, 2), dtype=torch.float32).pin_memory() for _ in range(T)] for _ in range(2)]
        # (2 x 8.1 GB per rank at the default size; umx_cuda_demix_async needs both sets for as long as two calls are in flight)
        ins_np = [h.numpy() for h in host_in]
        outs_np = [[h.numpy() for h in hs] for hs in host_out]
        sess.demix_into(ins_np, outs_np[0])
        # (a) one call at a time: umx_cuda_demix returns when the stems are in the host buffers
        k_sync = max(1, min(args.steps, 3))
        barrier()
        t0 = time.perf_counter()
        for _ in range(k_sync):
            sess.demix_into(ins_np, outs_np[0])
        torch.cuda.synchronize()
        sync_s = (time.perf_counter() - t0) / k_sync
        copy_ms = read_stages()  # busy time of the two copy streams in the last of these steps
        if args.timeline and rank == 0:
            json.dump(sess.timeline(), open(args.timeline + ".e2e.json", "w"))
        _, e2e_sync_ms = sharding.aggregate(T * args.track_seconds, sync_s * 1e3, dist if world > 1 else None, dev)
        e2e_sync_value = audio_s / (e2e_sync_ms / 1e3)
        # (b) a queue of batches, as a serving host feeds them: umx_cuda_demix_async with two calls in flight.  Every step's inputs
        # go up from pinned host memory and its stems come down inside the timed region; a step's result is read on the host as
        # soon as its ticket is done, while the next step is already running.
        k_e2e = max(4, min(args.steps, 8))  # (the first call of the queue has nothing to overlap with)
        sess.wait(sess.demix_async(ins_np, outs_np[1]))  # warm the second set of tables and buffers
        checksum = 0.0
        barrier()
        t0 = time.perf_counter()
        tickets = []
        for i in range(k_e2e):
            tickets.append(sess.demix_async(ins_np, outs_np[i & 1]))
            if i >= 1:
                sess.wait(tickets[i - 1])
                checksum += float(np.abs(outs_np[(i - 1) & 1][0][:, ::4096]).sum())
        sess.wait(tickets[-1])
        checksum += float(np.abs(outs_np[(k_e2e - 1) & 1][0][:, ::4096]).sum())  # every step's result was read back
        torch.cuda.synchronize()
        e2e_s = (time.perf_counter() - t0) / k_e2e
        checksum /= k_e2e
        _, e2e_ms_max = sharding.aggregate(T * args.track_seconds, e2e_s * 1e3, dist if world > 1 else None, dev)
        e2e_value = audio_s / (e2e_ms_max / 1e3)

    if rank == 0:
        pk = peaks()
        # Work actually performed per track and step.  The reference runs every frame of every zero-extended 60 s window; the
        # frames beyond a chunk are computed once (STFT, fc1, first-layer W_ih: fz + 1 of F frames) or not at all (fc2, fc3,
        # Wiener, iSTFT: fz of F) here, so a short last segment counts for what is really done.
        plan = umx.segment_plan(n)
        fz = [min(FRAMES, (ln + 2048 + 1023) // 1024 if ln > 0 else 0) for _, ln in plan]
        pre = sum(min(FRAMES, z + 1) / FRAMES for z in fz)   # segment-equivalents of the stages before the recurrences
        post = sum(z / FRAMES for z in fz)                   # ... and after the network
        gemm_1 = 4 * ((15.74e9 + 65.03e9 / 3) * pre + (65.03e9 * 2 / 3) * segs + (10.84e9 + 21.69e9) * post)
        work_per_track = {"stft": ALGO["stft"][1] * pre, "istft": ALGO["istft"][1] * post, "wiener": ALGO["wiener"][1] * post,
                          "gemm": gemm_1, "lstm": ALGO["lstm"][1] * segs}
        terms = args.gemm_terms if args.gemm_terms else 3
        # MMA work actually issued: every fp16 split pass counts (fc1/fc2/fc3: `terms` passes, W_ih: min(terms, 2) on the exact
        # integers); SURVEY 8(d) asks for both the algorithmic and the issued utilisation of the tensor pipe
        issued = 4 * (15.74e9 * terms * pre + (65.03e9 / 3) * min(terms, 2) * pre + (65.03e9 * 2 / 3) * min(terms, 2) * segs
                      + (10.84e9 + 21.69e9) * terms * post)

        def stage_table(ms_by_stage):
            tab = {}
            for k, (bound, _) in ALGO.items():
                t_ms = ms_by_stage.get(k, 0.0)
                if t_ms <= 0:
                    continue
                ach = work_per_track[k] * T / (t_ms / 1e3) / (1e9 if bound == "hbm" else 1e12)
                tab[k] = {"bound": bound, "ms_per_step": t_ms, "achieved": ach, "peak": pk[bound],
                          "unit": "GB/s" if bound == "hbm" else "TFLOP/s", "frac": ach / pk[bound]}
            if "gemm" in tab:
                g = tab["gemm"]
                g["achieved_issued"] = issued * T / (g["ms_per_step"] / 1e3) / 1e12
                g["frac_issued"] = g["achieved_issued"] / pk["tensor"]
            if "lstm" in tab:
                # the recurrence is a latency problem: time per dependent step is the figure of merit; the
                # tensor-pipe fraction is reported because the schema asks for one
                tab["lstm"]["us_per_dependent_step"] = tab["lstm"]["ms_per_step"] * 1e3 / (3 * segs * FRAMES)
                tab["lstm"]["note"] = "latency-bound: 8-CTA clusters, one MMA batch + DSMEM exchange per step"
            return tab

        # Stage times of the LAST TIMED STEP: for every stage the time during which at least one of its kernels was running
        # (union of the CUDA-event intervals over all streams).  In the default schedule the stages share the GPU, so these
        # overlap and each stage's rate is what it got while sharing; `roofline_in_order` below has the same kernels alone.
        per_stage = stage_table(stage_ms)

        def roofline_of(tab, step_ms):
            if "gemm" not in tab:  # dominant kernel: gemm_tc_kernel (profiles/: largest share of the GPU time in the launch list)
                return None
            r = tab["gemm"]
            return {"kernel": "gemm_tc_kernel", "bound": r["bound"], "achieved": r["achieved"], "peak": r["peak"], "unit": r["unit"],
                    "frac": r["frac"], "achieved_issued": r.get("achieved_issued"), "frac_issued": r.get("frac_issued"),
                    "traffic": TRAFFIC_32.get("gemm") if T == 32 else None,
                    "traffic_unit": "DRAM bytes per launch (ncu, 32-track segment batch; see profiles/)", "peak_source": pk["src"],
                    "share_of_step": r["ms_per_step"] / step_ms}

        roofline = roofline_of(per_stage, ms)
        in_order = None
        if stage_ms_in_order:
            tab = stage_table(stage_ms_in_order)
            in_order = roofline_of(tab, ms_in_order)
            if in_order:
                in_order["ms_per_step_in_order"] = ms_in_order
                in_order["stages"] = tab
                in_order["note"] = ("one extra step with option pipeline = 0 (every kernel in order on one stream) after the timed "
                                    "region: stage times do not overlap and match an ncu launch list of the step")
        schedules = {0: "every kernel in order on one stream",
                     3: "segment wavefront: LSTM layer l of segment s+1 beside layer l+1 of segment s, one stream per (layer, target "
                        "group), post stage on a low-priority stream; stage times overlap"}
        GB = 1e9
        out = {
            "metric": "audio-sec demixed/sec (UMX-L, 4 stems, 44.1kHz stereo)", "value": value, "unit": "audio-s/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32 (fp16 hi/lo split tensor-core operands, fp32 accumulate)",
            "data": "synthetic",
            "config": {"workload": f"UMX-L 4-stem full pipeline, batch of {T} synthetic {args.track_seconds:.0f} s stereo "
                                   f"tracks per GPU sharded with no collective (configs[3]: 256 x 3 min at 8 GPUs), "
                                   f"60 s segments, streaming LSTM, Wiener-EM",
                       "tracks_per_gpu": T, "segments_per_track": segs, "weights": "synthetic UMX-L u8/u16 ggml (seed 1234)",
                       "l2": "inputs and intermediates per step are far larger than L2 (>1 GB per track-segment)",
                       "schedule": schedules[0 if pipeline == 0 else 3], "gemm_terms": terms},
            "e2e": {"value": e2e_value, "unit": "audio-s/s", "h2d_bytes_per_step": T * n * 8, "d2h_bytes_per_step": T * n * 32,
                    "steps": k_e2e, "ms_per_step": e2e_ms_max, "h2d_ms": copy_ms.get("h2d"), "d2h_ms": copy_ms.get("d2h"),
                    "h2d_gbs": (T * n * 8 / GB) / (copy_ms["h2d"] / 1e3) if copy_ms.get("h2d") else None,
                    "d2h_gbs": (T * n * 32 / GB) / (copy_ms["d2h"] / 1e3) if copy_ms.get("d2h") else None,
                    "mode": "umx_cuda_demix_async on pinned host buffers, two calls in flight (a queue of batches): every step's "
                            "inputs go up and its stems come down inside the timed region, each step's result is read on the host",
                    "one_call_at_a_time": {"value": e2e_sync_value, "ms_per_step": e2e_sync_ms,
                                           "note": "umx_cuda_demix, returns when the stems are in the host buffers"},
                    "note": "h2d_ms / d2h_ms: busy time of this rank's two copy streams in the last one-call-at-a-time step",
                    "checksum": checksum},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "roofline_in_order": in_order, "stages": per_stage,
            "realtime_factor_per_gpu": value / world,
        }
        if world == 1 and not args.no_cpu_baseline:
            with tempfile.TemporaryDirectory() as td:
                dt, frames = reference_sample(synth_model_file(td), 0, 3)
            rate, t_window, rsegs = stock_rate(dt, frames, args.track_seconds)
            out["cpu_baseline"] = {"value": rate, "unit": "audio-s/s", "cores": os.cpu_count(), "kind": "reference",
                                   "sample": f"compiled unmodified reference (oracle/_ref): umx_inference on a {REF_SAMPLE_S} s window "
                                             f"({frames} frames) = {dt:.2f} s, 3 repeats; value = {args.track_seconds:.0f} s track / "
                                             f"({rsegs} stock 60 s windows x {t_window:.1f} s, cost linear in frames); "
                                             f"{os.cpu_count()} host threads available, single-threaded in practice (SURVEY F6)"}
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
