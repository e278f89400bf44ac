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


def reference_rate(model_path: str, window_s: float, repeats: int = 1):
    """The compiled, unmodified reference (oracle/_ref) on one bounded window through umx_inference.
    Returns (audio-seconds advanced per wall second, seconds per window, threads)."""
    from oracle import ref, synth

    rm = ref.Model(model_path)
    n_buf = int(window_s * SR)
    st = ref.LstmState(512, ref.nb_frames(n_buf))
    audio = synth.make_audio(n_buf, 7)
    ts = []
    for _ in range(repeats):
        t0 = time.time()
        ref.umx_inference(rm, st, audio, n_buf)
        ts.append(time.time() - t0)
    t = float(np.mean(ts))
    # a window advances 0.75 of its length (25 % overlap, umx.cpp:180)
    return 0.75 * window_s / t, t, os.cpu_count()


def run_reference(args, rank):
    """--impl reference: the reference's own CPU implementation of the path, rank 0 only."""
    if rank != 0:
        return
    window_s = 7.5  # bounded sample: the reference's cost is linear in frames (323 of 2584 per step)
    with tempfile.TemporaryDirectory() as td:
        mp = synth_model_file(td)
        from oracle import ref, synth

        rm = ref.Model(mp)
        n_buf = int(window_s * SR)
        st = ref.LstmState(512, ref.nb_frames(n_buf))
        audio = synth.make_audio(n_buf, 7)
        for _ in range(args.warmup):
            ref.umx_inference(rm, st, audio, n_buf)
        t0 = time.time()
        for _ in range(args.steps):
            ref.umx_inference(rm, st, audio, n_buf)
        dt = (time.time() - t0) / args.steps
    value = 0.75 * window_s / dt
    sample = (f"compiled unmodified reference (oracle/_ref, Eigen GEMM/GEMV, -O3 -march=x86-64-v3 -ffast-math): "
              f"umx_inference on one {window_s} s window ({ref.nb_frames(n_buf)} frames) per step, synthetic UMX-L "
              f"u8/u16 weights; a window advances 0.75 of its length")
    print(json.dumps({
        "impl": "reference", "metric": "audio-sec demixed/sec (UMX-L, 4 stems, 44.1kHz stereo)", "value": value,
        "unit": "audio-s/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "UMX-L 4-stem full pipeline, synthetic stereo 44.1 kHz (bounded CPU sample)"},
        "cpu_baseline": {"value": value, "unit": "audio-s/s", "cores": os.cpu_count(), "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--tracks", type=int, default=32, help="tracks per GPU (32 x 8 GPUs = configs[3])")
    ap.add_argument("--track-seconds", type=float, default=180.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--gemm-terms", type=int, default=0)
    ap.add_argument("--pipeline", type=int, default=-1, help="session option: post stage of segment s on a second stream under the network of s+1")
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
    if args.pipeline >= 0:
        sess.set_option("pipeline", args.pipeline)
    stream = torch.cuda.current_stream()
    sess.set_stream(stream.cuda_stream)
    sess.set_profiling(True)

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

    for _ in range(args.warmup):
        step_device()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stage_ms = {k: 0.0 for k in ALGO}
    launches = 0
    ev0.record(stream)
    for _ in range(args.steps):
        step_device()
        for k in stage_ms:
            stage_ms[k] += max(sess.stage_ms(k), 0.0)
        launches += sess.launches()
    ev1.record(stream)
    barrier()
    clocks = sampler.stop()
    ms = ev0.elapsed_time(ev1) / args.steps
    audio_s, ms_max = sharding.aggregate(T * args.track_seconds, ms, dist if world > 1 else None, dev)
    value = audio_s / (ms_max / 1e3)

    # ---- e2e: host buffers through umx_cuda_demix, copies inside the timed region ----
    host_out = [torch.empty((4, n, 2), dtype=torch.float32).pin_memory() for _ in range(T)]
    ins_np = [h.numpy() for h in host_in]
    outs_np = [h.numpy() for h in host_out]
    sess.demix_into(ins_np, outs_np)
    k_e2e = max(1, min(args.steps, 3))
    barrier()
    t0 = time.perf_counter()
    for _ in range(k_e2e):
        sess.demix_into(ins_np, outs_np)
    torch.cuda.synchronize()
    e2e_s = (time.perf_counter() - t0) / k_e2e
    _, e2e_ms_max = sharding.aggregate(T * args.track_seconds, e2e_s * 1e3, dist if world > 1 else None, dev)
    e2e_value = audio_s / (e2e_ms_max / 1e3)
    checksum = float(np.abs(outs_np[0][:, ::4096]).sum())  # the result was read back

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
        per_stage = {}
        for k, (bound, _) in ALGO.items():
            t_ms = stage_ms[k] / args.steps
            if t_ms <= 0:
                continue
            ach = work_per_track[k] * T / (t_ms / 1e3) / (1e9 if bound == "hbm" else 1e12)
            per_stage[k] = {"bound": bound, "ms_per_step": t_ms, "achieved": ach, "peak": pk[bound],
                            "unit": "GB/s" if bound == "hbm" else "TFLOP/s", "frac": ach / pk[bound]}
        if "gemm" in per_stage:
            # MMA work actually issued: every fp16 split pass counts (fc1/fc2/fc3: `terms` passes, W_ih: min(terms, 2) on the
            # exact integers); SURVEY 8(d) asks for both the algorithmic and the issued utilisation of the tensor pipe
            terms = args.gemm_terms if args.gemm_terms else 3
            issued = 4 * (15.74e9 * terms * pre + (65.03e9 / 3) * min(terms, 2) * pre + (65.03e9 * 2 / 3) * min(terms, 2) * segs
                          + (10.84e9 + 21.69e9) * terms * post)
            g = per_stage["gemm"]
            g["achieved_issued"] = issued * T / (g["ms_per_step"] / 1e3) / 1e12
            g["frac_issued"] = g["achieved_issued"] / pk["tensor"]
        if "lstm" in per_stage:
            # the recurrence is a latency problem: time per dependent step is the figure of merit; the
            # tensor-pipe fraction above is reported because the schema asks for one
            launches_l = 3 * segs
            per_stage["lstm"]["us_per_dependent_step"] = per_stage["lstm"]["ms_per_step"] * 1e3 / (launches_l * FRAMES)
            per_stage["lstm"]["note"] = "latency-bound: 8-CTA clusters on 64 SMs, one MMA batch + DSMEM exchange per step"
        # dominant kernel: gemm_tc_kernel (its two instantiations are 43 % of the GPU time in the ncu launch list of this command,
        # profiles/r01_launches_bench_steps1_summary.txt; the recurrence kernel is 34 %).  Under the default schedule the
        # recurrence's stage time is inflated by the post-stage kernels sharing the GPU with it, so it is not picked by time.
        top = "gemm" if "gemm" in per_stage else (max(per_stage, key=lambda k: per_stage[k]["ms_per_step"]) if per_stage else None)
        roofline = None
        if top:
            r = per_stage[top]
            roofline = {"kernel": top, "bound": r["bound"], "achieved": r["achieved"], "peak": r["peak"], "unit": r["unit"],
                        "frac": r["frac"], "achieved_issued": r.get("achieved_issued"), "frac_issued": r.get("frac_issued"),
                        "us_per_dependent_step": r.get("us_per_dependent_step"), "note": r.get("note"),
                        "traffic": TRAFFIC_32.get(top) if T == 32 else None,
                        "traffic_unit": "DRAM bytes per launch (ncu, 32-track segment batch; see profiles/)", "peak_source": pk["src"],
                        "share_of_step": r["ms_per_step"] / ms}
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
                       "schedule": ("in order on one stream" if args.pipeline == 0 else
                                    "Wiener/iSTFT/finalize of segment s run on a second stream under the recurrence kernels of "
                                    "segment s+1, so the per-stage times overlap and add up to more than the step")},
            "e2e": {"value": e2e_value, "unit": "audio-s/s", "h2d_bytes_per_step": T * n * 8, "d2h_bytes_per_step": T * n * 32,
                    "steps": k_e2e, "checksum": checksum},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "stages": per_stage,
            "realtime_factor_per_gpu": value / world,
        }
        if world == 1 and not args.no_cpu_baseline:
            with tempfile.TemporaryDirectory() as td:
                rate, t_win, cores = reference_rate(synth_model_file(td), 15.0)
            out["cpu_baseline"] = {"value": rate, "unit": "audio-s/s", "cores": 1, "kind": "reference",
                                   "host_cores": cores,
                                   "sample": f"compiled unmodified reference (oracle/_ref): umx_inference on one 15 s window "
                                             f"(646 frames) = {t_win:.1f} s; single-threaded in practice (SURVEY F6)"}
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
