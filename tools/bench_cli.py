"""Wall-clock of the two command-line front ends on synthetic WAVs (model load from the gzipped ggml file included):
  umx.cpp.main          cli/umx_main.cpp over the C ABI (umx_cuda_demix: whole track, segment wavefront)
  umx.cpp.main.dropin   the reference's UNMODIFIED umx.cpp linked against shim/umx_dropin.cpp (umxcpp::umx_inference per
                        segment -> umx_cuda_infer_segment; its own shift / split / recombine loop on the host)
    python tools/bench_cli.py [seconds ...]     (default 180 420)"""
import gzip
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import synth  # noqa: E402

CLI = os.path.join(ROOT, "umx.cpp_b200", "umx.cpp.main")
DROPIN = os.path.join(ROOT, "oracle", "_ref", "umx.cpp.main.dropin")
secs = [float(a) for a in sys.argv[1:]] or [180.0, 420.0]
with tempfile.TemporaryDirectory() as td:
    mp = os.path.join(td, "m.bin.gz")
    with open(mp, "wb") as f:
        f.write(gzip.compress(synth.make_ggml(hidden=1024, seed=1234), 1))
    for s in secs:
        wp = os.path.join(td, f"in_{int(s)}.wav")
        synth.write_wav_f32(wp, synth.make_audio(int(s * 44100), 7))
        for name, exe in (("umx.cpp.main", CLI), ("umx.cpp.main.dropin", DROPIN)):
            if not os.path.exists(exe):
                continue
            best, line = 1e9, ""
            for rep in range(2):
                t = time.perf_counter()
                r = subprocess.run([exe, mp, wp, os.path.join(td, "out")], capture_output=True, text=True, cwd=td)
                dt = time.perf_counter() - t
                if r.returncode != 0:
                    print(name, "failed:", r.stderr[-400:])
                    break
                if dt < best:
                    best = dt
                    line = next((l for l in r.stdout.splitlines() if "demixed" in l or "Loaded model" in l), "")
            print(f"{name:20s} {s:5.0f} s WAV: {best:6.2f} s wall -> {s / best:6.0f} audio-s/s end to end   [{line.strip()}]", flush=True)
