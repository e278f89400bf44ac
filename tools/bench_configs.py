"""Side measurements for the other BASELINE.json configs (the bench line itself is configs[3]):
#1 10 s clip, #2 STFT -> iSTFT round trip of a 7-minute track, #3 one 7-minute track through the full pipeline,
and the CLI end to end (model load included).  Host-buffer API calls, wall clock around the call, best of 3."""
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402
from oracle import synth  # noqa: E402

umx = g.load_package()
raw = synth.make_ggml(hidden=1024, seed=1234)
t0 = time.perf_counter()
model = umx.load_umx_model(raw)
print(f"model load (138 MB ggml in memory -> device operands): {time.perf_counter() - t0:.2f} s")


def best(fn, n=3):
    fn()
    ts = []
    for _ in range(n):
        t = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t)
    return min(ts)


clip = synth.make_audio(441000, 7)
sess1 = umx.Session(model, 1, umx.SEGMENT_SAMPLES)
t = best(lambda: sess1.demix([clip]))
print(f"config 1: 10 s clip, 1 track: {t * 1e3:.1f} ms -> {10.0 / t:.0f} audio-s/s")

long = synth.make_audio(18522000, 7)
t = best(lambda: sess1.demix([long]))
print(f"config 3: 7 min track, 1 track, 10 segments: {t * 1e3:.1f} ms -> {420.0 / t:.0f} audio-s/s")

import ctypes as C  # noqa: E402

n7 = long.shape[1]
F7 = n7 // 1024 + 1
wave_il = np.ascontiguousarray(long.T.astype(np.float32))       # interleaved, the reference's MatrixXf(2, n) memory
spec_ref = np.empty(2 * F7 * 2049, dtype=np.complex64)            # Tensor3dXcf(2, F, 2049) memory
back_il = np.empty((n7, 2), dtype=np.float32)
L = umx.lib()


def roundtrip():
    assert L.umx_cuda_stft(wave_il.ctypes.data_as(C.c_void_p), n7, n7, spec_ref.ctypes.data_as(C.c_void_p)) == 0
    assert L.umx_cuda_istft(spec_ref.ctypes.data_as(C.c_void_p), n7, n7, back_il.ctypes.data_as(C.c_void_p)) == 0


t = best(roundtrip, 2)
err = float(np.abs(back_il - wave_il).max())
print(f"config 2: STFT -> iSTFT of a 7 min track through the C ABI (pageable host buffers both ways, Eigen memory order): "
      f"{t * 1e3:.1f} ms -> {420.0 / t:.0f} audio-s/s, round-trip max-abs error {err:.2e}")

cli = os.path.join(ROOT, "umx.cpp_b200", "umx.cpp.main")
if os.path.exists(cli):
    import gzip
    import wave
    with tempfile.TemporaryDirectory() as td:
        mp = os.path.join(td, "m.bin.gz")
        with gzip.open(mp, "wb", compresslevel=1) as f:
            f.write(raw)
        wp = os.path.join(td, "in.wav")
        a = synth.make_audio(7938000, 7)
        pcm = (np.clip(a.T if a.shape[0] == 2 else a, -1, 1) * 32767).astype("<i2")
        with wave.open(wp, "wb") as w:
            w.setnchannels(2); w.setsampwidth(2); w.setframerate(44100); w.writeframes(pcm.tobytes())
        t = time.perf_counter()
        r = subprocess.run([cli, mp, wp, os.path.join(td, "out")], capture_output=True)
        dt = time.perf_counter() - t
        print(f"CLI end to end, 3 min WAV, gz model load included (exit {r.returncode}): {dt:.2f} s -> {180.0 / dt:.0f} audio-s/s")
