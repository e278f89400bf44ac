"""One demix of `--tracks` synthetic tracks of `--seconds` (single 60 s segment by default) for ncu captures."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402
from oracle import synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--tracks", type=int, default=16)
ap.add_argument("--seconds", type=float, default=50.0)
ap.add_argument("--reps", type=int, default=2)
ap.add_argument("--terms", type=int, default=0)
a = ap.parse_args()
umx = g.load_package()
model = umx.load_umx_model(synth.make_ggml(hidden=1024, seed=1234))
sess = umx.Session(model, a.tracks, umx.SEGMENT_SAMPLES)
if a.terms:
    sess.set_option("gemm_terms", a.terms)
sess.set_profiling(True)
n = int(a.seconds * 44100)
base = synth.make_audio(n, 7)
tracks = [base * (1.0 - 0.01 * i) for i in range(a.tracks)]
for r in range(a.reps):
    out = sess.demix(tracks)
    print({k: round(sess.stage_ms(k), 2) for k in ("stft", "gemm", "lstm", "wiener", "istft", "finalize", "h2d", "d2h")},
          "launches", sess.launches(), flush=True)
