"""Seeded inputs shared by tests/golden/make_golden.py (which ran the compiled reference on them) and the
tests that compare against the stored outputs."""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
from oracle import synth  # noqa: E402

HIDDEN = 32
N_BUF = 8192
N = 6000
F = N_BUF // 1024 + 1
WIENER_BIN_STEP = 8
GOLDEN = os.path.join(ROOT, "tests", "golden", "ref_small.npz")

_model = None


def model_bytes() -> bytes:
    global _model
    if _model is None:
        _model = synth.make_ggml(hidden=HIDDEN, seed=4321)
    return _model


def audio() -> np.ndarray:
    return synth.make_audio(N, seed=11)


def lstm_input() -> np.ndarray:
    return np.random.default_rng(0).uniform(-1, 1, (F, HIDDEN)).astype(np.float32)


def wiener_mags(spec: np.ndarray) -> np.ndarray:
    rng = np.random.default_rng(1)
    return (rng.uniform(0, 1, (4,) + spec.shape).astype(np.float32) * np.abs(spec)[None]).astype(np.float32)


def gtest_rand_waveform() -> np.ndarray:
    """test/test_dsp.cpp:49-55: audio(0,i) = audio(1,i) = rand()/RAND_MAX with the unseeded glibc rand()."""
    libc = ctypes.CDLL("libc.so.6")
    libc.srand(1)  # an unseeded process behaves as srand(1)
    v = np.array([libc.rand() for _ in range(4096)], dtype=np.float64) / 2147483647.0
    w = v.astype(np.float32)
    return np.stack([w, w])


def load_golden():
    return np.load(GOLDEN)


# ---- multi-segment whole-track case: the reference CLI (umx.cpp:26-295) on a 100 s track = 3 segments, the last one short ----
MS_HIDDEN = 32
MS_SECONDS = 100
MS_N = MS_SECONDS * 44100
MS_GOLDEN = os.path.join(ROOT, "tests", "golden", "ref_cli_multiseg.npz")


def multiseg_model_bytes() -> bytes:
    return synth.make_ggml(hidden=MS_HIDDEN, seed=4321)


def multiseg_audio() -> np.ndarray:
    return synth.make_audio(MS_N, seed=7)


def multiseg_indices(n: int = MS_N, shift: int = 4033) -> np.ndarray:
    """Output samples kept in the golden file: every 499th, plus 256-sample windows around every place where the segment
    recombination changes regime (start / peak / end of each triangular window, in trimmed output coordinates) and the track ends."""
    seg, stride = 2646000, 1984500
    marks = [0, n - 1]
    for k in range(0, (n + 22050 - shift + stride - 1) // stride):
        for p in (k * stride, k * stride + seg // 2, k * stride + seg):
            marks.append(p - shift)
    idx = set(range(0, n, 499))
    for m in marks:
        idx.update(i for i in range(m - 128, m + 128) if 0 <= i < n)
    return np.array(sorted(idx), dtype=np.int64)


def parse_cli_plan(stdout: str):
    """[(offset, chunk_length)] from the lines the reference prints at umx.cpp:220-222."""
    import re

    return [(int(a), int(b)) for a, b in re.findall(r"offset: (\d+), chunk shape: \(2, (\d+)\)", stdout)]


# ---- upstream-faithful variants: inputs for tests/golden/make_golden_upstream.py (which ran vendor/open-unmix-pytorch on them) ----
UP_F = 12          # frames
UP_WIN = 5         # frames per Wiener window (upstream default is 300; 5 gives three windows, the last one short)
UP_BIN_STEP = 16   # bins kept in the golden file
UP_GOLDEN = os.path.join(ROOT, "tests", "golden", "upstream_variants.npz")


def upstream_wiener_inputs():
    """(mix (2, F, 2049) c64, mags (4, 2, F, 2049) f32): a loud and a quiet stretch so the per-window max|X| differs."""
    rng = np.random.default_rng(21)
    mix = (rng.standard_normal((2, UP_F, 2049)) + 1j * rng.standard_normal((2, UP_F, 2049))).astype(np.complex64)
    mix[:, :UP_WIN] *= 40.0
    mix[:, UP_WIN:] *= 0.5
    mags = (rng.uniform(0, 1, (4, 2, UP_F, 2049)) * np.abs(mix)[None]).astype(np.float32)
    return mix, mags
