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
