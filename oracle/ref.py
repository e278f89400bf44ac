"""TEST INFRASTRUCTURE ONLY — ctypes view of oracle/_ref/libumxref.so (the compiled, unmodified
reference; see oracle/Makefile and oracle/ref_harness.cpp).

numpy conventions used here (converted to/from the reference's Eigen column-major memory):
    audio      (2, n) float32
    spec       (2, F, B) complex64          B = 2049
    mags       (4, 2, F, B) float32
"""
from __future__ import annotations

import ctypes as C
import os
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_ref", "libumxref.so")
NB_BINS = 2049

_lib = None


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        l = C.CDLL(LIB_PATH)
        vp, ip, fp = C.c_void_p, C.c_int, C.POINTER(C.c_float)
        l.ref_model_load.restype = vp
        l.ref_model_load.argtypes = [C.c_char_p]
        l.ref_model_free.argtypes = [vp]
        l.ref_model_hidden.argtypes = [vp]
        l.ref_model_tensor.restype = C.c_long
        l.ref_model_tensor.argtypes = [vp, C.c_char_p, ip, ip, ip, vp, C.POINTER(ip), C.POINTER(ip)]
        l.ref_stft.argtypes = [vp, ip, ip, vp]
        l.ref_istft.argtypes = [vp, ip, ip, vp]
        l.ref_lstm_create.restype = vp
        l.ref_lstm_create.argtypes = [ip, ip]
        l.ref_lstm_free.argtypes = [vp]
        l.ref_lstm_forward.argtypes = [vp, vp, ip, vp, ip, vp]
        l.ref_lstm_get_state.argtypes = [vp, ip, ip, ip, vp, vp]
        l.ref_wiener.argtypes = [vp, vp, ip, vp]
        l.ref_umx_inference.argtypes = [vp, vp, vp, ip, ip, vp]
        l.ref_main.argtypes = [ip, C.POINTER(C.c_char_p)]
        l.ref_nb_frames.argtypes = [ip]
        _lib = l
    return _lib


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def nb_frames(n_buf: int) -> int:
    return n_buf // 1024 + 1


def _spec_to_np(raw: np.ndarray, F: int) -> np.ndarray:
    # Eigen (2,F,B) col-major == numpy (B,F,2) C-order
    return np.ascontiguousarray(raw.reshape(NB_BINS, F, 2).transpose(2, 1, 0))


def _spec_from_np(spec: np.ndarray) -> np.ndarray:
    return np.ascontiguousarray(spec.transpose(2, 1, 0))


def stft(audio: np.ndarray, n_buf: int | None = None) -> np.ndarray:
    """src/dsp.cpp:141 — audio (2,n) -> spec (2,F,2049), buffers sized for n_buf (default n)."""
    n = audio.shape[1]
    n_buf = n if n_buf is None else n_buf
    F = nb_frames(n_buf)
    a = np.ascontiguousarray(audio.T.astype(np.float32))  # interleaved
    out = np.empty(2 * F * NB_BINS, dtype=np.complex64)
    lib().ref_stft(_p(a), n, n_buf, _p(out))
    return _spec_to_np(out, F)


def istft(spec: np.ndarray, n_buf: int, n_wave: int | None = None) -> np.ndarray:
    """src/dsp.cpp:178 — spec (2,F,2049) -> (2,n_wave)."""
    n_wave = n_buf if n_wave is None else n_wave
    s = _spec_from_np(spec.astype(np.complex64))
    out = np.empty((n_wave, 2), dtype=np.float32)
    lib().ref_istft(_p(s), n_buf, n_wave, _p(out))
    return np.ascontiguousarray(out.T)


def wiener(mix: np.ndarray, mags: np.ndarray):
    """src/wiener.cpp:92 — mix (2,F,B) c64, mags (4,2,F,B) f32 -> (y (4,2,F,B) c64, mix_after)."""
    F = mix.shape[1]
    m = _spec_from_np(mix.astype(np.complex64))
    g = np.ascontiguousarray(mags.astype(np.float32).transpose(0, 3, 2, 1))
    out = np.empty((4, NB_BINS, F, 2), dtype=np.complex64)
    lib().ref_wiener(_p(m), _p(g), F, _p(out))
    return np.ascontiguousarray(out.transpose(0, 3, 2, 1)), _spec_to_np(m.reshape(-1), F)


class Model:
    """src/model.cpp:42 load_umx_model on a .gz ggml file."""

    def __init__(self, path: str):
        cwd = os.getcwd()
        with tempfile.TemporaryDirectory() as td:  # the loader drops ./temp.decompressed
            os.chdir(td)
            try:
                self.h = lib().ref_model_load(os.path.abspath(os.path.join(cwd, path)).encode())
            finally:
                os.chdir(cwd)
        if not self.h:
            raise RuntimeError("reference load_umx_model failed: " + path)
        self.hidden = lib().ref_model_hidden(self.h)

    def tensor(self, name: str, target: int = 0, layer: int = 0, direction: int = 0) -> np.ndarray:
        """Dequantised tensor as numpy (rows, cols) (values identical to the Eigen matrix)."""
        r, c = C.c_int(), C.c_int()
        n = lib().ref_model_tensor(self.h, name.encode(), target, layer, direction, None, C.byref(r), C.byref(c))
        if n == 0:
            raise KeyError(name)
        buf = np.empty(n, dtype=np.float32)
        lib().ref_model_tensor(self.h, name.encode(), target, layer, direction, _p(buf), None, None)
        return np.ascontiguousarray(buf.reshape(c.value, r.value).T)

    def close(self):
        if self.h:
            lib().ref_model_free(self.h)
            self.h = None


class LstmState:
    def __init__(self, lstm_hidden: int, seq_len: int):
        self.h = lib().ref_lstm_create(lstm_hidden, seq_len)
        self.hid = lstm_hidden

    def get(self, target: int, layer: int, direction: int):
        h = np.empty(self.hid, np.float32)
        c = np.empty(self.hid, np.float32)
        lib().ref_lstm_get_state(self.h, target, layer, direction, _p(h), _p(c))
        return h, c

    def close(self):
        if self.h:
            lib().ref_lstm_free(self.h)
            self.h = None


def lstm_forward(model: Model, state: LstmState, target: int, x: np.ndarray) -> np.ndarray:
    """src/lstm.cpp:101 — x (F, hidden) -> (F, hidden); state carried in `state`."""
    F = x.shape[0]
    xin = np.asfortranarray(x.astype(np.float32))
    out = np.empty((F, model.hidden), dtype=np.float32, order="F")
    lib().ref_lstm_forward(model.h, state.h, target, _p(xin), F, _p(out))
    return np.ascontiguousarray(out)


def umx_inference(model: Model, state: LstmState, audio: np.ndarray, n_buf: int) -> np.ndarray:
    """src/inference.cpp:12 — audio (2,n) -> (4,2,n); buffers/state sized for n_buf."""
    n = audio.shape[1]
    a = np.ascontiguousarray(audio.T.astype(np.float32))
    out = np.empty((4, n, 2), dtype=np.float32)
    lib().ref_umx_inference(model.h, state.h, _p(a), n, n_buf, _p(out))
    return np.ascontiguousarray(out.transpose(0, 2, 1))


MAIN_PATH = os.path.join(HERE, "_ref", "umx.cpp.main.ref")


def main(model_gz: str, wav: str, out_dir: str) -> int:
    """The reference CLI (umx.cpp:26) as a separate process.  Only for tracks >= 30 s, and even then
    its istft writes past the end of the chunk's waveform matrix whenever a chunk is shorter than 60 s
    (src/dsp.cpp:203-205 under -DNDEBUG), so the process may die in free() AFTER the four WAVs are written:
    callers check the files, not the exit status."""
    import subprocess

    with tempfile.TemporaryDirectory() as td:  # the loader drops ./temp.decompressed
        r = subprocess.run([MAIN_PATH, os.path.abspath(model_gz), os.path.abspath(wav), os.path.abspath(out_dir)],
                           cwd=td, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return r.returncode
