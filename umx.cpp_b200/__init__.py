"""Host-side mirror (Python) of the reference's C++ interface for the UMX-L demixing hot path, on top
of the C ABI in include/umx_cuda.h (libumx_cuda.so, hand-written sm_100a CUDA).

Names follow the reference: load_umx_model (src/model.hpp:58), umx_inference (src/inference.hpp:20),
stft / istft (src/dsp.hpp:114-115), wiener_filter (src/wiener.hpp:19), umx_lstm_forward
(src/lstm.hpp:22), and demix = shift_inference + split_inference (umx.cpp:99-295).

numpy conventions (converted to the reference's Eigen column-major order at the C boundary):
    audio (2, n) float32 ; spec (2, F, 2049) complex64 ; mags (4, 2, F, 2049) float32 ; stems (4, 2, n)

There is NO CPU fallback and nothing here imports oracle/: if the CUDA library is missing or no GPU is
present, every call raises.
"""
from __future__ import annotations

import ctypes as C
import os

# The segment wavefront runs on up to 18 CUDA streams; with the driver's default of 8 hardware work queues several streams share
# a queue and their kernels serialise behind each other's event waits.  Each queue costs ~50 ms of context creation, so 16 here
# (as the C library does when it is loaded, csrc/capi.cu); bench.py asks for 32.  Only takes effect before the CUDA context exists.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "16")

import numpy as np  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libumx_cuda.so")

NB_BINS = 2049
SEGMENT_SAMPLES = 2646000
STRIDE_SAMPLES = 1984500
MAX_SHIFT = 22050
DEFAULT_SHIFT_OFFSET = 4033
SAMPLE_RATE = 44100

EXPORTS = [
    "umx_cuda_last_error", "umx_cuda_device_count", "umx_cuda_nb_frames", "umx_cuda_model_create",
    "umx_cuda_model_create_from_file", "umx_cuda_model_destroy", "umx_cuda_model_hidden_size",
    "umx_cuda_model_device", "umx_cuda_model_tensor", "umx_cuda_session_create", "umx_cuda_session_destroy",
    "umx_cuda_session_reset", "umx_cuda_session_get_state", "umx_cuda_infer_segment", "umx_cuda_demix",
    "umx_cuda_demix_async", "umx_cuda_session_wait",
    "umx_cuda_demix_device", "umx_cuda_host_alloc", "umx_cuda_host_free", "umx_cuda_segment_plan", "umx_cuda_stft", "umx_cuda_istft", "umx_cuda_lstm_forward",
    "umx_cuda_wiener", "umx_cuda_wiener_opts", "umx_cuda_stft_opts", "umx_cuda_target_mags", "umx_cuda_session_set_profiling", "umx_cuda_session_stage_ms",
    "umx_cuda_session_launches", "umx_cuda_session_timeline", "umx_cuda_session_set_option", "umx_cuda_session_set_stream",
    "umx_cuda_wiener_noalloc", "umx_cuda_lstm_max_active_clusters", "umx_cuda_nccl_unique_id",
    "umx_cuda_session_init_target_parallel",
]

_lib = None


class UmxCudaError(RuntimeError):
    pass


def lib():
    """The C-ABI library; raises if it has not been built (python __graft_entry__.py build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise UmxCudaError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
                           "There is no CPU fallback.")
    l = C.CDLL(LIB_PATH)
    vp, ip, fpp = C.c_void_p, C.c_int, C.POINTER(C.c_void_p)
    l.umx_cuda_last_error.restype = C.c_char_p
    l.umx_cuda_model_create.argtypes = [vp, C.c_size_t, ip, fpp]
    l.umx_cuda_model_create_from_file.argtypes = [C.c_char_p, ip, fpp]
    l.umx_cuda_model_destroy.argtypes = [vp]
    l.umx_cuda_model_destroy.restype = None
    l.umx_cuda_model_hidden_size.argtypes = [vp]
    l.umx_cuda_model_device.argtypes = [vp]
    l.umx_cuda_model_tensor.restype = C.c_long
    l.umx_cuda_model_tensor.argtypes = [vp, C.c_char_p, ip, ip, ip, vp, C.c_long, C.POINTER(ip), C.POINTER(ip)]
    l.umx_cuda_session_create.argtypes = [vp, ip, ip, fpp]
    l.umx_cuda_session_destroy.argtypes = [vp]
    l.umx_cuda_session_destroy.restype = None
    l.umx_cuda_session_reset.argtypes = [vp]
    l.umx_cuda_session_get_state.argtypes = [vp, ip, ip, ip, ip, vp, vp]
    l.umx_cuda_infer_segment.argtypes = [vp, vp, ip, vp]
    l.umx_cuda_demix.argtypes = [vp, vp, vp, ip, ip, vp]
    l.umx_cuda_demix_device.argtypes = [vp, vp, vp, ip, ip, vp]
    l.umx_cuda_demix_async.argtypes = [vp, vp, vp, ip, ip, vp, vp]
    l.umx_cuda_session_wait.argtypes = [vp, C.c_int64]
    l.umx_cuda_host_alloc.argtypes = [C.c_size_t, fpp]
    l.umx_cuda_host_free.argtypes = [vp]
    l.umx_cuda_host_free.restype = None
    l.umx_cuda_segment_plan.argtypes = [C.c_int64, ip, ip, vp, vp, ip]
    l.umx_cuda_stft.argtypes = [vp, ip, ip, vp]
    l.umx_cuda_istft.argtypes = [vp, ip, ip, vp]
    l.umx_cuda_lstm_forward.argtypes = [vp, ip, vp, ip, vp]
    l.umx_cuda_wiener.argtypes = [vp, vp, ip, vp]
    l.umx_cuda_wiener_opts.argtypes = [vp, vp, ip, vp, ip, ip, ip, ip]
    l.umx_cuda_stft_opts.argtypes = [vp, ip, ip, vp, ip]
    l.umx_cuda_target_mags.argtypes = [vp, vp, vp]
    l.umx_cuda_session_set_profiling.argtypes = [vp, ip]
    l.umx_cuda_session_stage_ms.restype = C.c_double
    l.umx_cuda_session_stage_ms.argtypes = [vp, C.c_char_p]
    l.umx_cuda_session_launches.restype = C.c_long
    l.umx_cuda_session_launches.argtypes = [vp]
    l.umx_cuda_session_timeline.argtypes = [vp, vp, vp, vp, ip]
    l.umx_cuda_session_set_option.argtypes = [vp, C.c_char_p, ip]
    l.umx_cuda_session_set_stream.argtypes = [vp, vp]
    l.umx_cuda_wiener_noalloc.argtypes = [vp, vp, ip, ip, vp, vp]
    l.umx_cuda_nccl_unique_id.argtypes = [vp]
    l.umx_cuda_session_init_target_parallel.argtypes = [vp, vp, ip, ip]
    _lib = l
    return l


def _check(rc: int, what: str):
    if rc < 0:
        raise UmxCudaError(f"{what} failed ({rc}): {lib().umx_cuda_last_error().decode(errors='replace')}")
    return rc


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def nb_frames(n_buf: int) -> int:
    return n_buf // 1024 + 1


def _spec_to_ref(spec: np.ndarray) -> np.ndarray:  # (2,F,B) -> Eigen col-major == C-order (B,F,2)
    return np.ascontiguousarray(spec.transpose(2, 1, 0))


def _spec_from_ref(raw: np.ndarray, F: int) -> np.ndarray:
    return np.ascontiguousarray(raw.reshape(NB_BINS, F, 2).transpose(2, 1, 0))


# ---------------------------------------------------------------------------------------------------
class UmxModel:
    """struct umx_model (src/model.hpp:12-56), weights resident on one GPU."""

    def __init__(self, handle, lib_):
        self._h, self._lib = handle, lib_
        self.hidden_size = lib_.umx_cuda_model_hidden_size(handle)
        self.is_initialized = True
        self.load_progress = 1.0
        self.inference_progress = 0.0

    def tensor(self, name: str, target: int = 0, layer: int = 0, direction: int = 0) -> np.ndarray:
        """Dequantised member `name` (reference member names, e.g. 'fc1_w'), numpy (rows, cols)."""
        r, c = C.c_int(), C.c_int()
        n = _check(self._lib.umx_cuda_model_tensor(self._h, name.encode(), target, layer, direction, None, 0,
                                                   C.byref(r), C.byref(c)), "umx_cuda_model_tensor")
        buf = np.empty(n, dtype=np.float32)
        _check(self._lib.umx_cuda_model_tensor(self._h, name.encode(), target, layer, direction, _p(buf), n, None, None),
               "umx_cuda_model_tensor")
        return np.ascontiguousarray(buf.reshape(c.value, r.value).T)

    def close(self):
        if self._h:
            self._lib.umx_cuda_model_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def load_umx_model(model_file, device: int = 0) -> UmxModel:
    """src/model.hpp:58 — `model_file` is a path to a (gzipped) ggml file, or its bytes."""
    l = lib()
    h = C.c_void_p()
    if isinstance(model_file, (bytes, bytearray)):
        buf = (C.c_char * len(model_file)).from_buffer_copy(model_file)
        _check(l.umx_cuda_model_create(buf, len(model_file), device, C.byref(h)), "load_umx_model")
    else:
        _check(l.umx_cuda_model_create_from_file(os.fspath(model_file).encode(), device, C.byref(h)), "load_umx_model")
    return UmxModel(h, l)


class Session:
    """Streaming LSTM state (lstm_data, src/lstm.hpp:10-16) + work buffers (stft_buffers) for up to
    `max_tracks` tracks demixed in lockstep."""

    def __init__(self, model: UmxModel, max_tracks: int = 1, n_buf: int = SEGMENT_SAMPLES):
        self._lib = lib()
        self.model, self.n_buf, self.max_tracks = model, n_buf, max_tracks
        self.F = nb_frames(n_buf)
        h = C.c_void_p()
        _check(self._lib.umx_cuda_session_create(model._h, max_tracks, n_buf, C.byref(h)), "umx_cuda_session_create")
        self._h = h

    def reset(self):
        _check(self._lib.umx_cuda_session_reset(self._h), "umx_cuda_session_reset")

    def set_option(self, name: str, value: int):
        _check(self._lib.umx_cuda_session_set_option(self._h, name.encode(), value), "umx_cuda_session_set_option")

    def set_stream(self, cuda_stream: int | None):
        """Run on the caller's CUDA stream (e.g. torch.cuda.current_stream().cuda_stream); None = own stream."""
        _check(self._lib.umx_cuda_session_set_stream(self._h, cuda_stream), "umx_cuda_session_set_stream")

    def init_target_parallel(self, unique_id: bytes, rank: int, world: int):
        """Target-parallel mode: this rank runs targets [rank*4/world, (rank+1)*4/world); masks are all-gathered
        over NCCL before the Wiener stage.  `unique_id`: the 128 bytes of nccl_unique_id() made on rank 0."""
        buf = (C.c_char * 128).from_buffer_copy(unique_id) if unique_id else None
        _check(self._lib.umx_cuda_session_init_target_parallel(self._h, buf, rank, world),
               "umx_cuda_session_init_target_parallel")

    def set_profiling(self, on: bool):
        _check(self._lib.umx_cuda_session_set_profiling(self._h, int(on)), "umx_cuda_session_set_profiling")

    def stage_ms(self, stage: str) -> float:
        return float(self._lib.umx_cuda_session_stage_ms(self._h, stage.encode()))

    def launches(self) -> int:
        return int(self._lib.umx_cuda_session_launches(self._h))

    def timeline(self):
        """[(stage, begin_ms, end_ms)] of the most recent profiled call."""
        n = _check(self._lib.umx_cuda_session_timeline(self._h, None, None, None, 0), "umx_cuda_session_timeline")
        names, b, e = np.zeros((n, 16), np.uint8), np.zeros(n, np.float32), np.zeros(n, np.float32)
        _check(self._lib.umx_cuda_session_timeline(self._h, _p(names), _p(b), _p(e), n), "umx_cuda_session_timeline")
        return [(bytes(names[i]).split(b"\0")[0].decode(), float(b[i]), float(e[i])) for i in range(n)]

    def state(self, track: int, target: int, layer: int, direction: int):
        hh = self.model.hidden_size // 2
        h, c = np.empty(hh, np.float32), np.empty(hh, np.float32)
        _check(self._lib.umx_cuda_session_get_state(self._h, track, target, layer, direction, _p(h), _p(c)),
               "umx_cuda_session_get_state")
        return h, c

    # -- src/inference.hpp:20-23
    def umx_inference(self, audio: np.ndarray) -> np.ndarray:
        """One chunk (2, n<=n_buf) -> (4, 2, n); the state streams to the next call."""
        n = audio.shape[1]
        a = np.ascontiguousarray(audio.T.astype(np.float32))
        out = np.empty((4, n, 2), dtype=np.float32)
        _check(self._lib.umx_cuda_infer_segment(self._h, _p(a), n, _p(out)), "umx_inference")
        return np.ascontiguousarray(out.transpose(0, 2, 1))

    # -- src/lstm.hpp:22-24
    def umx_lstm_forward(self, target: int, x: np.ndarray) -> np.ndarray:
        F = x.shape[0]
        xin = np.asfortranarray(x.astype(np.float32))
        out = np.empty((F, self.model.hidden_size), dtype=np.float32, order="F")
        _check(self._lib.umx_cuda_lstm_forward(self._h, target, _p(xin), F, _p(out)), "umx_lstm_forward")
        return np.ascontiguousarray(out)

    # -- src/inference.cpp:29-186
    def target_mags(self, spec: np.ndarray) -> np.ndarray:
        F = spec.shape[1]
        s = _spec_to_ref(spec.astype(np.complex64))
        out = np.empty((4, NB_BINS, F, 2), dtype=np.float32)
        _check(self._lib.umx_cuda_target_mags(self._h, _p(s), _p(out)), "umx_cuda_target_mags")
        return np.ascontiguousarray(out.transpose(0, 3, 2, 1))

    # -- umx.cpp:99-295
    def demix(self, tracks, shift_offset: int = DEFAULT_SHIFT_OFFSET, interleaved: bool = False):
        """tracks: list of (2, n_i) arrays (or (n_i, 2) when interleaved). Returns list of (4, 2, n_i)
        (or (4, n_i, 2))."""
        ins = [np.ascontiguousarray(t.astype(np.float32) if interleaved else t.T.astype(np.float32)) for t in tracks]
        outs = [np.empty((4, a.shape[0], 2), dtype=np.float32) for a in ins]
        self.demix_into(ins, outs, shift_offset)
        return outs if interleaved else [np.ascontiguousarray(o.transpose(0, 2, 1)) for o in outs]

    def demix_into(self, ins, outs, shift_offset: int = DEFAULT_SHIFT_OFFSET):
        """Interleaved (n_i, 2) float32 inputs and preallocated (4, n_i, 2) outputs (e.g. pinned)."""
        k = len(ins)
        pin = (C.c_void_p * k)(*[a.ctypes.data for a in ins])
        pout = (C.c_void_p * k)(*[o.ctypes.data for o in outs])
        ns = (C.c_int64 * k)(*[a.shape[0] for a in ins])
        _check(self._lib.umx_cuda_demix(self._h, pin, ns, k, shift_offset, pout), "umx_cuda_demix")

    def demix_async(self, ins, outs, shift_offset: int = DEFAULT_SHIFT_OFFSET) -> int:
        """demix_into that only enqueues the work: returns a ticket for wait().  ins / outs must stay alive (and should be
        pinned) until then; up to two calls are in flight, so give consecutive calls different outs."""
        k = len(ins)
        pin = (C.c_void_p * k)(*[a.ctypes.data for a in ins])
        pout = (C.c_void_p * k)(*[o.ctypes.data for o in outs])
        ns = (C.c_int64 * k)(*[a.shape[0] for a in ins])
        ticket = C.c_int64(0)
        _check(self._lib.umx_cuda_demix_async(self._h, pin, ns, k, shift_offset, pout, C.byref(ticket)), "umx_cuda_demix_async")
        return ticket.value

    def wait(self, ticket: int = 0):
        """Blocks until the call with this ticket (0: every call enqueued so far) has delivered its outputs."""
        _check(self._lib.umx_cuda_session_wait(self._h, ticket), "umx_cuda_session_wait")

    def demix_device(self, in_ptrs, n_samples, out_ptrs, shift_offset: int = DEFAULT_SHIFT_OFFSET):
        """Device pointers (ints) of interleaved tracks / (4, n, 2) outputs."""
        k = len(in_ptrs)
        pin = (C.c_void_p * k)(*in_ptrs)
        pout = (C.c_void_p * k)(*out_ptrs)
        ns = (C.c_int64 * k)(*n_samples)
        _check(self._lib.umx_cuda_demix_device(self._h, pin, ns, k, shift_offset, pout), "umx_cuda_demix_device")

    def close(self):
        if self._h:
            self._lib.umx_cuda_session_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def nccl_unique_id() -> bytes:
    buf = (C.c_char * 128)()
    _check(lib().umx_cuda_nccl_unique_id(buf), "umx_cuda_nccl_unique_id")
    return bytes(buf)


def segment_plan(n_samples: int, shift_offset: int = DEFAULT_SHIFT_OFFSET, n_buf: int = SEGMENT_SAMPLES):
    """[(offset, chunk_length)] of split_inference (umx.cpp:214-218)."""
    l = lib()
    cnt = _check(l.umx_cuda_segment_plan(n_samples, shift_offset, n_buf, None, None, 0), "umx_cuda_segment_plan")
    offs = (C.c_int64 * cnt)()
    lens = (C.c_int * cnt)()
    _check(l.umx_cuda_segment_plan(n_samples, shift_offset, n_buf, offs, lens, cnt), "umx_cuda_segment_plan")
    return [(int(offs[i]), int(lens[i])) for i in range(cnt)]


# -- src/dsp.hpp:114-115
def stft(audio: np.ndarray, n_buf: int | None = None, reflect_pad: bool = False) -> np.ndarray:
    """reflect_pad: true reflect padding (upstream torch.stft) instead of the reference's symmetric edge copy."""
    n = audio.shape[1]
    n_buf = n if n_buf is None else n_buf
    F = nb_frames(n_buf)
    a = np.ascontiguousarray(audio.T.astype(np.float32))
    out = np.empty(2 * F * NB_BINS, dtype=np.complex64)
    _check(lib().umx_cuda_stft_opts(_p(a), n, n_buf, _p(out), int(reflect_pad)), "stft")
    return _spec_from_ref(out, F)


def istft(spec: np.ndarray, n_buf: int, n_wave: int | None = None) -> np.ndarray:
    n_wave = n_buf if n_wave is None else n_wave
    s = _spec_to_ref(spec.astype(np.complex64))
    out = np.empty((n_wave, 2), dtype=np.float32)
    _check(lib().umx_cuda_istft(_p(s), n_buf, n_wave, _p(out)), "istft")
    return np.ascontiguousarray(out.T)


# -- src/wiener.hpp:19-21
def wiener_filter(mix: np.ndarray, mags: np.ndarray, psd_abs2: bool = False, reg_once: bool = False, win_len: int = 0, niter: int = 1):
    """-> (y (4,2,F,B) complex64, mix/max_abs as the reference leaves its first argument).  The keyword arguments select the
    upstream-faithful variants (session options "wiener_psd", "wiener_reg", "wiener_win_len", "wiener_niter")."""
    F = mix.shape[1]
    m = _spec_to_ref(mix.astype(np.complex64))
    g = np.ascontiguousarray(mags.astype(np.float32).transpose(0, 3, 2, 1))
    out = np.empty((4, NB_BINS, F, 2), dtype=np.complex64)
    _check(lib().umx_cuda_wiener_opts(_p(m), _p(g), F, _p(out), int(psd_abs2), int(reg_once), int(win_len), int(niter)), "wiener_filter")
    return np.ascontiguousarray(out.transpose(0, 3, 2, 1)), _spec_from_ref(m.reshape(-1), F)


def wiener_noalloc(d_mix: int, d_mags: int, F: int, n_tracks: int, d_out: int, stream: int | None = None):
    """umx_cuda_wiener_noalloc: device pointers in the internal [frame][channel][bin] order; nothing is
    allocated (see include/umx_cuda.h for who needs that and what it costs)."""
    _check(lib().umx_cuda_wiener_noalloc(d_mix, d_mags, F, n_tracks, d_out, stream), "umx_cuda_wiener_noalloc")
