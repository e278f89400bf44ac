"""GPU: size-independent properties at BASELINE.json's full sizes, where the CPU oracle would take minutes to hours.

* STFT -> iSTFT round trip of a 7-minute track (the reference's own bar in test/test_dsp.cpp: 1e-4);
* linearity of the STFT;
* Wiener-EM: the four filtered stems add up to the mixture (sum_j W_j = I up to the regulariser, src/wiener.cpp:268-325);
* tracks demixed in lockstep do not see each other: a track alone and the same track inside a batch give the same stems;
* demixing is deterministic: two sessions, same input, identical bits.
All through the C ABI."""
import ctypes as C

import numpy as np
import pytest

from oracle import synth

pytestmark = pytest.mark.gpu

SEVEN_MIN = 18522000


def _vp(a):
    return a.ctypes.data_as(C.c_void_p)


def test_stft_istft_round_trip_seven_minutes(umx):
    L = umx.lib()
    a = synth.make_audio(SEVEN_MIN, 7)
    n = a.shape[1]
    F = n // 1024 + 1
    wave = np.ascontiguousarray(a.T.astype(np.float32))  # the reference's MatrixXf(2, n) memory
    spec = np.empty(2 * F * 2049, dtype=np.complex64)
    back = np.empty((n, 2), dtype=np.float32)
    assert L.umx_cuda_stft(_vp(wave), n, n, _vp(spec)) == 0
    assert L.umx_cuda_istft(_vp(spec), n, n, _vp(back)) == 0
    assert float(np.abs(back - wave).max()) < 1e-4
    # linearity on the same buffers: stft(2.5 x) = 2.5 stft(x)
    spec2 = np.empty_like(spec)
    wave2 = wave * np.float32(2.5)
    assert L.umx_cuda_stft(_vp(wave2), n, n, _vp(spec2)) == 0
    scale = float(np.abs(spec).max())
    assert float(np.abs(spec2 - np.complex64(2.5) * spec).max()) < 1e-5 * 2.5 * scale


def test_wiener_stems_sum_to_mixture_full_segment(umx):
    rng = np.random.default_rng(5)
    F = 2584
    mix = (50.0 * (rng.standard_normal((2, F, 2049)) + 1j * rng.standard_normal((2, F, 2049)))).astype(np.complex64)
    mags = (np.abs(mix)[None] * rng.uniform(0.05, 1.0, size=(4, 2, F, 2049))).astype(np.float32)
    y, mix_scaled = umx.wiener_filter(mix, mags)
    assert np.isfinite(y.real).all() and np.isfinite(y.imag).all()
    # the reference leaves the mixture divided by max(1, max |X| / 10) (src/wiener.cpp:17-30,118-129) and returns the stems at full scale
    max_abs = max(1.0, float(np.abs(mix).max()) / 10.0)
    assert max_abs > 20.0
    assert np.allclose(mix_scaled * np.float32(max_abs), mix, rtol=1e-5, atol=1e-5 * max_abs)
    # sum_j W_j = (Cxx - reg I) Cxx^-1: the identity wherever the model covariance dominates the regulariser; judged on the
    # bulk of the (channel, frame, bin) cells, since single cells with a near-singular covariance may deviate
    rel = (np.abs(y.sum(axis=0) - mix) / np.abs(mix).max()).ravel()
    med, p99 = float(np.median(rel)), float(np.percentile(rel, 99.0))
    print(f"wiener sum-to-mixture residual / peak: median {med:.3e}, 99th percentile {p99:.3e}, max {float(rel.max()):.3e}")
    assert med < 2e-5 and p99 < 1e-3  # measured 1.4e-6 / 7.6e-5


def test_lockstep_tracks_are_independent_and_deterministic(umx):
    gm = umx.load_umx_model(synth.make_ggml(hidden=1024, seed=1234))
    n_buf = 65536  # 64 frames + 1 per window: the cluster recurrence kernel and the tensor-core GEMMs, small windows
    tracks = [synth.make_audio(n, 90 + i) for i, n in enumerate((150000, 90000, 200000, 61000, 120000))]
    alone = umx.Session(gm, 1, n_buf).demix([tracks[2]])[0]
    batch_a = umx.Session(gm, len(tracks), n_buf).demix(tracks)
    batch_b = umx.Session(gm, len(tracks), n_buf).demix(tracks)
    for x, y in zip(batch_a, batch_b):
        assert np.array_equal(x, y)
    assert np.array_equal(batch_a[2], alone)


@pytest.mark.parametrize("hidden,n_buf", [(1024, 65536), (64, 16384)])
def test_skipping_frames_beyond_the_chunk_changes_nothing(umx, hidden, n_buf):
    """The reference runs every frame of the zero-extended 60 s window; here the frames beyond a chunk are computed once (first
    layer) or not at all (after the network).  Option skip_zero_frames=0 runs them all: the stems must be identical bits.
    Ragged lockstep batch: short last chunks, and tracks that run out of segments early."""
    gm = umx.load_umx_model(synth.make_ggml(hidden=hidden, seed=1234))
    tracks = [synth.make_audio(n, 70 + i) for i, n in enumerate((3 * n_buf + 900, n_buf // 3, 2 * n_buf + n_buf // 2, 5000))]
    outs = []
    for skip in (1, 0):
        sess = umx.Session(gm, len(tracks), n_buf)
        sess.set_option("skip_zero_frames", skip)
        outs.append(sess.demix(tracks))
    for x, y in zip(*outs):
        assert np.isfinite(x).all()
        assert np.array_equal(x, y)


def test_stage_timeline_and_pinned_host_buffers(umx):
    """umx_cuda_session_timeline (the raw stage intervals behind stage_ms) and umx_cuda_host_alloc (page-locked buffers for hosts
    without the CUDA runtime): a demix through pinned buffers gives the bits of the pageable path, and the timeline of the call
    holds every stage, ordered and inside the call."""
    import ctypes as C

    gm = umx.load_umx_model(synth.make_ggml(hidden=64, seed=81))
    sess = umx.Session(gm, 2, 16384)
    tracks = [synth.make_audio(n, 90 + i) for i, n in enumerate((40000, 9000))]
    want = [w.transpose(0, 2, 1) for w in sess.demix(tracks)]  # (4, n, 2) like the interleaved output buffers
    L = umx.lib()
    ins, outs, raw = [], [], []
    for t in tracks:
        n = t.shape[1]
        for nbytes, shape, dst in ((8 * n, (n, 2), ins), (32 * n, (4, n, 2), outs)):
            p = C.c_void_p()
            assert L.umx_cuda_host_alloc(nbytes, C.byref(p)) == 0 and p.value
            raw.append(p)
            dst.append(np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(nbytes // 4,)).reshape(shape))
        ins[-1][:] = t.T
    sess.set_profiling(True)
    sess.demix_into(ins, outs)
    for o, w in zip(outs, want):
        assert np.array_equal(o, w)
    tl = sess.timeline()
    names = {s for s, _, _ in tl}
    assert {"stft", "gemm", "lstm", "wiener", "istft", "h2d", "d2h"} <= names
    assert all(0.0 <= b <= e for _, b, e in tl)
    assert abs(sess.stage_ms("gemm") - sum(e - b for s, b, e in tl if s == "gemm")) >= -1e-3  # union <= sum of intervals
    assert sess.stage_ms("gemm") <= sum(e - b for s, b, e in tl if s == "gemm") + 1e-3
    for p in raw:
        L.umx_cuda_host_free(p)


def test_demix_async_two_calls_in_flight(umx):
    """umx_cuda_demix_async / umx_cuda_session_wait: three different batches enqueued back to back (two in flight at a time, the
    third call waits for the first) give, each in its own output buffers, the bits of umx_cuda_demix one call at a time; the
    synchronous entry points in between wait for what is in flight."""
    gm = umx.load_umx_model(synth.make_ggml(hidden=64, seed=82))
    sess = umx.Session(gm, 3, 16384)
    batches = [[synth.make_audio(n, 100 + 10 * b + i) for i, n in enumerate(lens)]
               for b, lens in enumerate(((40000, 9000, 23000), (12000, 52000), (30000, 30000, 1)))]
    want = [[w.transpose(0, 2, 1) for w in sess.demix(tr)] for tr in batches]
    for pipeline in (3, 0):
        sess.set_option("pipeline", pipeline)
        ins = [[np.ascontiguousarray(t.T) for t in tr] for tr in batches]
        outs = [[np.full((4, t.shape[1], 2), np.nan, np.float32) for t in tr] for tr in batches]
        tickets = [sess.demix_async(i, o) for i, o in zip(ins, outs)]
        assert tickets == sorted(tickets) and len(set(tickets)) == 3
        sess.wait(tickets[1])
        for o, w in zip(outs[0] + outs[1], want[0] + want[1]):
            assert np.array_equal(o, w)
        sess.wait()
        for o, w in zip(outs[2], want[2]):
            assert np.array_equal(o, w)
        # a synchronous call right behind an enqueued one
        outs0 = [np.full_like(o, np.nan) for o in outs[0]]
        sess.demix_async(ins[0], outs0)
        again = [w.transpose(0, 2, 1) for w in sess.demix(batches[1])]
        for o, w in zip(again, want[1]):
            assert np.array_equal(o, w)
        for o, w in zip(outs0, want[0]):
            assert np.array_equal(o, w)
