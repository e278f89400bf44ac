"""GPU parity: every stage of the CUDA path, called through the C ABI, against the oracle (numpy
restatement pinned to the compiled reference) and the golden vectors of the compiled reference.

Tolerances are BASELINE.json's: dequantised weights and segment indexing bit-exact; spectrograms
<= 1e-4 relative; stem waveforms <= 1e-3 max-abs and SDR >= 60 dB against the reference."""
import os

import numpy as np
import pytest

import golden_inputs as gi
from oracle import synth, umx_oracle as O

pytestmark = pytest.mark.gpu

SPEC_RTOL = 1e-4
WAVE_ATOL = 1e-3
SDR_MIN = 60.0


def rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def sdr(got, want):
    return float(10 * np.log10((want.astype(np.float64) ** 2).sum() / max(((got - want).astype(np.float64) ** 2).sum(), 1e-30)))


@pytest.fixture(scope="module")
def G():
    return gi.load_golden()


@pytest.fixture(scope="module")
def small(umx):
    raw = gi.model_bytes()
    return umx.load_umx_model(raw), O.load_model(raw)


# ---- model --------------------------------------------------------------------------------------------
def test_dequant_bit_exact(umx, small, G):
    gm, om = small
    assert gm.hidden_size == gi.HIDDEN
    assert np.array_equal(G["w_fc1_w"], gm.tensor("fc1_w", 2)[:64, :64])
    assert np.array_equal(G["w_lstm_hh_w"], gm.tensor("lstm_hh_w", 3, 1, 1))
    assert np.array_equal(G["w_input_mean"], gm.tensor("input_mean", 2))
    for t in range(4):
        for ref_name, name in (("fc1_w", "fc1.weight"), ("fc2_w", "fc2.weight"), ("fc3_w", "fc3.weight")):
            assert np.array_equal(gm.tensor(ref_name, t).view(np.uint32), np.ascontiguousarray(om.t[t][name].T).view(np.uint32))
        for ref_name, name in (("bn1_rv", "bn1.running_var"), ("bn3_w", "bn3.weight"), ("bn2_b", "bn2.bias")):
            assert np.array_equal(gm.tensor(ref_name, t).ravel(), om.t[t][name])
        for l in range(3):
            for d in range(2):
                sfx = f"_l{l}" + ("_reverse" if d else "")
                assert np.array_equal(gm.tensor("lstm_ih_w", t, l, d), om.t[t]["lstm.weight_ih" + sfx].T)
                assert np.array_equal(gm.tensor("lstm_hh_b", t, l, d).ravel(), om.t[t]["lstm.bias_hh" + sfx])


def test_model_load_errors(umx):
    raw = bytearray(gi.model_bytes())
    bad = bytes([0]) + bytes(raw[1:])
    with pytest.raises(umx.UmxCudaError, match="bad magic"):
        umx.load_umx_model(bad)
    with pytest.raises(umx.UmxCudaError):
        umx.load_umx_model(bytes(raw[: len(raw) // 2]))
    with pytest.raises(umx.UmxCudaError, match="failed to open"):
        umx.load_umx_model("/nonexistent/model.gz")
    import gzip

    m = umx.load_umx_model(gzip.compress(bytes(raw), 1))  # gz path, src/model.cpp:56-84
    assert m.hidden_size == gi.HIDDEN


# ---- STFT / iSTFT -------------------------------------------------------------------------------------
def test_stft_golden(umx, G):
    assert rel(umx.stft(gi.audio(), gi.N_BUF), G["spec"]) < SPEC_RTOL


def test_istft_golden(umx, G):
    assert np.abs(umx.istft(G["spec"], gi.N_BUF, gi.N) - G["istft"]).max() < 1e-5


def test_gtest_roundtrip_rand(umx, G):
    """test/test_dsp.cpp:41-80"""
    w = gi.gtest_rand_waveform()
    s = umx.stft(w, 4096)
    assert s.shape == (2, 5, 2049)
    assert rel(s, G["gtest_rand_spec"]) < SPEC_RTOL
    assert np.abs(umx.istft(s, 4096, 4096) - w).max() < 1e-4


@pytest.mark.parametrize("n,n_buf", [(4096, 4096), (13000, 20000), (1, 8192), (0, 4096), (50000, 50000), (44100 * 10, 2646000)])
def test_stft_istft_vs_oracle(umx, n, n_buf):
    audio = synth.make_audio(max(n, 1), 5)[:, :n]
    s = umx.stft(audio, n_buf)
    want = O.stft(audio, n_buf)
    assert s.shape == want.shape
    assert rel(s, want) < SPEC_RTOL if n else np.abs(s).max() == 0
    if n:
        back = umx.istft(s, n_buf, n)
        assert back.shape == (2, n)
        assert np.abs(back - O.istft(want, n_buf, n)).max() < 1e-5
        assert np.abs(back - audio).max() < 1e-4  # round trip, test_dsp.cpp tolerance


def test_stft_true_stereo_channels_do_not_mix(umx):
    """The reference never tests L != R (its stereo fixture is dual mono); the packed L+iR FFT must not leak."""
    a = synth.make_audio(20000, 1)
    a[1] = 0
    s = umx.stft(a, 20000)
    assert np.abs(s[1]).max() < 1e-4 * np.abs(s[0]).max()
    assert rel(s[0], O.stft(a, 20000)[0]) < SPEC_RTOL


# ---- LSTM -----------------------------------------------------------------------------------------------
def test_lstm_streaming_golden(umx, small, G):
    gm, _ = small
    sess = umx.Session(gm, 1, gi.N_BUF)
    x = gi.lstm_input()
    assert np.abs(sess.umx_lstm_forward(1, x) - G["lstm_call1"]).max() < 1e-4
    assert np.abs(sess.umx_lstm_forward(1, x) - G["lstm_call2"]).max() < 1e-4
    h, c = sess.state(0, 1, 2, 1)
    assert np.abs(h - G["lstm_h_l2_bwd"]).max() < 1e-4 and np.abs(c - G["lstm_c_l2_bwd"]).max() < 1e-4
    sess.reset()
    assert np.abs(sess.umx_lstm_forward(1, x) - G["lstm_call1"]).max() < 1e-4


# ---- Wiener ---------------------------------------------------------------------------------------------
def test_wiener_golden(umx, G):
    y, mix = umx.wiener_filter(G["spec"], gi.wiener_mags(G["spec"]))
    assert rel(y[..., :: gi.WIENER_BIN_STEP], G["wiener_y_sub"]) < 1e-4
    assert rel(mix[..., :: gi.WIENER_BIN_STEP], G["wiener_mix_sub"]) < 1e-5


def test_wiener_zero_mix_and_loud_mix(umx):
    F = 23
    rng = np.random.default_rng(3)
    spec = (rng.standard_normal((2, F, 2049)) + 1j * rng.standard_normal((2, F, 2049))).astype(np.complex64) * 40
    spec[:, 5:9] = 0  # silent frames: polar(mag, arg 0)
    mags = rng.uniform(0, 50, (4, 2, F, 2049)).astype(np.float32)
    y, mix = umx.wiener_filter(spec, mags)
    wy, wmix = O.wiener(spec, mags)
    assert rel(y, wy) < 1e-4 and rel(mix, wmix) < 1e-5


# ---- network / inference ------------------------------------------------------------------------------
PAIRS = [3]  # gemm_impl: 0 = tcgen05 (CTA pairs in order, single-CTA tiles in the wavefront), 2 = pairs always, 3 = single-CTA always


@pytest.mark.parametrize("gemm_impl", [1, 0] + PAIRS)
def test_umx_inference_golden(umx, small, G, gemm_impl):
    gm, _ = small
    sess = umx.Session(gm, 1, gi.N_BUF)
    sess.set_option("gemm_impl", gemm_impl)
    a = gi.audio()
    for call, audio in (("inference_call1", a), ("inference_call2", a[:, : gi.N // 2])):
        got = sess.umx_inference(audio)
        assert np.abs(got - G[call]).max() < WAVE_ATOL
        assert min(sdr(got[j], G[call][j]) for j in range(4)) > SDR_MIN


@pytest.mark.parametrize("gemm_impl", [1, 0] + PAIRS)
def test_target_mags_vs_oracle(umx, small, G, gemm_impl):
    gm, om = small
    sess = umx.Session(gm, 1, gi.N_BUF)
    sess.set_option("gemm_impl", gemm_impl)
    got = sess.target_mags(G["spec"])
    mix_mag = np.abs(G["spec"])
    for j in range(4):
        want = O.masked_mags(O.target_mask(om.t[j], mix_mag, O.LstmState(gi.HIDDEN // 2)), mix_mag)
        assert rel(got[j], want) < 1e-3


@pytest.mark.parametrize("gemm_impl,terms", [(1, 3), (0, 3), (0, 2), (0, 1)] + [(3, 3), (3, 2), (3, 1)])
def test_demix_multi_segment_multi_track(umx, gemm_impl, terms):
    """shift + split + recombine (umx.cpp:99-295) on three tracks of different lengths in lockstep."""
    hidden, n_buf = 64, 16384
    raw = synth.make_ggml(hidden=hidden, seed=77)
    gm, om = umx.load_umx_model(raw), O.load_model(raw)
    sess = umx.Session(gm, 3, n_buf)
    sess.set_option("gemm_impl", gemm_impl)
    sess.set_option("gemm_terms", terms)
    tracks = [synth.make_audio(n, 20 + i) for i, n in enumerate((30000, 52000, 9000))]
    got = sess.demix(tracks, shift_offset=4033)
    saved = (O.SEGMENT_SAMPLES, O.STRIDE_SAMPLES)
    O.SEGMENT_SAMPLES, O.STRIDE_SAMPLES = n_buf, int(np.float32(0.75) * np.float32(n_buf))
    try:
        for t, g in zip(tracks, got):
            want = O.demix(om, t, shift_offset=4033)
            assert g.shape == want.shape
            assert np.abs(g - want).max() < WAVE_ATOL
            assert min(sdr(g[j], want[j]) for j in range(4)) > SDR_MIN
    finally:
        O.SEGMENT_SAMPLES, O.STRIDE_SAMPLES = saved
    # the same session demixes again identically (state is reset per call)
    again = sess.demix(tracks[:1], shift_offset=4033)[0]
    assert np.abs(again - got[0]).max() < 1e-6


@pytest.mark.parametrize("hidden", [64, 1024])  # validation recurrence kernel / the tcgen05 cluster kernel
@pytest.mark.parametrize("mode,lane_targets", [(3, 0), (3, 1), (3, 2), (3, 4)])  # (3, 0) is the default
def test_demix_pipeline_modes_identical(umx, mode, lane_targets, hidden):
    """Session option "pipeline": the segment wavefront (3: layer l of segment s+1 beside layer l+1 of segment s, one stream per
    layer and target group) must give bit-identical stems to the in-order schedule (0).  Up to 7 segments per track, i.e. deeper
    than every buffer ring of the wavefront."""
    n_buf = 16384
    gm = umx.load_umx_model(synth.make_ggml(hidden=hidden, seed=78))
    tracks = [synth.make_audio(n, 30 + i) for i, n in enumerate((70000, 41000, 9000))]
    ref_sess = umx.Session(gm, 3, n_buf)
    ref_sess.set_option("pipeline", 0)
    want = ref_sess.demix(tracks, shift_offset=4033)
    sess = umx.Session(gm, 3, n_buf)
    sess.set_option("pipeline", mode)
    sess.set_option("lane_targets", lane_targets)
    for _ in range(2):
        got = sess.demix(tracks, shift_offset=4033)
        for g, w in zip(got, want):
            assert np.array_equal(g, w)


@pytest.mark.parametrize("hidden", [64, 1024])
def test_demix_direct_final_identical(umx, hidden):
    """Session option "direct_final": output samples that only one segment covers are written finalised by the inverse STFT
    kernel itself, (w y)/w, instead of going through the accumulator and finalize_kernel.  Same bits, for every track length
    (short last segments, tracks shorter than one stride, a track of one sample) and through the host and device copies."""
    n_buf = 16384
    gm = umx.load_umx_model(synth.make_ggml(hidden=hidden, seed=79))
    tracks = [synth.make_audio(n, 50 + i) for i, n in enumerate((70000, 12288 - 18017 + 22050, 9000, 1, 40961))]
    a, b = umx.Session(gm, len(tracks), n_buf), umx.Session(gm, len(tracks), n_buf)
    a.set_option("direct_final", 0)
    for shift in (4033, 0, 11025):
        want, got = a.demix(tracks, shift_offset=shift), b.demix(tracks, shift_offset=shift)
        for g, w in zip(got, want):
            assert np.array_equal(g, w)


@pytest.mark.parametrize("terms", [3, 2, 1])
def test_gemm_cta_pairs_identical_to_single_cta(umx, terms):
    """CTA pairs (tcgen05.mma.cta_group::2 on 256 x 256 tiles, B split across the pair; what gemm_impl 0 uses in the in-order schedule)
    accumulate every output element over the same k-blocks in the same order as the single-CTA kernel (gemm_impl 3): identical bits,
    at the UMX-L hidden size, ragged row counts."""
    n_buf = 16384
    gm = umx.load_umx_model(synth.make_ggml(hidden=1024, seed=80))
    tracks = [synth.make_audio(n, 60 + i) for i, n in enumerate((40000, 9000, 23000, 30000, 12000))]
    a, b = umx.Session(gm, len(tracks), n_buf), umx.Session(gm, len(tracks), n_buf)
    for sess, impl in ((a, 3), (b, 0)):
        sess.set_option("gemm_impl", impl)
        sess.set_option("gemm_terms", terms)
        sess.set_option("pipeline", 0)
    want = a.demix(tracks)
    for g, w in zip(b.demix(tracks), want):
        assert np.array_equal(g, w)
    if os.environ.get("UMX_TEST_PAIRS_WAVEFRONT") == "1":  # pairs drawing from the wavefront's tile queues (opt-in: gemm_impl 2)
        b.set_option("gemm_impl", 2)
        b.set_option("pipeline", 3)
        for g, w in zip(b.demix(tracks), want):
            assert np.array_equal(g, w)


def test_demix_shift_offsets_and_trim(umx, small):
    gm, om = small
    n_buf = 8192
    sess = umx.Session(gm, 1, n_buf)
    sess.set_option("gemm_impl", 1)
    t = synth.make_audio(20000, 4)
    saved = (O.SEGMENT_SAMPLES, O.STRIDE_SAMPLES)
    O.SEGMENT_SAMPLES, O.STRIDE_SAMPLES = n_buf, int(np.float32(0.75) * np.float32(n_buf))
    try:
        for shift in (0, 4033, 11025):  # offsets > 11025 overrun the reference's own buffer (umx.cpp:120-122)
            got = sess.demix([t], shift_offset=shift)[0]
            want = O.demix(om, t, shift_offset=shift)
            assert np.abs(got - want).max() < WAVE_ATOL
    finally:
        O.SEGMENT_SAMPLES, O.STRIDE_SAMPLES = saved
