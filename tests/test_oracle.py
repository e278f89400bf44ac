"""CPU: the numpy restatement (oracle/umx_oracle.py) against golden vectors produced by the compiled,
unmodified reference (tests/golden/make_golden.py), against the reference's own gtest cases
(test/test_dsp.cpp) restated on the same inputs, and live against oracle/_ref when it is present."""
import numpy as np
import pytest

import golden_inputs as gi
from oracle import ref, synth, umx_oracle as O

G = gi.load_golden()


def rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


@pytest.fixture(scope="module")
def model():
    return O.load_model(gi.model_bytes())


def test_dequant_bit_exact(model):
    # src/model.cpp:614,660; reference matrices are (in, out) = W^T
    t = model.t[2]
    assert np.array_equal(G["w_fc1_w"], t["fc1.weight"].T[:64, :64])
    assert np.array_equal(G["w_fc3_w"], t["fc3.weight"].T[:64, :64])
    assert np.array_equal(G["w_bn2_rv"].ravel(), t["bn2.running_var"])
    assert np.array_equal(G["w_input_mean"].ravel(), np.concatenate([t["input_mean"]] * 2))
    assert np.array_equal(G["w_lstm_hh_w"], model.t[3]["lstm.weight_hh_l1_reverse"].T)


def test_stft_matches_reference():
    assert rel(O.stft(gi.audio(), gi.N_BUF), G["spec"]) < 1e-5  # BASELINE tolerance is 1e-4


def test_istft_matches_reference():
    assert np.abs(O.istft(G["spec"], gi.N_BUF, gi.N) - G["istft"]).max() < 1e-5


def test_gtest_stft_roundtrip_rand_waveform():
    """test/test_dsp.cpp:41-80: 4096 rand() samples, 5 frames x 2049 bins, round trip within 1e-4."""
    w = gi.gtest_rand_waveform()
    s = O.stft(w, 4096)
    assert s.shape == (2, 5, 2049)
    assert rel(s, G["gtest_rand_spec"]) < 1e-5
    back = O.istft(s, 4096, 4096)
    assert np.abs(back - w).max() < 1e-4  # NEAR_TOLERANCE, test_dsp.cpp:7
    assert np.abs(G["gtest_rand_roundtrip"] - w).max() < 1e-4


def test_gtest_magphase_roundtrip():
    """test/test_dsp.cpp:118-273: abs/arg -> polar reproduces the spectrogram within 1e-4, mags >= 0."""
    s = G["spec"]
    mag, ph = np.abs(s), np.angle(s)
    assert (mag >= 0).all()
    back = (mag * np.exp(1j * ph)).astype(np.complex64)
    assert np.abs(back - s).max() < 1e-4 * max(1.0, np.abs(s).max())


def test_lstm_streaming_matches_reference(model):
    st = O.LstmState(gi.HIDDEN // 2)
    x = gi.lstm_input()
    assert np.abs(O.lstm_forward(model.t[1], x, st) - G["lstm_call1"]).max() < 1e-5
    assert np.abs(O.lstm_forward(model.t[1], x, st) - G["lstm_call2"]).max() < 1e-5  # state carried
    assert np.abs(st.h[2, 1] - G["lstm_h_l2_bwd"]).max() < 1e-5
    assert np.abs(st.c[2, 1] - G["lstm_c_l2_bwd"]).max() < 1e-5


def test_wiener_matches_reference():
    y, mix = O.wiener(G["spec"], gi.wiener_mags(G["spec"]))
    assert rel(y[..., :: gi.WIENER_BIN_STEP], G["wiener_y_sub"]) < 1e-4
    assert rel(mix[..., :: gi.WIENER_BIN_STEP], G["wiener_mix_sub"]) < 1e-5


def test_umx_inference_matches_reference(model):
    states = [O.LstmState(gi.HIDDEN // 2) for _ in range(4)]
    a = gi.audio()
    assert np.abs(O.umx_inference(model, a, gi.N_BUF, states) - G["inference_call1"]).max() < 1e-5
    assert np.abs(O.umx_inference(model, a[:, : gi.N // 2], gi.N_BUF, states) - G["inference_call2"]).max() < 1e-5


def test_segment_plan_constants():
    # umx.cpp:112-113,156-157,180 and SURVEY 8(a) a2
    assert (O.SEGMENT_SAMPLES, O.STRIDE_SAMPLES, O.MAX_SHIFT) == (2646000, 1984500, 22050)
    L = O.shifted_length(18522000, 4033)
    plan = O.segment_plan(L)
    assert L == 18540017 and len(plan) == 10 and plan[-1] == (9 * 1984500, 679517)
    plan3 = O.segment_plan(O.shifted_length(7938000, 4033))
    assert len(plan3) == 5 and plan3[-1][1] == 18017
    assert O.segment_plan(O.shifted_length(441000, 4033)) == [(0, 459017)]
    w = O.transition_weight()
    assert w[0] == np.float32(1) / np.float32(1323000) and w[1322999] == 1.0 and w[1323000] == 1.0 and w[-1] == w[0]


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built (needs /root/reference)")
def test_live_against_compiled_reference(tmp_path):
    """A second seed/shape than the golden file, straight against the compiled reference."""
    import gzip

    raw = synth.make_ggml(hidden=64, seed=99)
    p = tmp_path / "m.gz"
    p.write_bytes(gzip.compress(raw, 1))
    rm, om = ref.Model(str(p)), O.load_model(raw)
    n_buf, n = 12288, 9000
    audio = synth.make_audio(n, 3)
    rs, os_ = ref.LstmState(32, O.nb_frames(n_buf)), [O.LstmState(32) for _ in range(4)]
    for _ in range(2):
        a = ref.umx_inference(rm, rs, audio, n_buf)
        b = O.umx_inference(om, audio, n_buf, os_)
        assert np.abs(a - b).max() < 1e-5
    for t in range(4):
        assert np.array_equal(rm.tensor("fc2_w", t), om.t[t]["fc2.weight"].T)
        assert np.array_equal(rm.tensor("lstm_ih_w", t, 2, 0), om.t[t]["lstm.weight_ih_l2"].T)


@pytest.mark.slow
def test_demix_multisegment_matches_reference_cli():
    """umx.cpp:99-295 end to end on THREE segments (the last one 459,017 samples): O.demix against the four WAVs the reference's
    own CLI wrote for the same 100 s track (tests/golden/make_golden_multiseg.py), and the integer segment table against the lines
    the CLI printed (umx.cpp:220-222).  ~1 minute of numpy."""
    g = np.load(gi.MS_GOLDEN)
    L = O.shifted_length(gi.MS_N, O.DEFAULT_SHIFT_OFFSET)
    assert O.segment_plan(L) == [tuple(int(v) for v in row) for row in g["plan"]]  # bit-exact indexing
    got = O.demix(O.load_model(gi.multiseg_model_bytes()), gi.multiseg_audio())
    assert got.shape == (4, 2, gi.MS_N)
    want = g["stems"]
    sub = got[:, :, g["idx"]]
    err = float(np.abs(sub - want).max())
    sdr = float(10 * np.log10((want.astype(np.float64) ** 2).sum() / ((sub - want).astype(np.float64) ** 2).sum()))
    assert err < 1e-5 and sdr > 100.0, (err, sdr)


def test_upstream_variants_match_open_unmix():
    """The option-switched variants of the oracle (what the session options "stft_pad", "wiener_psd", "wiener_reg",
    "wiener_win_len", "wiener_niter" select) against UPSTREAM Open-Unmix itself: golden vectors made by
    tests/golden/make_golden_upstream.py from vendor/open-unmix-pytorch (torch.stft with reflect padding;
    openunmix.filtering.wiener per window, float64)."""
    g = np.load(gi.UP_GOLDEN)
    assert rel(O.stft(gi.audio(), gi.N_BUF, reflect=True), g["stft_reflect"]) < 1e-5
    assert rel(O.stft(gi.audio(), gi.N_BUF, reflect=False), g["stft_reflect"]) > 1e-3  # the reference's padding IS different
    mix, mags = gi.upstream_wiener_inputs()
    for it in (0, 1, 2):
        y = O.wiener_windows(mix, mags, gi.UP_WIN, psd_abs2=True, reg_once=True, niter=it)
        assert rel(y[..., :: gi.UP_BIN_STEP], g[f"wiener_win_niter{it}"]) < 2e-4, it
    # and each switch matters: the reference's own arithmetic does not reproduce upstream
    y_ref = O.wiener_windows(mix, mags, 0)
    assert rel(y_ref[..., :: gi.UP_BIN_STEP], g["wiener_win_niter1"]) > 1e-3
