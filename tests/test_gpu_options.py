"""GPU: the upstream-faithful session options ("stft_pad", "wiener_psd", "wiener_reg", "wiener_win_len", "wiener_niter",
"lstm_reset": each switches one of the reference's deviations from Open-Unmix back) against
 (a) UPSTREAM itself: golden vectors made from vendor/open-unmix-pytorch (tests/golden/make_golden_upstream.py) for the
     kernel-level STFT and Wiener entry points, and
 (b) the numpy oracle with the same options for whole tracks through the fused pipeline (several segments, ragged batch)."""
import numpy as np
import pytest

import golden_inputs as gi
from oracle import synth, umx_oracle as O

pytestmark = pytest.mark.gpu


def rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def sdr(got, want):
    return float(10 * np.log10((want.astype(np.float64) ** 2).sum() / max(((got - want).astype(np.float64) ** 2).sum(), 1e-30)))


def test_stft_reflect_padding_matches_torch_stft(umx):
    g = np.load(gi.UP_GOLDEN)
    got = umx.stft(gi.audio(), gi.N_BUF, reflect_pad=True)
    assert rel(got, g["stft_reflect"]) < 1e-4  # BASELINE's spectrogram tolerance
    assert rel(umx.stft(gi.audio(), gi.N_BUF), g["stft_reflect"]) > 1e-3  # the default (reference) padding differs


@pytest.mark.parametrize("niter", [0, 1, 2])
def test_wiener_upstream_variant_matches_open_unmix(umx, niter):
    """PSD |y|^2, regulariser once, per-window statistics and max|X| (three windows, the last one short), 0, 1 or 2 EM iterations."""
    g = np.load(gi.UP_GOLDEN)
    mix, mags = gi.upstream_wiener_inputs()
    got, _ = umx.wiener_filter(mix, mags, psd_abs2=True, reg_once=True, win_len=gi.UP_WIN, niter=niter)
    assert rel(got[..., :: gi.UP_BIN_STEP], g[f"wiener_win_niter{niter}"]) < 2e-4


OPTION_SETS = [
    {"stft_pad": 1},
    {"wiener_psd": 1},
    {"wiener_reg": 1},
    {"wiener_win_len": 5},
    {"wiener_niter": 0},
    {"wiener_niter": 2},
    {"wiener_niter": 3, "wiener_win_len": 5, "wiener_psd": 1},
    {"lstm_reset": 1},
    {"stft_pad": 1, "wiener_psd": 1, "wiener_reg": 1, "wiener_win_len": 6, "lstm_reset": 1},  # "as upstream"
]


@pytest.mark.parametrize("opts", OPTION_SETS, ids=lambda o: "+".join(f"{k}={v}" for k, v in o.items()))
def test_demix_with_options_matches_oracle(umx, opts):
    """Whole tracks (several segments, three ragged tracks in lockstep, default schedule) with every option and with all of
    them against the oracle restated with the same options; and the option changes the result (it is not a no-op)."""
    hidden, n_buf = 64, 16384
    raw = synth.make_ggml(hidden=hidden, seed=91)
    gm, om = umx.load_umx_model(raw), O.load_model(raw)
    tracks = [synth.make_audio(n, 70 + i) for i, n in enumerate((45000, 30000, 9000))]
    sess = umx.Session(gm, 3, n_buf)
    base = sess.demix(tracks)
    for k, v in opts.items():
        sess.set_option(k, v)
    got = sess.demix(tracks)
    saved = (O.SEGMENT_SAMPLES, O.STRIDE_SAMPLES)
    O.SEGMENT_SAMPLES, O.STRIDE_SAMPLES = n_buf, int(np.float32(0.75) * np.float32(n_buf))
    try:
        for i in (0, 2):
            want = O.demix(om, tracks[i], opts=opts)
            err, s = float(np.abs(got[i] - want).max()), min(sdr(got[i][j], want[j]) for j in range(4))
            assert err < 1e-3 and s > 60.0, (opts, i, err, s)
    finally:
        O.SEGMENT_SAMPLES, O.STRIDE_SAMPLES = saved
    assert float(np.abs(got[0] - base[0]).max()) > 1e-5


def test_option_errors(umx):
    gm = umx.load_umx_model(synth.make_ggml(hidden=16, seed=3))
    sess = umx.Session(gm, 1, 16384)  # 17 frames
    with pytest.raises(umx.UmxCudaError, match="19 windows"):
        sess.set_option("wiener_win_len", -3)
    with pytest.raises(umx.UmxCudaError):
        sess.set_option("wiener_niter", 9)
    with pytest.raises(umx.UmxCudaError):
        sess.set_option("no_such_option", 1)
