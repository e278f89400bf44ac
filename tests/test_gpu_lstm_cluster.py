"""GPU: the tensor-core cluster recurrence (lstm_cluster.cu: UMX-L's 512 units per direction on clusters of 8 CTAs, UMX-HQ's 256
on clusters of 4) against the simple recurrence kernel and the oracle, single and multi-track, streaming across calls."""
import numpy as np
import pytest

from oracle import synth, umx_oracle as O

pytestmark = pytest.mark.gpu

N_BUF = 16384
F = N_BUF // 1024 + 1


@pytest.fixture(scope="module", params=[1024, 512], ids=["umxl-hidden1024", "umxhq-hidden512"])
def umxl(umx, request):
    raw = synth.make_ggml(hidden=request.param, seed=1234)
    return umx.load_umx_model(raw), O.load_model(raw)


def test_lstm_forward_cluster_vs_oracle_streaming(umx, umxl):
    gm, om = umxl
    H = gm.hidden_size
    x = np.random.default_rng(0).uniform(-1, 1, (F, H)).astype(np.float32)
    fast, slow = umx.Session(gm, 1, N_BUF), umx.Session(gm, 1, N_BUF)
    fast.set_option("lstm_impl", 0)
    slow.set_option("lstm_impl", 1)
    st = O.LstmState(H // 2)
    for call in range(3):  # state streams from call to call (lstm.cpp:160-161)
        want = O.lstm_forward(om.t[2], x, st)
        a, b = fast.umx_lstm_forward(2, x), slow.umx_lstm_forward(2, x)
        assert np.abs(b - want).max() < 2e-5, f"simple kernel, call {call}"
        assert np.abs(a - want).max() < 2e-5, f"cluster kernel, call {call}"
    for l in range(3):
        for d in range(2):
            h, c = fast.state(0, 2, l, d)
            assert np.abs(h - st.h[l, d]).max() < 2e-5 and np.abs(c - st.c[l, d]).max() < 1e-4


@pytest.mark.parametrize("n_tracks", [1, 3, 8, 16, 19, 32, 35])
def test_demix_cluster_vs_simple_multi_track(umx, umxl, n_tracks):
    """Whole pipeline, several tracks in lockstep (a cluster takes up to 32 tracks as two halves, then a second group starts)."""
    gm, _ = umxl
    fast, slow = umx.Session(gm, n_tracks, N_BUF), umx.Session(gm, n_tracks, N_BUF)
    fast.set_option("lstm_impl", 0)
    slow.set_option("lstm_impl", 1)
    tracks = [synth.make_audio(20000 + 1500 * i, 40 + i) for i in range(n_tracks)]
    a, b = fast.demix(tracks), slow.demix(tracks)
    for x, y in zip(a, b):
        assert np.isfinite(x).all()
        assert np.abs(x - y).max() < 2e-5
