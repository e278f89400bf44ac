"""GPU parity at the real UMX-L size (hidden 1024, 60 s windows, 2584 frames) and at UMX-HQ's (hidden 512) against the
compiled, unmodified reference (oracle/_ref/libumxref.so, which travels to the GPU box).  Two streamed
umx_inference calls each: ~2 + 1 minutes of single-core reference time."""
import gzip
import os

import numpy as np
import pytest

from oracle import ref, synth

pytestmark = [pytest.mark.gpu, pytest.mark.slow]

N_BUF = 2646000


def sdr(got, want):
    return float(10 * np.log10((want.astype(np.float64) ** 2).sum() / max(((got - want).astype(np.float64) ** 2).sum(), 1e-30)))


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref/libumxref.so not built")
@pytest.mark.parametrize("terms", [int(os.environ.get("UMX_TEST_TERMS", "0"))])
@pytest.mark.parametrize("hidden", [1024, 512], ids=["umxl", "umxhq"])
def test_umxl_two_streamed_segments_vs_reference(umx, tmp_path, terms, hidden):
    raw = synth.make_ggml(hidden=hidden, seed=1234)
    if hidden == 1024:
        assert len(raw) == 138340380  # 172 tensors, 131.93 MB
    p = tmp_path / "umxl-synth.gz"
    p.write_bytes(gzip.compress(raw, 1))
    rm = ref.Model(str(p))
    gm = umx.load_umx_model(str(p))
    # bit-exact dequantisation at full size (one big u8 and one big u16 tensor per target)
    for t in range(4):
        assert np.array_equal(rm.tensor("fc1_w", t), gm.tensor("fc1_w", t))
        assert np.array_equal(rm.tensor("fc3_w", t), gm.tensor("fc3_w", t))
        assert np.array_equal(rm.tensor("lstm_hh_w", t, 2, 1), gm.tensor("lstm_hh_w", t, 2, 1))
    sess = umx.Session(gm, 1, N_BUF)
    if terms:
        sess.set_option("gemm_terms", terms)
    rs = ref.LstmState(hidden // 2, ref.nb_frames(N_BUF))
    audio = synth.make_audio(10 * 44100 + 18017, 7)
    for chunk in (audio[:, :441000], audio[:, 441000 - 300000:]):  # 10 s, then a different 7 s chunk (state streams)
        want = ref.umx_inference(rm, rs, chunk, N_BUF)
        got = sess.umx_inference(chunk)
        err = float(np.abs(got - want).max())
        sdrs = [sdr(got[j], want[j]) for j in range(4)]
        print(f"{'UMX-L' if hidden == 1024 else 'UMX-HQ size'} chunk {chunk.shape[1]}: max-abs {err:.3e}, SDR {['%.1f' % s for s in sdrs]}, "
              f"stems rms {np.sqrt((want ** 2).mean()):.4f}")
        assert err < 1e-3 and min(sdrs) > 60.0
    for l in range(3):
        for d in range(2):
            h, c = sess.state(0, 3, l, d)
            rh, rc = rs.get(3, l, d)
            assert np.abs(h - rh).max() < 1e-3 and np.abs(c - rc).max() < 5e-3
