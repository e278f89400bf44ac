"""GPU: the configuration bench.py times — several ragged tracks in one lockstep batch, multi-segment, default options
(post stage of segment s under the recurrences of s+1, zero frames skipped, cluster recurrence, 3-term GEMMs) — against THE
REFERENCE ITSELF: oracle/_ref/umx.cpp.main.ref (the unmodified umx.cpp + src/*.cpp as its own process) demixes a 100 s synthetic
WAV with synthetic UMX-L weights: three segments (2,646,000 / 2,443,517 / 459,017 samples), two triangular cross-fades, the
sum_weight division and the trim (umx.cpp:197-273).  The segment table is compared bit-exactly with the lines the reference
prints (umx.cpp:220-222); the four stems within BASELINE's 1e-3 / 60 dB.  About 1.5 minutes of single-core reference time.
A second, CPU-free case checks the same batch against the golden file of the reference CLI at a small hidden size."""
import gzip
import os
import subprocess
import tempfile

import numpy as np
import pytest

import golden_inputs as gi
from oracle import ref, synth

pytestmark = [pytest.mark.gpu, pytest.mark.slow]


def sdr(got, want):
    return float(10 * np.log10((want.astype(np.float64) ** 2).sum() / max(((got - want).astype(np.float64) ** 2).sum(), 1e-30)))


def ragged_batch(track, seeds=(21, 22), seconds=(130, 40)):
    """The track under test in the middle of a ragged lockstep batch: a longer neighbour (4 segments) and a shorter one (1)."""
    a, b = (synth.make_audio(int(s * 44100), seed) for s, seed in zip(seconds, seeds))
    return [a, track, b]


def test_multisegment_batch_matches_reference_cli_golden(umx, have_gpu):
    """Small hidden size (validation recurrence kernel), stored output of the reference CLI."""
    assert have_gpu
    g = np.load(gi.MS_GOLDEN)
    assert umx.segment_plan(gi.MS_N) == [tuple(int(v) for v in row) for row in g["plan"]]
    model = umx.load_umx_model(gi.multiseg_model_bytes())
    sess = umx.Session(model, max_tracks=3)
    got = sess.demix(ragged_batch(gi.multiseg_audio()))[1]
    sub, want = got[:, :, g["idx"]], g["stems"]
    err, s = float(np.abs(sub - want).max()), sdr(sub, want)
    print(f"multiseg hidden {gi.MS_HIDDEN}: max-abs {err:.3e} SDR {s:.1f} dB")
    assert err < 1e-3 and s > 60.0


@pytest.mark.skipif(not os.path.exists(ref.MAIN_PATH), reason="oracle/_ref/umx.cpp.main.ref not built")
def test_multisegment_batch_matches_reference_cli_umxl(umx, have_gpu, tmp_path):
    assert have_gpu
    raw = synth.make_ggml(hidden=1024, seed=1234)
    model_gz = tmp_path / "ggml-model-umxl-u8-synth.bin.gz"
    model_gz.write_bytes(gzip.compress(raw, 1))
    audio = synth.make_audio(gi.MS_N, 7)
    wav = tmp_path / "mix.wav"
    synth.write_wav_f32(str(wav), audio)
    out_ref = tmp_path / "ref"
    with tempfile.TemporaryDirectory() as td:  # the reference's loader drops ./temp.decompressed
        r = subprocess.run([ref.MAIN_PATH, str(model_gz), str(wav), str(out_ref)], cwd=td, capture_output=True, text=True)
    # exit status unreliable (the reference's istft overruns the chunk matrix for chunks < 60 s, oracle/ref.py); the files count
    assert all((out_ref / f"target_{k}.wav").exists() for k in range(4)), r.stderr[-2000:]
    plan_ref = gi.parse_cli_plan(r.stdout)
    assert len(plan_ref) == 3
    assert umx.segment_plan(gi.MS_N) == plan_ref  # indexing: bit-exact against what the reference printed
    want = np.stack([synth.read_wav_f32(str(out_ref / f"target_{k}.wav")) for k in range(4)])

    model = umx.load_umx_model(raw)
    sess = umx.Session(model, max_tracks=3)  # default options: this is the path bench.py times
    got = sess.demix(ragged_batch(audio))[1]
    assert got.shape == want.shape == (4, 2, gi.MS_N)
    for k in range(4):
        err, s = float(np.abs(got[k] - want[k]).max()), sdr(got[k], want[k])
        print(f"multiseg UMX-L target_{k}: max-abs {err:.3e} SDR {s:.1f} dB")
        assert err < 1e-3 and s > 60.0
    # the cross-fade regions on their own (where two segments overlap): same bar
    for a, b in ((1984500 - 4033, 2646000 - 4033), (3969000 - 4033, gi.MS_N)):
        e = float(np.abs(got[:, :, a:b] - want[:, :, a:b]).max())
        print(f"  overlap [{a}, {b}): max-abs {e:.3e}")
        assert e < 1e-3
