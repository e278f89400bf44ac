"""GPU: the CLI contract (umx.cpp:26-97) end to end.  The reference's own, unmodified CLI (oracle/_ref/umx.cpp.main.ref, its own
process) demixes a 32 s synthetic WAV with synthetic UMX-L weights; the same files go through
(a) our C++ CLI (cli/umx_main.cpp over the C ABI) and (b) the reference's unmodified driver linked against the
drop-in shim (shim/umx_dropin.cpp) instead of its own library.  target_0..3.wav must agree within the
BASELINE tolerances.  ~1 minute of single-core reference time."""
import gzip
import os
import subprocess

import numpy as np
import pytest

from oracle import ref, synth

pytestmark = [pytest.mark.gpu, pytest.mark.slow]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "umx.cpp_b200", "umx.cpp.main")
DROPIN = os.path.join(ROOT, "oracle", "_ref", "umx.cpp.main.dropin")


def sdr(got, want):
    return float(10 * np.log10((want.astype(np.float64) ** 2).sum() / max(((got - want).astype(np.float64) ** 2).sum(), 1e-30)))


def test_cli_usage_and_load_errors(tmp_path):
    r = subprocess.run([CLI], capture_output=True, text=True)
    assert r.returncode == 1 and "Usage" in r.stderr  # umx.cpp:28-33
    wav = tmp_path / "a.wav"
    synth.write_wav_f32(str(wav), synth.make_audio(44100, 1))
    r = subprocess.run([CLI, str(tmp_path / "missing.gz"), str(wav), str(tmp_path / "o")], capture_output=True, text=True)
    assert r.returncode == 1 and "Error loading model" in r.stderr  # umx.cpp:66-70
    bad = tmp_path / "b.wav"
    synth.write_wav_f32(str(bad), synth.make_audio(8000, 1), rate=8000)
    r = subprocess.run([CLI, str(tmp_path / "missing.gz"), str(bad), str(tmp_path / "o")], capture_output=True, text=True)
    assert r.returncode == 1 and "sample rate" in r.stderr  # src/dsp.cpp:27-33


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref/libumxref.so not built")
def test_cli_and_dropin_match_reference_cli(tmp_path):
    model = tmp_path / "ggml-model-umxl-u8-synth.bin.gz"
    model.write_bytes(gzip.compress(synth.make_ggml(hidden=1024, seed=1234), 1))
    wav = tmp_path / "mix.wav"
    audio = synth.make_audio(32 * 44100, 7)  # >= 30 s: shorter tracks overflow a buffer in the reference CLI
    synth.write_wav_f32(str(wav), audio)
    out_ref = tmp_path / "ref"
    ref.main(str(model), str(wav), str(out_ref))  # exit status unreliable (heap overrun in the reference), files count
    assert all((out_ref / f"target_{k}.wav").exists() for k in range(4))
    want = [synth.read_wav_f32(str(out_ref / f"target_{k}.wav")) for k in range(4)]
    runs = {"cli": [CLI, str(model), str(wav), str(tmp_path / "cli")]}
    if os.path.exists(DROPIN):
        runs["dropin"] = [DROPIN, str(model), str(wav), str(tmp_path / "dropin")]
    for name, cmd in runs.items():
        r = subprocess.run(cmd, capture_output=True, text=True, cwd=str(tmp_path))
        assert r.returncode == 0, r.stderr[-2000:]
        for k in range(4):
            got = synth.read_wav_f32(os.path.join(cmd[3], f"target_{k}.wav"))
            assert got.shape == want[k].shape == (2, audio.shape[1])
            err, s = float(np.abs(got - want[k]).max()), sdr(got, want[k])
            print(f"{name} target_{k}: max-abs {err:.3e} SDR {s:.1f} dB")
            assert err < 1e-3 and s > 60.0
