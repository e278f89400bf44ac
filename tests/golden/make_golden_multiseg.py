"""Generates tests/golden/ref_cli_multiseg.npz: the reference's own CLI (oracle/_ref/umx.cpp.main.ref = the UNMODIFIED
/root/reference/umx.cpp + src/*.cpp, oracle/Makefile) run as a process on a 100 s synthetic WAV, i.e. THREE segments
(2,646,000 / 2,443,517 / 459,017 samples) with two triangular cross-fades, the sum_weight division and the trim
(umx.cpp:99-295).  Stored: the (offset, chunk length) lines it prints (umx.cpp:220-222) and its four stems at the sample
positions of golden_inputs.multiseg_indices().  Run here (the reference tree is not on the GPU box):
    python tests/golden/make_golden_multiseg.py            (~35 s)"""
import gzip
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import ref, synth  # noqa: E402
import golden_inputs as gi  # noqa: E402


def main():
    td = tempfile.mkdtemp()
    gz, wav, out = os.path.join(td, "m.gz"), os.path.join(td, "a.wav"), os.path.join(td, "out")
    with open(gz, "wb") as f:
        f.write(gzip.compress(gi.multiseg_model_bytes(), 1))
    synth.write_wav_f32(wav, gi.multiseg_audio())
    r = subprocess.run([ref.MAIN_PATH, gz, wav, out], cwd=td, capture_output=True, text=True)
    plan = gi.parse_cli_plan(r.stdout)
    stems = np.stack([synth.read_wav_f32(os.path.join(out, f"target_{k}.wav")) for k in range(4)])
    assert stems.shape == (4, 2, gi.MS_N) and len(plan) == 3
    idx = gi.multiseg_indices()
    np.savez_compressed(gi.MS_GOLDEN, plan=np.array(plan, dtype=np.int64), idx=idx, stems=stems[:, :, idx])
    print("plan", plan, "kept", idx.size, "of", gi.MS_N, "samples;", os.path.getsize(gi.MS_GOLDEN), "bytes")


if __name__ == "__main__":
    main()
