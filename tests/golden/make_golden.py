"""Generates tests/golden/ref_small.npz by running the compiled, UNMODIFIED reference
(oracle/_ref/libumxref.so, built from /root/reference by oracle/Makefile) on seeded synthetic inputs.
Run here (the reference tree does not exist on the GPU box):  python tests/golden/make_golden.py
All inputs are regenerated from seeds by tests/golden_inputs.py, so only outputs are stored."""
import os
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
    gz = os.path.join(td, "m.gz")
    with open(gz, "wb") as f:
        import gzip
        f.write(gzip.compress(gi.model_bytes(), 1))
    rm = ref.Model(gz)
    audio = gi.audio()
    out = {}
    out["spec"] = ref.stft(audio, gi.N_BUF)
    out["istft"] = ref.istft(out["spec"], gi.N_BUF, gi.N)
    st = ref.LstmState(gi.HIDDEN // 2, gi.F)
    x = gi.lstm_input()
    out["lstm_call1"] = ref.lstm_forward(rm, st, 1, x)
    out["lstm_call2"] = ref.lstm_forward(rm, st, 1, x)
    h, c = st.get(1, 2, 1)
    out["lstm_h_l2_bwd"], out["lstm_c_l2_bwd"] = h, c
    mags = gi.wiener_mags(out["spec"])
    y, mix_after = ref.wiener(out["spec"], mags)
    out["wiener_y_sub"] = y[..., :: gi.WIENER_BIN_STEP]
    out["wiener_mix_sub"] = mix_after[..., :: gi.WIENER_BIN_STEP]
    st2 = ref.LstmState(gi.HIDDEN // 2, gi.F)
    out["inference_call1"] = ref.umx_inference(rm, st2, audio, gi.N_BUF)
    out["inference_call2"] = ref.umx_inference(rm, st2, audio[:, : gi.N // 2], gi.N_BUF)
    for name in ("fc1_w", "fc3_w", "bn2_rv", "input_mean"):
        out["w_" + name] = rm.tensor(name, 2)[:64, :64] if name.endswith("_w") else rm.tensor(name, 2)
    out["w_lstm_hh_w"] = rm.tensor("lstm_hh_w", 3, 1, 1)
    # gtest DSP_STFT.STFTRoundtripRandWaveform (test/test_dsp.cpp:41-80): 4096 rand()/RAND_MAX samples, L == R
    rnd = gi.gtest_rand_waveform()
    s = ref.stft(rnd, 4096)
    out["gtest_rand_spec"] = s
    out["gtest_rand_roundtrip"] = ref.istft(s, 4096, 4096)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ref_small.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
