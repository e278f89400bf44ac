"""Generates tests/golden/upstream_variants.npz from UPSTREAM Open-Unmix itself (vendor/open-unmix-pytorch in the reference tree,
imported here; PyTorch, float64): the behaviours the reference deviates from and that the session options
"stft_pad", "wiener_psd", "wiener_reg", "wiener_win_len", "wiener_niter" switch back on.
    python tests/golden/make_golden_upstream.py
Stored: torch.stft with reflect padding of golden_inputs.audio(); openunmix.filtering.wiener (0, 1 and 2 iterations)
applied per window of UP_WIN frames exactly as openunmix.model.Separator.forward does (model.py:297-311)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, "/root/reference/vendor/open-unmix-pytorch")
import golden_inputs as gi  # noqa: E402
from openunmix import filtering  # noqa: E402


def main():
    out = {}
    # --- STFT with true reflect padding: the chunk sits in a zero-extended N_BUF window like everywhere else in the path
    a = np.zeros((2, gi.N_BUF), np.float64)
    a[:, : gi.N] = gi.audio()
    win = torch.hann_window(4096, periodic=True, dtype=torch.float64)
    s = torch.stft(torch.from_numpy(a), n_fft=4096, hop_length=1024, window=win, center=True, pad_mode="reflect",
                   normalized=False, onesided=True, return_complex=True)  # (2, 2049, F)
    out["stft_reflect"] = s.permute(0, 2, 1).numpy().astype(np.complex64)

    # --- Wiener per window, upstream arithmetic
    mix, mags = gi.upstream_wiener_inputs()
    mix_t = torch.view_as_real(torch.from_numpy(mix.astype(np.complex128)).permute(1, 2, 0).contiguous())  # (F, B, C, 2)
    mags_t = torch.from_numpy(mags.astype(np.float64)).permute(2, 3, 1, 0).contiguous()                   # (F, B, C, sources)
    for it in (0, 1, 2):
        y = torch.zeros(mix_t.shape + (4,), dtype=torch.float64)
        pos = 0
        while pos < gi.UP_F:  # model.py:301-311
            t = torch.arange(pos, min(gi.UP_F, pos + gi.UP_WIN))
            pos = int(t[-1]) + 1
            y[t] = filtering.wiener(mags_t[t], mix_t[t], it, softmask=False, residual=False)
        yc = torch.view_as_complex(y.permute(4, 2, 0, 1, 3).contiguous()).numpy()  # (sources, C, F, B)
        out[f"wiener_win_niter{it}"] = yc[..., :: gi.UP_BIN_STEP].astype(np.complex64)
    np.savez_compressed(gi.UP_GOLDEN, **out)
    print({k: v.shape for k, v in out.items()}, os.path.getsize(gi.UP_GOLDEN), "bytes")


if __name__ == "__main__":
    main()
