"""TEST INFRASTRUCTURE ONLY — synthetic inputs for the parity oracle, the tests and bench.py.

The real UMX-L weights are not available offline (ggml-models/ggml-model-umxl-u8.bin.gz in the
reference tree is a git-LFS pointer), so parity runs on synthetic weights written in the exact
on-disk format the reference loader parses (src/model.cpp:93-553), as produced by
scripts/convert-umx-pth-to-ggml.py:96-160:

    int32 magic 0x756d7867 ; int32 hidden_size
    per tensor: float32 scale, float32 offset, int32 n_dims, int32 name_len,
                int32 dims[n_dims] (reversed PyTorch shape), name bytes, payload (u8 | u16, C order)

uint16 for tensors whose name contains bn2/bn3/fc2/fc3, uint8 otherwise
(convert-umx-pth-to-ggml.py:146-150).  Distributions follow SURVEY.md Appendix D.
"""
from __future__ import annotations

import gzip
import io
import struct

import numpy as np

MAGIC = 0x756D7867
NB_BINS = 2049
NB_BINS_CROPPED = 1487
SAMPLE_RATE = 44100
TARGETS = ("bass", "drums", "other", "vocals")  # file order, convert-umx-pth-to-ggml.py:104


def quantize(a: np.ndarray, qtype) -> tuple[np.ndarray, np.float32, np.float32]:
    """convert-umx-pth-to-ggml.py:13-26 (float32 arithmetic, as numpy does on float32 data)."""
    a = a.astype(np.float32)
    lo, hi = np.min(a), np.max(a)
    scale = (hi - lo) / float(np.iinfo(qtype).max - 1)
    offset = lo
    q = np.round((a - offset) / scale).astype(qtype)
    return q, np.float32(scale), np.float32(offset)


def tensor_specs(hidden: int) -> list[tuple[str, tuple[int, ...], str]]:
    """(name, PyTorch shape, distribution key) in state_dict order."""
    H = hidden
    specs: list[tuple[str, tuple[int, ...], str]] = [
        ("input_mean", (NB_BINS_CROPPED,), "in_mean"),
        ("input_scale", (NB_BINS_CROPPED,), "in_scale"),
        ("output_scale", (NB_BINS,), "out_scale"),
        ("output_mean", (NB_BINS,), "out_mean"),
        ("fc1.weight", (H, 2 * NB_BINS_CROPPED), "fc1"),
    ]

    def bn(prefix: str, n: int):
        return [
            (f"{prefix}.weight", (n,), "bn_w"),
            (f"{prefix}.bias", (n,), "bn_b"),
            (f"{prefix}.running_mean", (n,), "bn_rm"),
            (f"{prefix}.running_var", (n,), "bn_rv"),
        ]

    specs += bn("bn1", H)
    for layer in range(3):
        for suffix in ("", "_reverse"):
            specs += [
                (f"lstm.weight_ih_l{layer}{suffix}", (2 * H, H), "lstm"),
                (f"lstm.weight_hh_l{layer}{suffix}", (2 * H, H // 2), "lstm"),
                (f"lstm.bias_ih_l{layer}{suffix}", (2 * H,), "lstm"),
                (f"lstm.bias_hh_l{layer}{suffix}", (2 * H,), "lstm"),
            ]
    specs += [("fc2.weight", (H, 2 * H), "fc2")]
    specs += bn("bn2", H)
    specs += [("fc3.weight", (2 * NB_BINS, H), "fc3")]
    specs += bn("bn3", 2 * NB_BINS)
    return specs


def _draw(rng: np.random.Generator, key: str, shape, hidden: int) -> np.ndarray:
    H = hidden
    u = lambda lo, hi: rng.uniform(lo, hi, size=shape).astype(np.float32)
    if key == "in_mean":
        return u(-0.6, -0.1)
    if key == "in_scale":
        return u(0.5, 2.0)
    if key == "out_scale":
        return u(0.5, 1.5)
    if key == "out_mean":
        return u(0.2, 0.8)
    if key == "fc1":
        b = 1.0 / np.sqrt(2 * NB_BINS_CROPPED)
        return u(-b, b)
    if key == "bn_w":
        return u(0.5, 1.5)
    if key in ("bn_b", "bn_rm"):
        return u(-0.1, 0.1)
    if key == "bn_rv":
        return u(0.5, 1.5)
    if key == "lstm":
        b = 1.0 / np.sqrt(H // 2)
        return u(-b, b)
    if key == "fc2":
        b = 1.0 / np.sqrt(2 * H)
        return u(-b, b)
    if key == "fc3":
        b = 1.0 / np.sqrt(H)
        return u(-b, b)
    raise KeyError(key)


def make_ggml(hidden: int = 1024, seed: int = 1234) -> bytes:
    """Uncompressed ggml byte stream with synthetic weights (4 targets x 43 tensors)."""
    rng = np.random.default_rng(seed)
    out = io.BytesIO()
    out.write(struct.pack("<i", MAGIC))
    out.write(struct.pack("<i", hidden))
    for _target in TARGETS:
        for name, shape, key in tensor_specs(hidden):
            data = _draw(rng, key, shape, hidden)
            qtype = np.uint16 if any(s in name for s in ("bn2", "bn3", "fc2", "fc3")) else np.uint8
            q, scale, offset = quantize(data, qtype)
            nm = name.encode("utf-8")
            out.write(struct.pack("<ffii", float(scale), float(offset), len(shape), len(nm)))
            for i in range(len(shape)):
                out.write(struct.pack("<i", shape[len(shape) - 1 - i]))
            out.write(nm)
            out.write(q.tobytes(order="C"))
    return out.getvalue()


def write_ggml_gz(path: str, hidden: int = 1024, seed: int = 1234, level: int = 1) -> int:
    raw = make_ggml(hidden, seed)
    with gzip.open(path, "wb", compresslevel=level) as f:
        f.write(raw)
    return len(raw)


def make_audio(n: int, seed: int = 7) -> np.ndarray:
    """(2, n) float32 synthetic stereo track: two sines + Gaussian noise per channel."""
    rng = np.random.default_rng(seed)
    t = np.arange(n, dtype=np.float64) / SAMPLE_RATE
    out = np.empty((2, n), dtype=np.float32)
    for c, (f1, f2) in enumerate(((220.0, 1333.0), (330.0, 977.0))):
        out[c] = (
            0.3 * np.sin(2 * np.pi * f1 * t)
            + 0.2 * np.sin(2 * np.pi * f2 * t)
            + 0.1 * rng.standard_normal(n)
        ).astype(np.float32)
    return out


def write_wav_f32(path: str, audio: np.ndarray, rate: int = SAMPLE_RATE) -> None:
    """Float32 RIFF/WAVE writer (channels, n) -> interleaved."""
    ch, n = audio.shape
    data = np.ascontiguousarray(audio.T.astype("<f4")).tobytes()
    with open(path, "wb") as f:
        f.write(b"RIFF" + struct.pack("<I", 36 + len(data)) + b"WAVE")
        f.write(b"fmt " + struct.pack("<IHHIIHH", 16, 3, ch, rate, rate * ch * 4, ch * 4, 32))
        f.write(b"data" + struct.pack("<I", len(data)))
        f.write(data)


def read_wav_f32(path: str) -> np.ndarray:
    """Minimal float32/PCM16 reader -> (channels, n) float32."""
    b = open(path, "rb").read()
    assert b[:4] == b"RIFF" and b[8:12] == b"WAVE"
    pos, fmt = 12, None
    while pos + 8 <= len(b):
        cid, ln = b[pos : pos + 4], struct.unpack("<I", b[pos + 4 : pos + 8])[0]
        body = b[pos + 8 : pos + 8 + ln]
        if cid == b"fmt ":
            fmt = struct.unpack("<HHIIHH", body[:16])
        elif cid == b"data":
            tag, ch, _rate, _br, _ba, bits = fmt
            if tag == 3 and bits == 32:
                a = np.frombuffer(body, dtype="<f4")
            elif tag == 1 and bits == 16:
                a = np.frombuffer(body, dtype="<i2").astype(np.float32) / 32767.0
            else:
                raise ValueError("unsupported wav")
            return np.ascontiguousarray(a.reshape(-1, ch).T)
        pos += 8 + ln + (ln & 1)
    raise ValueError("no data chunk")
