"""CPU: host-side logic and the C-ABI surface (no compute calls)."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

from oracle import umx_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_fft_core_and_padding_on_host(tmp_path):
    """The exact radix-16 FFT code of the kernels (csrc/fft4096.cuh), emulated thread by thread."""
    exe = tmp_path / "host_fft_check"
    subprocess.run(["/usr/bin/g++", "-O2", "-std=c++17", "-o", str(exe),
                    os.path.join(ROOT, "tests", "host", "host_fft_check.cpp")], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr


def test_library_exports_every_declared_symbol(umx):
    hdr = open(os.path.join(ROOT, "include", "umx_cuda.h")).read()
    declared = sorted(set(re.findall(r"\b(umx_cuda_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 20
    lib = ctypes.CDLL(umx.LIB_PATH)
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, f"not exported: {missing}"
    assert sorted(umx.EXPORTS) == declared


@pytest.mark.parametrize("n,shift", [(441000, 4033), (7938000, 4033), (18522000, 4033), (1, 0), (1984500, 11025),
                                     (2646000 - 22050 + 4033, 4033), (5 * 1984500, 0)])
def test_segment_plan_bit_exact(umx, n, shift):
    """umx.cpp:214-218 through the C ABI vs the oracle's integer math."""
    assert umx.segment_plan(n, shift) == O.segment_plan(O.shifted_length(n, shift))


def test_segment_plan_small_buffers(umx):
    n_buf = 16384
    stride = int(np.float32(0.75) * np.float32(n_buf))
    for n in (100, 30000, 16384 * 3):
        L = n + 22050 - 4033
        want = [(off, min(n_buf, L - off)) for off in range(0, L, stride)]
        assert umx.segment_plan(n, 4033, n_buf) == want


def test_nb_frames(umx):
    assert umx.lib().umx_cuda_nb_frames(2646000) == 2584 == O.nb_frames(2646000)
    assert umx.lib().umx_cuda_nb_frames(4096) == 5


def test_no_cpu_fallback(umx, have_gpu):
    """Without a device every compute entry point must fail loudly."""
    if have_gpu:
        pytest.skip("GPU present")
    from oracle import synth

    with pytest.raises(umx.UmxCudaError):
        umx.load_umx_model(synth.make_ggml(hidden=8))
    with pytest.raises(umx.UmxCudaError):
        umx.stft(np.zeros((2, 4096), np.float32))


def test_product_does_not_import_oracle():
    src = open(os.path.join(ROOT, "umx.cpp_b200", "__init__.py")).read()
    assert "import oracle" not in src and "from oracle" not in src
    for f in os.listdir(os.path.join(ROOT, "umx.cpp_b200", "csrc")):
        if not f.endswith((".cu", ".cuh", ".h", ".cpp")):
            continue
        assert "oracle" not in open(os.path.join(ROOT, "umx.cpp_b200", "csrc", f)).read().lower()


def _frames_reaching(n_wave: int, F: int) -> int:
    """umx.cpp_b200/csrc/common.cuh: frames_reaching (mirrored here; the oracle decides whether the bound is right)."""
    return min(F, (n_wave + 2048 + 1023) // 1024 if n_wave > 0 else 0)


@pytest.mark.parametrize("n_buf,n_wave", [(16384, 1), (16384, 5000), (16384, 6144), (16384, 6145), (16384, 14000),
                                           (16384, 14336), (16384, 14337), (16384, 16384), (32768, 18017)])
def test_frames_beyond_a_chunk_are_exactly_zero_in_the_oracle(n_buf, n_wave):
    """The product skips every frame f >= frames_reaching(chunk length) of a zero-extended window after the network and computes
    only the first of them before it.  That is only right if the oracle's STFT of such a window is EXACTLY zero there (so the
    masks cannot matter in the Wiener sums) and the bound is tight."""
    rng = np.random.default_rng(n_wave)
    audio = rng.uniform(0.2, 1.0, size=(2, n_wave)).astype(np.float32)  # no zero samples inside the chunk
    spec = O.stft(audio, n_buf)  # (2, F, 2049), buffers sized for n_buf, chunk zero-extended
    F = spec.shape[1]
    fz = _frames_reaching(n_wave, F)
    assert np.all(spec[:, fz:, :] == 0)
    # tight up to the window's zero first tap: at most one frame more than strictly necessary is computed
    last_nonzero = max(f for f in range(F) if np.any(spec[:, f, :] != 0))
    assert fz - 1 - last_nonzero in (0, 1)


def test_cpp_ggml_writer_round_trip(tmp_path):
    """cli/umx_ggml_writer.hpp (the C++ counterpart of scripts/convert-umx-pth-to-ggml.py:96-160): the file it writes parses
    with the oracle's reader (= src/model.cpp:93-553 restated), has the 4 x 43 tensors in state_dict order with the reference's
    shapes and payload widths, and dequantises to within half a quantisation step of the fp32 values it was given."""
    from oracle import ref, synth

    exe = tmp_path / "umx.make-ggml"
    subprocess.run(["/usr/bin/g++", "-O2", "-std=c++17", "-o", str(exe), os.path.join(ROOT, "cli", "umx_make_ggml.cpp"), "-lz"], check=True)
    out, dump = tmp_path / "m.bin.gz", tmp_path / "m.f32"
    r = subprocess.run([str(exe), str(out), "16", "5", "--dump-f32", str(dump)], capture_output=True, text=True)
    assert r.returncode == 0 and "172 tensors" in r.stdout, r.stdout + r.stderr
    hidden, targets = O.parse_ggml(O.gzip.decompress(out.read_bytes()))
    assert hidden == 16 and len(targets) == 4 and all(len(t) == 43 for t in targets)
    f32 = np.fromfile(dump, dtype=np.float32)
    pos = 0
    for tensors in targets:
        for qt, (name, shape, _) in zip(tensors, synth.tensor_specs(16)):
            assert qt.name == name and tuple(qt.dims) == tuple(reversed(shape))
            assert qt.q.dtype == (np.uint16 if any(k in name for k in ("bn2", "bn3", "fc2", "fc3")) else np.uint8)
            n = int(np.prod(shape))
            want = f32[pos:pos + n].reshape(shape)
            pos += n
            assert np.abs(qt.dequant().reshape(shape) - want).max() <= 0.5001 * float(qt.scale) + 1e-7
    assert pos == f32.size
    if ref.available():  # the compiled reference's own loader accepts it too
        rm = ref.Model(str(out))
        assert rm.hidden == 16
        assert np.array_equal(rm.tensor("fc1_w", 1), targets[1][4].dequant().reshape(16, 2974).T)


def test_header_is_plain_c_and_links_from_c(umx, tmp_path):
    """include/umx_cuda.h compiles as C99 with strict warnings, and a C program linked against libumx_cuda.so gets the reference's
    segment table (umx.cpp:214-218: 5 segments for 3 minutes, the last one 18,017 samples) through the ABI."""
    inc = os.path.join(ROOT, "include")
    subprocess.run(["/usr/bin/gcc", "-std=c99", "-pedantic", "-Wall", "-Wextra", "-Wstrict-prototypes", "-Werror", "-fsyntax-only",
                    "-x", "c", os.path.join(inc, "umx_cuda.h")], check=True)
    exe = tmp_path / "abi_from_c"
    libdir = os.path.dirname(umx.LIB_PATH)
    subprocess.run(["/usr/bin/gcc", "-std=c99", "-Wall", "-Werror", "-I", inc, "-o", str(exe), os.path.join(ROOT, "tests", "host", "abi_from_c.c"),
                    "-L", libdir, "-l:libumx_cuda.so", f"-Wl,-rpath,{libdir}"], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split("\n")
    assert out[0] == "frames 2584 segments 5"
    rows = [tuple(int(x) for x in l.split()) for l in out[1:] if l]
    total = 7938000 + 22050 - 4033  # umx.cpp:120-121
    assert rows == [(k * 1984500, min(2646000, total - k * 1984500)) for k in range(5)] and rows[-1][1] == 18017
