"""GPU: the allocation-free Wiener entry point (umx_cuda_wiener_noalloc) against the ordinary one.

The no-alloc variant is for callers that must not call cudaMalloc while filtering: a real-time audio
thread (cudaMalloc synchronises the whole device and has no latency bound) or a plugin whose host
process has given all GPU memory to another allocator.  Its scratch is a static array inside the
library that every call shares and that keeps the previous call's contents, so this test makes TWO calls
in a row, with different inputs and different frame counts, and checks each against umx_cuda_wiener.
It lives in its own file, collected last, so a failure here cannot hide the other tests' results, and
it must pass both alone in a fresh process and after everything else has run.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NB = 2049


def _internal(x):  # (..., 2, F, B) numpy  ->  [F][2][B] order
    return np.ascontiguousarray(np.moveaxis(x, -2, -3))


def rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def _case(seed, F, scale):
    rng = np.random.default_rng(seed)
    spec = ((rng.standard_normal((2, F, NB)) + 1j * rng.standard_normal((2, F, NB))) * scale).astype(np.complex64)
    spec[:, F // 3] = 0  # a silent frame: polar(mag, arg 0) path
    mags = (rng.uniform(0, 1, (4, 2, F, NB)) * np.abs(spec)[None] + rng.uniform(0, 0.1, (4, 2, F, NB))).astype(np.float32)
    return spec, mags


def test_wiener_noalloc_matches_wiener_twice_in_a_row(umx):
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda", 0)
    # second call: fewer frames and a quieter mix, so stale partial sums or a stale max|X| would show
    for seed, F, scale in ((5, 41, 60.0), (6, 17, 0.5)):
        spec, mags = _case(seed, F, scale)
        want, _ = umx.wiener_filter(spec, mags)  # existing entry point (host buffers, reference order)
        d_mix = torch.from_numpy(_internal(spec)).to(dev)
        d_mags = torch.from_numpy(_internal(mags)).to(dev)
        d_out = torch.full((4, F, 2, NB), float("nan"), dtype=torch.complex64, device=dev)
        torch.cuda.synchronize()
        umx.wiener_noalloc(d_mix.data_ptr(), d_mags.data_ptr(), F, 1, d_out.data_ptr(), None)
        torch.cuda.synchronize()
        got = np.moveaxis(d_out.cpu().numpy(), -3, -2)  # back to (4, 2, F, B)
        assert np.isfinite(got).all()
        assert rel(got, want) < 1e-4  # the tolerance of the other Wiener parity tests
        assert torch.equal(d_mix.cpu(), torch.from_numpy(_internal(spec)))  # the mix is left untouched


def test_wiener_noalloc_two_tracks_and_limit(umx):
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda", 0)
    F = 12
    cases = [_case(9, F, 30.0), _case(10, F, 2.0)]  # different loudness: max|X| is per track
    d_mix = torch.from_numpy(np.concatenate([_internal(c[0]) for c in cases], axis=0)).to(dev)
    d_mags = torch.from_numpy(np.concatenate([_internal(c[1]) for c in cases], axis=1)).to(dev)
    d_out = torch.empty((4, 2 * F, 2, NB), dtype=torch.complex64, device=dev)
    umx.wiener_noalloc(d_mix.data_ptr(), d_mags.data_ptr(), F, 2, d_out.data_ptr(), None)
    torch.cuda.synchronize()
    got = d_out.cpu().numpy()
    for i, (spec, mags) in enumerate(cases):
        want, _ = umx.wiener_filter(spec, mags)
        assert rel(np.moveaxis(got[:, i * F:(i + 1) * F], -3, -2), want) < 1e-4
    with pytest.raises(umx.UmxCudaError, match="at most"):
        umx.wiener_noalloc(d_mix.data_ptr(), d_mags.data_ptr(), F, 3, d_out.data_ptr(), None)
    with pytest.raises(umx.UmxCudaError):
        umx.wiener_noalloc(d_mix.data_ptr(), d_mags.data_ptr(), 0, 1, d_out.data_ptr(), None)


def test_wiener_noalloc_own_stream_back_to_back(umx):
    """Two calls of different sizes queued on the caller's own stream with no host synchronisation between them: the second
    starts from the first one's scratch contents and is ordered behind it by the stream alone."""
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda", 0)
    st = torch.cuda.Stream(device=dev)
    cases = [_case(15, 33, 40.0), _case(16, 9, 0.25)]
    bufs = []
    for spec, mags in cases:
        F = spec.shape[1]
        bufs.append((torch.from_numpy(_internal(spec)).to(dev), torch.from_numpy(_internal(mags)).to(dev),
                     torch.full((4, F, 2, NB), float("nan"), dtype=torch.complex64, device=dev)))
    torch.cuda.synchronize()
    for (spec, _), (d_mix, d_mags, d_out) in zip(cases, bufs):
        umx.wiener_noalloc(d_mix.data_ptr(), d_mags.data_ptr(), spec.shape[1], 1, d_out.data_ptr(), st.cuda_stream)
    st.synchronize()
    for (spec, mags), (_, _, d_out) in zip(cases, bufs):
        want, _ = umx.wiener_filter(spec, mags)
        got = np.moveaxis(d_out.cpu().numpy(), -3, -2)
        assert np.isfinite(got).all() and rel(got, want) < 1e-4


def test_wiener_noalloc_rejects_foreign_memory_and_huge_F(umx):
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda", 0)
    spec, mags = _case(17, 4, 1.0)
    d_mix = torch.from_numpy(_internal(spec)).to(dev)
    d_mags = torch.from_numpy(_internal(mags)).to(dev)
    d_out = torch.empty((4, 4, 2, NB), dtype=torch.complex64, device=dev)
    host = np.zeros(16, np.float32)
    with pytest.raises(umx.UmxCudaError, match="device memory"):  # a host pointer where a device buffer is expected
        umx.wiener_noalloc(host.ctypes.data, d_mags.data_ptr(), 4, 1, d_out.data_ptr(), None)
    with pytest.raises(umx.UmxCudaError, match="1,048,560"):  # rejected before anything is queued
        umx.wiener_noalloc(d_mix.data_ptr(), d_mags.data_ptr(), 16 * 65535 + 1, 1, d_out.data_ptr(), None)
