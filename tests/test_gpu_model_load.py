"""GPU: load_umx_model (src/model.cpp:42-665) on damaged model files.  The reference compares whole tensor names and answers
anything else with "failed to load" (src/model.cpp:240-546); the C ABI promises a clean negative status, never a crash."""
import numpy as np
import pytest

from oracle import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def raw():
    return synth.make_ggml(hidden=16, seed=3)


@pytest.mark.parametrize("old,new,what", [
    (b"lstm.weight_ih_l1", b"lstm.weight_ih_l9", "failed to load"),   # layer index out of range
    (b"lstm.weight_hh_l2", b"lstm.weight_hh_X2", "failed to load"),   # no "_l<digit>" at all
    (b"lstm.weight_ih_l0_reverse", b"lstm.weight_ih_l0_reversf", "failed to load"),  # unknown suffix
    (b"bn2.running_var", b"bn2.running_vbr", "failed to load"),
    (b"fc1.weight", b"fc1.weighs", "failed to load"),
    (b"lstm.bias_hh_l0", b"lstm.bias_ih_l0", "appears twice"),        # 43 tensors, one of them a duplicate
])
def test_crafted_tensor_names_are_rejected(umx, raw, old, new, what):
    assert raw.count(old) >= 1 and len(old) == len(new)
    bad = raw.replace(old, new, 1)
    with pytest.raises(umx.UmxCudaError, match=what):
        umx.load_umx_model(bad)


def test_truncated_and_garbage_files(umx, raw):
    for cut in (7, 8, 40, len(raw) // 2, len(raw) - 1):
        with pytest.raises(umx.UmxCudaError):
            umx.load_umx_model(raw[:cut])
    with pytest.raises(umx.UmxCudaError, match="magic"):
        umx.load_umx_model(b"\x00" * 64)
    umx.load_umx_model(raw).close()  # the intact file still loads


def test_model_written_by_the_cpp_writer_loads_bit_exact(umx, tmp_path):
    """A model file made by the C++ ggml writer (cli/umx_make_ggml.cpp) loads, and its dequantised tensors equal the oracle's."""
    import os
    import subprocess

    from oracle import umx_oracle as O

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "umx.cpp_b200", "umx.make-ggml")
    if not os.path.exists(exe):
        pytest.skip("umx.make-ggml not built")
    out = tmp_path / "m.bin.gz"
    assert subprocess.run([exe, str(out), "32", "9"], capture_output=True).returncode == 0
    gm, om = umx.load_umx_model(str(out)), O.load_model(str(out))
    assert gm.hidden_size == 32
    for t in (0, 3):
        assert np.array_equal(gm.tensor("fc1_w", t), om.t[t]["fc1.weight"].T)
        assert np.array_equal(gm.tensor("fc3_w", t), om.t[t]["fc3.weight"].T)
        assert np.array_equal(gm.tensor("lstm_hh_w", t, 1, 1), om.t[t]["lstm.weight_hh_l1_reverse"].T)
