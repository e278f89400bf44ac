"""GPU: load_umx_model (src/model.cpp:42-665) on damaged model files.  The reference compares whole tensor names and answers
anything else with "failed to load" (src/model.cpp:240-546); the C ABI promises a clean negative status, never a crash."""
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
