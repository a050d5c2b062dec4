"""poz_forward_noalloc -- the entry point for callers that may not allocate device memory (their
pool already owns the GPU, or they cannot afford an allocation's stall) -- must return what the
context-based poz_forward returns for the same float32 parameters.

Collected last on purpose (file name), so a run that stops at the first failure has already
exercised everything else.  Passes on its own in a fresh process as well as after the other tests:
it builds every input itself and compares two entry points of the same library.
"""
import numpy as np
import pytest
import torch

from oracle import mlp_oracle as orc
from oracle.detgen import det_normalish, det_params

pytestmark = pytest.mark.gpu

RTOL, ATOL_REL = 1e-5, 1e-5      # the float32 tolerance of tests/test_gpu_parity.py


def _pair(sizes, algos, batch, salt, has_bias=True):
    """(output of poz_forward on a context, output of poz_forward_noalloc) for the same model."""
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    from penr_oz_neural_network_torch_b200.engine import forward_noalloc
    spec = StackSpec.from_algos(algos)
    params = det_params(sizes, has_bias, salt)
    per = 2 if has_bias else 1
    eng = DeviceMLP(sizes, spec, has_bias, "float32", 0)
    for l in range(spec.n_linear):
        eng.upload_params(l, params[per * l], params[per * l + 1] if has_bias else None)
    x = torch.tensor(det_normalish((batch, sizes[0]), salt + 1), dtype=torch.float32, device="cuda:0")
    want = eng.forward(x)
    ws = [torch.tensor(params[per * l], dtype=torch.float32, device="cuda:0") for l in range(spec.n_linear)]
    bs = ([torch.tensor(params[per * l + 1], dtype=torch.float32, device="cuda:0")
           for l in range(spec.n_linear)] if has_bias else None)
    out = torch.full((batch, sizes[-1]), float("nan"), dtype=torch.float32, device="cuda:0")
    torch.cuda.synchronize()
    got = forward_noalloc(sizes, spec.acts, ws, bs, x, out)
    torch.cuda.synchronize()
    eng.close()
    # and directly against the float64 oracle (neural_net_model.py:384-396 restated), not only against the
    # library's other entry point
    ref = orc.forward(orc.parse_stack(algos), params, x.cpu().numpy().astype(np.float64))[0][-1].reshape(batch, -1)
    res = got.cpu().numpy().astype(np.float64)
    bad = np.abs(res - ref) > 1e-5 * np.abs(ref) + 1e-7
    assert not bad.any(), (sizes, int(bad.sum()))
    return want.cpu().numpy().astype(np.float64), res


def _check(want, got):
    assert np.isfinite(got).all()
    np.testing.assert_allclose(got, want, rtol=RTOL, atol=ATOL_REL * np.abs(want).max())


def test_two_calls_in_a_row_match_the_context_entry_point():
    # first call: wide hidden layers, several row slices through the shared scratch
    want, got = _pair([784, 256, 128, 10], ["linear", "relu", "linear", "relu", "linear", "softmax"], 9000, 31)
    _check(want, got)
    assert (want.argmax(1) == got.argmax(1)).all()
    # second call straight after, smaller and differently shaped: whatever the first call left in
    # the scratch must not leak into it
    want, got = _pair([20, 33, 7, 5], ["linear", "tanh", "linear", "sigmoid", "linear"], 257, 32)
    _check(want, got)
    # and a third without biases / with a single layer
    want, got = _pair([12, 6], ["linear", "sigmoid"], 3, 33, has_bias=False)
    _check(want, got)


def test_sizes_beyond_the_fixed_scratch_are_refused():
    from penr_oz_neural_network_torch_b200 import PozError
    from penr_oz_neural_network_torch_b200 import _native as nat
    from penr_oz_neural_network_torch_b200.engine import forward_noalloc
    wide = nat.NOALLOC_MAX_WIDTH + 1
    w = [torch.zeros((4, wide), device="cuda:0"), torch.zeros((wide, 2), device="cuda:0")]
    x = torch.zeros((1, 4), device="cuda:0")
    out = torch.zeros((1, 2), device="cuda:0")
    with pytest.raises(PozError) as e:
        forward_noalloc([4, wide, 2], ["relu", "none"], w, None, x, out)
    assert "POZ_NOALLOC_MAX_WIDTH" in str(e.value)
