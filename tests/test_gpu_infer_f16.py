"""Large-batch inference on the fp16-split tensor-core kernels (two fp16 planes per operand, three MMAs per product;
csrc/poz_tc3_gemm.cuh) against the float64 oracle: the same contract as the split-bf16 kernels (rtol 1e-5 on every
output, neural_net_model.py:117-119,188), the fall-back to those kernels when an operand leaves the range in which two
fp16 planes carry fp32 precision, and the switch that turns the path off."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from oracle import mlp_oracle as orc
from oracle.detgen import det_params

pytestmark = pytest.mark.gpu

SIZES = [784, 256, 128, 10]
ALGOS = ["linear", "relu", "linear", "relu", "linear", "softmax"]
ROWS = 32768                      # >= POZ_TC3_F16_MIN_ROWS (16384): the fp16-split path is taken


def _engine(params=None):
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    params = params if params is not None else det_params(SIZES, True, 5)
    eng = DeviceMLP(SIZES, StackSpec.from_algos(ALGOS), True, "float32", 0)
    for l in range(3):
        eng.upload_params(l, params[2 * l], params[2 * l + 1])
    return eng, params


def _check(out, xd, params, rows, rtol=1e-5):
    ref = orc.forward(orc.parse_stack(ALGOS), params, xd[rows].cpu().numpy().astype(np.float64))[0][-1]
    got = out[rows].cpu().numpy().astype(np.float64).reshape(ref.shape)
    bad = np.abs(got - ref) > rtol * np.abs(ref) + 1e-7
    assert not bad.any(), (int(bad.sum()), float((np.abs(got - ref) / np.abs(ref)).max()))


def _launches(eng, xd):
    from penr_oz_neural_network_torch_b200 import launch_count
    eng.forward(xd)                                  # (the first call also makes the weight planes)
    n0 = launch_count()
    out = eng.forward(xd)
    torch.cuda.synchronize()
    return out, launch_count() - n0


@pytest.mark.parametrize("scale", [1.0, 1e-2, 3.0])
def test_fp16_split_pass_against_the_oracle(scale):
    eng, params = _engine()
    g = torch.Generator(device="cuda:0").manual_seed(7)
    xd = torch.randn((ROWS, 784), generator=g, device="cuda:0", dtype=torch.float32) * scale
    out, launches = _launches(eng, xd)
    # two fp16-split launches and the two split-bf16 launches that return at once (the flag's memset is no kernel)
    assert launches == 4
    rows = torch.arange(0, ROWS, 7, device="cuda:0")
    _check(out, xd, params, rows)
    assert torch.allclose(out.sum(1), torch.ones(ROWS, device="cuda:0"), atol=1e-5)
    eng.close()


@pytest.mark.parametrize("scale", [1e-3, 30.0, 1000.0])
def test_fp16_split_pass_over_the_input_range(scale):
    """Inputs of other magnitudes, on a stack that ends in a plain Linear (a softmax over logits in the hundreds turns
    fp32 rounding of the logits into 1e-4 relative error of the small probabilities, whatever the kernel): the suite's
    float32 tolerance, rtol 1e-5 + 1e-5 max|ref|."""
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    algos = ["linear", "relu", "linear", "relu", "linear"]
    params = det_params(SIZES, True, 5)
    eng = DeviceMLP(SIZES, StackSpec.from_algos(algos), True, "float32", 0)
    for l in range(3):
        eng.upload_params(l, params[2 * l], params[2 * l + 1])
    g = torch.Generator(device="cuda:0").manual_seed(9)
    xd = torch.randn((ROWS, 784), generator=g, device="cuda:0", dtype=torch.float32) * scale
    out, launches = _launches(eng, xd)
    assert launches == 4
    rows = torch.arange(0, ROWS, 11, device="cuda:0")
    ref = orc.forward(orc.parse_stack(algos), params, xd[rows].cpu().numpy().astype(np.float64))[0][-1]
    got = out[rows].cpu().numpy().astype(np.float64).reshape(ref.shape)
    assert np.allclose(got, ref, rtol=1e-5, atol=1e-5 * np.abs(ref).max()), float(np.abs(got - ref).max() / np.abs(ref).max())
    eng.close()


@pytest.mark.parametrize("what", ["huge", "tiny", "inf"])
def test_operands_outside_the_fp16_range_fall_back(what):
    """One row with |x| >= 2^15 (or a row whose entries are all below 2^-16, or an infinity) makes the pass repeat on the
    split-bf16 kernels: every row of the result still meets the contract."""
    eng, params = _engine()
    g = torch.Generator(device="cuda:0").manual_seed(11)
    xd = torch.randn((ROWS, 784), generator=g, device="cuda:0", dtype=torch.float32)
    if what == "huge":
        xd[12345, 17] = 1.0e5
        xd[20000] *= 4.0e4
    elif what == "tiny":
        xd[777] *= 1e-7
        xd[30000] *= 1e-9
    else:
        xd[5, 5] = float("inf")
    out = eng.forward(xd)
    rows = torch.cat([torch.arange(0, ROWS, 13, device="cuda:0"),
                      torch.tensor([5, 777, 12345, 20000, 30000], device="cuda:0")]).unique()
    ref = orc.forward(orc.parse_stack(ALGOS), params, xd[rows].cpu().numpy().astype(np.float64))[0][-1]
    got = out[rows].cpu().numpy().astype(np.float64).reshape(ref.shape)
    finite = np.isfinite(ref).all(axis=1)
    bad = np.abs(got[finite] - ref[finite]) > 1e-5 * np.abs(ref[finite]) + 1e-7
    assert not bad.any(), (what, int(bad.sum()))
    if what == "inf":           # inf * w sums to NaN, and relu keeps a NaN (torch.relu does): no finite result in either
        assert not np.isfinite(ref[~finite]).any() and not np.isfinite(got[~finite]).any()
    eng.close()


def test_weights_outside_the_fp16_range_fall_back():
    params = det_params(SIZES, True, 5)
    params = [p.copy() for p in params]
    params[0] = params[0] * 1e-6              # the whole first weight matrix below 2^-16
    eng, _ = _engine(params)
    g = torch.Generator(device="cuda:0").manual_seed(3)
    xd = torch.randn((ROWS, 784), generator=g, device="cuda:0", dtype=torch.float32)
    out = eng.forward(xd)
    _check(out, xd, params, torch.arange(0, ROWS, 17, device="cuda:0"))
    eng.close()


def test_new_parameters_reach_the_fp16_planes():
    """The fp16 planes are rebuilt after an upload (and after training steps): a second forward with other parameters
    follows them."""
    eng, params = _engine()
    g = torch.Generator(device="cuda:0").manual_seed(5)
    xd = torch.randn((ROWS, 784), generator=g, device="cuda:0", dtype=torch.float32)
    eng.forward(xd)
    params2 = det_params(SIZES, True, 99)
    for l in range(3):
        eng.upload_params(l, params2[2 * l], params2[2 * l + 1])
    out = eng.forward(xd)
    _check(out, xd, params2, torch.arange(0, ROWS, 19, device="cuda:0"))
    eng.close()


CHILD = r"""
import sys, numpy as np, torch
sys.path.insert(0, {root!r}); sys.path.insert(0, {root!r} + "/tests")
from test_gpu_infer_f16 import _engine, _check, ROWS
eng, params = _engine()
g = torch.Generator(device="cuda:0").manual_seed(7)
xd = torch.randn((ROWS, 784), generator=g, device="cuda:0", dtype=torch.float32)
from penr_oz_neural_network_torch_b200 import launch_count
eng.forward(xd)
n0 = launch_count(); out = eng.forward(xd); torch.cuda.synchronize()
print("launches", launch_count() - n0)
_check(out, xd, params, torch.arange(0, ROWS, 7, device="cuda:0"))
print("ok")
"""


def test_switch_keeps_the_split_bf16_kernels():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, POZ_TC3_F16="0")
    r = subprocess.run([sys.executable, "-c", CHILD.format(root=root)], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout[-500:] + r.stderr[-1500:]
    assert "launches 2" in r.stdout, r.stdout


def test_fp16_planes_follow_training_steps_replayed_from_a_graph():
    """train x4 (the last two replayed from the CUDA graph), infer 32,768 rows, train x4, infer again: each inference must
    see the parameters as the steps before it left them (the fp16 weight planes are rebuilt whenever a step -- launched or
    replayed -- has changed the parameters)."""
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    params = det_params(SIZES, True, 321)
    eng = DeviceMLP(SIZES, StackSpec.from_algos(ALGOS), True, "float32", 0)
    for l in range(3):
        eng.upload_params(l, params[2 * l], params[2 * l + 1])
    from oracle.detgen import det_normalish, det_labels
    xt = det_normalish((1024, 784), 322)
    yt = det_labels(1024, 10, 323)
    xtd = torch.tensor(xt, dtype=torch.float32, device="cuda:0")
    ytd = torch.tensor(yt, dtype=torch.int32, device="cuda:0")
    g = torch.Generator(device="cuda:0").manual_seed(324)
    xi = torch.randn((ROWS, 784), generator=g, device="cuda:0", dtype=torch.float32)
    rows = torch.arange(0, ROWS, 23, device="cuda:0")
    stack = orc.parse_stack(ALGOS)
    adam = orc.AdamState()
    p = params
    tgt = [[int(v)] for v in yt]
    eng.train_begin(optimizer="adam", learning_rate=0.01, decay_rate=0.9, l2_lambda=0.001, epochs=8)
    side = torch.cuda.Stream()
    step = 0
    for _round in range(2):
        with torch.cuda.stream(side):
            for _ in range(4):
                eng.train_step(xtd, ytd)
        side.synchronize()
        for _ in range(4):
            p, _c, *_ = orc.train_step(stack, p, xt, tgt, lr=0.01 * 0.9 ** step, optimizer="adam", adam=adam, l2_lambda=0.001)
            step += 1
        out = eng.forward(xi)
        ref = orc.forward(stack, p, xi[rows].cpu().numpy().astype(np.float64))[0][-1]
        got = out[rows].cpu().numpy().astype(np.float64).reshape(ref.shape)
        # the parameters themselves carry the float32 training tolerance (relu-flip outliers): compare loosely, but a stale
        # set of planes would be off by the whole of four Adam steps (~4e-2 in the weights)
        assert np.abs(got - ref).max() < 5e-3, float(np.abs(got - ref).max())
    eng.close()
