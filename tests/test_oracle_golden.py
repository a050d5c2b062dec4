"""Pins the numpy oracle against outputs of the unmodified reference (tests/golden/step_*.npz).

fp64 vs fp64: the only differences are summation order inside BLAS / torch reductions, so the
tolerance is rtol 1e-11 (atol 1e-13).
"""
import numpy as np
import pytest

from conftest import check_against, golden_names, load_golden
from oracle import mlp_oracle as orc
from oracle.detgen import det_labels, det_normalish, det_params, det_uniform

RTOL, ATOL = 1e-11, 1e-13


def case_inputs(meta):
    algos = meta["expanded_algos"]
    stack = orc.parse_stack(algos)
    params = det_params(meta["sizes"], bool(meta["bias"]), meta["salt"])
    x = det_normalish((meta["batch"], meta["sizes"][0]), meta["salt"] + 3)
    if stack.loss == "ce":
        target = [[int(v)] for v in det_labels(meta["batch"], meta["sizes"][-1], meta["salt"] + 7)]
    else:
        target = det_uniform((meta["batch"], meta["sizes"][-1]), meta["salt"] + 9, 0.0, 1.0).tolist()
    return stack, params, x, target


def masks_for(arrays, step):
    sfx = "" if step == 0 else f"_s{step}"
    out = {}
    for k, v in arrays.items():
        if k.startswith("mask") and ((step == 0 and "_s" not in k) or (step > 0 and k.endswith(sfx))):
            idx = int(k[4:].split("_")[0])
            out[idx] = v.astype(np.float64)
    return out


@pytest.mark.parametrize("name", golden_names("step_"))
def test_expand_algos_matches_reference(name):
    _, meta = load_golden(name)
    algos = meta["algos"] or ["relu"] * (len(meta["sizes"]) - 1)
    assert orc.expand_algos(algos) == meta["expanded_algos"]


@pytest.mark.parametrize("name", golden_names("step_"))
def test_oracle_train_steps_match_reference(name):
    arrays, meta = load_golden(name)
    stack, params, x, target = case_inputs(meta)
    p = meta.get("dropout", 0.0)
    adam = orc.AdamState()
    costs = []
    for step in range(meta["steps"]):
        lr_t = meta["lr"] * meta["decay"] ** step
        masks = masks_for(arrays, step) if p > 0 else None
        params_new, cost, outs, grads, out_grads = orc.train_step(
            stack, params, x, target, lr=lr_t, optimizer=meta["optimizer"] or "sgd", adam=adam,
            l2_lambda=meta["l2"], dropout_rate=p, masks=masks)
        if step == 0:
            for i, o in enumerate(outs):
                check_against(arrays, f"act{i}", o, RTOL, ATOL)
            for i, g in enumerate(out_grads):
                has = (f"actgrad{i}" in arrays) or (f"actgrad{i}__sub" in arrays)
                assert has == (g is not None), f"actgrad{i}"
                if g is not None:
                    check_against(arrays, f"actgrad{i}", g, RTOL, ATOL)
            for i, g in enumerate(grads):
                check_against(arrays, f"grad{i}", g, RTOL, ATOL)
        params = params_new
        costs.append(cost)
    np.testing.assert_allclose(costs, arrays["costs"], rtol=RTOL, atol=ATOL)
    for i, q in enumerate(params):
        # Adam's m/(sqrt(v)+eps) amplifies last-bit differences of tiny gradients
        check_against(arrays, f"param{i}", q, 1e-9, 1e-12)
    if meta["optimizer"] == "adam":
        assert adam.step == int(arrays["adam_step"])
        for i in range(len(params)):
            check_against(arrays, f"adam_m{i}", adam.m[i], RTOL, ATOL)
            check_against(arrays, f"adam_v{i}", adam.v[i], RTOL, 1e-20)
    # inference on the updated parameters
    outs, into_last = orc.forward(stack, params, x, training=True)
    check_against(arrays, "infer_out", outs[-1], RTOL, ATOL)
    np.testing.assert_allclose(orc.loss_value(stack, outs, into_last, target), arrays["infer_cost"],
                               rtol=RTOL, atol=ATOL)
    o1, _ = orc.forward(stack, params, x[0])
    check_against(arrays, "infer_out_1d", o1[-1], RTOL, ATOL)
    o1t, il = orc.forward(stack, params, x[0], training=True)
    np.testing.assert_allclose(orc.loss_value(stack, o1t, il, target[0]), arrays["infer_cost_1d_t"],
                               rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize("name", [n for n in golden_names("step_") if "dropout" not in n])
def test_torch_cpu_port_matches_reference(name):
    """The torch-operator port used as the timed CPU baseline gives the reference's numbers."""
    import torch
    from oracle.torch_cpu_port import TorchCpuMLP
    arrays, meta = load_golden(name)
    stack, params, x, target = case_inputs(meta)
    m = TorchCpuMLP(meta["expanded_algos"], params, meta["optimizer"] or "sgd")
    xt = torch.tensor(x)
    tt = (torch.tensor([t[0] for t in target], dtype=torch.int64) if stack.loss == "ce"
          else torch.tensor(target, dtype=torch.float64))
    costs = [m.train_step(xt, tt, meta["lr"] * meta["decay"] ** s, l2_lambda=meta["l2"])
             for s in range(meta["steps"])]
    np.testing.assert_allclose(costs, arrays["costs"], rtol=RTOL, atol=ATOL)
    for i, q in enumerate(m.params):
        check_against(arrays, f"param{i}", q.detach().numpy(), 1e-9, 1e-12)
