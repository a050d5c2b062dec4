"""Pins the numpy oracle against outputs of the unmodified reference (tests/golden/step_*.npz).

fp64 vs fp64: the only differences are summation order inside BLAS / torch reductions, so the
tolerance is rtol 1e-11 (atol 1e-13).
"""
import numpy as np
import pytest

from conftest import check_against, check_means_within_standard_error, golden_names, load_golden
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


# ---- general per-entry oracle (embedding / flatten / batchnorm stacks) ------------------------------
def net_case_inputs(meta):
    from oracle import net_oracle as no
    algos = meta["expanded_algos"]
    entries = no.build_entries(meta["sizes"], algos, has_bias=bool(meta["bias"]))
    B = meta["batch"]
    if meta["tokens"]:
        x = det_labels(B * meta["tokens"], meta["sizes"][0], meta["salt"] + 3).reshape(B, meta["tokens"]).astype(np.float64)
    else:
        x = det_normalish((B, meta["sizes"][0]), meta["salt"] + 3)
    n_out = meta["sizes"][-1]
    if algos[-1] == "softmax":
        target = [[int(v)] for v in det_labels(B, n_out, meta["salt"] + 7)]
    else:
        target = det_uniform((B, n_out), meta["salt"] + 9, 0.0, 1.0).tolist()
    return entries, x, target


def net_det_params(entries, salt):
    """Same construction as oracle/make_golden.py:net_params."""
    out = []
    for k, e in enumerate(entries):
        if e.algo == "embedding":
            out.append(det_uniform(e.weights.shape, salt * 97 + k))
        elif e.algo == "linear":
            out.append((det_uniform(e.weights.shape, salt * 97 + k) * np.sqrt(3.0 / e.weights.shape[0]))
                       .astype(np.float32).astype(np.float64))
            if e.bias is not None:
                out.append(det_uniform(e.bias.shape, salt * 97 + 50 + k) * 0.1)
        elif e.algo == "batchnorm":
            out.append(1.0 + det_uniform(e.gain.shape, salt * 97 + k) * 0.25)
            out.append(det_uniform(e.bn_bias.shape, salt * 97 + 50 + k) * 0.1)
    return [np.asarray(a, dtype=np.float32).astype(np.float64) for a in out]


@pytest.mark.parametrize("name", golden_names("net_"))
def test_net_oracle_matches_reference(name):
    from oracle import net_oracle as no
    arrays, meta = load_golden(name)
    algos = meta["algos"]
    assert no.expand_algos(algos) == meta["expanded_algos"]
    entries, x, target = net_case_inputs(meta)
    assert [e.hidden for e in entries] == meta["hidden"]
    assert [e.ratio if e.algo == "flatten" else 0 for e in entries] == [r or 0 for r in meta["flatten_ratios"]]
    params = net_det_params(entries, meta["salt"])
    no.set_params(entries, params)
    assert sum(p.size for p in no.get_params(entries)) == meta["num_params"]
    p = meta.get("dropout", 0.0)
    adam = orc.AdamState()
    costs = []
    for step in range(meta["steps"]):
        masks = None
        if p > 0:
            masks = {int(k[4:].split("_")[0]): v.astype(np.float64) for k, v in arrays.items()
                     if k.startswith("mask") and k.endswith(f"_s{step}")}
        outs, into_last, tape = no.forward(entries, x, training=True, dropout_rate=p, masks=masks)
        cost = no.loss_value(entries, outs, into_last, target) + no.l2_penalty(entries, meta["l2"])
        grads, out_grads = no.backward(entries, outs, into_last, tape, target, dropout_rate=p, masks=masks,
                                       l2_lambda=meta["l2"])
        if step == 0:
            for i, o in enumerate(outs):
                check_against(arrays, f"act{i}", o, RTOL, ATOL)
            for i, g in enumerate(out_grads):
                assert (f"actgrad{i}" in arrays) == (g is not None), f"actgrad{i}"
                if g is not None:
                    check_against(arrays, f"actgrad{i}", g, 1e-10, 1e-13)
            for i, g in enumerate(grads):
                check_against(arrays, f"grad{i}", g, 1e-10, 1e-13)
        lr_t = meta["lr"] * meta["decay"] ** step
        cur = no.get_params(entries)
        new = (orc.adam_update(cur, grads, adam, lr_t) if meta["optimizer"] == "adam" else orc.sgd_update(cur, grads, lr_t))
        no.set_params(entries, new)
        costs.append(cost)
    np.testing.assert_allclose(costs, arrays["costs"], rtol=1e-10)
    for i, q in enumerate(no.get_params(entries)):
        check_against(arrays, f"param{i}", q, 1e-8, 1e-9)
    for i, e in enumerate(entries):
        if e.algo == "batchnorm":
            # the bias in front of a batchnorm has a mathematically zero gradient; Adam turns its
            # round-off noise into ~1e-11 drifts of that bias, which the running mean follows
            np.testing.assert_allclose(e.mean.reshape(-1), arrays[f"bn_mean{i}"], rtol=1e-9, atol=1e-9)
            np.testing.assert_allclose(e.variance.reshape(-1), arrays[f"bn_var{i}"], rtol=1e-10, atol=1e-13)
    outs, _, _ = no.forward(entries, x, training=False)
    check_against(arrays, "infer_out", outs[-1], 1e-9, 1e-12)
    outs1, _, _ = no.forward(entries, x[0], training=False)
    check_against(arrays, "infer_out_1d", outs1[-1], 1e-9, 1e-12)


@pytest.mark.parametrize("name", [n for n in golden_names("net_") if "dropout" not in n])
def test_torch_cpu_port_general_stacks_match_reference(name):
    import torch
    from oracle import net_oracle as no
    from oracle.torch_cpu_port import TorchCpuMLP
    arrays, meta = load_golden(name)
    entries, x, target = net_case_inputs(meta)
    params = net_det_params(entries, meta["salt"])
    ratios = [e.ratio for e in entries if e.algo == "flatten"]
    m = TorchCpuMLP(meta["expanded_algos"], params, meta["optimizer"] or "sgd", flatten_ratios=ratios)
    xt = torch.tensor(x)
    tt = (torch.tensor([t[0] for t in target], dtype=torch.int64) if meta["expanded_algos"][-1] == "softmax"
          else torch.tensor(target, dtype=torch.float64))
    costs = [m.train_step(xt, tt, meta["lr"] * meta["decay"] ** s, l2_lambda=meta["l2"]) for s in range(meta["steps"])]
    np.testing.assert_allclose(costs, arrays["costs"], rtol=1e-10)
    for i, q in enumerate(m.params):
        check_against(arrays, f"param{i}", q.detach().numpy(), 1e-8, 1e-9)


@pytest.mark.parametrize("name", golden_names("dropstat_"))
def test_oracle_dropout_statistics_match_reference(name):
    """Dropout in distribution: the reference's cost and mean |dcost/dW| over 64 F.dropout seeds
    (tests/golden/dropstat_*.npz) against the oracle's over 64 numpy-drawn keep masks placed on the same
    hidden entries -- the means must agree within 3 standard errors (the same check the device's Philox
    masks get in tests/test_gpu_round2.py)."""
    arrays, meta = load_golden(name)
    stack, params, x, target = case_inputs(meta)
    p, n = meta["dropout"], meta["seeds"]
    algos = meta["expanded_algos"]
    hidden = [i for i in range(len(algos)) if 0 < i < len(algos) - 1]
    widths = {}
    for spec, (w, _) in zip(stack.linears, orc.split_params(stack, params)):
        widths[spec.z_index] = w.shape[1]
        if spec.a_index is not None:
            widths[spec.a_index] = w.shape[1]
    rng = np.random.default_rng(7)
    costs, gabs = [], []
    for _ in range(n):
        masks = {i: (rng.random((meta["batch"], widths[i])) < 1.0 - p).astype(np.float64) for i in hidden}
        _, cost, _, grads, _ = orc.train_step(stack, params, x, target, lr=0.0, optimizer="sgd",
                                              dropout_rate=p, masks=masks)
        costs.append(cost)
        per = 2 if meta["bias"] else 1
        gabs.append([np.abs(grads[per * l]).mean() for l in range(len(stack.linears))])
    check_means_within_standard_error(np.asarray(costs), arrays["cost"], "cost")
    gabs = np.asarray(gabs)
    for l in range(gabs.shape[1]):
        check_means_within_standard_error(gabs[:, l], arrays["grad_abs_mean"][:, l], f"mean|dW{l}|")
