"""Generate tests/golden/*.npz by running the UNMODIFIED reference in the build container.

TEST INFRASTRUCTURE ONLY.  Run once, here, with ``/root/reference`` mounted:

    python oracle/make_golden.py

The reference is imported from where it lies (nothing is copied).  Each fixture stores the outputs
of the reference's own methods -- ``NeuralNetworkModel._forward`` (neural_net_model.py:384),
``compute_output`` (:371), torch autograd ``backward`` (:495), ``optimizer.step`` (:497) / the SGD
line (:499-500) and the whole list-level ``train`` (:410) -- for deterministic inputs built by
``oracle/detgen.py``, so the tests can rebuild the inputs without storing them.
"""
from __future__ import annotations

import json
import os
import sys
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

from oracle.detgen import det_labels, det_normalish, det_params, det_uniform  # noqa: E402
from oracle.golden_cases import STEP_CASES, TRAIN_API_CASES, INIT_CASES, NET_CASES, DROPSTAT_CASES  # noqa: E402

import neural_net_model as ref  # noqa: E402  (the reference, unmodified)

OUT = os.path.join(ROOT, "tests", "golden")


def build_ref_model(case) -> "ref.NeuralNetworkModel":
    m = ref.NeuralNetworkModel("golden", case["sizes"], "xavier", case["bias"], case["algos"],
                               case["optimizer"])
    n_lin = sum(1 for a in m.algos if a == "linear")
    sizes = case["sizes"]
    assert n_lin == len(sizes) - 1, (m.algos, sizes)
    params = det_params(sizes, bool(case["bias"]), case["salt"])
    it = iter(params)
    for layer in m.layers:
        if layer.algo == "linear":
            st = {"params": [next(it).tolist()] + ([next(it).tolist()] if case["bias"] else [])}
            layer.state_dict = st           # the reference's own setter (neural_net_model.py:39-45)
    if case["optimizer"] == "adam":
        m.optimizer = torch.optim.Adam(m.params)   # as deserialize does (neural_net_model.py:353)
    return m


def make_target(case, stack_last_is_softmax: bool):
    b, out = case["batch"], case["sizes"][-1]
    if stack_last_is_softmax:
        return [[int(v)] for v in det_labels(b, out, case["salt"] + 7)]
    return det_uniform((b, out), case["salt"] + 9, 0.0, 1.0).tolist()


class MaskRecorder:
    """Replaces torch.nn.functional.dropout by its definition (mask ~ Bernoulli(1-p) drawn with
    ``empty_like(x).bernoulli_(1-p)``, result ``x*mask/(1-p)``) so the masks can be stored.  The
    generator asserts that, on the same seed, the substituted forward reproduces the stock
    forward bit-for-bit before any fixture with dropout is written."""

    def __init__(self):
        self.masks = []

    def __call__(self, x, p=0.5, training=True, inplace=False):
        if p == 0.0:
            return x
        mask = torch.empty_like(x).bernoulli_(1.0 - p)
        self.masks.append(mask.detach().clone())
        return x * mask.div(1.0 - p)


def run_step_case(case):
    m = build_ref_model(case)
    algos = list(m.algos)
    is_ce = algos[-1] == "softmax"
    x = det_normalish((case["batch"], case["sizes"][0]), case["salt"] + 3)
    target = make_target(case, is_ce)
    xt = torch.tensor(x, dtype=torch.float64)
    p = case.get("dropout", 0.0)
    arrays = {}
    costs = []
    rec = None
    for step in range(case["steps"]):
        lr_t = case["lr"] * (case["decay"] ** step)              # neural_net_model.py:464
        if m.optimizer is not None:
            for g in m.optimizer.param_groups:
                g["lr"] = lr_t                                   # neural_net_model.py:467-469
        for q in m.params:
            q.requires_grad_()                                   # :477-478
        if p > 0.0:
            seed = 1234 + step
            torch.manual_seed(seed)
            stock_acts, stock_cost = m._forward(xt, target, p)
            rec = MaskRecorder()
            saved = torch.nn.functional.dropout
            torch.nn.functional.dropout = rec
            try:
                torch.manual_seed(seed)
                acts, cost = m._forward(xt, target, p)
            finally:
                torch.nn.functional.dropout = saved
            for a, b in zip(stock_acts, acts):
                assert torch.equal(a, b), "mask recorder diverged from stock F.dropout"
            assert torch.equal(stock_cost, cost)
        else:
            acts, cost = m._forward(xt, target, p)               # :480
        if case["l2"] > 0.0:
            cost = cost + case["l2"] * sum((w ** 2).sum() for w in m.weights)   # :482-483
        for q in m.params:
            q.grad = None                                        # :485-486
        for a in acts:
            a.retain_grad()                                      # :491-493
        cost.backward()                                          # :495
        if step == 0:
            for i, a in enumerate(acts):
                arrays[f"act{i}"] = a.detach().numpy().copy()
                if a.grad is not None:
                    arrays[f"actgrad{i}"] = a.grad.numpy().copy()
            for i, q in enumerate(m.params):
                arrays[f"grad{i}"] = q.grad.numpy().copy()
            if rec is not None:
                hidden_idx = [i for i, l in enumerate(m.layers) if l.hidden]
                assert len(hidden_idx) == len(rec.masks)
                for i, mk in zip(hidden_idx, rec.masks):
                    arrays[f"mask{i}"] = mk.numpy().astype(np.uint8)
        elif p > 0.0:
            hidden_idx = [i for i, l in enumerate(m.layers) if l.hidden]
            for i, mk in zip(hidden_idx, rec.masks):
                arrays[f"mask{i}_s{step}"] = mk.numpy().astype(np.uint8)
        if m.optimizer is not None:
            m.optimizer.step()                                   # :497
        else:
            for q in m.params:
                q.data -= lr_t * q.grad                          # :499-500
        costs.append(cost.item())                                # :503
    arrays["costs"] = np.asarray(costs)
    for i, q in enumerate(m.params):
        arrays[f"param{i}"] = q.detach().numpy().copy()
    if m.optimizer is not None:
        sd = m.optimizer.state_dict()["state"]
        for i in sorted(sd):
            arrays[f"adam_m{i}"] = sd[i]["exp_avg"].numpy().copy()
            arrays[f"adam_v{i}"] = sd[i]["exp_avg_sq"].numpy().copy()
        arrays["adam_step"] = np.asarray(float(sd[0]["step"]))
    # inference through the list-level API on the *updated* model (neural_net_model.py:371-382)
    out, c = m.compute_output(x.tolist(), target)
    arrays["infer_out"] = np.asarray(out)
    arrays["infer_cost"] = np.asarray(c)
    out1, c1 = m.compute_output(x[0].tolist(), None)
    assert c1 is None
    arrays["infer_out_1d"] = np.asarray(out1)
    out1t, c1t = m.compute_output(x[0].tolist(), target[0])
    arrays["infer_out_1d_t"] = np.asarray(out1t)
    arrays["infer_cost_1d_t"] = np.asarray(c1t)
    sub = case.get("subsample")
    if sub:
        for k in list(arrays):
            if arrays[k].size > 4096 and not k.startswith("mask"):
                arrays[k + "__sub"] = arrays[k].reshape(-1)[::sub].copy()
                arrays[k + "__sum"] = np.asarray(arrays[k].sum())
                del arrays[k]
    meta = dict(case)
    meta["expanded_algos"] = algos
    return arrays, meta


def run_train_api_case(case, workdir):
    """Whole list-level train() (neural_net_model.py:410-530) with dropout 0 on a fixed seed."""
    os.chdir(workdir)
    m = build_ref_model(case)
    algos = list(m.algos)
    is_ce = algos[-1] == "softmax"
    n = m.training_buffer_size + case.get("extra", 0)
    xs = det_uniform((n, case["sizes"][0]), case["salt"] + 3).tolist()
    if is_ce:
        ts = [[int(v)] for v in det_labels(n, case["sizes"][-1], case["salt"] + 7)]
    else:
        ts = det_uniform((n, case["sizes"][-1]), case["salt"] + 9, 0.0, 1.0).tolist()
    data = list(zip(xs, ts))
    torch.manual_seed(case["seed"])
    m.train(data, epochs=case["epochs"], learning_rate=case["lr"], batch_size=case["batch"],
            decay_rate=case["decay"], dropout_rate=0.0, l2_lambda=case["l2"])
    arrays = {}
    for i, q in enumerate(m.params):
        arrays[f"param{i}"] = q.detach().numpy().copy()
    arrays["progress_cost"] = np.asarray([pp["cost"] for pp in m.progress])
    arrays["progress_epoch"] = np.asarray([pp["epoch"] for pp in m.progress])
    wur = [[v for v in pp["weight_upd_ratio"] if v is not None] for pp in m.progress]
    arrays["progress_wur"] = np.asarray(wur)
    arrays["avg_cost"] = np.asarray(m.avg_cost)
    st = m.stats
    arrays["stats_act_mean"] = np.asarray([l["activation"]["mean"] for l in st["layers"]])
    arrays["stats_act_std"] = np.asarray([l["activation"]["std"] for l in st["layers"]])
    arrays["stats_act_sat"] = np.asarray([l["activation"]["saturated"] for l in st["layers"]])
    arrays["stats_grad_mean"] = np.asarray([np.nan if l["gradient"] is None else l["gradient"]["mean"]
                                            for l in st["layers"]])
    arrays["stats_grad_std"] = np.asarray([np.nan if l["gradient"] is None else l["gradient"]["std"]
                                           for l in st["layers"]])
    arrays["stats_wgrad_mean"] = np.asarray([w["gradient"]["mean"] for w in st["weights"] if w])
    arrays["stats_wgrad_std"] = np.asarray([w["gradient"]["std"] for w in st["weights"] if w])
    arrays["stats_hist_y0"] = np.asarray(st["layers"][0]["activation"]["histogram"]["y"])
    arrays["stats_hist_x0"] = np.asarray(st["layers"][0]["activation"]["histogram"]["x"])
    if m.optimizer is not None:
        sd = m.optimizer.state_dict()["state"]
        arrays["adam_step"] = np.asarray(float(sd[0]["step"]))
        arrays["adam_m0"] = sd[0]["exp_avg"].numpy().copy()
    meta = dict(case)
    meta["expanded_algos"] = algos
    meta["status"] = m.status
    meta["n_data"] = n
    return arrays, meta


def run_init_case(case):
    """Construction on a fixed torch seed (neural_net_model.py:190-290): the drop-in class must
    draw the same initial parameters."""
    torch.manual_seed(case["seed"])
    m = ref.NeuralNetworkModel("golden", case["sizes"], case["weight_algo"], case["bias"],
                               case["algos"], case["optimizer"], confidence=case["confidence"])
    arrays = {f"param{i}": q.detach().numpy().copy() for i, q in enumerate(m.params)}
    meta = dict(case)
    meta["expanded_algos"] = list(m.algos)
    meta["hidden"] = [bool(l.hidden) for l in m.layers]
    meta["num_params"] = m.num_params
    return arrays, meta


def net_params(m, salt):
    """Deterministic float32-exact parameters for every entry of a reference model, in `params` order."""
    out = []
    for k, layer in enumerate(m.layers):
        if layer.algo == "embedding":
            out.append(det_uniform(tuple(layer.weights.shape), salt * 97 + k))
        elif layer.algo == "linear":
            fan_in = layer.weights.shape[0]
            out.append((det_uniform(tuple(layer.weights.shape), salt * 97 + k) * np.sqrt(3.0 / fan_in))
                       .astype(np.float32).astype(np.float64))
            if layer.bias is not None:
                out.append(det_uniform(tuple(layer.bias.shape), salt * 97 + 50 + k) * 0.1)
        elif layer.algo == "batchnorm":
            out.append(1.0 + det_uniform(tuple(layer.gain.shape), salt * 97 + k) * 0.25)
            out.append(det_uniform(tuple(layer.bias.shape), salt * 97 + 50 + k) * 0.1)
    return [np.asarray(a, dtype=np.float32).astype(np.float64) for a in out]


def run_net_case(case):
    """Stacks with embedding / flatten / batchnorm entries, driven through the reference's own methods."""
    m = ref.NeuralNetworkModel("golden", case["sizes"], "xavier", case["bias"], case["algos"], case["optimizer"])
    params = net_params(m, case["salt"])
    it = iter(params)
    for layer in m.layers:
        if layer.algo == "batchnorm":
            layer.state_dict = {"params": [next(it).tolist(), next(it).tolist()], "eps": layer.eps,
                                "momentum": layer.momentum}
        elif layer.algo == "embedding":
            layer.state_dict = {"params": [next(it).tolist()]}
        elif layer.algo == "linear":
            layer.state_dict = {"params": [next(it).tolist()] + ([next(it).tolist()] if layer.bias is not None else [])}
    if case["optimizer"] == "adam":
        m.optimizer = torch.optim.Adam(m.params)
    algos = list(m.algos)
    is_ce = algos[-1] == "softmax"
    B = case["batch"]
    if case["tokens"]:
        x = det_labels(B * case["tokens"], case["sizes"][0], case["salt"] + 3).reshape(B, case["tokens"]).astype(np.float64)
    else:
        x = det_normalish((B, case["sizes"][0]), case["salt"] + 3)
    n_out = case["sizes"][-1]
    target = ([[int(v)] for v in det_labels(B, n_out, case["salt"] + 7)] if is_ce
              else det_uniform((B, n_out), case["salt"] + 9, 0.0, 1.0).tolist())
    xt = torch.tensor(x, dtype=torch.float64)
    p = case.get("dropout", 0.0)
    arrays, costs = {}, []
    for step in range(case["steps"]):
        lr_t = case["lr"] * (case["decay"] ** step)
        if m.optimizer is not None:
            for g in m.optimizer.param_groups:
                g["lr"] = lr_t
        for q in m.params:
            q.requires_grad_()
        rec = None
        if p > 0.0:
            rec = MaskRecorder()
            saved = torch.nn.functional.dropout
            torch.nn.functional.dropout = rec
            try:
                torch.manual_seed(4321 + step)
                acts, cost = m._forward(xt, target, p)
            finally:
                torch.nn.functional.dropout = saved
        else:
            acts, cost = m._forward(xt, target, p)
        if case["l2"] > 0.0:
            cost = cost + case["l2"] * sum((w ** 2).sum() for w in m.weights)
        for q in m.params:
            q.grad = None
        for a in acts:
            a.retain_grad()
        cost.backward()
        if step == 0:
            for i, a in enumerate(acts):
                arrays[f"act{i}"] = a.detach().numpy().copy()
                if a.grad is not None:
                    arrays[f"actgrad{i}"] = a.grad.numpy().copy()
            for i, q in enumerate(m.params):
                arrays[f"grad{i}"] = q.grad.numpy().copy()
        if rec is not None:
            hidden_idx = [i for i, l in enumerate(m.layers) if l.hidden]
            assert len(hidden_idx) == len(rec.masks)
            for i, mk in zip(hidden_idx, rec.masks):
                arrays[f"mask{i}_s{step}"] = mk.numpy().astype(np.uint8)
        if m.optimizer is not None:
            m.optimizer.step()
        else:
            for q in m.params:
                q.data -= lr_t * q.grad
        costs.append(cost.item())
    arrays["costs"] = np.asarray(costs)
    for i, q in enumerate(m.params):
        arrays[f"param{i}"] = q.detach().numpy().copy()
    for i, layer in enumerate(m.layers):
        if layer.algo == "batchnorm":
            arrays[f"bn_mean{i}"] = layer.mean.detach().numpy().reshape(-1).copy()
            arrays[f"bn_var{i}"] = layer.variance.detach().numpy().reshape(-1).copy()
    if m.optimizer is not None:
        sd = m.optimizer.state_dict()["state"]
        for i in sorted(sd):
            arrays[f"adam_m{i}"] = sd[i]["exp_avg"].numpy().copy()
            arrays[f"adam_v{i}"] = sd[i]["exp_avg_sq"].numpy().copy()
    # inference on the updated model: without a target batchnorm uses its running statistics (:167-168)
    out, c = m.compute_output(x.tolist(), None)
    arrays["infer_out"] = np.asarray(out)
    out1, _ = m.compute_output(x[0].tolist(), None)
    arrays["infer_out_1d"] = np.asarray(out1)
    meta = dict(case)
    meta["expanded_algos"] = algos
    meta["hidden"] = [bool(l.hidden) for l in m.layers]
    meta["flatten_ratios"] = [getattr(l, "ratio", 0) for l in m.layers]
    meta["num_params"] = m.num_params
    return arrays, meta


def run_dropstat_case(case):
    """Cost and mean |dcost/dW| of ONE step of the stock reference (its own F.dropout, neural_net_model.py:395)
    for `seeds` different torch seeds; parameters are not updated between the draws."""
    m = build_ref_model(case)
    x = torch.tensor(det_normalish((case["batch"], case["sizes"][0]), case["salt"] + 3), dtype=torch.float64)
    target = make_target(case, m.algos[-1] == "softmax")
    costs, gabs = [], []
    for seed in range(case["seeds"]):
        torch.manual_seed(1000 + seed)
        for q in m.params:
            q.requires_grad_()
            q.grad = None
        _, cost = m._forward(x, target, case["dropout"])
        cost.backward()
        costs.append(cost.item())
        gabs.append([w.grad.abs().mean().item() for w in m.weights])
    meta = dict(case)
    meta["expanded_algos"] = list(m.algos)
    return {"cost": np.asarray(costs), "grad_abs_mean": np.asarray(gabs)}, meta


def save(name, arrays, meta):
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, __meta__=np.asarray(json.dumps(meta)), **arrays)
    print(f"{name}: {os.path.getsize(path) / 1024:.1f} KiB, {len(arrays)} arrays")


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(1)     # fixed reduction order inside the reference's torch kernels
    for case in STEP_CASES:
        arrays, meta = run_step_case(case)
        save("step_" + case["name"], arrays, meta)
    with tempfile.TemporaryDirectory() as tmp:
        cwd = os.getcwd()
        try:
            for case in TRAIN_API_CASES:
                arrays, meta = run_train_api_case(case, tmp)
                save("train_" + case["name"], arrays, meta)
        finally:
            os.chdir(cwd)
    for case in NET_CASES:
        arrays, meta = run_net_case(case)
        save("net_" + case["name"], arrays, meta)
    for case in INIT_CASES:
        arrays, meta = run_init_case(case)
        save("init_" + case["name"], arrays, meta)
    for case in DROPSTAT_CASES:
        arrays, meta = run_dropstat_case(case)
        save("dropstat_" + case["name"], arrays, meta)


if __name__ == "__main__":
    main()
