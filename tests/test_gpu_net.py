"""Stacks with embedding / flatten / batchnorm entries through the C ABI,
against the outputs of the unmodified reference (tests/golden/net_*.npz).  Same stated tolerances
as tests/test_gpu_parity.py (float32: rtol 1e-5 + atol 1e-5*max|ref|; float64: rtol 1e-10)."""
import numpy as np
import pytest
import torch

from conftest import golden_names, load_golden
from oracle import net_oracle as no
from test_gpu_parity import ADAM_PARAM_ATOL, close, close_golden
from test_oracle_golden import net_case_inputs, net_det_params

pytestmark = pytest.mark.gpu


def build_net(meta, dtype):
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    algos = meta["expanded_algos"]
    entries, x, target = net_case_inputs(meta)
    ratios = {i: e.ratio for i, e in enumerate(entries) if e.algo == "flatten"}
    spec = StackSpec.from_algos(algos, ratios)
    dims = [entries[i].weights.shape for i in spec.z_index]
    emb = entries[0].weights.shape if spec.emb_index is not None else None
    bn = [(entries[i].eps, entries[i].momentum) if i is not None else None for i in spec.bn_index]
    eng = DeviceMLP(dims, spec, bool(meta["bias"]), dtype, 0, embedding=emb, batchnorm=bn)
    params = net_det_params(entries, meta["salt"])
    no.set_params(entries, params)
    upload_entries(eng, spec, entries)
    tdt = torch.float32 if dtype == "float32" else torch.float64
    xd = (torch.tensor(x, dtype=torch.int32, device="cuda:0") if emb is not None
          else torch.tensor(x, dtype=tdt, device="cuda:0"))
    if spec.loss == "ce":
        td = torch.tensor([t[0] for t in target], dtype=torch.int32, device="cuda:0")
    else:
        td = torch.tensor(target, dtype=tdt, device="cuda:0")
    return eng, spec, entries, x, xd, target, td


def upload_entries(eng, spec, entries):
    if spec.emb_index is not None:
        eng.upload_tensor("embedding", 0, entries[0].weights)
    for l in range(spec.n_linear):
        e = entries[spec.z_index[l]]
        eng.upload_tensor("weight", l, e.weights)
        if e.bias is not None:
            eng.upload_tensor("bias", l, e.bias)
        if spec.bn_index[l] is not None:
            b = entries[spec.bn_index[l]]
            eng.upload_tensor("bn_gain", l, b.gain)
            eng.upload_tensor("bn_bias", l, b.bn_bias)
            eng.upload_tensor("bn_mean", l, np.asarray(b.mean).reshape(-1))
            eng.upload_tensor("bn_var", l, np.asarray(b.variance).reshape(-1))


def entry_locations(spec):
    """entry index -> (layer, which) of the device tensor that holds its recorded output."""
    loc = {}
    if spec.emb_index is not None:
        loc[spec.emb_index] = (-1, 0)
        for i in spec.emb_flat:
            loc[i] = (-1, 0)
    for l in range(spec.n_linear):
        loc[spec.z_index[l]] = (l, 0)
        last = (l, 0)
        if spec.bn_index[l] is not None:
            loc[spec.bn_index[l]] = last = (l, 2)
        if spec.a_index[l] is not None:
            loc[spec.a_index[l]] = last = (l, 1)
        for i in spec.flat_after[l]:
            loc[i] = last
    return loc


def param_locations(spec, entries):
    """flat `params` index -> (kind, layer)."""
    out = []
    for i, e in enumerate(entries):
        if e.algo == "embedding":
            out.append(("embedding", 0))
        elif e.algo == "linear":
            l = spec.z_index.index(i)
            out.append(("weight", l))
            if e.bias is not None:
                out.append(("bias", l))
        elif e.algo == "batchnorm":
            l = spec.bn_index.index(i)
            out += [("bn_gain", l), ("bn_bias", l)]
    return out


def device_masks(arrays, spec, step):
    """Keep masks of the hidden entries -> one (multiplied) mask per device dropout point."""
    def m(i):
        return arrays.get(f"mask{i}_s{step}")
    slots = {}

    def put(slot, idxs):
        ms = [m(i) for i in idxs if m(i) is not None]
        if ms:
            prod = np.ones(ms[0].size, dtype=np.uint8)
            for a in ms:
                prod = prod * a.reshape(-1)
            slots[slot] = torch.tensor(prod, dtype=torch.uint8, device="cuda:0").contiguous()

    nl = spec.n_linear
    if spec.emb_index is not None:
        put(3 * nl, spec.emb_flat)
    for l in range(nl):
        z, b, a = [spec.z_index[l]], [], []
        if spec.bn_index[l] is not None:
            b = [spec.bn_index[l]]
        if spec.a_index[l] is not None:
            a = [spec.a_index[l]]
        tail = spec.flat_after[l]
        if a:
            a += tail
        elif b:
            b += tail
        else:
            z += tail
        put(3 * l, z); put(3 * l + 2, b); put(3 * l + 1, a)
    return slots


@pytest.mark.parametrize("dtype", ["float32", "float64"])
@pytest.mark.parametrize("name", golden_names("net_"))
def test_general_stacks_match_reference_outputs(name, dtype):
    arrays, meta = load_golden(name)
    eng, spec, entries, x, xd, target, td = build_net(meta, dtype)
    p = meta.get("dropout", 0.0)
    B = meta["batch"]
    adam = meta["optimizer"] == "adam"
    eng.train_begin(optimizer=meta["optimizer"] or "sgd", learning_rate=meta["lr"], decay_rate=meta["decay"],
                    dropout_rate=p, l2_lambda=meta["l2"], epochs=meta["steps"], progress_stride=1)
    loc = entry_locations(spec)
    plocs = param_locations(spec, entries)
    for step in range(meta["steps"]):
        masks = device_masks(arrays, spec, step) if p > 0 else None
        eng.train_step(xd, td, masks=masks, diagnostics=(step == 0))
        if step == 0:
            folded = set(spec.emb_flat) | {i for fl in spec.flat_after for i in fl}
            for i in range(len(entries)):
                layer, which = loc[i]
                # with dropout, an entry that precedes a folded flatten is recorded with the flatten's mask as well
                precise = p == 0 or (i in folded) or not any(j in folded for j in range(i + 1, len(entries)) if loc[j] == loc[i])
                got = eng.diag_activation(layer, which, B)
                want = arrays[f"act{i}"]
                if precise:
                    close(got.reshape(want.shape), want, dtype, f"act{i}")
                g = eng.diag_activation_grad(layer, which, B)
                assert (g is not None) == (f"actgrad{i}" in arrays), f"actgrad{i}"
                if g is not None and precise:
                    close(g.reshape(arrays[f"actgrad{i}"].shape), arrays[f"actgrad{i}"], dtype, f"actgrad{i}", outliers=True)
            wscale = max(np.abs(arrays[f"grad{k}"]).max() for k in range(len(plocs)))
            for k, (kind, layer) in enumerate(plocs):
                got = eng.download_tensor(kind, layer, "diag_grad")
                if kind == "bias" and spec.bn_index[layer] is not None and p == 0:
                    # mathematically zero (the batchnorm removes any bias): round-off noise on both sides
                    assert np.abs(arrays[f"grad{k}"]).max() < 1e-12 and np.abs(got).max() < 1e-5 * wscale
                    continue
                close(got, arrays[f"grad{k}"], dtype, f"grad{k} {kind}{layer}", outliers=True)
    close(eng.costs(0, meta["steps"]), arrays["costs"], dtype, "costs")
    for k, (kind, layer) in enumerate(plocs):
        # a bias in front of a batchnorm has a zero gradient up to round-off, which Adam normalises to +-lr:
        # its value is a gauge (the batchnorm removes it again), so it is not compared
        gauge = adam and kind == "bias" and spec.bn_index[layer] is not None and p == 0
        if not gauge:
            close(eng.download_tensor(kind, layer), arrays[f"param{k}"], dtype, f"param{k} {kind}{layer}",
                  atol=ADAM_PARAM_ATOL[dtype] if adam else None, outliers=True, loose_atol=2 * meta["lr"] if adam else None)
    for l in range(spec.n_linear):
        i = spec.bn_index[l]
        if i is not None and not (adam and meta["bias"]):
            close(eng.download_tensor("bn_mean", l), arrays[f"bn_mean{i}"], dtype, f"bn_mean{i}")
            close(eng.download_tensor("bn_var", l), arrays[f"bn_var{i}"], dtype, f"bn_var{i}")
    out = eng.forward(xd).cpu().numpy()
    if not (adam and meta["bias"] and any(b is not None for b in spec.bn_index)):
        close(out, arrays["infer_out"], dtype, "infer_out", rtol=1e-4 if dtype == "float32" else None)
        one = eng.forward(xd[:1].contiguous()).cpu().numpy()[0]
        close(one, arrays["infer_out_1d"].reshape(-1), dtype, "infer_out_1d", rtol=1e-4 if dtype == "float32" else None)
