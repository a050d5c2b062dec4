"""`_forward` as the reference defines it (neural_net_model.py:384-408): every entry's output, the cost, and
autograd that delivers the CUDA backward pass's gradients; plus the layers' own `forward` operators.

The gradient case restates the reference's test_torch_backward.py (embedding -> linear -> batchnorm -> tanh ->
linear -> softmax + cross-entropy, hand-derived backward, `:121-204`) at tolerances a different summation order
can meet: float64 context rtol 1e-9, float32 context rtol 2e-4 + 2e-5 * max|tensor|.
"""
import random

import numpy as np
import pytest
import torch

from oracle import mlp_oracle as orc
from oracle.detgen import det_labels, det_normalish

pytestmark = pytest.mark.gpu


def closed_form_check(dtype, rtol, atol_rel):
    from penr_oz_neural_network_torch_b200 import NeuralNetworkModel
    torch.manual_seed(3)
    random.seed(3)
    block, emb_sz, hidden, vocab = 3, 10, 64, 27
    model = NeuralNetworkModel("autograd", [vocab, emb_sz, emb_sz * block, hidden, vocab], "xavier", "random",
                               ["embedding", "linear", "batchnorm", "tanh", "linear", "softmax"], None, dtype=dtype)
    batch = 512
    x = torch.randint(low=0, high=vocab, size=(batch, block))
    target = [[random.randint(0, vocab - 1)] for _ in range(batch)]
    for p in model.params:
        p.requires_grad_()
    activations, cost = model._forward(x, target)
    for p in model.params:
        p.grad = None
    for a in activations:
        a.retain_grad()
    cost.backward()
    emb, emb_cat, h_pre_bn, h_pre_act, h, logits, probs = tuple(activations)
    c, w1, b1, bn_gain, bn_bias, w2, b2 = tuple(model.params)
    assert emb.shape == (batch, block, emb_sz) and emb_cat.shape == (batch, block * emb_sz)
    assert probs.grad is None                                   # the trailing softmax is not on the loss graph
    y = torch.tensor([t[0] for t in target], dtype=torch.int64)

    def close(got, want, what):
        scale = want.abs().max().item()
        torch.testing.assert_close(got, want, rtol=rtol, atol=atol_rel * scale, msg=lambda m: what + ": " + m)

    with torch.no_grad():
        # forward values against the same formulas in float64 torch on the host
        close(emb, c[x], "embedding output")
        close(h_pre_bn, emb_cat @ w1 + b1, "linear 1")
        bn_raw = (h_pre_act - bn_bias) / bn_gain
        mean = h_pre_bn.mean(0, keepdim=True)
        var = h_pre_bn.var(0, keepdim=True)
        close(h_pre_act, bn_gain * (h_pre_bn - mean) / torch.sqrt(var + 1e-5) + bn_bias, "batchnorm")
        close(h, h_pre_act.tanh(), "tanh")
        close(logits, h @ w2 + b2, "linear 2")
        close(probs, logits.softmax(1), "softmax")
        close(cost.detach(), torch.nn.functional.cross_entropy(logits, y), "cost")
        # the derived backward of test_torch_backward.py:160-187
        bn_var_inv = (var + 1e-5) ** -0.5
        d_logits = probs / batch
        d_logits[range(batch), y] -= 1.0 / batch
        d_w2 = h.T @ d_logits
        d_b2 = d_logits.sum(0)
        d_h = d_logits @ w2.T
        d_h_pre_act = (1.0 - h ** 2) * d_h
        d_bn_gain = (bn_raw * d_h_pre_act).sum(0)
        d_bn_bias = d_h_pre_act.sum(0)
        d_h_pre_bn = bn_gain * bn_var_inv / batch * (batch * d_h_pre_act - d_h_pre_act.sum(0) -
                                                     batch / (batch - 1) * bn_raw * (d_h_pre_act * bn_raw).sum(0))
        d_w1 = emb_cat.T @ d_h_pre_bn
        d_emb_cat = d_h_pre_bn @ w1.T
        d_emb = d_emb_cat.view(emb.shape)
        d_c = torch.zeros_like(c)
        d_c.index_put_((x.reshape(-1),), d_emb.reshape(-1, emb_sz), accumulate=True)
    close(logits.grad, d_logits, "d logits")
    close(w2.grad, d_w2, "d w2")
    close(b2.grad, d_b2, "d b2")
    close(h.grad, d_h, "d h")
    close(h_pre_act.grad, d_h_pre_act, "d h_pre_act")
    close(bn_gain.grad, d_bn_gain, "d bn gain")
    close(bn_bias.grad, d_bn_bias, "d bn bias")
    close(h_pre_bn.grad, d_h_pre_bn, "d h_pre_bn")
    close(w1.grad, d_w1, "d w1")
    close(emb_cat.grad, d_emb_cat, "d emb_cat")
    close(emb.grad, d_emb, "d emb")
    close(c.grad, d_c, "d embedding table")
    # b1 sits in front of a batchnorm: its gradient is mathematically zero (round-off only)
    assert b1.grad.abs().max().item() <= 1e-6 * d_w1.abs().max().item() + 1e-12
    # scaling the cost scales every gradient (the autograd node passes the upstream gradient on)
    g_before = w2.grad.clone()
    for p in model.params:
        p.grad = None
    acts2, cost2 = model._forward(x, target)
    (3.0 * cost2 + 0.5 * (w2 ** 2).sum()).backward()
    close(w2.grad, 3.0 * g_before + w2.detach(), "scaled cost + L2 term")


def test_forward_returns_every_entry_and_autograd_delivers_device_gradients_float64():
    closed_form_check("float64", 1e-9, 1e-11)


def test_forward_returns_every_entry_and_autograd_delivers_device_gradients_float32():
    closed_form_check("float32", 2e-4, 2e-5)


def test_forward_without_target_chains_the_unit_operators():
    """neural_net_model.py:389-396 with `layer.training = False`: six entry outputs for the C2-shaped stack, 1-D
    and 2-D inputs, against the numpy oracle; cost is `torch.empty(0)`."""
    from penr_oz_neural_network_torch_b200 import NeuralNetworkModel
    torch.manual_seed(1)
    model = NeuralNetworkModel("fwd", [20, 16, 12, 5], activation_algos=["relu", "tanh", "softmax"], bias_algo="random",
                               dtype="float64")
    x = det_normalish((33, 20), 9)
    acts, cost = model._forward(torch.tensor(x), None)
    assert cost.numel() == 0 and len(acts) == 6
    params = [p.detach().numpy() for p in model.params]
    ref, _ = orc.forward(orc.parse_stack(model.algos), params, x)
    for a, r in zip(acts, ref):
        np.testing.assert_allclose(a.numpy(), r, rtol=1e-12, atol=1e-14)
    one, _ = model._forward(torch.tensor(x[0]), None)
    assert [tuple(a.shape) for a in one] == [(16,), (16,), (12,), (12,), (5,), (5,)]
    np.testing.assert_allclose(one[-1].numpy(), ref[-1][0], rtol=1e-12)
    # compute_output (fused kernels) agrees with the chained operators
    out, c = model.compute_output(x.tolist())
    np.testing.assert_allclose(np.asarray(out), ref[-1], rtol=1e-12)
    assert c is None


def test_forward_applies_dropout_rate():
    """`dropout_rate` reaches the kernels: with p = 0.5 about half of every hidden entry is zero and the
    survivors are scaled by 2 (neural_net_model.py:393-395); p = 0 leaves nothing zeroed."""
    from penr_oz_neural_network_torch_b200 import NeuralNetworkModel
    torch.manual_seed(2)
    model = NeuralNetworkModel("drop", [32, 64, 48, 6], activation_algos=["sigmoid", "tanh", "softmax"])
    x = torch.tensor(det_normalish((256, 32), 5))
    target = [[int(v)] for v in det_labels(256, 6, 6)]
    plain, c0 = model._forward(x, target, 0.0)
    dropped, c1 = model._forward(x, target, 0.5)
    assert all((a != 0).all() for a in plain[1:5])
    # hidden entries: sigmoid, linear, tanh, linear (logits).  tanh(0) = 0, so the tanh entry shows the zeros of
    # the dropped linear entry in front of it plus its own: 1 - 0.5 * 0.5
    for i, want in ((1, 0.5), (2, 0.5), (3, 0.75), (4, 0.5)):
        zero = (dropped[i] == 0).double().mean().item()
        assert want - 0.08 < zero < want + 0.08, (i, zero)
    assert (dropped[0] != 0).all()                              # the first entry is not hidden
    kept = dropped[1] != 0
    sig = torch.sigmoid(plain[0])
    torch.testing.assert_close(dropped[1][kept], 2.0 * sig[kept], rtol=1e-4, atol=1e-6)
    assert abs(c0.item() - c1.item()) > 1e-6


def test_layer_forward_operators_match_the_reference_formulas():
    from penr_oz_neural_network_torch_b200 import (BatchNormLayer, EmbeddingLayer, FlattenLayer, Layer, LinearLayer,
                                                   ReluLayer, SigmoidLayer, SoftmaxLayer, TanhLayer)
    from penr_oz_neural_network_torch_b200 import launch_count
    torch.manual_seed(0)
    x = torch.randn(7, 9, dtype=torch.float64)
    assert Layer().forward(x) is x
    lin = LinearLayer(9, 5, "random")
    before = launch_count()
    y = lin.forward(x)
    assert launch_count() > before, "LinearLayer.forward must run the CUDA operator"
    torch.testing.assert_close(y, x @ lin.weights + lin.bias, rtol=1e-12, atol=1e-14)
    torch.testing.assert_close(lin.forward(x[0]), x[0] @ lin.weights + lin.bias, rtol=1e-12, atol=1e-14)
    nobias = LinearLayer(9, 5, "")
    torch.testing.assert_close(nobias.forward(x), x @ nobias.weights, rtol=1e-12, atol=1e-14)
    x3 = torch.randn(4, 3, 9, dtype=torch.float64)
    torch.testing.assert_close(lin.forward(x3), x3 @ lin.weights + lin.bias, rtol=1e-12, atol=1e-14)
    lin.weights[0, 0] += 1.0                                      # an in-place edit is picked up
    torch.testing.assert_close(lin.forward(x), x @ lin.weights + lin.bias, rtol=1e-12, atol=1e-14)
    for cls, fn in ((ReluLayer, torch.relu), (SigmoidLayer, torch.sigmoid), (TanhLayer, torch.tanh),
                    (SoftmaxLayer, lambda t: t.softmax(-1))):
        torch.testing.assert_close(cls().forward(x), fn(x), rtol=1e-12, atol=1e-14)
        torch.testing.assert_close(cls().forward(x[0]), fn(x[0]), rtol=1e-12, atol=1e-14)
    emb = EmbeddingLayer(27, 4)
    idx = torch.randint(0, 27, (10, 3))
    assert torch.equal(emb.forward(idx), emb.weights[idx])
    assert torch.equal(emb.forward(torch.tensor([0.0, 5.0, 13.0])), emb.weights[[0, 5, 13]])
    with pytest.raises(IndexError):
        emb.forward(torch.tensor([27]))
    flat = FlattenLayer(3)
    assert flat.forward(emb.forward(idx)).shape == (10, 12)
    bn = BatchNormLayer(9)
    bn.gain = torch.randn(9, dtype=torch.float64)
    bn.bias = torch.randn(9, dtype=torch.float64)
    bn.training = True
    got = bn.forward(x)
    mean, var = x.mean(0, keepdim=True), x.var(0, keepdim=True)
    torch.testing.assert_close(got, bn.gain * (x - mean) / torch.sqrt(var + 1e-5) + bn.bias, rtol=1e-12, atol=1e-13)
    torch.testing.assert_close(bn.mean, 0.1 * mean, rtol=1e-12, atol=1e-15)
    torch.testing.assert_close(bn.variance, 0.9 + 0.1 * var, rtol=1e-12, atol=1e-15)
    bn.training = False
    run_m, run_v = bn.mean.clone(), bn.variance.clone()
    got = bn.forward(x)
    torch.testing.assert_close(got, bn.gain * (x - run_m) / torch.sqrt(run_v + 1e-5) + bn.bias, rtol=1e-12, atol=1e-13)
    assert torch.equal(bn.mean, run_m)
    # the 1-D training quirk: dim = () reduces over every element (neural_net_model.py:160-162)
    bn1 = BatchNormLayer(9)
    bn1.training = True
    v = x[0]
    got = bn1.forward(v)
    torch.testing.assert_close(got, (v - v.mean()) / torch.sqrt(v.var() + 1e-5), rtol=1e-12, atol=1e-13)
    assert tuple(bn1.mean.shape) == (9,)
    torch.testing.assert_close(bn1.mean, torch.full((9,), 0.1 * v.mean().item(), dtype=torch.float64), rtol=1e-12, atol=1e-15)
