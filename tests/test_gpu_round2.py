"""Parity cases added in round 2 (all through the C ABI, against the numpy oracle / reference fixtures):

* parameters changed by a train step that was REPLAYED from a CUDA graph must reach the tensor-core
  inference path's weight planes (train, infer, train, infer -- with and without graphs);
* the bf16 / tcgen05 context at the C4 / C5 layer width (4096, K = 4096): one SGD step against the float64
  oracle with a tolerance derived from K, and against a float64 CPU emulation of the mixed-precision recipe
  (tight); three Adam steps against the oracle;
* Philox dropout in distribution against the reference's F.dropout statistics (tests/golden/dropstat_*.npz).
"""
import os

import numpy as np
import pytest
import torch

from conftest import check_means_within_standard_error, golden_names, load_golden
from oracle import mlp_oracle as orc
from oracle.detgen import det_labels, det_normalish, det_params, det_uniform

pytestmark = pytest.mark.gpu

MNIST = ([784, 256, 128, 10], ["linear", "relu", "linear", "relu", "linear", "softmax"])


# ---- graph replay vs weight planes -------------------------------------------------------------------------
def _train_infer_sequence(batch, infer_rows, graphs):
    """train x3, infer, train x3, infer on a side stream; returns the two inference outputs and the parameters."""
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    old = os.environ.get("POZ_GRAPHS")
    os.environ["POZ_GRAPHS"] = "1" if graphs else "0"       # read when the context is created
    try:
        sizes, algos = MNIST
        eng = DeviceMLP(sizes, StackSpec.from_algos(algos), True, "float32", 0)
    finally:
        if old is None:
            os.environ.pop("POZ_GRAPHS", None)
        else:
            os.environ["POZ_GRAPHS"] = old
    params = det_params(sizes, True, 77)
    for l in range(3):
        eng.upload_params(l, params[2 * l], params[2 * l + 1])
    x = torch.tensor(det_normalish((batch, 784), 78), dtype=torch.float32, device="cuda:0")
    y = torch.tensor(det_labels(batch, 10, 79), dtype=torch.int32, device="cuda:0")
    xi = torch.tensor(det_normalish((infer_rows, 784), 80), dtype=torch.float32, device="cuda:0")
    side = torch.cuda.Stream()
    outs = []
    torch.cuda.synchronize()
    with torch.cuda.stream(side):
        eng.train_begin(optimizer="adam", learning_rate=0.01, decay_rate=0.9, l2_lambda=0.001, epochs=8)
        eng.train_step(x, y)                 # first sighting: eager
        outs.append(eng.forward(xi).clone())  # an inference between the first and the second sighting of the key
        eng.train_step(x, y)                 # second sighting: captured
        eng.train_step(x, y)                 # replayed
        outs.append(eng.forward(xi).clone())
        for _ in range(3):
            eng.train_step(x, y)             # replayed
        outs.append(eng.forward(xi).clone())
    side.synchronize()
    ps = [eng.download_params(l) for l in range(3)]
    costs = eng.costs(0, 6)
    eng.close()
    return [o.cpu().numpy() for o in outs], ps, costs


@pytest.mark.parametrize("batch,infer_rows", [(1024, 4096), (8192, 8192)], ids=["simt-step", "tcgen05-layer0-step"])
def test_inference_after_replayed_train_steps_sees_the_new_weights(batch, infer_rows):
    """The large-batch inference path keeps bf16 planes of the weights; a train step replayed from a CUDA graph
    changes the weights without passing through the host code that used to mark those planes stale.  The same
    train / infer sequence with graphs disabled is the reference, and the float64 oracle pins the last output."""
    with_g, p_g, c_g = _train_infer_sequence(batch, infer_rows, True)
    without, p_n, c_n = _train_infer_sequence(batch, infer_rows, False)
    np.testing.assert_allclose(c_g, c_n, rtol=1e-6)
    for (wg, bg), (wn, bn) in zip(p_g, p_n):
        np.testing.assert_allclose(wg, wn, rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(bg, bn, rtol=1e-5, atol=1e-7)
    for a, b in zip(with_g, without):
        np.testing.assert_allclose(a, b, rtol=2e-5, atol=1e-7)
    # the outputs really moved between the inferences (a stale plane set would reproduce the first one)
    assert np.abs(with_g[2] - with_g[0]).max() > 1e-3
    sizes, algos = MNIST
    flat = [t for wb in p_g for t in wb]
    xi = det_normalish((infer_rows, 784), 80).astype(np.float32).astype(np.float64)
    ref = orc.forward(orc.parse_stack(algos), flat, xi)[0][-1]
    bad = np.abs(with_g[2] - ref) > 1e-5 * np.abs(ref) + 1e-7
    assert not bad.any(), int(bad.sum())


# ---- bf16 context at the C4 / C5 width -----------------------------------------------------------------------
WIDE = [4096] * 5
WIDE_ALGOS = ["linear", "relu"] * 4
WB = 2048


def bf16_round(a):
    """float64 -> nearest bfloat16 (ties to even) -> float64, through float32 as the device does."""
    f = np.ascontiguousarray(a, dtype=np.float32)
    u = f.view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000
    return u.astype(np.uint32).view(np.float32).astype(np.float64).reshape(np.shape(a))


def wide_setup():
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    rng = np.random.default_rng(404)
    params = []
    for l in range(4):                        # "he" scaling for relu (neural_net_model.py:243-253), small random biases
        params += [rng.standard_normal((4096, 4096)) * np.sqrt(2.0 / 4096), 0.1 * rng.standard_normal(4096)]
    eng = DeviceMLP(WIDE, StackSpec.from_algos(WIDE_ALGOS), True, "bfloat16", 0)
    for l in range(4):
        eng.upload_params(l, params[2 * l], params[2 * l + 1])
    x = bf16_round(rng.standard_normal((WB, 4096)))
    t = bf16_round(rng.standard_normal((WB, 4096)))
    xd = torch.tensor(x, dtype=torch.float32, device="cuda:0").to(torch.bfloat16)
    td = torch.tensor(t, dtype=torch.float32, device="cuda:0").to(torch.bfloat16)
    return eng, params, x, t, xd, td


def rms(v):
    return float(np.sqrt(np.mean(np.square(v))))


def _tc_gemm(variant, a, b, m, n, k, out, bias=None, aux=None, act=0, splits=1):
    from penr_oz_neural_network_torch_b200 import _native as nat
    torch.cuda.synchronize()
    nat.check(nat.lib().poz_tc_gemm(0, variant, a.data_ptr(), b.data_ptr(), m, n, k, out.data_ptr(),
                                    None if bias is None else bias.data_ptr(), None if aux is None else aux.data_ptr(),
                                    act, splits, int(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()


def _dev_bf16(a):
    return torch.tensor(a, dtype=torch.float32, device="cuda:0").to(torch.bfloat16).contiguous()


def _check_bf16_result(got, exact, what):
    """`exact` = the float64 value of what the kernel rounds to bf16.  The kernel sums in fp32 (relative error of
    a K = 4096 sum ~1e-6 of the row's magnitude), so a value that close to a bf16 tie may round the other way:
    >= 99 % of the elements bit-equal to round(exact), everything within one bf16 ulp (2^-7 relative) plus the
    fp32 summation error."""
    want = bf16_round(exact)
    got = got.float().cpu().numpy().astype(np.float64)
    same = got == want
    assert same.mean() >= 0.99, (what, same.mean())
    tol = np.abs(exact) * 2.0 ** -7 + 4e-6 * np.abs(exact).max()
    assert (np.abs(got - exact) <= tol).all(), (what, float((np.abs(got - exact) / tol).max()))


def test_tc_gemm_variants_at_the_c4_layer_shape_against_cpu_float64():
    """The three tcgen05 GEMMs of a bf16 train step at the C4 / C5 layer shape (N = K = 4096; 2048 rows) against a
    float64 CPU product of the same bf16 operands -- not a GPU library product.  A k-block or a tile fetched from
    the wrong place shows up as whole 128 x 256 blocks outside one ulp."""
    rng = np.random.default_rng(11)
    m, n, k = 2048, 4096, 4096
    a = bf16_round(rng.standard_normal((m, k)))
    w = bf16_round(rng.standard_normal((k, n)) * np.sqrt(2.0 / k))
    bias = rng.standard_normal(n).astype(np.float32)
    out = torch.full((m, n), float("nan"), device="cuda:0", dtype=torch.bfloat16)
    _tc_gemm(0, _dev_bf16(a), _dev_bf16(w), m, n, k, out, bias=torch.tensor(bias, device="cuda:0"), act=1)
    z = a @ w + bias.astype(np.float64)
    _check_bf16_result(out, np.maximum(z, 0.0), "forward (bias + relu epilogue)")
    # backward-x: dz [m][k'] (k' = outputs of the layer), W [n'][k'] -> [m][n'], masked by the activation in front
    dz = bf16_round(rng.standard_normal((m, k)) * 1e-6)
    aprev = bf16_round(np.maximum(rng.standard_normal((m, n)), 0.0))
    out = torch.full((m, n), float("nan"), device="cuda:0", dtype=torch.bfloat16)
    _tc_gemm(1, _dev_bf16(dz), _dev_bf16(w), m, n, k, out, aux=_dev_bf16(aprev), act=1)
    _check_bf16_result(out, (dz @ w.T) * (aprev > 0), "backward-x (relu-gradient epilogue)")
    # backward-w: x [rows][m'], dz [rows][n'] -> fp32 [m'][n'], summed over the 2048 rows
    x2 = bf16_round(rng.standard_normal((m, 4096)))
    g = torch.full((1, 4096, 4096), float("nan"), device="cuda:0", dtype=torch.float32)
    _tc_gemm(2, _dev_bf16(x2), _dev_bf16(dz), 4096, 4096, m, g, splits=1)
    want = x2.T @ dz
    got = g[0].cpu().numpy().astype(np.float64)
    assert np.abs(got - want).max() <= 4e-6 * np.abs(want).max() + 1e-6 * np.abs(want).max()


def test_wide_bf16_step_at_width_4096_against_oracle():
    """One SGD step (lr 1, so parameter delta = -gradient) and the forward pass of 4 x Linear(4096 -> 4096) + relu
    against the float64 oracle, on bf16-representable inputs and unrounded float64 weights.

    Stated bf16 tolerance, derived from K = 4096.  Forward: every product a*w carries two independent bf16
    roundings (relative 2^-9 each, uniform: sigma = 2^-9 / sqrt(3)); over a K-long sum of random-sign terms the
    relative error of a pre-activation is sqrt(2) * 2^-9 / sqrt(3) = 1.6e-3 of its rms whatever K is, the stored
    activation adds 1.1e-3, and four layers compound to ~4e-3 of the output rms: rms error <= 8e-3 * rms(ref),
    max error <= 6 times that.  Backward: relu' is discontinuous -- a pre-activation inside its own error band
    (|z| < ~4e-3 rms, a fraction f ~ 3e-3 of the elements) switches that element's gradient on or off, an rms
    error of sqrt(f) ~ 5.5e-2 that no amount of care in a bf16 kernel removes: rms error <= 8e-2 * rms(ref) for
    gradients.  The tight check of the same kernels (one bf16 ulp against float64 CPU products at this layer
    shape) is test_tc_gemm_variants_at_the_c4_layer_shape_against_cpu_float64."""
    eng, params, x, t, xd, td = wide_setup()
    out = eng.forward(xd).float().cpu().numpy().astype(np.float64)
    eng.train_begin(optimizer="sgd", learning_rate=1.0, decay_rate=1.0, l2_lambda=0.0, epochs=1)
    eng.train_step(xd, td)
    cost = float(eng.costs(0, 1)[0])
    got_grads = []
    for l in range(4):
        w, b = eng.download_params(l)
        got_grads += [params[2 * l] - w, params[2 * l + 1] - b]
    eng.close()
    stack = orc.parse_stack(WIDE_ALGOS)
    _, ref_cost, outs, ref_grads, _ = orc.train_step(stack, params, x, t, lr=1.0, optimizer="sgd")
    ref_out = outs[-1]
    assert rms(out - ref_out) <= 8e-3 * rms(ref_out), rms(out - ref_out) / rms(ref_out)
    assert np.abs(out - ref_out).max() <= 6 * 8e-3 * rms(ref_out)
    np.testing.assert_allclose(cost, ref_cost, rtol=5e-3)
    for i, (g, r) in enumerate(zip(got_grads, ref_grads)):
        # parameters are float32 masters of magnitude ~2e-2: a delta of ~1e-5 resolves to ~2e-9 (float32 ulp)
        assert rms(g - r) <= 8e-2 * rms(r) + 2e-9, (i, rms(g - r) / rms(r))
        cos = float((g * r).sum() / np.sqrt((g * g).sum() * (r * r).sum()))
        assert cos > 0.995, (i, cos)


def test_wide_bf16_three_adam_steps_at_width_4096_against_oracle():
    eng, params, x, t, xd, td = wide_setup()
    eng.train_begin(optimizer="adam", learning_rate=1e-4, decay_rate=0.9, l2_lambda=1e-6, epochs=3)
    for _ in range(3):
        eng.train_step(xd, td)
    costs = eng.costs(0, 3)
    assert eng.get_adam_step() == 3
    out, cost = eng.forward_loss(xd, td)
    got = out.float().cpu().numpy().astype(np.float64)
    deltas = []
    for l in range(4):
        w, _ = eng.download_params(l)
        deltas.append(w - params[2 * l])
    eng.close()
    stack = orc.parse_stack(WIDE_ALGOS)
    adam = orc.AdamState()
    p, ref_costs = params, []
    for s in range(3):
        p, c, *_ = orc.train_step(stack, p, x, t, lr=1e-4 * 0.9 ** s, optimizer="adam", adam=adam,
                                  l2_lambda=1e-6)
        ref_costs.append(c)
    np.testing.assert_allclose(costs, ref_costs, rtol=1e-2)
    ref = orc.forward(stack, p, x)[0][-1]
    # Adam's first steps are ~lr * sign(g): the ~4 % of the elements whose gradient is smaller than its own bf16 /
    # relu-flip noise (rms 5.5e-2, see the one-step test) move the other way, 2 * lr = 1 % of |w| ~ 2e-2 each:
    # ~2e-3 of output rms per step and layer on top of the forward tolerance of 8e-3 -> 2.5e-2; the direction of
    # the accumulated update agrees to cos > 0.85 (1 - 2 * 0.044 = 0.91 expected)
    assert rms(got - ref) <= 2.5e-2 * rms(ref), rms(got - ref) / rms(ref)
    np.testing.assert_allclose(cost, np.mean((ref - t) ** 2), rtol=2e-2)
    for l in range(4):
        want = p[2 * l] - params[2 * l]
        cos = float((deltas[l] * want).sum() / np.sqrt((deltas[l] ** 2).sum() * (want ** 2).sum()))
        assert cos > 0.85, (l, cos)


# ---- Philox dropout in distribution -----------------------------------------------------------------------------
@pytest.mark.parametrize("name", golden_names("dropstat_"))
def test_philox_dropout_matches_reference_in_distribution(name):
    """The perf-mode default (dropout_rate 0.2, masks drawn in-kernel from Philox and regenerated in the
    backward) cannot agree with torch's mask stream step by step; over 64 seeds the cost and the mean
    |weight gradient| of one step must agree with the reference's 64 F.dropout seeds within 3 standard errors
    (and have a comparable spread) -- i.e. same placement (every hidden entry, the logits included,
    neural_net_model.py:393-395), same keep probability, same 1/(1-p) scaling, in forward AND backward."""
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    arrays, meta = load_golden(name)
    sizes, algos = meta["sizes"], meta["expanded_algos"]
    spec = StackSpec.from_algos(algos)
    has_bias = bool(meta["bias"])
    params = det_params(sizes, has_bias, meta["salt"])
    per = 2 if has_bias else 1
    eng = DeviceMLP(sizes, spec, has_bias, "float32", 0)
    for l in range(spec.n_linear):
        eng.upload_params(l, params[per * l], params[per * l + 1] if has_bias else None)
    x = torch.tensor(det_normalish((meta["batch"], sizes[0]), meta["salt"] + 3), dtype=torch.float32, device="cuda:0")
    y = torch.tensor(det_labels(meta["batch"], sizes[-1], meta["salt"] + 7), dtype=torch.int32, device="cuda:0")
    costs, gabs = [], []
    for seed in range(meta["seeds"]):
        # learning rate 0: the step leaves the parameters where they are
        eng.train_begin(optimizer="sgd", learning_rate=0.0, decay_rate=1.0, dropout_rate=meta["dropout"],
                        l2_lambda=0.0, epochs=1, dropout_seed=5000 + seed)
        eng.train_step(x, y, diagnostics=True)
        costs.append(float(eng.costs(0, 1)[0]))
        gabs.append([np.abs(eng.diag_weight_grad(l)[0]).mean() for l in range(spec.n_linear)])
    eng.close()
    check_means_within_standard_error(np.asarray(costs), arrays["cost"], "cost")
    gabs = np.asarray(gabs)
    for l in range(gabs.shape[1]):
        check_means_within_standard_error(gabs[:, l], arrays["grad_abs_mean"][:, l], f"mean|dW{l}|")
