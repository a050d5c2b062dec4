"""Parity of the CUDA path (through the C ABI) with the reference.

Every case is checked twice: against the committed outputs of the unmodified reference
(tests/golden/step_*.npz) and, where useful, against the numpy oracle run live on the same inputs.

Stated tolerances
  float32 context: rtol 1e-5 (BASELINE.json north_star) plus atol = 1e-5 * max|reference tensor|,
                   which covers elements that are small because of cancellation in a K- or
                   batch-long fp32 sum.
  float64 context: rtol 1e-10, atol = 1e-12 * max|reference tensor|.
  argmax of the network output: identical.
  relu outliers (float32 only): the reference function is discontinuous in its own rounding -- a
                   pre-activation within ~1e-6 of zero gets relu'(z) = 0 or 1 depending on the last
                   bit, which switches one sample's contribution to a gradient on or off.  For
                   gradient-derived tensors of float32 runs at most max(2, 0.02 %) of the elements
                   may miss the tolerance above, and those must still agree to 2 % of max|tensor|
                   (after Adam: to 2*lr).  float64 runs get no such allowance.
"""
import numpy as np
import pytest
import torch

from conftest import golden_names, load_golden
from oracle import mlp_oracle as orc
from oracle.detgen import det_labels, det_normalish, det_params, det_uniform

pytestmark = pytest.mark.gpu

TOL = {"float32": (1e-5, 1e-5), "float64": (1e-10, 1e-12)}
# parameters after Adam steps: m/(sqrt(v)+eps) turns a relative gradient error into an absolute
# update error of lr * that error, so the absolute term is tied to lr rather than to |w|
ADAM_PARAM_ATOL = {"float32": 2e-6, "float64": 1e-12}


def launch_count():
    from penr_oz_neural_network_torch_b200 import _native
    return int(_native.lib().poz_launch_count())


def close(got, want, dtype, what, rtol=None, atol=None, outliers=False, loose_atol=None):
    r, a = TOL[dtype]
    want = np.asarray(want, dtype=np.float64)
    got = np.asarray(got, dtype=np.float64)
    scale = float(np.max(np.abs(want))) if want.size else 0.0
    rtol = rtol or r
    atol = atol if atol is not None else a * max(scale, 1e-30)
    if outliers and dtype == "float32" and want.size:
        bad = ~np.isclose(got, want, rtol=rtol, atol=atol)
        allowed = max(2, int(np.ceil(2e-4 * want.size)))
        assert bad.sum() <= allowed, f"{what}: {bad.sum()} elements outside tolerance (allowed {allowed})"
        loose = loose_atol if loose_atol is not None else 2e-2 * max(scale, 1e-30)
        np.testing.assert_allclose(got[bad], want[bad], rtol=1e-2, atol=loose, err_msg=what + " (relu outliers)")
        return
    np.testing.assert_allclose(got, want, rtol=rtol, atol=atol, err_msg=what)


def expand(arrays, key, want_shape):
    """Golden entry -> (full array | None, flat sub-sample, stride)."""
    if key in arrays:
        return arrays[key], None, None
    sub = arrays[key + "__sub"]
    n = int(np.prod(want_shape))
    stride = next(s for s in range(1, 4096) if len(range(0, n, s)) == sub.size)
    return None, sub, stride


def close_golden(arrays, key, got, dtype, what=None, **kw):
    got = np.asarray(got, dtype=np.float64)
    full, sub, stride = expand(arrays, key, got.shape)
    if full is not None:
        close(got.reshape(full.shape), full, dtype, what or key, **kw)
    else:
        close(got.reshape(-1)[::stride], sub, dtype, (what or key) + "__sub", **kw)


def build_case(meta, dtype):
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    algos = meta["expanded_algos"]
    spec = StackSpec.from_algos(algos)
    sizes = meta["sizes"]
    has_bias = bool(meta["bias"])
    params = det_params(sizes, has_bias, meta["salt"])
    eng = DeviceMLP(sizes, spec, has_bias, dtype, 0)
    per = 2 if has_bias else 1
    for l in range(spec.n_linear):
        eng.upload_params(l, params[per * l], params[per * l + 1] if has_bias else None)
    x = det_normalish((meta["batch"], sizes[0]), meta["salt"] + 3)
    tdt = torch.float32 if dtype == "float32" else torch.float64
    xd = torch.tensor(x, dtype=tdt, device="cuda:0")
    if spec.loss == "ce":
        y = det_labels(meta["batch"], sizes[-1], meta["salt"] + 7)
        td = torch.tensor(y, dtype=torch.int32, device="cuda:0")
        target = [[int(v)] for v in y]
    else:
        t = det_uniform((meta["batch"], sizes[-1]), meta["salt"] + 9, 0.0, 1.0)
        td = torch.tensor(t, dtype=tdt, device="cuda:0")
        target = t.tolist()
    return eng, spec, params, x, xd, target, td


def device_masks(arrays, spec, step):
    out = {}
    for l in range(spec.n_linear):
        for which, idx in ((0, spec.z_index[l]), (1, spec.a_index[l])):
            if idx is None:
                continue
            key = f"mask{idx}" if step == 0 else f"mask{idx}_s{step}"
            if key in arrays:
                out[2 * l + which] = torch.tensor(arrays[key], dtype=torch.uint8, device="cuda:0").contiguous()
    return out or None


@pytest.mark.parametrize("dtype", ["float32", "float64"])
@pytest.mark.parametrize("name", golden_names("step_"))
def test_train_steps_match_reference_outputs(name, dtype):
    arrays, meta = load_golden(name)
    eng, spec, params, x, xd, target, td = build_case(meta, dtype)
    p = meta.get("dropout", 0.0)
    batch = meta["batch"]
    has_bias = bool(meta["bias"])
    per = 2 if has_bias else 1
    eng.train_begin(optimizer=meta["optimizer"] or "sgd", learning_rate=meta["lr"],
                    decay_rate=meta["decay"], dropout_rate=p, l2_lambda=meta["l2"], epochs=meta["steps"],
                    progress_stride=1)
    for step in range(meta["steps"]):
        masks = device_masks(arrays, spec, step) if p > 0 else None
        eng.train_step(xd, td, masks=masks, diagnostics=(step == 0))
        if step == 0:
            for l in range(spec.n_linear):
                zi, ai = spec.z_index[l], spec.a_index[l]
                close_golden(arrays, f"act{zi}", eng.diag_activation(l, 0, batch), dtype)
                g = eng.diag_activation_grad(l, 0, batch)
                assert g is not None
                close_golden(arrays, f"actgrad{zi}", g, dtype, outliers=True)
                if ai is not None:
                    close_golden(arrays, f"act{ai}", eng.diag_activation(l, 1, batch), dtype)
                    g = eng.diag_activation_grad(l, 1, batch)
                    has = (f"actgrad{ai}" in arrays) or (f"actgrad{ai}__sub" in arrays)
                    assert (g is not None) == has     # trailing softmax: .grad is None
                    if g is not None:
                        close_golden(arrays, f"actgrad{ai}", g, dtype, outliers=True)
                wg, bg = eng.diag_weight_grad(l)
                close_golden(arrays, f"grad{per * l}", wg, dtype, outliers=True)
                if has_bias:
                    close_golden(arrays, f"grad{per * l + 1}", bg, dtype, outliers=True)
    close(eng.costs(0, meta["steps"]), arrays["costs"], dtype, "costs")
    adam = meta["optimizer"] == "adam"
    patol = ADAM_PARAM_ATOL[dtype] if adam else None
    for l in range(spec.n_linear):
        w, b = eng.download_params(l)
        lo = 2 * meta["lr"] if adam else None
        close_golden(arrays, f"param{per * l}", w, dtype, atol=patol, outliers=True, loose_atol=lo)
        if has_bias:
            close_golden(arrays, f"param{per * l + 1}", b, dtype, atol=patol, outliers=True, loose_atol=lo)
    if adam:
        assert eng.get_adam_step() == int(arrays["adam_step"])
        for l in range(spec.n_linear):
            m_w, v_w, m_b, v_b = eng.download_adam(l)
            close_golden(arrays, f"adam_m{per * l}", m_w, dtype, outliers=True)
            close_golden(arrays, f"adam_v{per * l}", v_w, dtype, outliers=True)
            if has_bias:
                close_golden(arrays, f"adam_m{per * l + 1}", m_b, dtype, outliers=True)
                close_golden(arrays, f"adam_v{per * l + 1}", v_b, dtype, outliers=True)
    # inference on the updated parameters, with and without a target
    out, cost = eng.forward_loss(xd, td)
    ptol = dict(rtol=1e-4) if (adam and dtype == "float32") else {}
    close_golden(arrays, "infer_out", out.cpu().numpy(), dtype, **ptol)
    close(cost, arrays["infer_cost"], dtype, "infer_cost", **ptol)
    out2 = eng.forward(xd)
    assert torch.equal(out, out2) or torch.allclose(out, out2, rtol=1e-6, atol=0)
    one = eng.forward(xd[:1].contiguous()).cpu().numpy()[0]
    close_golden(arrays, "infer_out_1d", one, dtype, **ptol)


@pytest.mark.parametrize("dtype", ["float32", "float64"])
def test_weight_update_ratio_matches_oracle(dtype):
    arrays, meta = load_golden("step_relu_tanh_adam")
    eng, spec, params, x, xd, target, td = build_case(meta, dtype)
    stack = orc.parse_stack(meta["expanded_algos"])
    eng.train_begin(optimizer="adam", learning_rate=meta["lr"], decay_rate=meta["decay"],
                    l2_lambda=meta["l2"], epochs=3, progress_stride=1)
    adam = orc.AdamState()
    want = []
    p = params
    for step in range(3):
        eng.train_step(xd, td)
        newp, *_ = orc.train_step(stack, p, x, target, lr=meta["lr"] * meta["decay"] ** step,
                                  optimizer="adam", adam=adam, l2_lambda=meta["l2"])
        want.append([orc.weight_update_ratio(newp[2 * l], p[2 * l]) for l in range(spec.n_linear)])
        p = newp
    got = eng.update_ratios(0, 3)
    np.testing.assert_allclose(got, np.asarray(want), rtol=2e-3 if dtype == "float32" else 1e-8)


def mnist_stack():
    sizes = [784, 256, 128, 10]
    algos = ["linear", "relu", "linear", "relu", "linear", "softmax"]
    return sizes, algos


@pytest.mark.parametrize("optimizer", ["adam", "sgd"])
def test_config2_full_size_against_live_oracle(optimizer):
    """BASELINE.json configs[1]: 784-256-128-10, softmax + CE, batch 1024, fp32; 3 steps."""
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    sizes, algos = mnist_stack()
    B = 1024
    params = det_params(sizes, True, 77)
    x = det_normalish((B, 784), 78)
    y = det_labels(B, 10, 79)
    eng = DeviceMLP(sizes, StackSpec.from_algos(algos), True, "float32", 0)
    for l in range(3):
        eng.upload_params(l, params[2 * l], params[2 * l + 1])
    xd = torch.tensor(x, dtype=torch.float32, device="cuda:0")
    yd = torch.tensor(y, dtype=torch.int32, device="cuda:0")
    eng.train_begin(optimizer=optimizer, learning_rate=0.01, decay_rate=0.9, l2_lambda=0.001, epochs=3)
    for _ in range(3):
        eng.train_step(xd, yd)
    stack = orc.parse_stack(algos)
    adam = orc.AdamState()
    p, costs = params, []
    tgt = [[int(v)] for v in y]
    for step in range(3):
        p, c, *_ = orc.train_step(stack, p, x, tgt, lr=0.01 * 0.9 ** step, optimizer=optimizer,
                                  adam=adam, l2_lambda=0.001)
        costs.append(c)
    close(eng.costs(0, 3), costs, "float32", "costs")
    for l in range(3):
        w, b = eng.download_params(l)
        atol = ADAM_PARAM_ATOL["float32"] if optimizer == "adam" else None
        lo = 0.02 if optimizer == "adam" else None
        close(w, p[2 * l], "float32", f"W{l}", atol=atol, outliers=True, loose_atol=lo)
        close(b, p[2 * l + 1], "float32", f"b{l}", atol=atol, outliers=True, loose_atol=lo)
    out = eng.forward(xd).cpu().numpy()
    ref = orc.forward(stack, p, x)[0][-1]
    close(out, ref, "float32", "probs", rtol=1e-4)
    # argmax identical wherever the reference's top-2 margin exceeds the stated tolerance
    srt = np.sort(ref, axis=1)
    decided = (srt[:, -1] - srt[:, -2]) > 1e-4
    assert decided.mean() > 0.99
    assert (out.argmax(1)[decided] == ref.argmax(1)[decided]).all()


@pytest.mark.parametrize("batch", [1, 8, 64, 512, 4096, 32768, 262144, 1048576])
def test_config3_inference_sweep(batch):
    """BASELINE.json configs[2]: batched inference on the 784-256-128-10 model, batch 1 to 1M.
    Every batch is compared with the float64 oracle at the contract's float32 tolerance -- rtol 1e-5 on every
    output (+ 1e-7 absolute, 1e-2 of a float32 ulp of a probability near 1) -- on a sample of rows spread over the
    WHOLE batch: every row up to 4,096 rows, else ~4,096 rows at a fixed stride plus the first and the last 256
    rows (first and last tiles of the first and last waves).  Every row must sum to 1; argmax identical wherever
    the reference's top-2 margin exceeds 1e-4; the tail is re-run as its own call (tiling independence)."""
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    sizes, algos = mnist_stack()
    params = det_params(sizes, True, 5)
    eng = DeviceMLP(sizes, StackSpec.from_algos(algos), True, "float32", 0)
    for l in range(3):
        eng.upload_params(l, params[2 * l], params[2 * l + 1])
    g = torch.Generator(device="cuda:0").manual_seed(batch)
    xd = torch.randn((batch, 784), generator=g, device="cuda:0", dtype=torch.float32)
    out = eng.forward(xd)
    assert out.shape == (batch, 10)
    assert torch.allclose(out.sum(1), torch.ones(batch, device="cuda:0"), atol=1e-5)
    if batch <= 4096:
        rows = torch.arange(batch, device="cuda:0")
    else:
        stride = batch // 4096 + 1                    # odd strides walk through every position inside a 128-row tile
        stride += 1 - stride % 2
        rows = torch.cat([torch.arange(0, 256, device="cuda:0"), torch.arange(0, batch, stride, device="cuda:0"),
                          torch.arange(batch - 256, batch, device="cuda:0")]).unique()
    ref = orc.forward(orc.parse_stack(algos), params, xd[rows].cpu().numpy().astype(np.float64))[0][-1]
    got = out[rows].cpu().numpy().astype(np.float64).reshape(ref.shape)
    bad = np.abs(got - ref) > 1e-5 * np.abs(ref) + 1e-7
    assert not bad.any(), (batch, int(bad.sum()), float((np.abs(got - ref) / np.abs(ref)).max()))
    ref2, got2 = ref.reshape(-1, 10), got.reshape(-1, 10)
    srt = np.sort(ref2, axis=1)
    decided = (srt[:, -1] - srt[:, -2]) > 1e-4
    assert (got2.argmax(1)[decided] == ref2.argmax(1)[decided]).all()
    if batch > 4096:
        tail = xd[batch - 4096:].contiguous()
        again = eng.forward(tail)
        # the re-run may take another kernel variant than the large call: both fp32-grade, not bit-identical
        diff = ((out[batch - 4096:] - again).abs() / again.abs().clamp_min(1e-6)).max().item()
        assert torch.allclose(out[batch - 4096:], again, rtol=1e-5, atol=1e-7), diff
    eng.close()


@pytest.mark.parametrize("sizes,algos,has_bias", [
    ([784, 256, 128, 10], ["linear", "relu", "linear", "relu", "linear", "softmax"], True),
    ([13, 37, 5], ["linear", "sigmoid", "linear", "tanh"], False),
    ([9, 70, 33, 21, 4], ["linear", "tanh", "linear", "relu", "linear", "sigmoid", "linear", "softmax"], True),
    ([300, 1000, 7], ["linear", "relu", "linear"], True),
])
@pytest.mark.parametrize("batch", [1, 3, 4])
def test_few_row_inference_in_one_cluster_launch(sizes, algos, has_bias, batch):
    """Up to 4 rows of a plain float32 stack go through infer_small_kernel (one cluster, the layers
    chained through distributed shared memory): against the oracle, and against the per-layer kernels
    (which a 17-row call of the same rows uses)."""
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    params = det_params(sizes, has_bias, 611)
    nl = len(sizes) - 1
    per = 2 if has_bias else 1
    eng = DeviceMLP(sizes, StackSpec.from_algos(algos), has_bias, "float32", 0)
    for l in range(nl):
        eng.upload_params(l, params[per * l], params[per * l + 1] if has_bias else None)
    x = det_normalish((17, sizes[0]), 612)
    xd = torch.tensor(x, dtype=torch.float32, device="cuda:0")
    before = launch_count()
    out = eng.forward(xd[:batch].contiguous())
    assert launch_count() - before == 1, "the few-row path is a single launch"
    ref = orc.forward(orc.parse_stack(algos), params, x[:batch])[0][-1]
    close(out.cpu().numpy(), ref.reshape(batch, -1), "float32", "output", rtol=1e-4)
    layered = eng.forward(xd)[:batch]
    assert torch.allclose(out, layered, rtol=2e-5, atol=1e-6)
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):                      # replayed from a graph on a side stream: same numbers
        outs = [eng.forward(xd[:batch].contiguous()) for _ in range(3)]
    side.synchronize()
    assert all(torch.equal(o, out) for o in outs)


@pytest.mark.parametrize("sizes,batch", [([100, 200, 70, 5], 8200), ([784, 256, 128, 10], 8192 + 77), ([132, 136, 64], 9000)])
def test_large_batch_inference_first_layer_from_fp32_rows(sizes, batch):
    """From 8192 rows the first layer of the tcgen05 inference path reads the fp32 rows itself and builds
    its bf16 planes in shared memory (tc3_fwd_x32_kernel): ragged row counts, a contraction that is not a
    multiple of the 64-wide k-block and an output width that is not a multiple of the 128-wide tile, against
    the oracle on every row."""
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    nl = len(sizes) - 1
    algos = []
    for l in range(nl):
        algos += ["linear", "softmax" if l == nl - 1 else ("relu" if l % 2 == 0 else "tanh")]
    params = det_params(sizes, True, 909)
    eng = DeviceMLP(sizes, StackSpec.from_algos(algos), True, "float32", 0)
    for l in range(nl):
        eng.upload_params(l, params[2 * l], params[2 * l + 1])
    x = det_normalish((batch, sizes[0]), 910)
    xd = torch.tensor(x, dtype=torch.float32, device="cuda:0")
    eng.profile_begin()
    out = eng.forward(xd)
    tags = eng.profile_end()
    assert "tc3_fwd_l0" in tags and "split_x" not in tags, tags      # no separate plane split of the input
    ref = orc.forward(orc.parse_stack(algos), params, x)[0][-1]
    close(out.cpu().numpy(), ref, "float32", "probs", rtol=1e-4)
    assert (out.cpu().numpy().argmax(1) == ref.argmax(1)).mean() > 0.999


@pytest.mark.parametrize("sizes,last_act,has_bias,batch", [
    ([784, 256, 128, 10], "softmax", True, 2048 + 5),
    ([90, 160, 100, 16], "sigmoid", True, 3000),
    ([64, 72, 1], "tanh", False, 2500),
    ([200, 96, 3], "softmax", False, 8192 + 130),          # the head in the fp32-input kernel itself
    ([784, 256, 128, 10], "softmax", True, 8192 + 300),    # fp32 rows from unit 0 to unit 1, head in unit 1
    ([96, 130, 72, 100, 12], "tanh", True, 9000),          # 130 % 4 != 0: that hand-over stays bf16 planes
])
def test_large_batch_inference_last_unit_in_the_epilogue(sizes, last_act, has_bias, batch):
    """A last unit of at most 16 outputs behind a 65..128-wide one is applied by the epilogue warps of that
    unit's tcgen05 GEMM (fp32 FMAs on the tile's rows, then the activation / softmax): no launch of its own,
    the last hidden activation is never written.  Against the oracle on every row."""
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    nl = len(sizes) - 1
    algos = []
    for l in range(nl):
        algos += ["linear", last_act if l == nl - 1 else ("relu" if l % 2 == 0 else "sigmoid")]
    params = det_params(sizes, has_bias, 1201)
    per = 2 if has_bias else 1
    eng = DeviceMLP(sizes, StackSpec.from_algos(algos), has_bias, "float32", 0)
    for l in range(nl):
        eng.upload_params(l, params[per * l], params[per * l + 1] if has_bias else None)
    x = det_normalish((batch, sizes[0]), 1202)
    xd = torch.tensor(x, dtype=torch.float32, device="cuda:0")
    eng.profile_begin()
    out = eng.forward(xd)
    tags = eng.profile_end()
    assert f"tc3_fwd_l{nl - 2}_head" in tags and f"tc3_fwd_l{nl - 1}" not in tags and "softmax" not in tags, tags
    ref = orc.forward(orc.parse_stack(algos), params, x)[0][-1]
    close(out.cpu().numpy(), ref.reshape(batch, -1), "float32", "output", rtol=1e-4)
    if sizes[-1] > 1:
        assert (out.cpu().numpy().argmax(1) == ref.reshape(batch, -1).argmax(1)).mean() > 0.999


@pytest.mark.parametrize("batch", [1, 4, 300, 8192 + 77])
def test_inference_writes_only_its_output_rows(batch):
    """Guard bands around the caller's output buffer stay untouched on every inference path (few-row
    cluster kernel, per-layer SIMT kernels, tcgen05 path with the fp32-input first layer)."""
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    sizes, algos = mnist_stack()
    params = det_params(sizes, True, 5)
    eng = DeviceMLP(sizes, StackSpec.from_algos(algos), True, "float32", 0)
    for l in range(3):
        eng.upload_params(l, params[2 * l], params[2 * l + 1])
    g = torch.Generator(device="cuda:0").manual_seed(batch)
    guard = 4096
    xbuf = torch.full((guard + batch * 784 + guard,), float("nan"), device="cuda:0", dtype=torch.float32)
    x = xbuf[guard:guard + batch * 784].view(batch, 784)
    x.copy_(torch.randn((batch, 784), generator=g, device="cuda:0"))
    obuf = torch.full((guard + batch * 10 + guard,), -7.0, device="cuda:0", dtype=torch.float32)
    out = obuf[guard:guard + batch * 10].view(batch, 10)
    eng.forward(x, out)
    torch.cuda.synchronize()
    assert torch.all(obuf[:guard] == -7.0) and torch.all(obuf[guard + batch * 10:] == -7.0)
    assert torch.isfinite(out).all()             # no NaN from the guard bands of the input leaked in
    assert torch.allclose(out.sum(1), torch.ones(batch, device="cuda:0"), atol=1e-5)
    ref = eng.forward(x.clone())
    assert torch.equal(out, ref)


def test_row_gather_equals_materialised_batch():
    """The in-kernel row gather (replacing neural_net_model.py:460-461,472) is bit-identical to
    training on the gathered copy."""
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    sizes, algos = [31, 20, 7], ["linear", "tanh", "linear", "softmax"]
    params = det_params(sizes, True, 9)
    pool = torch.tensor(det_normalish((500, 31), 10), dtype=torch.float32, device="cuda:0")
    labels = torch.tensor(det_labels(500, 7, 11), dtype=torch.int32, device="cuda:0")
    idx = torch.tensor(det_labels(257, 500, 12), dtype=torch.int32, device="cuda:0")
    res = []
    for gather in (True, False):
        eng = DeviceMLP(sizes, StackSpec.from_algos(algos), True, "float32", 0)
        for l in range(2):
            eng.upload_params(l, params[2 * l], params[2 * l + 1])
        eng.train_begin(optimizer="adam", learning_rate=0.01, l2_lambda=0.001, epochs=2)
        for _ in range(2):
            if gather:
                eng.train_step(pool, labels, row_idx=idx)
            else:
                eng.train_step(pool[idx.long()].contiguous(), labels[idx.long()].contiguous())
        res.append([eng.download_params(l) for l in range(2)] + [eng.costs(0, 2)])
    for a, b in zip(res[0][:2], res[1][:2]):
        np.testing.assert_array_equal(a[0], b[0])
        np.testing.assert_array_equal(a[1], b[1])
    np.testing.assert_allclose(res[0][2], res[1][2], rtol=1e-12)


def test_in_kernel_dropout_statistics_and_determinism():
    """Philox masks: keep-rate ~ 1-p on every hidden entry, kept values scaled by 1/(1-p),
    backward sees the same mask as forward (zero gradient exactly where the output was dropped)."""
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    sizes, algos = [64, 96, 80, 10], ["linear", "sigmoid", "linear", "tanh", "linear", "softmax"]
    spec = StackSpec.from_algos(algos)
    params = det_params(sizes, True, 21)
    B, p = 2048, 0.3
    xd = torch.tensor(det_normalish((B, 64), 22), dtype=torch.float32, device="cuda:0")
    yd = torch.tensor(det_labels(B, 10, 23), dtype=torch.int32, device="cuda:0")
    runs = []
    for seed in (1, 1, 2):
        eng = DeviceMLP(sizes, spec, True, "float32", 0)
        for l in range(3):
            eng.upload_params(l, params[2 * l], params[2 * l + 1])
        eng.train_begin(optimizer="sgd", learning_rate=0.01, dropout_rate=p, epochs=1, dropout_seed=seed)
        eng.train_step(xd, yd, diagnostics=True)
        a1 = eng.diag_activation(0, 1, B)       # sigmoid output, hidden -> dropped
        g1 = eng.diag_activation_grad(0, 1, B)
        z2 = eng.diag_activation(1, 0, B)       # second linear output, hidden -> dropped
        runs.append((a1, z2, eng.costs(0, 1)[0]))
        keep = a1 != 0.0
        assert abs(keep.mean() - (1 - p)) < 0.01
        assert a1[keep].min() >= 0.0 and a1[keep].max() <= 1.0 / (1 - p) + 1e-5
        assert abs((z2 != 0.0).mean() - (1 - p)) < 0.01
        assert g1 is not None
    np.testing.assert_array_equal(runs[0][0], runs[1][0])       # same seed -> same masks
    assert (runs[0][0] != runs[2][0]).any()                     # different seed -> different masks
    assert np.isfinite(runs[0][2])


@pytest.mark.parametrize("B", [2304 + 40, 8192])
def test_large_batch_training_with_first_layer_on_tensor_cores(B):
    """Batches from the threshold (2,304 rows) at which the first layer's forward / weight-gradient GEMMs run on the
    persistent split-bf16 tcgen05 kernels: three steps against the live oracle, same float32 tolerance."""
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    sizes, algos = mnist_stack()
    params = det_params(sizes, True, 177)
    x = det_normalish((B, 784), 178)
    y = det_labels(B, 10, 179)
    eng = DeviceMLP(sizes, StackSpec.from_algos(algos), True, "float32", 0)
    for l in range(3):
        eng.upload_params(l, params[2 * l], params[2 * l + 1])
    xd = torch.tensor(x, dtype=torch.float32, device="cuda:0")
    yd = torch.tensor(y, dtype=torch.int32, device="cuda:0")
    eng.train_begin(optimizer="sgd", learning_rate=0.05, decay_rate=0.9, l2_lambda=0.001, epochs=3)
    eng.profile_begin()
    for _ in range(3):
        eng.train_step(xd, yd)
    assert "tc3_bwd_w_l0" in eng.profile_end()
    stack = orc.parse_stack(algos)
    p, costs = params, []
    tgt = [[int(v)] for v in y]
    for step in range(3):
        p, c, *_ = orc.train_step(stack, p, x, tgt, lr=0.05 * 0.9 ** step, optimizer="sgd", l2_lambda=0.001)
        costs.append(c)
    close(eng.costs(0, 3), costs, "float32", "costs")
    for l in range(3):
        w, b = eng.download_params(l)
        close(w, p[2 * l], "float32", f"W{l}", outliers=True)
        close(b, p[2 * l + 1], "float32", f"b{l}", outliers=True)


@pytest.mark.parametrize("dtype", ["float32", "float64"])
def test_device_statistics_equal_host_torch_statistics(dtype):
    """The device-side reductions behind `stats` (neural_net_model.py:542-583) against the reference's
    own host formulas (mean / unbiased std / saturation rule / torch.histogram(density=True)) applied to
    the downloaded tensors: histogram densities identical, moments to 1e-10."""
    arrays, meta = load_golden("step_odd_sizes_adam")          # tanh, sigmoid, softmax; batch 67
    eng, spec, params, x, xd, target, td = build_case(meta, dtype)
    eng.train_begin(optimizer="adam", learning_rate=0.01, l2_lambda=0.001, epochs=1)
    eng.train_step(xd, td, diagnostics=True)
    B = meta["batch"]
    rules = {"tanh": (1, 0.97), "sigmoid": (1, 0.97), "relu": (2, 0.0), "softmax": (3, 0.97), "none": (1, 5.0)}

    def host(t, rule, thr):
        t = torch.from_numpy(t)
        h = torch.histogram(t, density=True)
        sat = ((t.abs() > thr) if rule == 1 else (t <= 0) if rule == 2 else (t.max(dim=-1).values > thr)).float().mean().item()
        return t.mean().item(), t.std().item(), sat, h.bin_edges[:-1].tolist(), h.hist.tolist()

    def check(dev, t, rule=1, thr=5.0, sat=True):
        m, s, st, hx, hy = host(t, rule, thr)
        np.testing.assert_allclose(dev["mean"], m, rtol=1e-10, atol=1e-14)
        np.testing.assert_allclose(dev["std"], s, rtol=1e-10)
        if sat:
            np.testing.assert_allclose(dev["saturated"], st, rtol=1e-6)
        np.testing.assert_allclose(dev["histogram"]["x"], hx, rtol=0, atol=0)
        np.testing.assert_allclose(dev["histogram"]["y"], hy, rtol=1e-12)

    for l in range(spec.n_linear):
        check(eng.diag_stats(l, 0, 0, 1, 5.0), eng.diag_activation(l, 0, B), 1, 5.0)
        check(eng.diag_stats(l, 0, 1), eng.diag_activation_grad(l, 0, B), sat=False)
        act = spec.acts[l]
        rule, thr = rules[act]
        check(eng.diag_stats(l, 1, 0, rule, thr), eng.diag_activation(l, 1, B), rule, thr)
        g = eng.diag_activation_grad(l, 1, B)
        d = eng.diag_stats(l, 1, 1)
        assert (g is None) == (d is None)
        if g is not None:
            check(d, g, sat=False)
        wg, _ = eng.diag_weight_grad(l)
        check(eng.diag_stats(l, 0, 3), wg, sat=False)
        w, _ = eng.download_params(l)
        dw = eng.diag_stats(l, 0, 2, histogram=False)
        np.testing.assert_allclose([dw["mean"], dw["std"]], [w.mean(), w.std(ddof=1)], rtol=1e-10, atol=1e-14)


def test_host_buffer_steps_equal_device_buffer_steps():
    """poz_train_step_host (pinned host batch in, the cost of the step two calls earlier out) trains
    exactly like poz_train_step on device-resident copies of the same batches."""
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    sizes, algos = [40, 24, 6], ["linear", "relu", "linear", "softmax"]
    params = det_params(sizes, True, 301)
    batches = [(torch.tensor(det_normalish((96, 40), 310 + i), dtype=torch.float32).pin_memory(),
                torch.tensor(det_labels(96, 6, 320 + i), dtype=torch.int32).pin_memory()) for i in range(5)]
    runs = []
    for host in (True, False):
        eng = DeviceMLP(sizes, StackSpec.from_algos(algos), True, "float32", 0)
        for l in range(2):
            eng.upload_params(l, params[2 * l], params[2 * l + 1])
        eng.train_begin(optimizer="adam", learning_rate=0.01, l2_lambda=0.001, epochs=5)
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            if host:
                got = [eng.train_step_host(x, y) for x, y in batches]
                assert all(np.isnan(v) for v in got[:eng.HOST_LAG])
                got = got[eng.HOST_LAG:] + eng.train_host_flush()
            else:
                for x, y in batches:
                    eng.train_step(x.cuda(), y.cuda())
                got = list(eng.costs(0, 5))
        side.synchronize()
        runs.append((got, eng.costs(0, 5), [eng.download_params(l) for l in range(2)]))
    np.testing.assert_allclose(runs[0][0], runs[1][0], rtol=1e-12)
    np.testing.assert_allclose(runs[0][1], runs[1][1], rtol=1e-12)


@pytest.mark.parametrize("batch", [3, 300, 4096 + 5])
def test_a_nan_input_row_gives_a_nan_output_row_as_in_the_reference(batch):
    """torch.relu keeps a NaN (neural_net_model.py:179): a row with a NaN in it comes out all NaN -- through the few-row
    cluster launch, the SIMT kernels and the tensor-core kernels alike -- and the other rows are untouched."""
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    sizes, algos = mnist_stack()
    params = det_params(sizes, True, 5)
    eng = DeviceMLP(sizes, StackSpec.from_algos(algos), True, "float32", 0)
    for l in range(3):
        eng.upload_params(l, params[2 * l], params[2 * l + 1])
    x = det_normalish((batch, 784), 61)
    bad = batch // 2
    x[bad, 17] = np.nan
    out = eng.forward(torch.tensor(x, dtype=torch.float32, device="cuda:0")).cpu().numpy()
    ref = orc.forward(orc.parse_stack(algos), params, x)[0][-1].reshape(batch, -1)
    assert np.isnan(ref[bad]).all() and np.isnan(out[bad]).all()
    ok = np.arange(batch) != bad
    close(out[ok], ref[ok], "float32", "probs", rtol=1e-4)
    eng.close()
