"""The bf16 / tcgen05 context (wide layers): one train step against
  (a) a plain PyTorch fp32 emulation of the same mixed-precision recipe (bf16 weights and
      activations, fp32 accumulation, fp32 master weights) -- tight, and
  (b) the float64 oracle -- within the stated bf16 tolerance: outputs rtol 5e-2 + atol 5e-2*max,
      cost rtol 2e-2.
"""
import numpy as np
import pytest
import torch

from oracle import mlp_oracle as orc
from oracle.detgen import det_normalish, det_params, det_uniform

pytestmark = pytest.mark.gpu

SIZES = [512, 768, 512, 256]
ALGOS = ["linear", "relu", "linear", "sigmoid", "linear"]
ACTS = ["relu", "sigmoid", "none"]
B = 1024


def bf(t):
    return t.to(torch.bfloat16).float()


def act(name, z):
    return z.relu() if name == "relu" else z.sigmoid() if name == "sigmoid" else z.tanh() if name == "tanh" else z


def dact(name, a):
    return (a > 0).float() if name == "relu" else a * (1 - a) if name == "sigmoid" else 1 - a * a if name == "tanh" else torch.ones_like(a)


def emulate_step(params, x, t, lr, l2):
    """fp32 torch emulation of the device recipe; returns (cost, grads, final output)."""
    ws = [p.clone() for p in params[0::2]]
    bs = [p.clone() for p in params[1::2]]
    a = [bf(x)]
    for l, name in enumerate(ACTS):
        a.append(bf(act(name, a[-1] @ bf(ws[l]) + bs[l])))
    tb = bf(t)
    cost = ((a[-1] - tb) ** 2).mean().item() + l2 * sum((w ** 2).sum().item() for w in ws)
    dz = bf(2.0 * (a[-1] - tb) / a[-1].numel() * dact(ACTS[-1], a[-1]))
    grads = [None] * 6
    for l in range(2, -1, -1):
        grads[2 * l] = a[l].T @ dz + 2 * l2 * ws[l]
        grads[2 * l + 1] = dz.sum(0)
        if l > 0:
            dz = bf((dz @ bf(ws[l]).T) * dact(ACTS[l - 1], a[l]))
    return cost, grads, a[-1]


def make(optimizer):
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    params = det_params(SIZES, True, 91)
    eng = DeviceMLP(SIZES, StackSpec.from_algos(ALGOS), True, "bfloat16", 0)
    for l in range(3):
        eng.upload_params(l, params[2 * l], params[2 * l + 1])
    x = det_normalish((B, SIZES[0]), 92)
    t = det_uniform((B, SIZES[-1]), 93, 0.0, 1.0)
    xd = torch.tensor(x, dtype=torch.bfloat16, device="cuda:0")
    td = torch.tensor(t, dtype=torch.bfloat16, device="cuda:0")
    return eng, params, x, t, xd, td


def test_sgd_step_matches_mixed_precision_emulation():
    eng, params, x, t, xd, td = make("sgd")
    lr, l2 = 0.5, 0.001
    eng.train_begin(optimizer="sgd", learning_rate=lr, decay_rate=1.0, l2_lambda=l2, epochs=1)
    eng.train_step(xd, td)
    p32 = [torch.tensor(p, dtype=torch.float32, device="cuda:0") for p in params]
    cost, grads, _ = emulate_step(p32, xd.float(), td.float(), lr, l2)
    np.testing.assert_allclose(eng.costs(0, 1)[0], cost, rtol=1e-3)
    for l in range(3):
        w, b = eng.download_params(l)
        for got, p, g in ((w, p32[2 * l], grads[2 * l]), (b, p32[2 * l + 1], grads[2 * l + 1])):
            delta = (torch.tensor(got, device="cuda:0", dtype=torch.float32) - p) / -lr     # = gradient
            scale = g.abs().max().item()
            # bf16 rounding of activations / gradients can differ by one ulp between the two
            # implementations (different fp32 summation order before the rounding)
            torch.testing.assert_close(delta, g, rtol=2e-2, atol=2e-2 * scale)


def test_adam_steps_and_inference_against_oracle_within_bf16_tolerance():
    eng, params, x, t, xd, td = make("adam")
    eng.train_begin(optimizer="adam", learning_rate=0.001, decay_rate=0.9, l2_lambda=0.001, epochs=3)
    for _ in range(3):
        eng.train_step(xd, td)
    stack = orc.parse_stack(ALGOS)
    adam = orc.AdamState()
    xb, tb = xd.float().cpu().numpy().astype(np.float64), td.float().cpu().numpy().astype(np.float64)
    p, costs = params, []
    for s in range(3):
        p, c, *_ = orc.train_step(stack, p, xb, tb.tolist(), lr=0.001 * 0.9 ** s, optimizer="adam", adam=adam,
                                  l2_lambda=0.001)
        costs.append(c)
    np.testing.assert_allclose(eng.costs(0, 3), costs, rtol=2e-2)
    assert eng.get_adam_step() == 3
    out, cost = eng.forward_loss(xd, td)
    ref = orc.forward(stack, p, xb)[0][-1]
    got = out.float().cpu().numpy()
    np.testing.assert_allclose(got, ref, rtol=5e-2, atol=5e-2 * np.abs(ref).max())
    np.testing.assert_allclose(cost, np.mean((ref - tb) ** 2), rtol=2e-2)
    out2 = eng.forward(xd)
    assert torch.equal(out, out2)


def test_unsupported_requests_fail_loudly():
    from penr_oz_neural_network_torch_b200 import DeviceMLP, PozError, StackSpec
    with pytest.raises(PozError):
        DeviceMLP([784, 256, 10], StackSpec.from_algos(["linear", "relu", "linear", "softmax"]), True, "bfloat16", 0)
    with pytest.raises(PozError):          # layer sizes must be multiples of 8 (16-byte TMA rows)
        DeviceMLP([100, 50, 20], StackSpec.from_algos(["linear", "relu", "linear"]), True, "bfloat16", 0)


def test_dropout_and_row_gather_in_the_bf16_context():
    """Philox dropout in the tcgen05 epilogues (forward masks regenerated in the backward) and the row gather of
    train()'s sampling: with p = 0 a gathered step equals the step on the materialised rows bit for bit; with
    p = 0.3 the step runs, is reproducible for a seed, differs between seeds, and the cost stays in range."""
    eng, params, x, t, xd, td = make("adam")
    idx = torch.randperm(B, device="cuda:0")[: B // 2].to(torch.int32).contiguous()
    eng.train_begin(optimizer="sgd", learning_rate=0.0, decay_rate=1.0, epochs=8)
    eng.train_step(xd, td, row_idx=idx)
    eng.train_step(xd[idx.long()].contiguous(), td[idx.long()].contiguous())
    c = eng.costs(0, 2)
    assert c[0] == c[1]
    costs = []
    for seed in (5, 5, 6):
        eng.train_begin(optimizer="sgd", learning_rate=0.0, decay_rate=1.0, dropout_rate=0.3, dropout_seed=seed, epochs=1)
        eng.train_step(xd, td)
        costs.append(float(eng.costs(0, 1)[0]))
    assert costs[0] == costs[1] and costs[0] != costs[2]
    assert 0.5 * c[0] < costs[0] < 3.0 * c[0]
