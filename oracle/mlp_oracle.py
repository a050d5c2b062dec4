"""CPU oracle for the NeuralNetworkModel hot path -- TEST INFRASTRUCTURE ONLY.

This file restates, in closed form and plain numpy float64, the arithmetic the reference
performs through torch on its train-step / inference path.  It is the checker the CUDA path is
compared with; nothing under the product package imports it.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may
import it.

Parity status: PINNED.  The reference's own tests hold no golden vectors for this path
(SURVEY.md section 8c), so the oracle is pinned against outputs of the reference itself, generated in
the build container by ``oracle/make_golden.py`` (which imports ``/root/reference``) and committed
under ``tests/golden/``; ``tests/test_oracle_golden.py`` re-checks them on every CPU run.

Reference citations are ``neural_net_model.py:<line>`` in derinworks/penr-oz-neural-network-torch.
The closed-form gradients follow the reference's own hand derivation in
``test_torch_backward.py:163-170``.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

ACTIVATIONS = ("relu", "sigmoid", "tanh", "softmax")


@dataclass
class LinearSpec:
    """One Linear layer of the stack plus the activation entry that follows it (if any)."""
    z_index: int                 # position of the "linear" entry in the algos list
    act: str = "none"            # "none" | "relu" | "sigmoid" | "tanh" | "softmax"
    a_index: int | None = None   # position of the activation entry, None if act == "none"
    drop_z: bool = False         # reference applies dropout to this linear's output
    drop_a: bool = False         # reference applies dropout to the activation's output


@dataclass
class Stack:
    algos: list[str]
    linears: list[LinearSpec] = field(default_factory=list)

    @property
    def loss(self) -> str:
        # neural_net_model.py:400-406 -- cross-entropy iff the last entry is softmax, else MSE
        return "ce" if self.algos[-1] == "softmax" else "mse"


def expand_algos(activation_algos: list[str]) -> list[str]:
    """The reference's "linear sandwich" (neural_net_model.py:202-212), Linear/activation part."""
    out: list[str] = []
    for i in range(len(activation_algos) - 1, -1, -1):
        algo = activation_algos[i]
        out.insert(0, algo)
        if algo in ACTIVATIONS and (i == 0 or activation_algos[i - 1] not in ("batchnorm", "linear")):
            out.insert(0, "linear")
    return out


def parse_stack(algos: list[str]) -> Stack:
    """Group an expanded algos list into Linear(+activation) units and derive the dropout points.

    ``hidden = 0 < i < n-1`` is neural_net_model.py:255; dropout is applied to the output of every
    hidden entry when a target is given (neural_net_model.py:393-395), which includes the
    pre-activation outputs of every Linear except the first.
    """
    n = len(algos)
    st = Stack(list(algos))
    i = 0
    while i < n:
        if algos[i] != "linear":
            raise ValueError(f"oracle covers Linear+activation stacks only, got {algos}")
        spec = LinearSpec(z_index=i, drop_z=(0 < i < n - 1))
        if i + 1 < n and algos[i + 1] in ACTIVATIONS:
            spec.act = algos[i + 1]
            spec.a_index = i + 1
            spec.drop_a = (0 < i + 1 < n - 1)
            i += 2
        else:
            i += 1
        st.linears.append(spec)
    return st


def split_params(stack: Stack, params: list[np.ndarray]):
    """Flat [W0,(b0),W1,(b1),...] (neural_net_model.py:262-267) -> per-linear (W, b|None)."""
    nl = len(stack.linears)
    if len(params) == nl:
        return [(params[i], None) for i in range(nl)]
    if len(params) == 2 * nl:
        return [(params[2 * i], params[2 * i + 1]) for i in range(nl)]
    raise ValueError("params do not match stack")


def _act(name: str, z: np.ndarray) -> np.ndarray:
    if name == "relu":      # neural_net_model.py:179
        return np.maximum(z, 0.0)
    if name == "sigmoid":   # neural_net_model.py:174
        return 1.0 / (1.0 + np.exp(-z))
    if name == "tanh":      # neural_net_model.py:184
        return np.tanh(z)
    if name == "softmax":   # neural_net_model.py:188
        m = z.max(axis=-1, keepdims=True)
        e = np.exp(z - m)
        return e / e.sum(axis=-1, keepdims=True)
    return z


def forward(stack: Stack, params: list[np.ndarray], x: np.ndarray, *, training: bool = False,
            dropout_rate: float = 0.0, masks: dict[int, np.ndarray] | None = None):
    """neural_net_model.py:384-396.  Returns the list of every entry's output (as the reference
    does) and the tensor that went *into* the last entry (the reference's ``logits``).

    ``masks`` maps an algos index to a 0/1 keep-mask; used only when training with
    ``dropout_rate > 0`` (F.dropout: ``x * mask / (1-p)``, neural_net_model.py:395).
    """
    x = np.asarray(x, dtype=np.float64)
    outs: list[np.ndarray] = []
    cur = x
    into_last = x
    scale = 1.0 / (1.0 - dropout_rate) if dropout_rate > 0.0 else 1.0

    def drop(idx: int, t: np.ndarray) -> np.ndarray:
        if training and dropout_rate > 0.0:
            return t * masks[idx] * scale
        return t

    for (w, b), spec in zip(split_params(stack, params), stack.linears):
        into_last = cur
        z = cur @ w                      # neural_net_model.py:117
        if b is not None:
            z = z + b                    # neural_net_model.py:119
        if spec.drop_z:
            z = drop(spec.z_index, z)
        outs.append(z)
        cur = z
        if spec.a_index is not None:
            into_last = cur
            a = _act(spec.act, cur)
            if spec.drop_a:
                a = drop(spec.a_index, a)
            outs.append(a)
            cur = a
    return outs, into_last


def labels_from_target(target) -> np.ndarray:
    """neural_net_model.py:401-402: label = first element of each target row, cast to int64."""
    t = np.asarray(target)
    if t.ndim == 1:
        return np.asarray([int(t[0])], dtype=np.int64)
    return t[:, 0].astype(np.int64)


def loss_value(stack: Stack, outs, into_last, target) -> float:
    if stack.loss == "ce":               # neural_net_model.py:400-403
        z = into_last if into_last.ndim == 2 else into_last[None, :]
        y = labels_from_target(target)
        m = z.max(axis=-1, keepdims=True)
        lse = m[:, 0] + np.log(np.exp(z - m).sum(axis=-1))
        return float(np.mean(lse - z[np.arange(z.shape[0]), y]))
    t = np.asarray(target, dtype=np.float64)   # neural_net_model.py:405-406
    return float(np.mean((outs[-1] - t) ** 2))


def l2_penalty(stack: Stack, params, l2_lambda: float) -> float:
    """neural_net_model.py:482-483: weights only, never biases."""
    if l2_lambda <= 0.0:
        return 0.0
    return l2_lambda * sum(float((w ** 2).sum()) for w, _ in split_params(stack, params))


def backward(stack: Stack, params, x, outs, target, *, dropout_rate: float = 0.0,
             masks: dict[int, np.ndarray] | None = None, l2_lambda: float = 0.0):
    """Closed-form gradient of ``loss + l2`` w.r.t. every parameter and every entry output.

    Returns (param_grads in flat params order, out_grads aligned with ``outs``; an entry that is
    not on the loss graph -- the trailing softmax under cross-entropy -- gets ``None``, matching
    ``a.grad is None`` at neural_net_model.py:545).
    """
    x = np.asarray(x, dtype=np.float64)
    x2 = x if x.ndim == 2 else x[None, :]
    pl = split_params(stack, params)
    nl = len(pl)
    batch = x2.shape[0]
    scale = 1.0 / (1.0 - dropout_rate) if dropout_rate > 0.0 else 1.0
    use_drop = dropout_rate > 0.0
    out_grads: list[np.ndarray | None] = [None] * len(outs)

    def as2(t):
        return t if t.ndim == 2 else t[None, :]

    # gradient w.r.t. the *recorded* (post-dropout) output of the last unit
    last = stack.linears[-1]
    if stack.loss == "ce":
        z = as2(outs[last.z_index])
        y = labels_from_target(target)
        probs = _act("softmax", z)
        g = probs / batch                           # test_torch_backward.py:163
        g[np.arange(batch), y] -= 1.0 / batch       # test_torch_backward.py:164
        out_grads[last.z_index] = g
        g_rec_z = g                                 # grad of recorded z of the last linear
        skip_act = True
    else:
        yhat = as2(outs[-1])
        t = as2(np.asarray(target, dtype=np.float64))
        g = 2.0 * (yhat - t) / yhat.size            # d mean((y-t)^2)
        skip_act = False
        g_rec_z = None

    grads: list[np.ndarray] = []
    upstream = g        # grad of the recorded output of the current unit's last entry
    for li in range(nl - 1, -1, -1):
        spec = stack.linears[li]
        w, b = pl[li]
        if li == nl - 1 and skip_act:
            gz_rec = g_rec_z
        else:
            if spec.a_index is not None:
                out_grads[spec.a_index] = upstream
                ga = upstream
                if spec.drop_a and use_drop:
                    ga = ga * masks[spec.a_index] * scale
                # recover the *pre-dropout* activation value for the derivative
                a_rec = as2(outs[spec.a_index])
                z_rec = as2(outs[spec.z_index])
                if spec.act == "relu":
                    gz_rec = ga * (z_rec > 0.0)
                elif spec.act == "sigmoid":
                    s = _act("sigmoid", z_rec)
                    gz_rec = ga * s * (1.0 - s)
                elif spec.act == "tanh":
                    th = np.tanh(z_rec)
                    gz_rec = ga * (1.0 - th * th)   # test_torch_backward.py:170
                elif spec.act == "softmax":
                    s = _act("softmax", z_rec)
                    gz_rec = s * (ga - (ga * s).sum(axis=-1, keepdims=True))
                else:
                    gz_rec = ga
                del a_rec
            else:
                gz_rec = upstream
        out_grads[spec.z_index] = gz_rec
        gz = gz_rec
        if spec.drop_z and use_drop:
            gz = gz * masks[spec.z_index] * scale
        inp = x2 if li == 0 else as2(outs[stack.linears[li - 1].a_index
                                          if stack.linears[li - 1].a_index is not None
                                          else stack.linears[li - 1].z_index])
        gw = inp.T @ gz                             # test_torch_backward.py:165
        if l2_lambda > 0.0:
            gw = gw + 2.0 * l2_lambda * w
        gb = gz.sum(axis=0) if b is not None else None   # test_torch_backward.py:166
        if b is not None:
            grads.insert(0, gb)
        grads.insert(0, gw)
        if li > 0:
            upstream = gz @ w.T                     # test_torch_backward.py:168
    if x.ndim == 1:
        out_grads = [None if g_ is None else g_[0] for g_ in out_grads]
    return grads, out_grads


def sgd_update(params, grads, lr: float):
    """neural_net_model.py:499-500."""
    return [p - lr * g for p, g in zip(params, grads)]


@dataclass
class AdamState:
    step: int = 0
    m: list[np.ndarray] = field(default_factory=list)
    v: list[np.ndarray] = field(default_factory=list)


def adam_update(params, grads, state: AdamState, lr: float, beta1=0.9, beta2=0.999, eps=1e-8):
    """torch.optim.Adam single-tensor CPU path as driven by neural_net_model.py:445-448,467-469,497
    (torch/optim/adam.py ``_single_tensor_adam``): lerp of exp_avg, mul+addcmul of exp_avg_sq,
    ``denom = sqrt(v)/sqrt(1-b2^t) + eps``, ``p -= (lr/(1-b1^t)) * m/denom``.  No weight decay,
    no amsgrad."""
    if not state.m:
        state.m = [np.zeros_like(p) for p in params]
        state.v = [np.zeros_like(p) for p in params]
    state.step += 1
    t = state.step
    bc1 = 1.0 - beta1 ** t
    bc2 = 1.0 - beta2 ** t
    step_size = lr / bc1
    bc2_sqrt = math.sqrt(bc2)
    new_params = []
    for i, (p, g) in enumerate(zip(params, grads)):
        state.m[i] = state.m[i] + (g - state.m[i]) * (1.0 - beta1)
        state.v[i] = state.v[i] * beta2 + (1.0 - beta2) * g * g
        denom = np.sqrt(state.v[i]) / bc2_sqrt + eps
        new_params.append(p - step_size * (state.m[i] / denom))
    return new_params


def train_step(stack: Stack, params, x, target, *, lr: float, optimizer: str = "adam",
               adam: AdamState | None = None, l2_lambda: float = 0.0, dropout_rate: float = 0.0,
               masks=None, beta1=0.9, beta2=0.999, eps=1e-8):
    """One iteration of the loop body neural_net_model.py:472-500 at tensor level.
    Returns (new_params, cost_including_l2, outs, grads, out_grads)."""
    outs, into_last = forward(stack, params, x, training=True, dropout_rate=dropout_rate, masks=masks)
    cost = loss_value(stack, outs, into_last, target) + l2_penalty(stack, params, l2_lambda)
    grads, out_grads = backward(stack, params, x, outs, target, dropout_rate=dropout_rate,
                                masks=masks, l2_lambda=l2_lambda)
    if optimizer == "adam":
        new_params = adam_update(params, grads, adam, lr, beta1, beta2, eps)
    else:
        new_params = sgd_update(params, grads, lr)
    return new_params, cost, outs, grads, out_grads


def weight_update_ratio(w_new: np.ndarray, w_old: np.ndarray) -> float:
    """neural_net_model.py:506: unbiased std of the delta over unbiased std of the new weights."""
    return float(np.std(w_new - w_old, ddof=1) / (np.std(w_new, ddof=1) + 1e-8))
