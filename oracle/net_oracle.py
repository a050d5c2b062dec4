"""General per-entry CPU oracle (numpy float64) -- TEST INFRASTRUCTURE ONLY.

Extends oracle/mlp_oracle.py to every entry type of the reference's layer list: embedding, flatten,
linear, batchnorm and the four activations, as a forward tape plus closed-form reverse pass.
Citations are neural_net_model.py:<line> / test_torch_backward.py:<line> of the reference.
Pinned like mlp_oracle: tests/test_oracle_golden.py holds it to outputs of the unmodified reference
recorded in tests/golden/net_*.npz (oracle/make_golden.py).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

ACTS = ("relu", "sigmoid", "tanh", "softmax")


@dataclass
class Entry:
    algo: str
    hidden: bool = False
    # parameters (float64 arrays) by role
    weights: np.ndarray | None = None     # linear W [in,out]; embedding C [vocab,emb]
    bias: np.ndarray | None = None        # linear b
    gain: np.ndarray | None = None        # batchnorm gain
    bn_bias: np.ndarray | None = None     # batchnorm bias
    mean: np.ndarray | None = None        # batchnorm running mean
    variance: np.ndarray | None = None    # batchnorm running variance
    eps: float = 1e-5
    momentum: float = 0.1
    ratio: int = 1                        # flatten

    @property
    def params(self):
        if self.algo == "batchnorm":
            return [p for p in (self.gain, self.bn_bias) if p is not None]
        return [p for p in (self.weights, self.bias) if p is not None]


def expand_algos(activation_algos):
    """neural_net_model.py:202-212, all entry types."""
    algos = list(activation_algos)
    out = []
    for i in range(len(algos) - 1, -1, -1):
        algo = algos[i]
        out.insert(0, algo)
        if algo in ACTS and (i == 0 or algos[i - 1] not in ("batchnorm", "linear")):
            out.insert(0, "linear")
        elif algo == "embedding" and algos[i + 1] != "flatten":
            out.insert(i + 1, "flatten")
    return out


def build_entries(layer_sizes, algos_expanded, has_bias=True, batchnorm=(1e-5, 0.1)):
    """Topology of MultiLayerPerceptron.__init__ (neural_net_model.py:213-260) with empty parameters."""
    entries, size_idx, n = [], 0, len(algos_expanded)
    ns = len(layer_sizes)
    for i, algo in enumerate(algos_expanded):
        in_sz = layer_sizes[size_idx] if size_idx < ns else 0
        out_sz = layer_sizes[size_idx + 1] if size_idx + 1 < ns else 0
        e = Entry(algo, hidden=(0 < i < n - 1))
        if algo == "flatten":
            e.ratio = out_sz // max(1, in_sz)
        elif algo == "batchnorm":
            e.eps, e.momentum = batchnorm
            e.gain, e.bn_bias = np.ones(in_sz), np.zeros(in_sz)
            e.mean, e.variance = np.zeros(in_sz), np.ones(in_sz)
        elif algo == "linear":
            e.weights = np.zeros((in_sz, out_sz))
            e.bias = np.zeros(out_sz) if has_bias else None
        elif algo == "embedding":
            e.weights = np.zeros((in_sz, out_sz))
        entries.append(e)
        if algo in ("embedding", "flatten", "linear"):
            size_idx += 1
    return entries


def set_params(entries, flat):
    """Assign a flat parameter list in the reference's `params` order."""
    it = iter(flat)
    for e in entries:
        if e.algo == "batchnorm":
            e.gain, e.bn_bias = np.array(next(it), dtype=np.float64), np.array(next(it), dtype=np.float64)
        elif e.algo in ("linear", "embedding"):
            e.weights = np.array(next(it), dtype=np.float64)
            if e.algo == "linear" and e.bias is not None:
                e.bias = np.array(next(it), dtype=np.float64)


def get_params(entries):
    return [p for e in entries for p in e.params]


def _softmax(z):
    m = z.max(axis=-1, keepdims=True)
    ez = np.exp(z - m)
    return ez / ez.sum(axis=-1, keepdims=True)


def forward(entries, x, *, training=False, dropout_rate=0.0, masks=None, update_running=True):
    """neural_net_model.py:384-396.  Returns (recorded outputs per entry, tensor into the last entry,
    tape of per-entry saved values for the reverse pass)."""
    cur = np.asarray(x)
    outs, tape = [], []
    into_last = cur
    scale = 1.0 / (1.0 - dropout_rate) if dropout_rate > 0 else 1.0
    for i, e in enumerate(entries):
        into_last = cur
        saved = {"in": cur}
        if e.algo == "embedding":
            idx = cur.astype(np.int64)                       # :68
            saved["idx"] = idx
            cur = e.weights[idx]
        elif e.algo == "flatten":                            # :82-91
            if cur.ndim > 1:
                shp = list(cur.shape)
                shp[-2] //= e.ratio
                shp[-1] *= e.ratio
                cur = cur.reshape(shp)
                if shp[-2] == 1:
                    cur = cur.squeeze(len(shp) - 2)
        elif e.algo == "linear":
            cur = cur @ e.weights
            if e.bias is not None:
                cur = cur + e.bias
        elif e.algo == "batchnorm":                          # :158-170
            if training:
                axes = tuple(range(cur.ndim - 1))
                mean = cur.mean(axis=axes, keepdims=True)
                var = cur.var(axis=axes, keepdims=True, ddof=1)
                if update_running:
                    e.mean = (1 - e.momentum) * e.mean + e.momentum * mean
                    e.variance = (1 - e.momentum) * e.variance + e.momentum * var
            else:
                mean, var = e.mean, e.variance
            inv = 1.0 / np.sqrt(var + e.eps)
            xhat = (cur - mean) * inv
            saved.update(xhat=xhat, inv=inv)
            cur = e.gain * xhat + e.bn_bias
        elif e.algo == "relu":
            cur = np.maximum(cur, 0.0)
        elif e.algo == "sigmoid":
            cur = 1.0 / (1.0 + np.exp(-cur))
        elif e.algo == "tanh":
            cur = np.tanh(cur)
        elif e.algo == "softmax":
            cur = _softmax(cur)
        saved["pre_drop"] = cur
        if e.hidden and training and dropout_rate > 0:
            cur = cur * masks[i] * scale
        outs.append(cur)
        tape.append(saved)
    return outs, into_last, tape


def loss_value(entries, outs, into_last, target):
    if entries[-1].algo == "softmax":                        # :400-403
        z = into_last if into_last.ndim == 2 else into_last[None, :]
        t = np.asarray(target)
        y = (np.asarray([int(t[0])]) if t.ndim == 1 else t[:, 0].astype(np.int64))
        m = z.max(axis=-1, keepdims=True)
        lse = m[:, 0] + np.log(np.exp(z - m).sum(axis=-1))
        return float(np.mean(lse - z[np.arange(z.shape[0]), y]))
    return float(np.mean((outs[-1] - np.asarray(target, dtype=np.float64)) ** 2))


def backward(entries, outs, into_last, tape, target, *, dropout_rate=0.0, masks=None, l2_lambda=0.0):
    """Reverse pass.  Returns (parameter grads in `params` order, gradient of every recorded entry
    output; None where the entry is not on the loss graph)."""
    n = len(entries)
    scale = 1.0 / (1.0 - dropout_rate) if dropout_rate > 0 else 1.0
    out_grads = [None] * n
    pgrads = [None] * n           # per entry: list of grads aligned with entry.params
    last = n - 1
    if entries[-1].algo == "softmax":
        z = into_last if into_last.ndim == 2 else into_last[None, :]
        t = np.asarray(target)
        y = (np.asarray([int(t[0])]) if t.ndim == 1 else t[:, 0].astype(np.int64))
        B = z.shape[0]
        g = _softmax(z) / B                                   # test_torch_backward.py:163-164
        g[np.arange(B), y] -= 1.0 / B
        g = g.reshape(into_last.shape)
        start = last - 1          # gradient w.r.t. the recorded output of entry last-1
    else:
        yhat = outs[-1]
        g = 2.0 * (yhat - np.asarray(target, dtype=np.float64)) / yhat.size
        start = last
    for i in range(start, -1, -1):
        e, sv = entries[i], tape[i]
        out_grads[i] = g
        if e.hidden and dropout_rate > 0:
            g = g * masks[i] * scale
        xin = sv["in"]
        if e.algo == "embedding":                             # test_torch_backward.py:97-100
            dc = np.zeros_like(e.weights)
            np.add.at(dc, sv["idx"].reshape(-1), g.reshape(-1, e.weights.shape[1]))
            if l2_lambda > 0:
                dc = dc + 2 * l2_lambda * e.weights
            pgrads[i] = [dc]
            g = None
        elif e.algo == "flatten":
            g = g.reshape(xin.shape)
        elif e.algo == "linear":
            x2 = xin.reshape(-1, xin.shape[-1])
            g2 = g.reshape(-1, g.shape[-1])
            gw = x2.T @ g2                                    # test_torch_backward.py:165
            if l2_lambda > 0:
                gw = gw + 2 * l2_lambda * e.weights
            pgrads[i] = [gw] + ([g2.sum(axis=0)] if e.bias is not None else [])
            g = (g2 @ e.weights.T).reshape(xin.shape)         # :168
        elif e.algo == "batchnorm":                           # test_torch_backward.py:172-176
            xhat, inv = sv["xhat"], sv["inv"]
            axes = tuple(range(g.ndim - 1))
            R = int(np.prod([g.shape[a] for a in axes]))
            dgain = (xhat * g).sum(axis=axes)
            dbias = g.sum(axis=axes)
            pgrads[i] = [dgain, dbias]
            g = e.gain * inv / R * (R * g - g.sum(axis=axes, keepdims=True)
                                    - R / (R - 1) * xhat * (g * xhat).sum(axis=axes, keepdims=True))
        elif e.algo == "relu":
            g = g * (sv["pre_drop"] > 0)
        elif e.algo == "sigmoid":
            s = sv["pre_drop"]
            g = g * s * (1 - s)
        elif e.algo == "tanh":
            th = sv["pre_drop"]
            g = g * (1 - th * th)                             # test_torch_backward.py:170
        elif e.algo == "softmax":
            s = sv["pre_drop"]
            g = s * (g - (g * s).sum(axis=-1, keepdims=True))
    flat = [gr for pg in pgrads if pg is not None for gr in pg]
    return flat, out_grads


def l2_penalty(entries, l2_lambda):
    """neural_net_model.py:482-483: every entry that has `.weights` (linear and embedding)."""
    if l2_lambda <= 0:
        return 0.0
    return l2_lambda * sum(float((e.weights ** 2).sum()) for e in entries if e.weights is not None)


def weight_indices(entries):
    """Positions, in the flat `params` list, of the tensors the reference calls `weights`."""
    out, k = [], 0
    for e in entries:
        if e.algo in ("linear", "embedding"):
            out.append(k)
        k += len(e.params)
    return out
