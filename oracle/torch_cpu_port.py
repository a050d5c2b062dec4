"""The reference's CPU execution path, restated with the same torch operators -- TEST / BASELINE
INFRASTRUCTURE ONLY (imported by tests/, bench.py's cpu_baseline leg and bench.py --impl reference).

Where oracle/mlp_oracle.py derives everything in closed form (and so can disagree with torch), this
file issues the very torch calls the reference issues -- ``x @ W``, ``+= b`` (neural_net_model.py:
117-119), ``.relu()/.sigmoid()/.tanh()/.softmax()`` (:172-188), ``F.dropout`` (:395),
``F.cross_entropy`` / ``F.mse_loss`` (:403,:406), the L2 sum (:483), autograd ``backward()`` (:495),
``torch.optim.Adam.step()`` (:497) or the SGD line (:500) -- in float64 on CPU tensors, so timing it
on the GPU box's host cores measures what the reference's hot path costs there.  The reference
itself cannot travel to the GPU box (it lives in /root/reference); tests/test_oracle_golden.py pins
this port to its recorded outputs.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


class TorchCpuMLP:
    def __init__(self, algos, params, optimizer="adam", flatten_ratios=None, batchnorm=(1e-5, 0.1)):
        """algos: expanded entry list; params: flat list in the reference's `params` order (embedding
        table, Linear W (b), batchnorm gain, bias ...) as float64 arrays/tensors; flatten_ratios: ratio
        per flatten entry, in order."""
        self.algos = list(algos)
        n = len(self.algos)
        self.hidden = [0 < i < n - 1 for i in range(n)]                     # :255
        self.params = [torch.as_tensor(p, dtype=torch.float64).clone() for p in params]
        n_lin = sum(1 for a in self.algos if a == "linear")
        n_other = sum(1 for a in self.algos if a == "embedding") + 2 * sum(1 for a in self.algos if a == "batchnorm")
        self.has_bias = len(self.params) - n_other == 2 * n_lin
        self.ratios = list(flatten_ratios or [])
        self.eps, self.momentum = batchnorm
        self.running = {}                                                   # entry index -> [mean, variance]
        self.weight_idx = []
        k = 0
        for i, a in enumerate(self.algos):
            if a == "embedding":
                self.weight_idx.append(k); k += 1
            elif a == "linear":
                self.weight_idx.append(k); k += 2 if self.has_bias else 1
            elif a == "batchnorm":
                dim = self.params[k].shape[0]
                self.running[i] = [torch.zeros(dim, dtype=torch.float64), torch.ones(dim, dtype=torch.float64)]
                k += 2
        self.optimizer = torch.optim.Adam(self.params) if optimizer == "adam" else None   # :281-282

    @property
    def weights(self):
        return [self.params[k] for k in self.weight_idx]

    def forward(self, x, target=None, dropout_rate=0.0):
        """:384-408 -- returns (entry outputs, cost)."""
        training = target is not None
        outs, cur, logits = [], x, x
        it = iter(self.params)
        ratios = iter(self.ratios)
        for i, (algo, hidden) in enumerate(zip(self.algos, self.hidden)):
            logits = cur
            if algo == "embedding":                                          # :68
                cur = next(it)[logits.long()]
            elif algo == "flatten":                                          # :82-91
                ratio = next(ratios)
                if logits.ndim > 1:
                    shp = list(logits.shape)
                    shp[-2] //= ratio
                    shp[-1] *= ratio
                    cur = logits.view(*shp)
                    if shp[-2] == 1:
                        cur = cur.squeeze(len(shp) - 2)
            elif algo == "batchnorm":                                        # :158-170
                gain, bias = next(it), next(it)
                if training:
                    dim = tuple(range(logits.ndim - 1))
                    mean = logits.mean(dim, keepdim=True)
                    var = logits.var(dim, keepdim=True)
                    with torch.no_grad():
                        run = self.running[i]
                        run[0] = (1 - self.momentum) * run[0] + self.momentum * mean
                        run[1] = (1 - self.momentum) * run[1] + self.momentum * var
                else:
                    mean, var = self.running[i]
                cur = gain * (logits - mean) / torch.sqrt(var + self.eps) + bias
            elif algo == "linear":
                cur = logits @ next(it)
                if self.has_bias:
                    cur += next(it)
            elif algo == "relu":
                cur = logits.relu()
            elif algo == "sigmoid":
                cur = logits.sigmoid()
            elif algo == "tanh":
                cur = logits.tanh()
            elif algo == "softmax":
                cur = logits.softmax(dim=logits.ndim - 1)
            if hidden and training:
                cur = F.dropout(cur, p=dropout_rate)
            outs.append(cur)
        if not training:
            cost = torch.empty(0)
        elif self.algos[-1] == "softmax":
            cost = F.cross_entropy(logits, target)
        else:
            cost = F.mse_loss(cur, target)
        return outs, cost

    def train_step(self, x, target, lr, dropout_rate=0.0, l2_lambda=0.0, beta1=0.9, beta2=0.999, eps=1e-8):
        """:464-500 at tensor level.  target: int64 labels (CE) or float64 [B, out] (MSE)."""
        if self.optimizer is not None:
            for g in self.optimizer.param_groups:
                g["lr"], g["betas"], g["eps"] = lr, (beta1, beta2), eps
        for p in self.params:
            p.requires_grad_()
        outs, cost = self.forward(x, target, dropout_rate)
        if l2_lambda > 0.0:
            cost += l2_lambda * sum((w ** 2).sum() for w in self.weights)
        for p in self.params:
            p.grad = None
        cost.backward()
        if self.optimizer is not None:
            self.optimizer.step()
        else:
            for p in self.params:
                p.data -= lr * p.grad
        return cost.item()
