"""Host-side logic of the data-parallel path, world_size 2 over gloo on CPU (no kernels run here:
the engine is replaced by a stand-in that produces per-rank gradients with numpy through the
ORACLE, which is allowed in tests)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import mlp_oracle as orc
from oracle.detgen import det_labels, det_normalish, det_params
from penr_oz_neural_network_torch_b200.dp import DataParallelTrainer, reduce_costs, shard_bounds


def test_shard_bounds_cover_rows_exactly_once():
    for n in (0, 1, 7, 8, 1024, 524288, 1025):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_bounds(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for a, b in zip(spans, spans[1:]):
                assert a[1] == b[0]
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


class FakeEngine:
    """DeviceMLP look-alike on CPU tensors: gradients of the local shard, scaled by 1/global_batch."""
    device = None

    def __init__(self, algos, params):
        self.stack = orc.parse_stack(algos)
        self.params = [p.copy() for p in params]
        self.n_linear = len(self.stack.linears)
        self.cost_log = []

    def train_begin(self, global_batch=0, learning_rate=0.01, l2_lambda=0.0, **_):
        self.gb, self.lr, self.l2 = global_batch, learning_rate, l2_lambda

    def train_backward(self, x, target, row_idx=None, diagnostics=False):
        x, y = x.numpy(), target.numpy()
        outs, into_last = orc.forward(self.stack, self.params, x, training=True)
        scale = x.shape[0] / self.gb
        tgt = [[int(v)] for v in y]
        grads, _ = orc.backward(self.stack, self.params, x, outs, tgt)
        self.buckets = []
        for l in range(self.n_linear):
            flat = np.concatenate([grads[2 * l].reshape(-1), grads[2 * l + 1]]) * scale
            self.buckets.append(torch.tensor(flat))
        self.local_cost = orc.loss_value(self.stack, outs, into_last, tgt) * scale + \
            orc.l2_penalty(self.stack, self.params, self.l2) * scale

    def grad_layer(self, layer, stream):
        return self.buckets[layer]

    def train_apply(self):
        for l in range(self.n_linear):
            w, b = self.params[2 * l], self.params[2 * l + 1]
            g = self.buckets[l].numpy()
            gw = g[:w.size].reshape(w.shape) + 2 * self.l2 * w
            self.params[2 * l] = w - self.lr * gw
            self.params[2 * l + 1] = b - self.lr * g[w.size:]
        self.cost_log.append(self.local_cost)

    def costs(self, first, count):
        return np.asarray(self.cost_log[first:first + count])


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sizes, algos = [12, 9, 5], ["linear", "tanh", "linear", "softmax"]
    params = det_params(sizes, True, 3)
    B = 37                                  # not divisible by 2: ragged shards
    x, y = det_normalish((B, 12), 4), det_labels(B, 5, 5)
    tr = DataParallelTrainer(FakeEngine(algos, params))
    tr.train_begin(global_batch=B, learning_rate=0.05, l2_lambda=0.01)
    lo, hi = tr.local_rows()
    for _ in range(3):
        tr.train_step(torch.tensor(x[lo:hi]), torch.tensor(y[lo:hi]))
    costs = tr.costs(0, 3)
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), costs=costs, *tr.eng.params)
    dist.destroy_process_group()


def test_two_ranks_equal_one_rank_on_the_global_batch(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = np.load(tmp_path / "r0.npz"), np.load(tmp_path / "r1.npz")
    sizes, algos = [12, 9, 5], ["linear", "tanh", "linear", "softmax"]
    stack = orc.parse_stack(algos)
    p = det_params(sizes, True, 3)
    x, y = det_normalish((37, 12), 4), det_labels(37, 5, 5)
    want_costs = []
    for _ in range(3):
        p, c, *_ = orc.train_step(stack, p, x, [[int(v)] for v in y], lr=0.05, optimizer="sgd", l2_lambda=0.01)
        want_costs.append(c)
    np.testing.assert_allclose(r0["costs"], want_costs, rtol=1e-12)
    np.testing.assert_allclose(r1["costs"], want_costs, rtol=1e-12)
    for i in range(4):
        np.testing.assert_allclose(r0[f"arr_{i}"], p[i], rtol=1e-11, atol=1e-14)
        np.testing.assert_array_equal(r0[f"arr_{i}"], r1[f"arr_{i}"])       # replicas stay identical
