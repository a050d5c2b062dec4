"""Data-parallel training of one model over the GPUs of a box (new: the reference is single-process).

One process per GPU; every rank holds the same parameters and optimizer state and draws the same
sample indices (same ``torch.manual_seed``, as ``torch.randint`` at neural_net_model.py:460 uses the
global generator).  Rank r trains on rows ``shard_bounds(B, r, N)`` of the sampled batch; the loss
is a mean over the *global* batch, so local gradients are pre-scaled by 1/global_batch in the
kernels and combined with one all-reduce(SUM) per Linear layer.  Layer l's bucket is reduced on a
communication stream as soon as its weight gradient is complete, while the backward of layers
l-1..0 keeps running; the (identical) update is applied on every rank once all buckets are in.
The L2 term is added after the reduction, once.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def shard_bounds(n_rows: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous, near-equal row ranges; the first ``n_rows % world`` ranks get one extra row."""
    base, extra = divmod(n_rows, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def reduce_costs(local_costs, group=None) -> np.ndarray:
    """Per-rank cost shares (local loss sum / global batch + this rank's share of the L2 penalty)
    add up to the global-batch cost of neural_net_model.py:480-483."""
    t = torch.as_tensor(np.asarray(local_costs, dtype=np.float64))
    if dist.get_backend(group) == "nccl":
        t = t.cuda()
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t.cpu().numpy()


class DataParallelTrainer:
    """Wraps a ``DeviceMLP`` (or anything with its train_* / grad_layer interface)."""

    def __init__(self, engine, group=None, comm_stream=None):
        if getattr(engine, "general", False):
            # the per-layer buckets are the Linear [W;b] images; an embedding table, batchnorm gain / bias and a
            # batchnorm's batch statistics would stay per shard and the ranks would drift apart
            raise ValueError("data-parallel training covers Linear + activation stacks; embedding / flatten / "
                             "batchnorm stacks are not sharded")
        self.eng = engine
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.comm_stream = comm_stream
        if comm_stream is None and torch.cuda.is_available() and getattr(engine, "device", None) is not None:
            self.comm_stream = torch.cuda.Stream(device=engine.device)

    def train_begin(self, global_batch: int, **hyper):
        self.global_batch = int(global_batch)
        self.eng.train_begin(global_batch=self.global_batch, **hyper)

    def local_rows(self, n_rows: int | None = None) -> tuple[int, int]:
        return shard_bounds(self.global_batch if n_rows is None else n_rows, self.rank, self.world)

    def train_step(self, x_local, target_local, row_idx_local=None, diagnostics=False):
        eng = self.eng
        eng.train_backward(x_local, target_local, row_idx=row_idx_local, diagnostics=diagnostics)
        if self.world > 1:
            for layer in range(eng.n_linear - 1, -1, -1):       # the order the backward finishes them
                bucket = eng.grad_layer(layer, self.comm_stream)
                self._all_reduce(bucket)
            self._join()
        eng.train_apply()

    def _all_reduce(self, bucket):
        if self.comm_stream is not None:
            with torch.cuda.stream(self.comm_stream):
                dist.all_reduce(bucket, op=dist.ReduceOp.SUM, group=self.group)
        else:
            dist.all_reduce(bucket, op=dist.ReduceOp.SUM, group=self.group)

    def _join(self):
        if self.comm_stream is not None:
            torch.cuda.current_stream(self.eng.device).wait_stream(self.comm_stream)

    def costs(self, first: int, count: int) -> np.ndarray:
        local = self.eng.costs(first, count)
        return reduce_costs(local, self.group) if self.world > 1 else local
