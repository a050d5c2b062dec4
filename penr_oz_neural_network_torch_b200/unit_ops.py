"""Unit operators on the device: one entry of the reference's layer list per call.

Host side of ``poz_op_linear / poz_op_activation / poz_op_embedding / poz_op_batchnorm``
(include/pozmlp.h).  The ``forward`` methods of the drop-in's Layer classes call these, so a test that
drives a single layer -- as the reference's own do (test_neural_net_model.py:13-66) -- runs the CUDA kernels,
not torch arithmetic.  PyTorch is plumbing here: it holds the buffers and moves them between host and device.
"""
from __future__ import annotations

import os

import torch

from . import _native as nat

# dtype of a stand-alone layer call (a layer inside a model uses the model's dtype): float64 is the reference's
UNIT_DTYPE = os.environ.get("POZ_UNIT_DTYPE", "float64")
_TORCH = {"float32": torch.float32, "float64": torch.float64}
_CODE = {"float32": nat.DTYPE_F32, "float64": nat.DTYPE_F64}


def _device(device=None) -> torch.device:
    if not torch.cuda.is_available():
        raise nat.PozError("no CUDA device: the layer operators have no CPU fallback")
    if device is None:
        return torch.device("cuda", torch.cuda.current_device())
    if isinstance(device, torch.device):
        return device if device.index is not None else torch.device("cuda", torch.cuda.current_device())
    return torch.device("cuda", int(device))


def _stream(dev) -> int:
    return int(torch.cuda.current_stream(dev).cuda_stream)


def to_device(t: torch.Tensor, dtype: str, dev: torch.device) -> torch.Tensor:
    return t.detach().to(device=dev, dtype=_TORCH[dtype]).contiguous()


def linear(x: torch.Tensor, w: torch.Tensor, b, dtype: str, dev: torch.device) -> torch.Tensor:
    """x [rows, in] (device, dtype), w [in, out], b [out] or None -> [rows, out]   (neural_net_model.py:117-119)"""
    rows, k = x.shape
    n = w.shape[1]
    if w.shape[0] != k:
        raise RuntimeError(f"mat1 and mat2 shapes cannot be multiplied ({rows}x{k} and {w.shape[0]}x{n})")
    y = torch.empty((rows, n), dtype=_TORCH[dtype], device=dev)
    nat.check(nat.lib().poz_op_linear(dev.index, _CODE[dtype], x.data_ptr(), w.data_ptr(),
                                      None if b is None else b.data_ptr(), rows, k, n, y.data_ptr(), _stream(dev)))
    return y


def activation(x: torch.Tensor, act: str, dtype: str, dev: torch.device) -> torch.Tensor:
    """x [rows, cols] -> act(x); softmax over the last dimension   (neural_net_model.py:172-188)"""
    rows, cols = x.shape
    y = torch.empty_like(x)
    nat.check(nat.lib().poz_op_activation(dev.index, _CODE[dtype], nat.ACT_CODES[act], x.data_ptr(), rows, cols,
                                          y.data_ptr(), _stream(dev)))
    return y


def embedding(table: torch.Tensor, idx: torch.Tensor, dtype: str, dev: torch.device) -> torch.Tensor:
    """table [vocab, dim], idx int32 [n] -> [n, dim]   (neural_net_model.py:68)"""
    n = idx.numel()
    y = torch.empty((n, table.shape[1]), dtype=_TORCH[dtype], device=dev)
    nat.check(nat.lib().poz_op_embedding(dev.index, _CODE[dtype], table.data_ptr(), table.shape[0], table.shape[1],
                                         idx.data_ptr(), n, y.data_ptr(), _stream(dev)))
    return y


def batchnorm(x: torch.Tensor, gain, bias, run_mean, run_var, eps: float, momentum: float, training: bool,
              scalar_stats: bool, dtype: str, dev: torch.device) -> torch.Tensor:
    """x [rows, cols]; run_mean / run_var [cols] are updated in place when training   (neural_net_model.py:158-170)"""
    rows, cols = x.shape
    y = torch.empty_like(x)
    nat.check(nat.lib().poz_op_batchnorm(dev.index, _CODE[dtype], int(training), int(scalar_stats), x.data_ptr(), rows, cols,
                                         gain.data_ptr(), bias.data_ptr(), run_mean.data_ptr(), run_var.data_ptr(),
                                         float(eps), float(momentum), y.data_ptr(), _stream(dev)))
    return y


class Chain:
    """Keeps an entry's input on the device while a model walks its layer list (``_forward`` without a
    target), and converts at the edges for a stand-alone ``layer.forward(cpu_tensor)`` call."""

    def __init__(self, dtype: str | None = None, device=None):
        self.dtype = dtype or UNIT_DTYPE
        if self.dtype not in _TORCH:
            raise ValueError("unit operators run in float32 or float64")
        self.dev = _device(device)

    def enter(self, t: torch.Tensor, as_index: bool = False) -> torch.Tensor:
        if as_index:
            return t.detach().long().to(torch.int32).to(self.dev).contiguous()
        return to_device(t, self.dtype, self.dev)

    @staticmethod
    def leave(y: torch.Tensor, like: torch.Tensor) -> torch.Tensor:
        """Result of a stand-alone call: where the input lived, float64 as the reference's tensors are."""
        return y.to(device=like.device, dtype=torch.float64)
