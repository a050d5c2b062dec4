"""Drop-in ``NeuralNetworkModel`` whose hot path runs in hand-written sm_100a CUDA.

Mirror of the reference's public class (derinworks/penr-oz-neural-network-torch
``neural_net_model.py:269-585``): same constructor, ``compute_output``, ``train``, ``_forward``,
``serialize`` / ``deserialize`` / ``delete`` and the same attributes, so ``main.py`` works by
changing its import line only.  What differs is underneath: the Linear +
activation stack, the loss, the backward pass and the Adam / SGD update run through the C ABI in
``include/pozmlp.h`` on ``cuda:<device>``; the host keeps float64 CPU tensors as the parameter
*view* the reference exposes (``params``, ``weights``, ``state_dict``, ``optimizer.state_dict()``)
and synchronises them with the device image at the boundaries of ``train`` / ``compute_output``.

Scope: an optional embedding head, then Linear layers each optionally followed
by a batchnorm and by relu / sigmoid / tanh, with flatten entries between them and an optional
trailing softmax.  Anything else raises ``ValueError`` -- there is deliberately no CPU fallback.
"""
from __future__ import annotations

import itertools
import json
import logging
import math
import os
import random
import time
from datetime import datetime as dt
from typing import Tuple

import numpy as np
import torch
from torch import Tensor

from . import unit_ops
from .engine import ACTIVATIONS, DeviceMLP, StackSpec

log = logging.getLogger(__name__)

_KNOWN_ALGOS = ("linear", "embedding", "flatten", "batchnorm") + ACTIVATIONS
_GAIN = {"relu": math.sqrt(2.0), "tanh": 5.0 / 3.0}     # neural_net_model.py:178,183

DEFAULT_DTYPE = os.environ.get("POZ_DTYPE", "float32")



def _list_to_f64(data):
    """torch.tensor(data, dtype=torch.float64) for the shape the API is fed with -- a list of equal-length lists of Python
    numbers (neural_net_model.py:379,472) -- at about twice the speed: one pass of numpy.fromiter over the chained rows instead
    of torch's generic nested-sequence walk, which is 97 % of a 1,024-row compute_output() call.  Anything else (a single
    vector, deeper nesting, ragged rows, non-numbers) goes to torch.tensor, whose result or exception stands."""
    try:
        if isinstance(data, list) and data and isinstance(data[0], list) and data[0] and not isinstance(data[0][0], (list, tuple)):
            width = len(data[0])
            if all(isinstance(r, list) and len(r) == width for r in data):
                flat = np.fromiter(itertools.chain.from_iterable(data), dtype=np.float64, count=len(data) * width)
                if not np.isnan(flat).any():        # (numpy reads None as NaN where torch raises: let torch decide those)
                    return torch.from_numpy(flat.reshape(len(data), width))
    except (TypeError, ValueError):
        pass
    return torch.tensor(data, dtype=torch.float64)


class Layer:
    """Parameter container + topology flags of one entry (neural_net_model.py:16-53).  ``train`` and
    ``compute_output`` execute the stack as a whole in the CUDA library; ``forward`` runs this one entry on
    the device through the unit operators (``poz_op_*``), for callers that drive a layer directly as the
    reference's tests do, and for ``_forward`` without a target."""
    algo = ""

    def __init__(self):
        self.hidden = False
        self.training = False
        self.weights: Tensor | None = None
        self.bias: Tensor | None = None

    def forward(self, input_tensor: Tensor) -> Tensor:
        """:param input_tensor: 1-D vector or 2-D batch matrix (any leading dimensions)
        :return: the forwarded tensor, float64 where the input lives (neural_net_model.py:48-53)"""
        if self._identity():
            return input_tensor                                       # Layer.forward is the identity (:53)
        chain = unit_ops.Chain()
        t = torch.as_tensor(input_tensor)
        return chain.leave(self._run(chain, chain.enter(t, as_index=self.algo == "embedding")), t)

    def _identity(self) -> bool:
        return type(self)._run is Layer._run

    def _run(self, chain: "unit_ops.Chain", x: Tensor) -> Tensor:
        """device tensor in, device tensor out"""
        return x

    def _on_device(self, name: str, chain: "unit_ops.Chain") -> Tensor | None:
        """Device copy of parameter ``name`` in the chain's dtype, refreshed when the host tensor changed."""
        host = getattr(self, name)
        if host is None:
            return None
        cache = self.__dict__.setdefault("_dev_cache", {})
        key = (name, chain.dtype, chain.dev.index)
        tag = (id(host), host._version)
        hit = cache.get(key)
        if hit is None or hit[0] != tag:
            hit = (tag, unit_ops.to_device(host, chain.dtype, chain.dev))
            cache[key] = hit
        return hit[1]

    @staticmethod
    def _rows(x: Tensor) -> Tensor:
        """[..., features] -> [rows, features]"""
        return x.reshape(-1, x.shape[-1]) if x.ndim != 2 else x

    @property
    def params(self) -> list[Tensor]:
        return [t for t in (self.weights, self.bias) if t is not None]

    @property
    def state_dict(self) -> dict:
        return {"params": [t.tolist() for t in self.params]}

    @state_dict.setter
    def state_dict(self, new_state: dict):
        values = new_state["params"]
        if len(values) > 0:
            self.weights = torch.tensor(values[0], dtype=torch.float64)
        if len(values) > 1:
            self.bias = torch.tensor(values[1], dtype=torch.float64)


class LinearLayer(Layer):
    algo = "linear"

    def __init__(self, input_size: int = 0, output_size: int = 0, bias_algo="zeros"):
        super().__init__()
        if input_size > 0 and output_size > 0:
            # same draws, in the same order, from torch's global CPU generator as the reference
            # (neural_net_model.py:104-108) so a given torch.manual_seed gives the same model
            self.weights = torch.randn(input_size, output_size).double()
            if bias_algo:
                self.bias = (torch.zeros(output_size) if bias_algo == "zeros"
                             else torch.randn(output_size)).double()

    def _run(self, chain, x):
        """x @ W, then += b (neural_net_model.py:110-120); 1-D and batched inputs alike"""
        w = self._on_device("weights", chain)
        y = unit_ops.linear(self._rows(x), w, self._on_device("bias", chain), chain.dtype, chain.dev)
        return y.reshape(*x.shape[:-1], w.shape[1])


class EmbeddingLayer(Layer):
    """Lookup table (neural_net_model.py:55-68); the gather / scatter-add run in emb_*_kernel."""
    algo = "embedding"

    def __init__(self, vocab_size: int = 0, embedding_size: int = 0):
        super().__init__()
        if vocab_size > 0 and embedding_size > 0:
            self.weights = torch.randn(vocab_size, embedding_size).double()

    def _run(self, chain, idx):
        """W[idx.long()] (neural_net_model.py:67-68): idx of any shape -> [..., embedding_size]"""
        table = self._on_device("weights", chain)
        vocab = table.shape[0]
        if idx.numel():
            lo, hi = int(idx.min()), int(idx.max())
            if lo < -vocab or hi >= vocab:
                raise IndexError(f"index {hi if hi >= vocab else lo} is out of bounds for dimension 0 with size {vocab}")
            if lo < 0:
                idx = torch.where(idx < 0, idx + vocab, idx)
        y = unit_ops.embedding(table, idx.reshape(-1).contiguous(), chain.dtype, chain.dev)
        return y.reshape(*idx.shape, table.shape[1])


class FlattenLayer(Layer):
    """Merges `ratio` rows into one (neural_net_model.py:70-91): a reinterpretation on the device."""
    algo = "flatten"

    def __init__(self, ratio: int):
        super().__init__()
        self.ratio = ratio

    @property
    def state_dict(self) -> dict:
        return {"ratio": self.ratio}

    @state_dict.setter
    def state_dict(self, new_state: dict):
        self.ratio = new_state["ratio"]

    def forward(self, input_tensor: Tensor) -> Tensor:
        return self._run(None, input_tensor)

    def _run(self, chain, x):
        """Merge `ratio` rows into one (neural_net_model.py:82-91): a view of the same row-major memory, on the
        host or on the device -- there is nothing to compute."""
        if x.ndim > 1:
            shape = list(x.shape)
            shape[-2] //= self.ratio
            shape[-1] *= self.ratio
            x = x.reshape(*shape)
            if shape[-2] == 1:
                x = x.squeeze(len(shape) - 2)
        return x


class BatchNormLayer(Layer):
    """Parameters and running statistics of a batchnorm entry (neural_net_model.py:122-170)."""
    algo = "batchnorm"

    def __init__(self, dim_size: int, eps=1e-5, momentum=0.1):
        super().__init__()
        self.eps = eps
        self.momentum = momentum
        self.gain = torch.ones(dim_size, dtype=torch.float64) if dim_size > 0 else None
        self.bias = torch.zeros(dim_size, dtype=torch.float64) if dim_size > 0 else None
        self.variance = torch.ones(dim_size, dtype=torch.float64) if dim_size > 0 else None
        self.mean = torch.zeros(dim_size, dtype=torch.float64) if dim_size > 0 else None

    @property
    def params(self) -> list[Tensor]:
        return [t for t in (self.gain, self.bias) if t is not None]

    @property
    def state_dict(self) -> dict:
        return {"params": [t.tolist() for t in self.params], "eps": self.eps, "momentum": self.momentum}

    @state_dict.setter
    def state_dict(self, new_state: dict):
        values = new_state["params"]
        if len(values) > 0:                      # running statistics are not persisted (:146-156)
            self.gain = torch.tensor(values[0], dtype=torch.float64)
            self.variance = torch.ones_like(self.gain)
        if len(values) > 1:
            self.bias = torch.tensor(values[1], dtype=torch.float64)
            self.mean = torch.zeros_like(self.bias)
        self.eps = new_state["eps"]
        self.momentum = new_state["momentum"]

    def _run(self, chain, x):
        """neural_net_model.py:158-170: batch statistics (unbiased variance) and a move of the running statistics
        when ``training``, the running statistics otherwise.  A 1-D input in training mode reduces over all of
        its elements (``dim = ()``), giving one mean / variance for every feature -- reproduced as is."""
        tdt = unit_ops._TORCH[chain.dtype]
        mean = self.mean.reshape(-1).to(device=chain.dev, dtype=tdt).contiguous().clone()
        var = self.variance.reshape(-1).to(device=chain.dev, dtype=tdt).contiguous().clone()
        scalar = bool(self.training and x.ndim == 1)
        x2 = x.reshape(1, -1) if x.ndim == 1 else self._rows(x)
        y = unit_ops.batchnorm(x2, self._on_device("gain", chain), self._on_device("bias", chain), mean, var,
                               self.eps, self.momentum, self.training, scalar, chain.dtype, chain.dev)
        if self.training:
            # the reference's running statistics take the keepdim shape of the batch statistics (:161-166)
            keep = [1] * (x.ndim - 1) + [x.shape[-1]] if x.ndim >= 2 else [x.shape[-1]]
            self.mean = mean.to("cpu", torch.float64).reshape(keep)
            self.variance = var.to("cpu", torch.float64).reshape(keep)
        return y.reshape(x.shape)


class _ActivationLayer(Layer):
    def _run(self, chain, x):
        """element-wise sigmoid / relu / tanh, softmax over the last dimension (neural_net_model.py:172-188)"""
        if x.ndim == 0:
            x = x.reshape(1)
        y = unit_ops.activation(x.reshape(1, -1) if x.ndim == 1 else self._rows(x), self.algo, chain.dtype, chain.dev)
        return y.reshape(x.shape)


class SigmoidLayer(_ActivationLayer):
    algo = "sigmoid"


class ReluLayer(_ActivationLayer):
    algo = "relu"
    weight_gain = _GAIN["relu"]


class TanhLayer(_ActivationLayer):
    algo = "tanh"
    weight_gain = _GAIN["tanh"]


class SoftmaxLayer(_ActivationLayer):
    algo = "softmax"


_LAYER_TYPES = {"sigmoid": SigmoidLayer, "relu": ReluLayer, "tanh": TanhLayer, "softmax": SoftmaxLayer}


def expand_algos(activation_algos: list[str]) -> list[str]:
    """The reference's entry-list expansion (neural_net_model.py:202-212): a Linear entry goes in
    front of every activation that does not already follow a Linear / batchnorm, and a flatten entry
    behind an embedding that has none.  Walks the list from the back, as the reference does, so
    that the insertion positions agree in every case."""
    requested = list(activation_algos)
    expanded: list[str] = []
    for i in range(len(requested) - 1, -1, -1):
        algo = requested[i]
        expanded.insert(0, algo)
        if algo in ACTIVATIONS and (i == 0 or requested[i - 1] not in ("batchnorm", "linear")):
            expanded.insert(0, "linear")
        elif algo == "embedding" and requested[i + 1] != "flatten":
            expanded.insert(i + 1, "flatten")
    return expanded


class _DeviceCost(torch.autograd.Function):
    """The scalar cost of a device pass as an autograd node whose inputs are the model's parameters: its
    backward hands out the parameter gradients the CUDA backward pass computed (scaled by the incoming
    gradient, so ``(2 * cost).backward()`` behaves)."""

    @staticmethod
    def forward(ctx, cost, grads, *params):
        ctx.grads = grads
        return cost.clone()

    @staticmethod
    def backward(ctx, upstream):
        return (None, None) + tuple(upstream * g for g in ctx.grads)


class MultiLayerPerceptron:
    def __init__(self, layer_sizes: list[int], weight_algo="xavier", bias_algo="zeros",
                 activation_algos: list[str] = None, batchnorm=(1e-5, 0.1)):
        requested = activation_algos or ["relu"] * (len(layer_sizes) - 1)
        for algo in requested:
            if algo not in _KNOWN_ALGOS:
                raise ValueError(f"Unsupported activation algorithm: {algo}")   # neural_net_model.py:241
        self.algos = expand_algos(requested)
        StackSpec.from_algos(self.algos)                    # structure check: ValueError if off the device path
        self.layers: list[Layer] = []
        n_entries = len(self.algos)
        n_sizes = len(layer_sizes)
        size_idx = 0
        pending_gain: LinearLayer | None = None
        for i, algo in enumerate(self.algos):
            in_sz = layer_sizes[size_idx] if size_idx < n_sizes else 0
            out_sz = layer_sizes[size_idx + 1] if size_idx + 1 < n_sizes else 0
            if algo == "embedding":
                layer = EmbeddingLayer(in_sz, out_sz)
            elif algo == "flatten":
                layer = FlattenLayer(out_sz // max(1, in_sz))                  # :227
            elif algo == "linear":
                layer = LinearLayer(in_sz, out_sz, bias_algo)
                if in_sz > 0 and layer.weights is not None and weight_algo in ("xavier", "he"):
                    layer.weights /= math.sqrt(in_sz)                          # :243-244
                pending_gain = layer if weight_algo == "he" else None          # :246-247
            elif algo == "batchnorm":
                layer = BatchNormLayer(in_sz, *batchnorm)                      # :231
            else:
                layer = _LAYER_TYPES[algo]()
                if pending_gain is not None and pending_gain.weights is not None and algo in _GAIN:
                    pending_gain.weights *= _GAIN[algo]                        # :249-253
            layer.hidden = 0 < i < n_entries - 1                               # :255
            self.layers.append(layer)
            if algo in ("embedding", "flatten", "linear"):                     # :259-260
                size_idx += 1

    @property
    def stack(self) -> StackSpec:
        """Units, dropout points and row multipliers, from the entries as they are now (flatten
        ratios are only known once the layers hold their sizes / their persisted state)."""
        ratios = {i: layer.ratio for i, layer in enumerate(self.layers) if layer.algo == "flatten"}
        return StackSpec.from_algos(self.algos, ratios)

    @property
    def params(self) -> list[Tensor]:
        return [p for layer in self.layers for p in layer.params]


class NeuralNetworkModel(MultiLayerPerceptron):
    def __init__(self, model_id, layer_sizes: list[int] = None, weight_algo="xavier", bias_algo="zeros",
                 activation_algos=None, optimizer_algo="adam", batchnorm=(1e-5, 0.1), confidence=1.0,
                 *, dtype: str | None = None, device=None):
        super().__init__(layer_sizes or [], weight_algo, bias_algo, activation_algos, batchnorm)
        self.model_id = model_id
        self.optimizer: torch.optim.Optimizer | None = None
        if len(self.params) > 0:
            if optimizer_algo == "adam":
                # the torch object is the *container* of Adam's state (state_dict()/.pth interop,
                # neural_net_model.py:338-341,351-354); the update itself runs in update_kernel
                self.optimizer = torch.optim.Adam(self.params)
            self.weights[-1] *= confidence                                      # :283
        self.progress = []
        self.training_data_buffer: list[Tuple] = []
        self.training_buffer_size: int = self.num_params
        self.avg_cost = None
        self.avg_cost_history = []
        self.stats = None
        self.status = "Created"
        self._dtype = dtype or DEFAULT_DTYPE
        self._device = device
        self._engine: DeviceMLP | None = None
        self._engine_signature = None
        self._synced_versions = None

    # ------------------------------------------------------------------ containers
    @property
    def weights(self) -> list[Tensor]:
        return [layer.weights for layer in self.layers if layer.weights is not None]

    @property
    def num_params(self) -> int:
        return sum(p.numel() for p in self.params)

    def get_model_data(self) -> dict:
        return {
            "algos": self.algos,
            "layers": [layer.state_dict for layer in self.layers],
            "progress": self.progress,
            "training_data_buffer": self.training_data_buffer,
            "average_cost": self.avg_cost,
            "average_cost_history": self.avg_cost_history,
            "stats": self.stats,
            "status": self.status,
        }

    def set_model_data(self, model_data: dict):
        for layer, state in zip(self.layers, model_data["layers"]):
            layer.state_dict = state
        self.progress = model_data["progress"]
        self.training_data_buffer = model_data["training_data_buffer"]
        self.training_buffer_size = self.num_params
        self.avg_cost = model_data["average_cost"]
        self.avg_cost_history = model_data["average_cost_history"]
        self.stats = model_data["stats"]
        self.status = model_data["status"]
        self._drop_engine()

    # ------------------------------------------------------------------ persistence (:330-369)
    @staticmethod
    def _paths(model_id):
        return (os.path.join("models", f"model_{model_id}.json"),
                os.path.join("models", f"optimizer_{model_id}.pth"))

    def serialize(self):
        os.makedirs("models", exist_ok=True)
        model_path, optimizer_path = self._paths(self.model_id)
        with open(model_path, "w", encoding="utf-8") as f:
            json.dump(self.get_model_data(), f, indent=4)
        log.info(f"Model saved successfully: {model_path}")
        if self.optimizer is not None:
            torch.save(self.optimizer.state_dict(), optimizer_path)
            log.info(f"Optimizer saved successfully: {optimizer_path}")

    @classmethod
    def deserialize(cls, model_id: str):
        model_path, optimizer_path = cls._paths(model_id)
        try:
            with open(model_path, "r", encoding="utf-8") as f:
                model_data = json.load(f)
        except FileNotFoundError as e:
            log.error(f"File not found error occurred: {str(e)}")
            raise KeyError(f"Model {model_id} not created yet.")
        model = cls(model_id, activation_algos=model_data["algos"])
        model.set_model_data(model_data)
        if os.path.exists(optimizer_path):
            model.optimizer = torch.optim.Adam(model.params)
            model.optimizer.load_state_dict(torch.load(optimizer_path))
        return model

    @classmethod
    def delete(cls, model_id: str):
        model_path, optimizer_path = cls._paths(model_id)
        try:
            os.remove(model_path)
            if os.path.exists(optimizer_path):
                os.remove(optimizer_path)
        except FileNotFoundError as e:
            log.warning(f"Failed to delete: {str(e)}")

    # ------------------------------------------------------------------ device image
    def _drop_engine(self):
        if self._engine is not None:
            self._engine.close()
        self._engine = None
        self._synced_versions = None

    def _param_slots(self, spec: StackSpec) -> list[tuple[Tensor, str, int]]:
        """Every parameter tensor in the reference's `params` order with its device location."""
        slots = []
        for i, layer in enumerate(self.layers):
            if layer.algo == "embedding":
                slots.append((layer.weights, "embedding", 0))
            elif layer.algo == "linear":
                l = spec.z_index.index(i)
                slots.append((layer.weights, "weight", l))
                if layer.bias is not None:
                    slots.append((layer.bias, "bias", l))
            elif layer.algo == "batchnorm":
                l = spec.bn_index.index(i)
                slots += [(layer.gain, "bn_gain", l), (layer.bias, "bn_bias", l)]
        return slots

    def _running_slots(self, spec: StackSpec):
        return [(layer, spec.bn_index.index(i)) for i, layer in enumerate(self.layers) if layer.algo == "batchnorm"]

    def _versions(self, spec: StackSpec):
        v = [(id(t), t._version) for t, _, _ in self._param_slots(spec)]
        for layer, _ in self._running_slots(spec):
            v += [(id(layer.mean), layer.mean._version), (id(layer.variance), layer.variance._version)]
        return v

    def _ensure_engine(self) -> DeviceMLP:
        spec = self.stack
        linears = [self.layers[i] for i in spec.z_index]
        if not linears or any(l.weights is None for l in linears):
            raise ValueError("model has no parameters")
        dims = [tuple(l.weights.shape) for l in linears]
        has_bias = linears[0].bias is not None
        if any((l.bias is not None) != has_bias for l in linears):
            raise ValueError("mixed bias / no-bias layers are not supported on the device path")
        emb = tuple(self.layers[0].weights.shape) if spec.emb_index is not None else None
        bn = [None if i is None else (self.layers[i].eps, self.layers[i].momentum) for i in spec.bn_index]
        signature = (dims, has_bias, emb, bn, spec.rmul, spec.kz, spec.kbn, spec.ka, spec.emb_drop)
        eng = self._engine
        if eng is None or self._engine_signature != signature:
            self._drop_engine()
            eng = DeviceMLP(dims, spec, has_bias, self._dtype, self._device, embedding=emb, batchnorm=bn)
            self._engine, self._engine_signature = eng, signature
        versions = self._versions(spec)
        if versions != self._synced_versions:
            for t, kind, l in self._param_slots(spec):
                eng.upload_tensor(kind, l, t)
            for layer, l in self._running_slots(spec):
                eng.upload_tensor("bn_mean", l, layer.mean.reshape(-1))
                eng.upload_tensor("bn_var", l, layer.variance.reshape(-1))
            self._synced_versions = versions
        return eng

    def _download_running_stats(self):
        eng, spec = self._engine, self.stack
        for layer, l in self._running_slots(spec):
            layer.mean = torch.from_numpy(eng.download_tensor("bn_mean", l))
            layer.variance = torch.from_numpy(eng.download_tensor("bn_var", l))
        self._synced_versions = self._versions(spec)

    def _download_params(self):
        eng, spec = self._engine, self.stack
        with torch.no_grad():
            for t, kind, l in self._param_slots(spec):
                t.copy_(torch.from_numpy(eng.download_tensor(kind, l)))
        self._download_running_stats()

    def _upload_adam_state(self):
        """torch.optim.Adam.state -> device (state persists across train() calls and .pth files)."""
        eng, opt = self._engine, self.optimizer
        step = 0
        for t, kind, l in self._param_slots(self.stack):
            st = opt.state.get(t, {})
            zero = torch.zeros_like(t)
            eng.upload_tensor(kind, l, st.get("exp_avg", zero), "adam_m")
            eng.upload_tensor(kind, l, st.get("exp_avg_sq", zero), "adam_v")
            if "step" in st:
                step = int(float(st["step"]))
        eng.set_adam_step(step)

    def _download_adam_state(self):
        eng, opt = self._engine, self.optimizer
        step = eng.get_adam_step()
        for t, kind, l in self._param_slots(self.stack):
            opt.state[t] = {"step": torch.tensor(float(step)),
                            "exp_avg": torch.from_numpy(eng.download_tensor(kind, l, "adam_m")).clone(),
                            "exp_avg_sq": torch.from_numpy(eng.download_tensor(kind, l, "adam_v")).clone()}

    def _inputs_to_device(self, eng: DeviceMLP, x: Tensor) -> Tensor:
        """float64 host tensor (features, or token indices for an embedding head) -> device batch."""
        if eng.embedding:
            idx = x.long()                                                   # :68 `input_tensor.long()`
            if idx.numel() and (int(idx.min()) < -eng.embedding[0] or int(idx.max()) >= eng.embedding[0]):
                raise IndexError(f"index out of range for an embedding of {eng.embedding[0]} rows")   # as W[idx] raises
            if idx.numel() and int(idx.min()) < 0:
                idx = idx % eng.embedding[0]                                 # negative indices wrap, as in W[idx]
            return idx.to(torch.int32).to(eng.device).contiguous()
        return x.to(eng.tdtype).to(eng.device).contiguous()

    # ------------------------------------------------------------------ forward (:371-408)
    def _targets_to_device(self, eng: DeviceMLP, target, batched: bool):
        if self.stack.loss == "ce":                                  # :401-402
            labels = [t[0] for t in target] if batched else [target[0]]
            lab = torch.tensor(labels, dtype=torch.int64)
            # F.cross_entropy raises on a class index outside [0, C) (its ignore_index -100 is not used here)
            if lab.numel() and (int(lab.min()) < 0 or int(lab.max()) >= eng.out_features):
                raise IndexError(f"Target {int(lab.max()) if int(lab.max()) >= eng.out_features else int(lab.min())} is out of bounds.")
            return lab.to(torch.int32).to(eng.device)
        t = torch.tensor(target, dtype=torch.float64)                # :405
        if not batched:
            t = t.unsqueeze(0)
        return t.to(eng.tdtype).to(eng.device).contiguous()

    def _forward(self, input_tensor: Tensor, target: list, dropout_rate=0.0) -> Tuple[list[Tensor], Tensor]:
        """Tensor-level forward (+ loss when a target is given): the output of EVERY entry of the layer list, as
        float64 CPU tensors of the reference's shapes, and the cost (``torch.empty(0)`` without a target, :399).

        Without a target the entries run one after the other on the device through their own ``forward``
        operators (neural_net_model.py:389-396, ``layer.training = False``).

        With a target the stack runs as ONE training-mode pass of the fused kernels -- dropout at
        ``dropout_rate`` on every hidden entry (masks drawn in the kernels), batchnorm on batch statistics
        (running statistics move, :161-166), the configured loss -- followed by its backward pass, and the
        results are tied into autograd: ``cost.backward()`` puts the kernels' gradients into ``p.grad`` of
        every parameter that requires grad and into ``a.grad`` of every returned activation (``None`` for a
        trailing softmax, which is not on the loss graph, :545); terms added to ``cost`` by the caller (the L2
        sum of :482-483) differentiate as usual.  Gradients reach the parameters through ``cost`` only: an
        objective the caller builds from the returned activations stops at those tensors."""
        target_specified = target is not None and all(t is not None for t in target)
        for layer in self.layers:
            layer.training = target_specified                        # :390
        x = torch.as_tensor(input_tensor)
        if not target_specified:
            chain = unit_ops.Chain(self._dtype if self._dtype in ("float32", "float64") else "float32", self._device)
            cur = chain.enter(x, as_index=self.layers[0].algo == "embedding")
            outs = []
            for layer in self.layers:
                cur = layer._run(chain, cur)
                outs.append(cur)
            return [o.to("cpu", torch.float64) for o in outs], torch.empty(0)
        return self._forward_with_loss(x, target, float(dropout_rate))

    def _forward_last(self, input_tensor: Tensor, target: list) -> Tuple[Tensor, Tensor]:
        """What ``compute_output`` needs of ``_forward``: the last entry's output and the cost, through the
        fused inference kernels (poz_forward / poz_forward_loss)."""
        target_specified = target is not None and all(t is not None for t in target)
        eng = self._ensure_engine()
        x = torch.as_tensor(input_tensor, dtype=torch.float64)
        batched = x.ndim == 2
        xd = self._inputs_to_device(eng, x if batched else x.unsqueeze(0))
        for layer in self.layers:
            layer.training = target_specified                        # :390
        if target_specified:
            td = self._targets_to_device(eng, target, batched)
            out, cost = eng.forward_loss(xd, td)
            cost_t = torch.tensor(cost, dtype=torch.float64)
            if eng.general:          # layer.training = True moves the batchnorm running statistics (:161-166)
                self._download_running_stats()
        else:
            out = eng.forward(xd)
            cost_t = torch.empty(0)
        out = out.to("cpu", torch.float64)
        return (out if batched else out[0]), cost_t

    def _entry_shapes(self, spec: StackSpec, batch: int, batched: bool) -> list[tuple]:
        """Shape of every entry's output as the reference produces it for a [batch, features] (or, with an
        embedding head, [batch, tokens]) input; the leading batch dimension is absent for a single sample."""
        shapes = []
        rows, width = 1, None
        for layer in self.layers:
            if layer.algo == "embedding":
                rows, width = spec.tokens, layer.weights.shape[1]
            elif layer.algo == "flatten":
                rows, width = rows // layer.ratio, width * layer.ratio
            elif layer.algo == "linear":
                width = layer.weights.shape[1]
            lead = (batch,) if batched else ()
            if layer.algo == "embedding" or rows > 1:
                shapes.append(lead + (rows, width))
            else:
                shapes.append(lead + (width,))
        return shapes

    def _forward_with_loss(self, x: Tensor, target: list, dropout_rate: float):
        eng = self._ensure_engine()
        spec = self.stack
        batched = x.ndim == 2
        xd = self._inputs_to_device(eng, (x if batched else x.unsqueeze(0)).to(torch.float64))
        td = self._targets_to_device(eng, target, batched)
        batch = xd.shape[0]
        # One step with learning rate 0 (plain SGD: p -= 0 * g leaves the parameters, and Adam's state is not
        # touched), recorded in diagnostics mode: every entry's output, its gradient and every parameter
        # gradient of exactly this pass stay on the device.
        eng.train_begin(optimizer="sgd", learning_rate=0.0, decay_rate=1.0, dropout_rate=dropout_rate, l2_lambda=0.0,
                        epochs=1, progress_stride=0, dropout_seed=random.getrandbits(63))
        side = torch.cuda.Stream(device=eng.device)
        side.wait_stream(torch.cuda.current_stream(eng.device))
        with torch.cuda.stream(side):
            eng.train_step(xd, td, diagnostics=True)
        side.synchronize()
        cost_value = float(eng.costs(0, 1)[0])
        if eng.general:
            self._download_running_stats()
        loc = self._entry_locations(spec)
        shapes = self._entry_shapes(spec, batch, batched)
        outs, out_grads = [], []
        for i in range(len(self.layers)):
            li, which = loc[i]
            a = torch.from_numpy(eng.diag_activation(li, which, batch)).reshape(shapes[i])
            g = eng.diag_activation_grad(li, which, batch)
            outs.append(a.clone().requires_grad_())
            out_grads.append(None if g is None else torch.from_numpy(g).reshape(shapes[i]).clone())
        slots = self._param_slots(spec)
        param_grads = [torch.from_numpy(eng.download_tensor(kind, l, "diag_grad")).reshape(t.shape).clone()
                       for t, kind, l in slots]
        cost = _DeviceCost.apply(torch.tensor(cost_value, dtype=torch.float64), param_grads, *[t for t, _, _ in slots])
        for a, g in zip(outs, out_grads):
            if g is not None:
                # exactly zero in value; its derivative hands the kernels' gradient of this entry to a.grad
                cost = cost + ((a - a.detach()) * g).sum()
        return outs, cost

    def compute_output(self, input_data: list, target: list = None) -> Tuple[list, float]:
        input_tensor = _list_to_f64(input_data)                       # :379
        last, cost = self._forward_last(input_tensor, target)
        return last.tolist(), cost.item() if cost.numel() > 0 else None

    # ------------------------------------------------------------------ train (:410-530)
    def train(self, training_data: list[Tuple[list[float], list[float]]], epochs=100, learning_rate=0.01,
              batch_size=None, decay_rate=0.9, dropout_rate=0.2, l2_lambda=0.001, beta1=0.9, beta2=0.999,
              epsilon=1e-8):
        self.training_data_buffer.extend(training_data)
        if len(self.training_data_buffer) < self.training_buffer_size:     # :429-434
            log.info(f"Model {self.model_id}: Insufficient training data. "
                     f"Current buffer size: {len(self.training_data_buffer)}, "
                     f"required: {self.training_buffer_size}")
            self.serialize()
            return
        training_data = self.training_data_buffer
        self.training_data_buffer = []
        sample_size = batch_size or int(len(training_data) / epochs)        # :441
        log.info(f"Training sample size: {sample_size}")
        if self.optimizer is not None:                                      # :445-448
            for group in self.optimizer.param_groups:
                group["betas"] = (beta1, beta2)
                group["eps"] = epsilon
        self.progress = []
        self.stats = None
        self.status = "Training"
        self.serialize()

        eng = self._ensure_engine()
        if self.optimizer is not None:
            self._upload_adam_state()
        # Data parallel (new: the reference is single-process): when this process is a rank of an initialised
        # torch.distributed group, every rank draws the same indices (same torch seed, as :460 uses the global
        # generator), trains on its contiguous share of them and all-reduces the layer gradients (dp.py); the
        # update and therefore the parameters are identical on every rank.
        dp = None
        if torch.distributed.is_available() and torch.distributed.is_initialized() and torch.distributed.get_world_size() > 1:
            from .dp import DataParallelTrainer
            dp = DataParallelTrainer(eng)
        n_data = len(training_data)
        # One upload of the whole buffer; each epoch's sample is a device-side row gather inside
        # the first-layer kernels instead of the reference's per-epoch list -> tensor copy (:472).
        x_all = self._inputs_to_device(eng, _list_to_f64([x for x, _ in training_data]))
        t_all = self._targets_to_device(eng, [t for _, t in training_data], True)
        stride = max(1, epochs // 100)                                      # :504
        hyper = dict(optimizer="adam" if self.optimizer is not None else "sgd",
                     learning_rate=learning_rate, decay_rate=decay_rate, dropout_rate=dropout_rate,
                     l2_lambda=l2_lambda, beta1=beta1, beta2=beta2, epsilon=epsilon, epochs=epochs,
                     progress_stride=stride, dropout_seed=random.getrandbits(63))
        lo, hi = 0, sample_size
        if dp is not None:
            dp.train_begin(global_batch=sample_size, **hyper)
            lo, hi = dp.local_rows(sample_size)
        else:
            eng.train_begin(**hyper)
        self._dp = dp
        for layer in self.layers:
            layer.training = True
        # The indices of K epochs are drawn in ONE torch.randint call -- the generator hands out the same numbers
        # whether asked for K x batch at once or batch at a time (tests/test_abi_and_host.py pins that), so the sampled
        # rows are exactly the reference's (:460) -- and converted into a pinned staging block.  Per epoch the host then
        # issues one small asynchronous copy, on a copy stream, into one of TWO device index buffers (fixed addresses:
        # the library replays the step from two CUDA graphs) -- the copy for epoch e + 1 runs under the step of epoch e --
        # and launches the step.
        m_local = hi - lo
        K = max(1, min(16, epochs))
        idx_bufs = [torch.empty((m_local,), dtype=torch.int32, device=eng.device) for _ in range(2)]
        ring = [torch.empty((K, m_local), dtype=torch.int32).pin_memory() for _ in range(3)]
        ring_done = [None] * len(ring)
        stamps = []
        fetched = 0
        last_serialized = time.time()
        # a side stream: the library replays the step from a CUDA graph, which the legacy default
        # stream does not allow
        side = torch.cuda.Stream(device=eng.device)
        side.wait_stream(torch.cuda.current_stream(eng.device))
        copier = torch.cuda.Stream(device=eng.device)
        copied = [torch.cuda.Event() for _ in idx_bufs]
        consumed = [None for _ in idx_bufs]
        loop_started = time.perf_counter()
        with torch.cuda.stream(side):
            for epoch in range(epochs):
                j = epoch % K
                blk = (epoch // K) % len(ring)
                if j == 0:
                    k_now = min(K, epochs - epoch)
                    idx = torch.randint(0, n_data, (k_now * sample_size,)).view(k_now, sample_size)   # :460, same generator
                    if ring_done[blk] is not None:
                        ring_done[blk].synchronize()                            # the copies out of this block, three blocks ago
                    ring[blk][:k_now].copy_(idx[:, lo:hi])                       # this rank's share (all of it on one GPU)
                buf = epoch & 1
                idx_d = idx_bufs[buf]
                with torch.cuda.stream(copier):
                    if consumed[buf] is not None:
                        copier.wait_event(consumed[buf])                        # the step of two epochs ago has read this buffer
                    idx_d.copy_(ring[blk][j], non_blocking=True)
                    copied[buf].record(copier)
                    if j == K - 1 or epoch + 1 == epochs:
                        ring_done[blk] = torch.cuda.Event()
                        ring_done[blk].record(copier)
                side.wait_event(copied[buf])
                if self.optimizer is not None:                                  # :467-469 (view only)
                    for group in self.optimizer.param_groups:
                        group["lr"] = learning_rate * (decay_rate ** epoch)
                long_training = time.time() - last_serialized >= 10             # :488
                diagnostics = epoch + 1 == epochs or long_training              # :491
                if dp is not None:
                    dp.train_step(x_all, t_all, row_idx_local=idx_d, diagnostics=diagnostics)
                else:
                    eng.train_step(x_all, t_all, row_idx=idx_d, diagnostics=diagnostics)
                if consumed[buf] is None:
                    consumed[buf] = torch.cuda.Event()
                consumed[buf].record(side)
                stamps.append(dt.now().isoformat())
                if long_training:  # pragma: no cover
                    fetched = self._collect_progress(eng, stamps, fetched, epoch + 1, stride)
                    self._download_params()
                    if self.optimizer is not None:
                        self._download_adam_state()
                    self._record_training_overall_progress(self._device_statistics(eng))
                    self.serialize()
                    last_serialized = time.time()
        torch.cuda.current_stream(eng.device).wait_stream(side)
        side.synchronize()
        # wall time of the epoch loop alone (sampling, index upload, the steps, until the last one has finished):
        # what bench.py reports as the class-level rate, apart from the list ingestion and the JSON saves around it
        self.train_loop_seconds = time.perf_counter() - loop_started
        self._collect_progress(eng, stamps, fetched, epochs, stride)
        self._download_params()
        if self.optimizer is not None:
            self._download_adam_state()
        self.status = "Trained"
        log.info(f"Model {self.model_id}: Done training for {epochs} epochs.")
        self._record_training_overall_progress(self._device_statistics(eng))
        self.serialize()

    def _collect_progress(self, eng: DeviceMLP, stamps, first: int, upto: int, stride: int) -> int:
        """Fetch cost.item() of epochs [first, upto) in one copy and append the progress points the
        reference would have recorded (:503-516)."""
        if upto <= first:
            return first
        dp = getattr(self, "_dp", None)
        costs = dp.costs(first, upto - first) if dp is not None else eng.costs(first, upto - first)
        rec_first = (first + stride - 1) // stride
        rec_count = (upto - 1) // stride + 1 - rec_first
        ratios = eng.update_ratios(rec_first, rec_count) if rec_count > 0 else np.empty((0, 0))
        for epoch in range(first, upto):
            cost = float(costs[epoch - first])
            if epoch % stride == 0:
                per_linear = list(ratios[epoch // stride - rec_first])
                self.progress.append({
                    "dt": stamps[epoch],
                    "epoch": epoch + 1,
                    "cost": cost,
                    "weight_upd_ratio": [per_linear.pop(0) if layer.weights is not None and per_linear
                                         else None for layer in self.layers],
                })
            log.info(f"Model {self.model_id}: Epoch {epoch + 1}, Cost: {cost:.4f}")
        return upto

    # saturation rule per entry type (neural_net_model.py:556-561): (rule code, threshold)
    _SATURATION_RULE = {"tanh": (1, 0.97), "sigmoid": (1, 0.97), "relu": (2, 0.0), "softmax": (3, 0.97),
                        "batchnorm": (1, 3.0), "embedding": (4, 5.0)}

    def _entry_locations(self, spec: StackSpec) -> dict:
        """entry index -> (layer, which) of the device tensor holding its recorded output.  A
        flatten entry is a view of the tensor in front of it."""
        loc = {}
        if spec.emb_index is not None:
            loc[spec.emb_index] = (-1, 0)
            for i in spec.emb_flat:
                loc[i] = (-1, 0)
        for l in range(spec.n_linear):
            loc[spec.z_index[l]] = last = (l, 0)
            if spec.bn_index[l] is not None:
                loc[spec.bn_index[l]] = last = (l, 2)
            if spec.a_index[l] is not None:
                loc[spec.a_index[l]] = last = (l, 1)
            for i in spec.flat_after[l]:
                loc[i] = last
        return loc

    def _device_statistics(self, eng: DeviceMLP) -> tuple[list, list]:
        """Statistics of the last diagnostics step (neural_net_model.py:542-583), reduced on the
        device: per entry the output's mean / std / saturation / density histogram and the same for
        its gradient; per weight tensor its mean / std and its gradient's mean / std / histogram."""
        spec = self.stack
        loc = self._entry_locations(spec)
        layer_stats, weight_stats = [], []
        for i, layer in enumerate(self.layers):
            li, which = loc[i]
            rule, thr = self._SATURATION_RULE.get(layer.algo, (1, 5.0))
            act = eng.diag_stats(li, which, 0, rule, thr)
            grad = eng.diag_stats(li, which, 1)
            layer_stats.append({
                "algo": layer.algo,
                "activation": {"mean": act["mean"], "std": act["std"], "saturated": act["saturated"],
                               "histogram": act["histogram"]},
                "gradient": None if grad is None else {"mean": grad["mean"], "std": grad["std"],
                                                       "histogram": grad["histogram"]},
            })
            if layer.weights is None:
                weight_stats.append(None)
            else:
                w = eng.diag_stats(li, 0, 2, histogram=False)
                wg = eng.diag_stats(li, 0, 3)
                weight_stats.append({
                    "shape": str(tuple(layer.weights.shape)),
                    "data": {"mean": w["mean"], "std": w["std"]},
                    "gradient": {"mean": wg["mean"], "std": wg["std"], "histogram": wg["histogram"]},
                })
        return layer_stats, weight_stats

    # ------------------------------------------------------------------ diagnostics (:532-585)
    def _record_training_overall_progress(self, statistics):
        costs = [p["cost"] for p in self.progress]
        avg_progress_cost = sum(costs) / len(self.progress)
        self.avg_cost = ((self.avg_cost or avg_progress_cost) + avg_progress_cost) / 2.0
        self.avg_cost_history.append(self.avg_cost)
        if len(self.avg_cost_history) > 100:
            self.avg_cost_history.pop(random.randint(1, 98))
        layer_stats, weight_stats = statistics
        self.stats = {"layers": layer_stats, "weights": weight_stats}
        log.info(f"Model {self.model_id} - Cost: {avg_progress_cost:.4f} Overall Cost: {self.avg_cost:.4f}")
