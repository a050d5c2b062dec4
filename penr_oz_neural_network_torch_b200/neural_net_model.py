"""Drop-in ``NeuralNetworkModel`` whose hot path runs in hand-written sm_100a CUDA.

Mirror of the reference's public class (derinworks/penr-oz-neural-network-torch
``neural_net_model.py:269-585``): same constructor, ``compute_output``, ``train``, ``_forward``,
``serialize`` / ``deserialize`` / ``delete`` and the same attributes, so ``main.py`` works by
changing its import line only.  What differs is underneath: the Linear +
activation stack, the loss, the backward pass and the Adam / SGD update run through the C ABI in
``include/pozmlp.h`` on ``cuda:<device>``; the host keeps float64 CPU tensors as the parameter
*view* the reference exposes (``params``, ``weights``, ``state_dict``, ``optimizer.state_dict()``)
and synchronises them with the device image at the boundaries of ``train`` / ``compute_output``.

Scope: stacks made of Linear layers each optionally followed by
relu / sigmoid / tanh, with an optional trailing softmax.  ``embedding`` / ``flatten`` /
``batchnorm`` entries raise ``ValueError`` -- there is deliberately no CPU fallback.
"""
from __future__ import annotations

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

from .engine import ACTIVATIONS, DeviceMLP, StackSpec

log = logging.getLogger(__name__)

_KNOWN_ALGOS = ("linear",) + ACTIVATIONS
_OFF_PATH_ALGOS = ("embedding", "flatten", "batchnorm")
_GAIN = {"relu": math.sqrt(2.0), "tanh": 5.0 / 3.0}     # neural_net_model.py:178,183

DEFAULT_DTYPE = os.environ.get("POZ_DTYPE", "float32")


class Layer:
    """Parameter container + topology flags of one entry (neural_net_model.py:16-53).  Entries
    hold no compute: the stack is executed as a whole by the CUDA library."""
    algo = ""

    def __init__(self):
        self.hidden = False
        self.training = False
        self.weights: Tensor | None = None
        self.bias: Tensor | None = None

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


class SigmoidLayer(Layer):
    algo = "sigmoid"


class ReluLayer(Layer):
    algo = "relu"
    weight_gain = _GAIN["relu"]


class TanhLayer(Layer):
    algo = "tanh"
    weight_gain = _GAIN["tanh"]


class SoftmaxLayer(Layer):
    algo = "softmax"


_LAYER_TYPES = {"sigmoid": SigmoidLayer, "relu": ReluLayer, "tanh": TanhLayer, "softmax": SoftmaxLayer}


def expand_algos(activation_algos: list[str]) -> list[str]:
    """The reference's "linear sandwich" (neural_net_model.py:202-212): a Linear entry is put in
    front of every activation that does not already follow one."""
    expanded: list[str] = []
    prev = None
    for algo in activation_algos:
        if algo in ACTIVATIONS and prev not in ("linear", "batchnorm"):
            expanded.append("linear")
        expanded.append(algo)
        prev = algo
    return expanded


class MultiLayerPerceptron:
    def __init__(self, layer_sizes: list[int], weight_algo="xavier", bias_algo="zeros",
                 activation_algos: list[str] = None, batchnorm=(1e-5, 0.1)):
        requested = activation_algos or ["relu"] * (len(layer_sizes) - 1)
        for algo in requested:
            if algo not in _KNOWN_ALGOS and algo not in _OFF_PATH_ALGOS:
                raise ValueError(f"Unsupported activation algorithm: {algo}")   # neural_net_model.py:241
        self.algos = expand_algos(requested)
        self.stack = StackSpec.from_algos(self.algos)       # raises ValueError for off-path entries
        self.layers: list[Layer] = []
        n_entries = len(self.algos)
        n_sizes = len(layer_sizes)
        size_idx = 0
        pending_gain: LinearLayer | None = None
        for i, algo in enumerate(self.algos):
            if algo == "linear":
                in_sz = layer_sizes[size_idx] if size_idx < n_sizes else 0
                out_sz = layer_sizes[size_idx + 1] if size_idx + 1 < n_sizes else 0
                layer = LinearLayer(in_sz, out_sz, bias_algo)
                if in_sz > 0 and layer.weights is not None and weight_algo in ("xavier", "he"):
                    layer.weights /= math.sqrt(in_sz)                          # :243-244
                pending_gain = layer if weight_algo == "he" else None          # :246-247
                size_idx += 1
            else:
                layer = _LAYER_TYPES[algo]()
                if pending_gain is not None and pending_gain.weights is not None and algo in _GAIN:
                    pending_gain.weights *= _GAIN[algo]                        # :249-253
            layer.hidden = 0 < i < n_entries - 1                               # :255
            self.layers.append(layer)

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
    def _linear_layers(self) -> list[Layer]:
        return [layer for layer in self.layers if layer.algo == "linear"]

    def _drop_engine(self):
        if self._engine is not None:
            self._engine.close()
        self._engine = None
        self._synced_versions = None

    def _versions(self):
        return [(id(p), p._version) for p in self.params]

    def _ensure_engine(self) -> DeviceMLP:
        linears = self._linear_layers()
        if not linears or any(l.weights is None for l in linears):
            raise ValueError("model has no parameters")
        sizes = [linears[0].weights.shape[0]] + [l.weights.shape[1] for l in linears]
        has_bias = linears[0].bias is not None
        if any((l.bias is not None) != has_bias for l in linears):
            raise ValueError("mixed bias / no-bias layers are not supported on the device path")
        eng = self._engine
        if eng is None or eng.sizes != sizes or eng.has_bias != has_bias:
            self._drop_engine()
            eng = DeviceMLP(sizes, self.stack, has_bias, self._dtype, self._device)
            self._engine = eng
        versions = self._versions()
        if versions != self._synced_versions:
            for i, layer in enumerate(linears):
                eng.upload_params(i, layer.weights, layer.bias)
            self._synced_versions = versions
        return eng

    def _download_params(self):
        eng = self._engine
        with torch.no_grad():
            for i, layer in enumerate(self._linear_layers()):
                w, b = eng.download_params(i)
                layer.weights.copy_(torch.from_numpy(w))
                if layer.bias is not None:
                    layer.bias.copy_(torch.from_numpy(b))
        self._synced_versions = self._versions()

    def _upload_adam_state(self):
        """torch.optim.Adam.state -> device (state persists across train() calls and .pth files)."""
        eng, opt = self._engine, self.optimizer
        step = 0
        for i, layer in enumerate(self._linear_layers()):
            sw = opt.state.get(layer.weights, {})
            sb = opt.state.get(layer.bias, {}) if layer.bias is not None else {}
            zw = torch.zeros_like(layer.weights)
            zb = torch.zeros_like(layer.bias) if layer.bias is not None else None
            eng.upload_adam(i, sw.get("exp_avg", zw), sw.get("exp_avg_sq", zw),
                            sb.get("exp_avg", zb) if zb is not None else None,
                            sb.get("exp_avg_sq", zb) if zb is not None else None)
            if "step" in sw:
                step = int(float(sw["step"]))
        eng.set_adam_step(step)

    def _download_adam_state(self):
        eng, opt = self._engine, self.optimizer
        step = eng.get_adam_step()
        for i, layer in enumerate(self._linear_layers()):
            m_w, v_w, m_b, v_b = eng.download_adam(i)
            pairs = [(layer.weights, m_w, v_w)]
            if layer.bias is not None:
                pairs.append((layer.bias, m_b, v_b))
            for p, m, v in pairs:
                opt.state[p] = {"step": torch.tensor(float(step)),
                                "exp_avg": torch.from_numpy(m).clone(),
                                "exp_avg_sq": torch.from_numpy(v).clone()}

    # ------------------------------------------------------------------ forward (:371-408)
    def _targets_to_device(self, eng: DeviceMLP, target, batched: bool):
        if self.stack.loss == "ce":                                  # :401-402
            labels = [t[0] for t in target] if batched else [target[0]]
            return torch.tensor(labels, dtype=torch.int64).to(torch.int32).to(eng.device)
        t = torch.tensor(target, dtype=torch.float64)                # :405
        if not batched:
            t = t.unsqueeze(0)
        return t.to(eng.tdtype).to(eng.device).contiguous()

    def _forward(self, input_tensor: Tensor, target: list, dropout_rate=0.0) -> Tuple[list[Tensor], Tensor]:
        """Tensor-level forward (+ loss when a target is given).  Returns ``([final_output], cost)``
        as float64 CPU tensors; ``cost`` is ``torch.empty(0)`` without a target (:399).  Gradients
        do not flow through this call -- the backward pass lives in ``train``."""
        target_specified = target is not None and all(t is not None for t in target)
        eng = self._ensure_engine()
        x = torch.as_tensor(input_tensor, dtype=torch.float64)
        batched = x.ndim == 2
        xd = (x if batched else x.unsqueeze(0)).to(eng.tdtype).to(eng.device).contiguous()
        for layer in self.layers:
            layer.training = target_specified                        # :390
        if target_specified:
            td = self._targets_to_device(eng, target, batched)
            out, cost = eng.forward_loss(xd, td)
            cost_t = torch.tensor(cost, dtype=torch.float64)
        else:
            out = eng.forward(xd)
            cost_t = torch.empty(0)
        out = out.to("cpu", torch.float64)
        return [out if batched else out[0]], cost_t

    def compute_output(self, input_data: list, target: list = None) -> Tuple[list, float]:
        input_tensor = torch.tensor(input_data, dtype=torch.float64)  # :379
        activations, cost = self._forward(input_tensor, target)
        return activations[-1].tolist(), cost.item() if cost.numel() > 0 else None

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
        n_data = len(training_data)
        # One upload of the whole buffer; each epoch's sample is a device-side row gather inside
        # the first-layer kernels instead of the reference's per-epoch list -> tensor copy (:472).
        x_all = torch.tensor([x for x, _ in training_data], dtype=torch.float64)
        x_all = x_all.to(eng.tdtype).to(eng.device).contiguous()
        t_all = self._targets_to_device(eng, [t for _, t in training_data], True)
        stride = max(1, epochs // 100)                                      # :504
        eng.train_begin(optimizer="adam" if self.optimizer is not None else "sgd",
                        learning_rate=learning_rate, decay_rate=decay_rate, dropout_rate=dropout_rate,
                        l2_lambda=l2_lambda, beta1=beta1, beta2=beta2, epsilon=epsilon, epochs=epochs,
                        progress_stride=stride, dropout_seed=random.getrandbits(63))
        for layer in self.layers:
            layer.training = True
        # one persistent index buffer: every epoch reuses the same device addresses, so the
        # library replays the step from a CUDA graph after the second epoch
        idx_d = torch.empty((sample_size,), dtype=torch.int32, device=eng.device)
        stamps = []
        fetched = 0
        last_serialized = time.time()
        # a side stream: the library replays the step from a CUDA graph, which the legacy default
        # stream does not allow
        side = torch.cuda.Stream(device=eng.device)
        side.wait_stream(torch.cuda.current_stream(eng.device))
        with torch.cuda.stream(side):
            for epoch in range(epochs):
                idx = torch.randint(0, n_data, (sample_size,))                  # :460, same generator
                idx_d.copy_(idx.to(torch.int32))
                if self.optimizer is not None:                                  # :467-469 (view only)
                    for group in self.optimizer.param_groups:
                        group["lr"] = learning_rate * (decay_rate ** epoch)
                long_training = time.time() - last_serialized >= 10             # :488
                diagnostics = epoch + 1 == epochs or long_training              # :491
                eng.train_step(x_all, t_all, row_idx=idx_d, diagnostics=diagnostics)
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
        costs = eng.costs(first, upto - first)
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

    _SATURATION_RULE = {"tanh": (1, 0.97), "sigmoid": (1, 0.97), "relu": (2, 0.0), "softmax": (3, 0.97)}

    def _device_statistics(self, eng: DeviceMLP) -> tuple[list, list]:
        """Statistics of the last diagnostics step (neural_net_model.py:542-583), reduced on the
        device: per entry the output's mean / std / saturation / density histogram and the same for
        its gradient; per weight matrix its mean / std and its gradient's mean / std / histogram."""
        spec = self.stack
        layer_stats, weight_stats = [], []
        for i, layer in enumerate(self.layers):
            if layer.algo == "linear":
                li, which = spec.z_index.index(i), 0
            else:
                li, which = spec.a_index.index(i), 1
            rule, thr = self._SATURATION_RULE.get(layer.algo, (1, 5.0))          # :556-561
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
