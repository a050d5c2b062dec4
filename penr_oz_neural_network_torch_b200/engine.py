"""Tensor-level host interface to the CUDA hot path (one ``DeviceMLP`` = one ``poz_ctx``).

PyTorch is used here for device memory, streams and ``torch.distributed`` only; every FLOP of the
forward / loss / backward / optimizer path runs in the kernels of ``libpozmlp.so``.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np
import torch

from . import _native as nat

ACTIVATIONS = ("relu", "sigmoid", "tanh", "softmax")
_TORCH_DTYPE = {"float32": torch.float32, "float64": torch.float64, "bfloat16": torch.bfloat16}
_DTYPE_CODE = {"float32": nat.DTYPE_F32, "float64": nat.DTYPE_F64, "bfloat16": nat.DTYPE_BF16}


@dataclass
class StackSpec:
    """Units derived from the reference's expanded ``algos`` list.

    A unit is one Linear entry with an optional BatchNorm and an optional activation after it.
    ``drop_z`` / ``drop_a`` (and ``kz`` / ``kbn`` / ``ka``) carry the reference's ``hidden`` flag
    (neural_net_model.py:255) of the entries: the dropout points of neural_net_model.py:393-395.
    Flatten entries (pure reshapes) fold their dropout into the tensor in front of them and divide
    the row count: ``rmul[l]`` = rows of unit l per sample.  An embedding entry may lead the stack.
    """
    algos: list[str]
    acts: list[str] = field(default_factory=list)
    z_index: list[int] = field(default_factory=list)
    a_index: list[int | None] = field(default_factory=list)
    bn_index: list[int | None] = field(default_factory=list)
    drop_z: list[bool] = field(default_factory=list)
    drop_a: list[bool] = field(default_factory=list)
    kz: list[int] = field(default_factory=list)
    kbn: list[int] = field(default_factory=list)
    ka: list[int] = field(default_factory=list)
    rmul: list[int] = field(default_factory=list)
    flat_after: list[list[int]] = field(default_factory=list)   # flatten entries folded into unit l
    emb_index: int | None = None
    emb_drop: int = 0
    emb_flat: list[int] = field(default_factory=list)           # flatten entries right after the embedding
    tokens: int = 1

    @property
    def n_linear(self) -> int:
        return len(self.acts)

    @property
    def loss(self) -> str:
        return "ce" if self.algos and self.algos[-1] == "softmax" else "mse"

    @property
    def general(self) -> bool:
        return (self.emb_index is not None or any(b is not None for b in self.bn_index)
                or any(m != 1 for m in self.rmul))

    @staticmethod
    def from_algos(algos: list[str], ratios: dict | None = None) -> "StackSpec":
        """``ratios`` maps the index of every flatten entry to its ratio (known once the layers
        hold their sizes); without it only the structure is validated."""
        spec = StackSpec(list(algos))
        n, i = len(algos), 0
        hidden = lambda k: 0 < k < n - 1
        ratio_of = lambda k: int((ratios or {}).get(k, 1)) or 1
        divs = []                     # cumulative flatten ratio in front of each unit
        div = 1
        if n and algos[0] == "embedding":
            spec.emb_index = 0
            i = 1
            while i < n and algos[i] == "flatten":
                spec.emb_flat.append(i)
                spec.emb_drop += int(hidden(i))
                div *= ratio_of(i)
                i += 1
        while i < n:
            if algos[i] != "linear":
                raise ValueError(f"Activation algorithm not on the B200 hot path here: {algos[i]} at entry {i} of "
                                 f"{algos} (a stack is [embedding] then Linear [batchnorm] [relu|sigmoid|tanh|softmax] "
                                 f"units, with flatten entries between them)")
            z = i
            kz, kbn, ka = int(hidden(z)), 0, 0
            bn_idx = a_idx = None
            act = "none"
            i += 1
            if i < n and algos[i] == "batchnorm":
                bn_idx, kbn = i, int(hidden(i))
                i += 1
            if i < n and algos[i] in ACTIVATIONS:
                act, a_idx, ka = algos[i], i, int(hidden(i))
                i += 1
            if act == "softmax" and a_idx != n - 1:
                raise ValueError("softmax is supported as the last activation only")
            flats = []
            divs.append(div)
            while i < n and algos[i] == "flatten":
                flats.append(i)
                k = int(hidden(i))
                if a_idx is not None:
                    ka += k
                elif bn_idx is not None:
                    kbn += k
                else:
                    kz += k
                div *= ratio_of(i)
                i += 1
            spec.acts.append(act)
            spec.z_index.append(z)
            spec.a_index.append(a_idx)
            spec.bn_index.append(bn_idx)
            spec.kz.append(kz); spec.kbn.append(kbn); spec.ka.append(ka)
            spec.drop_z.append(kz > 0)
            spec.drop_a.append(ka > 0)
            spec.flat_after.append(flats)
        if not spec.acts and n:
            raise ValueError(f"stack {algos} has no Linear layer")
        spec.tokens = div
        spec.rmul = [div // d for d in divs]
        if ratios is not None:
            if spec.emb_index is None and div != 1:
                raise ValueError("flatten entries need an embedding head on the device path")
            if spec.rmul and spec.rmul[-1] != 1:
                raise ValueError("the stack must end with one row per sample (flatten ratios do not use up the tokens)")
            if spec.bn_index and spec.bn_index[-1] is not None:
                raise ValueError("batchnorm on the last Linear layer is not supported on the device path")
        return spec


def _arr(ctype, values):
    return (ctype * len(values))(*values)


def _ptr(t) -> int:
    return 0 if t is None else int(t.data_ptr())


def _stream_ptr(device) -> int:
    return int(torch.cuda.current_stream(device).cuda_stream)


class DeviceMLP:
    """Device image of one model: parameters, Adam state, workspace; tensor-level fwd/train."""

    def __init__(self, sizes, spec: StackSpec, has_bias: bool, dtype: str = "float32", device=None, *,
                 embedding: tuple[int, int] | None = None, batchnorm: list | None = None):
        """``sizes``: the chained layer widths of a plain stack, or a list of (in, out) pairs when
        flatten entries make consecutive widths differ.  ``embedding`` = (vocab, dim);
        ``batchnorm[l]`` = (eps, momentum) or None per Linear layer."""
        if not torch.cuda.is_available():
            raise nat.PozError("no CUDA device: the NeuralNetworkModel hot path has no CPU fallback")
        if dtype not in _DTYPE_CODE:
            raise ValueError(f"dtype must be one of {sorted(_DTYPE_CODE)}")
        if sizes and isinstance(sizes[0], (tuple, list)):
            dims = [(int(a), int(b)) for a, b in sizes]
        else:
            dims = [(int(a), int(b)) for a, b in zip(sizes[:-1], sizes[1:])]
        if len(dims) != spec.n_linear:
            raise ValueError("layer sizes do not match the stack")
        self.lib = nat.lib()
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else
                                   (device.index if isinstance(device, torch.device) else int(device)))
        self.dims = dims
        self.sizes = [dims[0][0]] + [d[1] for d in dims]      # chained widths (plain stacks)
        self.spec = spec
        self.has_bias = bool(has_bias)
        self.dtype_name = dtype
        self.tdtype = _TORCH_DTYPE[dtype]
        # gradients / master parameters: float32 in bfloat16 contexts
        self.gdtype = torch.float64 if dtype == "float64" else torch.float32
        self.n_linear = spec.n_linear
        self.embedding = tuple(embedding) if embedding else None
        self.batchnorm = list(batchnorm) if batchnorm else [None] * self.n_linear
        self.general = spec.general
        self.tokens = spec.tokens if self.embedding else 1
        self.in_features = self.tokens if self.embedding else dims[0][0]
        self.out_features = dims[-1][1]
        handle = C.c_void_p()
        loss = nat.LOSS_CE if spec.loss == "ce" else nat.LOSS_MSE
        with torch.cuda.device(self.device):
            if not self.general:
                nat.check(self.lib.poz_ctx_create(
                    self.device.index, self.n_linear, _arr(C.c_int, self.sizes),
                    _arr(C.c_int, [nat.ACT_CODES[a] for a in spec.acts]),
                    _arr(C.c_ubyte, [int(b) for b in spec.drop_z]),
                    _arr(C.c_ubyte, [int(b) for b in spec.drop_a]),
                    int(self.has_bias), loss, _DTYPE_CODE[dtype], C.byref(handle)))
            else:
                layers = (nat.LayerDesc * self.n_linear)()
                for l, (k_in, k_out) in enumerate(dims):
                    bn = self.batchnorm[l]
                    layers[l] = nat.LayerDesc(k_in, k_out, nat.ACT_CODES[spec.acts[l]], int(bn is not None),
                                              float(bn[0]) if bn else 1e-5, float(bn[1]) if bn else 0.1,
                                              spec.kz[l], spec.kbn[l], spec.ka[l], spec.rmul[l])
                desc = nat.StackDesc(self.n_linear, layers, int(self.has_bias), loss, _DTYPE_CODE[dtype],
                                     self.embedding[0] if self.embedding else 0,
                                     self.embedding[1] if self.embedding else 0, spec.emb_drop, self.tokens)
                nat.check(self.lib.poz_ctx_create_ex(self.device.index, C.byref(desc), C.byref(handle)))
        self.ctx = handle
        self._epochs_declared = 0

    # ---- lifetime ----
    def close(self):
        if getattr(self, "ctx", None):
            self.lib.poz_ctx_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- parameters / optimizer state (float64 host images) ----
    @staticmethod
    def _f64(a):
        if a is None:
            return None
        if isinstance(a, torch.Tensor):
            a = a.detach().cpu().numpy()
        return np.ascontiguousarray(a, dtype=np.float64)

    def upload_params(self, layer: int, w, b=None):
        w, b = self._f64(w), self._f64(b)
        assert w.shape == self.dims[layer], w.shape
        nat.check(self.lib.poz_params_upload(self.ctx, layer, w.ctypes.data,
                                             b.ctypes.data if b is not None else None))

    def download_params(self, layer: int):
        w = np.empty(self.dims[layer], dtype=np.float64)
        b = np.empty((self.dims[layer][1],), dtype=np.float64) if self.has_bias else None
        nat.check(self.lib.poz_params_download(self.ctx, layer, w.ctypes.data,
                                               b.ctypes.data if b is not None else None))
        return w, b

    def upload_adam(self, layer: int, m_w, v_w, m_b=None, v_b=None):
        m_w, v_w, m_b, v_b = (self._f64(t) for t in (m_w, v_w, m_b, v_b))
        nat.check(self.lib.poz_adam_upload(
            self.ctx, layer, m_w.ctypes.data, v_w.ctypes.data,
            m_b.ctypes.data if m_b is not None else None, v_b.ctypes.data if v_b is not None else None))

    def download_adam(self, layer: int):
        shp = self.dims[layer]
        m_w, v_w = np.empty(shp), np.empty(shp)
        m_b = np.empty((shp[1],)) if self.has_bias else None
        v_b = np.empty((shp[1],)) if self.has_bias else None
        nat.check(self.lib.poz_adam_download(
            self.ctx, layer, m_w.ctypes.data, v_w.ctypes.data,
            m_b.ctypes.data if m_b is not None else None, v_b.ctypes.data if v_b is not None else None))
        return m_w, v_w, m_b, v_b

    # generic tensor <-> float64 host image (poz_tensor_io): kinds and buffers as in include/pozmlp.h
    KIND = {"weight": 0, "bias": 1, "embedding": 2, "bn_gain": 3, "bn_bias": 4, "bn_mean": 5, "bn_var": 6}
    BUFFER = {"value": 0, "adam_m": 1, "adam_v": 2, "diag_grad": 3}

    def tensor_shape(self, kind: str, layer: int):
        if kind == "embedding":
            return tuple(self.embedding)
        if kind == "weight":
            return self.dims[layer]
        return (self.dims[layer][1],)

    def upload_tensor(self, kind: str, layer: int, values, buffer: str = "value"):
        a = self._f64(values)
        assert a.shape == self.tensor_shape(kind, layer), (kind, layer, a.shape)
        nat.check(self.lib.poz_tensor_io(self.ctx, self.KIND[kind], layer, self.BUFFER[buffer], 0, a.ctypes.data))

    def download_tensor(self, kind: str, layer: int, buffer: str = "value") -> np.ndarray:
        out = np.empty(self.tensor_shape(kind, layer), dtype=np.float64)
        nat.check(self.lib.poz_tensor_io(self.ctx, self.KIND[kind], layer, self.BUFFER[buffer], 1, out.ctypes.data))
        return out

    def set_adam_step(self, step: int):
        nat.check(self.lib.poz_adam_set_step(self.ctx, int(step)))

    def get_adam_step(self) -> int:
        v = C.c_longlong()
        nat.check(self.lib.poz_adam_get_step(self.ctx, C.byref(v)))
        return int(v.value)

    # ---- helpers ----
    def _check_x(self, x: torch.Tensor):
        want = torch.int32 if self.embedding else self.tdtype
        if x.device != self.device or x.dtype != want or not x.is_contiguous():
            raise ValueError(f"x must be a contiguous {want} tensor on {self.device}")
        if x.ndim != 2 or x.shape[1] != self.in_features:
            raise ValueError(f"x must be [batch, {self.in_features}], got {tuple(x.shape)}")

    def _check_target(self, t: torch.Tensor, rows: int | None):
        if t.device != self.device or not t.is_contiguous():
            raise ValueError("target must be a contiguous tensor on the model's device")
        if self.spec.loss == "ce":
            if t.dtype != torch.int32 or t.ndim != 1:
                raise ValueError("cross-entropy target must be int32 labels [batch]")
        else:
            if t.dtype != self.tdtype or t.ndim != 2 or t.shape[1] != self.out_features:
                raise ValueError(f"MSE target must be {self.tdtype} [batch, {self.out_features}]")
        if rows is not None and t.shape[0] != rows:
            raise ValueError("target rows do not match the batch")

    # ---- inference (neural_net_model.py:371-396) ----
    def forward(self, x: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
        self._check_x(x)
        batch = x.shape[0]
        if out is None:
            out = torch.empty((batch, self.out_features), dtype=self.tdtype, device=self.device)
        if batch:
            nat.check(self.lib.poz_forward(self.ctx, _ptr(x), batch, _ptr(out), _stream_ptr(self.device)))
        return out

    def forward_loss(self, x: torch.Tensor, target: torch.Tensor):
        self._check_x(x)
        self._check_target(target, x.shape[0])
        out = torch.empty((x.shape[0], self.out_features), dtype=self.tdtype, device=self.device)
        cost = C.c_double()
        nat.check(self.lib.poz_forward_loss(self.ctx, _ptr(x), _ptr(target), x.shape[0], _ptr(out),
                                            C.byref(cost), _stream_ptr(self.device)))
        return out, float(cost.value)

    # ---- training (neural_net_model.py:444-500) ----
    def train_begin(self, *, optimizer: str = "adam", learning_rate=0.01, decay_rate=0.9,
                    dropout_rate=0.0, l2_lambda=0.0, beta1=0.9, beta2=0.999, epsilon=1e-8,
                    epochs=1, progress_stride=0, dropout_seed=0, global_batch=0):
        cfg = nat.TrainConfig(
            nat.OPT_ADAM if optimizer == "adam" else nat.OPT_SGD, float(learning_rate),
            float(decay_rate), float(dropout_rate), float(l2_lambda), float(beta1), float(beta2),
            float(epsilon), int(epochs), int(progress_stride), int(dropout_seed), int(global_batch))
        nat.check(self.lib.poz_train_begin(self.ctx, C.byref(cfg)))
        self._epochs_declared = int(epochs)

    def _mask_array(self, masks):
        if masks is None:
            return None, None
        keep = []
        # slots: plain stacks 2*l + {0: z, 1: a}; general stacks 3*l + {0: z, 1: a, 2: bn} and 3*n for the
        # embedding output
        n_slots = 3 * self.n_linear + 1 if self.general else 2 * self.n_linear
        ptrs = (C.c_void_p * n_slots)()
        for i in range(n_slots):
            m = masks.get(i) if isinstance(masks, dict) else masks[i]
            if m is None:
                ptrs[i] = None
            else:
                if m.dtype != torch.uint8 or m.device != self.device or not m.is_contiguous():
                    raise ValueError("keep masks must be contiguous uint8 tensors on the device")
                keep.append(m)
                ptrs[i] = m.data_ptr()
        return ptrs, keep

    def train_step(self, x, target, row_idx=None, masks=None, diagnostics=False):
        batch = self._check_step_args(x, target, row_idx)
        ptrs, _keep = self._mask_array(masks)
        nat.check(self.lib.poz_train_step(self.ctx, _ptr(x), _ptr(target), _ptr(row_idx), batch, ptrs,
                                          int(bool(diagnostics)), _stream_ptr(self.device)))

    def train_step_host(self, x_host: torch.Tensor, target_host: torch.Tensor) -> float:
        """One step from (pinned) HOST tensors through poz_train_step_host: copy, step and the
        read-back of an earlier step's cost in a single C call.  Returns the cost of the host step issued
        HOST_LAG (= 2, include/pozmlp.h POZ_HOST_LAG) calls earlier, NaN on the first HOST_LAG calls;
        ``train_host_flush`` returns the costs still outstanding.  The host tensors of a call may be
        overwritten once HOST_LAG further calls (or the flush) have returned."""
        if x_host.is_cuda or target_host.is_cuda or not x_host.is_contiguous() or not target_host.is_contiguous():
            raise ValueError("train_step_host takes contiguous CPU tensors (pinned for asynchronous copies)")
        want = torch.int32 if self.embedding else self.tdtype
        if x_host.dtype != want or x_host.ndim != 2 or x_host.shape[1] != self.in_features:
            raise ValueError(f"x_host must be {want} [batch, {self.in_features}]")
        prev = C.c_double()
        nat.check(self.lib.poz_train_step_host(self.ctx, x_host.data_ptr(), target_host.data_ptr(), x_host.shape[0],
                                               C.byref(prev), _stream_ptr(self.device)))
        return float(prev.value)

    HOST_LAG = 2

    def train_host_flush(self) -> list:
        """Waits for the outstanding host steps; their costs, oldest first (HOST_LAG entries, NaN where
        fewer steps were outstanding)."""
        last = (C.c_double * self.HOST_LAG)()
        nat.check(self.lib.poz_train_host_flush(self.ctx, last))
        return [float(v) for v in last]

    def train_backward(self, x, target, row_idx=None, masks=None, diagnostics=False):
        batch = self._check_step_args(x, target, row_idx)
        ptrs, _keep = self._mask_array(masks)
        nat.check(self.lib.poz_train_backward(self.ctx, _ptr(x), _ptr(target), _ptr(row_idx), batch, ptrs,
                                              int(bool(diagnostics)), _stream_ptr(self.device)))

    def train_apply(self):
        nat.check(self.lib.poz_train_apply(self.ctx, _stream_ptr(self.device)))

    def _check_step_args(self, x, target, row_idx) -> int:
        self._check_x(x)
        if row_idx is not None:
            if row_idx.dtype != torch.int32 or row_idx.device != self.device or row_idx.ndim != 1:
                raise ValueError("row_idx must be an int32 vector on the device")
            self._check_target(target, None)
            return int(row_idx.shape[0])
        self._check_target(target, x.shape[0])
        return int(x.shape[0])

    def grad_buffer(self) -> torch.Tensor:
        """Zero-copy torch view of the flat local gradient (valid until the next step)."""
        p, n, dt = C.c_void_p(), C.c_longlong(), C.c_int()
        nat.check(self.lib.poz_grad_buffer(self.ctx, C.byref(p), C.byref(n), C.byref(dt),
                                           _stream_ptr(self.device)))
        return _wrap_device_pointer(p.value, int(n.value), self.gdtype, self.device)

    def grad_layer(self, layer: int, stream: "torch.cuda.Stream") -> torch.Tensor:
        """Layer ``layer``'s slice of the flat gradient, ready on ``stream`` (see poz_grad_layer)."""
        p, n = C.c_void_p(), C.c_longlong()
        nat.check(self.lib.poz_grad_layer(self.ctx, layer, C.byref(p), C.byref(n), int(stream.cuda_stream)))
        return _wrap_device_pointer(p.value, int(n.value), self.gdtype, self.device)

    def costs(self, first: int, count: int, stream=None) -> np.ndarray:
        """cost.item() of epochs [first, first+count); synchronises `stream` (default: current)."""
        out = np.empty((count,), dtype=np.float64)
        sp = _stream_ptr(self.device) if stream is None else int(stream.cuda_stream)
        nat.check(self.lib.poz_train_costs(self.ctx, first, count, out.ctypes.data, sp))
        return out

    def update_ratios(self, first_record: int, count: int) -> np.ndarray:
        out = np.empty((count, self.n_linear + (1 if self.embedding else 0)), dtype=np.float64)
        nat.check(self.lib.poz_train_update_ratios(self.ctx, first_record, count, out.ctypes.data,
                                                   _stream_ptr(self.device)))
        return out

    # ---- per-kernel timing (bench.py roofline leg) ----
    def profile_begin(self):
        nat.check(self.lib.poz_profile_begin(self.ctx))

    def profile_end(self) -> dict:
        import json
        buf = C.create_string_buffer(1 << 16)
        nat.check(self.lib.poz_profile_end(self.ctx, buf, len(buf)))
        return json.loads(buf.value.decode())

    # ---- diagnostics of the last diagnostics step (neural_net_model.py:491-493,532-583) ----
    def _diag_shape(self, layer: int, batch: int):
        if layer < 0:
            return (batch * self.tokens, self.embedding[1])
        return (batch * self.spec.rmul[layer], self.dims[layer][1])

    def diag_activation(self, layer: int, which: int, batch: int) -> np.ndarray:
        """Recorded output of one entry of the last diagnostics step: layer l, which 0 = Linear,
        1 = activation, 2 = batchnorm; layer -1 = the embedding head."""
        out = np.empty(self._diag_shape(layer, batch), dtype=np.float64)
        nat.check(self.lib.poz_diag_activation(self.ctx, layer, which, out.ctypes.data))
        return out

    def diag_activation_grad(self, layer: int, which: int, batch: int):
        out = np.empty(self._diag_shape(layer, batch), dtype=np.float64)
        has = C.c_int()
        nat.check(self.lib.poz_diag_activation_grad(self.ctx, layer, which, out.ctypes.data, C.byref(has)))
        return out if has.value else None

    def diag_stats(self, layer: int, which: int, source: int, rule: int = 0, threshold: float = 0.0,
                   histogram: bool = True):
        """mean / unbiased std / saturated fraction / torch.histogram(density=True) of one diagnostics
        tensor, computed on the device (neural_net_model.py:542-583).  Returns None if the tensor does
        not exist (e.g. the gradient of the trailing softmax).  source: 0 output, 1 its gradient,
        2 weights, 3 weight gradient."""
        m = np.zeros(6)
        nat.check(self.lib.poz_diag_moments(self.ctx, layer, which, source, rule, float(threshold), m.ctypes.data))
        n = m[0]
        if n == 0:
            return None
        mean = m[1] / n
        var = (m[2] - n * mean * mean) / (n - 1) if n > 1 else float("nan")
        out = {"mean": float(mean), "std": float(np.sqrt(max(var, 0.0))) if n > 1 else float("nan"),
               "saturated": float(m[5])}
        if histogram:
            lo, hi = float(m[3]), float(m[4])
            if lo == hi:                       # torch.histogram widens a degenerate range by 0.5 each side
                lo, hi = lo - 0.5, hi + 0.5
            edges = torch.linspace(lo, hi, 101, dtype=torch.float64)
            e = np.ascontiguousarray(edges.numpy())
            counts = np.zeros(100)
            nat.check(self.lib.poz_diag_histogram(self.ctx, layer, which, source, e.ctypes.data, 100,
                                                  counts.ctypes.data))
            widths = np.diff(e)
            out["histogram"] = {"x": edges[:-1].tolist(), "y": (counts / (n * widths)).tolist()}
        return out

    def diag_weight_grad(self, layer: int):
        w = np.empty(self.dims[layer], dtype=np.float64)
        b = np.empty((self.dims[layer][1],), dtype=np.float64) if self.has_bias else None
        nat.check(self.lib.poz_diag_weight_grad(self.ctx, layer, w.ctypes.data,
                                                b.ctypes.data if b is not None else None))
        return w, b


def forward_noalloc(sizes, acts, weights, biases, x: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    """Inference through ``poz_forward_noalloc``: no context, no allocation inside the call.

    Meant for code that already owns all of its device memory and may not ask for more (a server
    whose pool has claimed the GPU, a loop that cannot stall on an allocation).  ``weights`` /
    ``biases`` are float32 CUDA tensors ``[in, out]`` / ``[out]`` (``biases`` or its items may be
    None), ``x`` ``[batch, sizes[0]]``, ``out`` a caller-provided ``[batch, sizes[-1]]`` tensor.
    Layers wider than ``_native.NOALLOC_MAX_WIDTH`` are refused.  Calls on one GPU must not overlap:
    they share one fixed scratch array inside the library.
    """
    lib = nat.lib()
    n = len(acts)
    sizes = [int(v) for v in sizes]
    if not x.is_cuda:
        raise ValueError("forward_noalloc takes CUDA tensors")
    dev = x.device if x.device.index is not None else torch.device("cuda", torch.cuda.current_device())
    if len(sizes) != n + 1 or len(weights) != n or (biases is not None and len(biases) != n):
        raise ValueError("forward_noalloc: sizes, acts, weights and biases do not describe the same stack")
    for t in list(weights) + [b for b in (biases or []) if b is not None] + [x, out]:
        if t.dtype != torch.float32 or not t.is_cuda or t.get_device() != dev.index or not t.is_contiguous():
            raise ValueError("forward_noalloc takes contiguous float32 tensors on one CUDA device")
    for l in range(n):
        if tuple(weights[l].shape) != (sizes[l], sizes[l + 1]):
            raise ValueError(f"forward_noalloc: weights[{l}] must be [{sizes[l]}, {sizes[l + 1]}]")
        if biases is not None and biases[l] is not None and tuple(biases[l].shape) != (sizes[l + 1],):
            raise ValueError(f"forward_noalloc: biases[{l}] must be [{sizes[l + 1]}]")
    if x.ndim != 2 or x.shape[1] != sizes[0] or tuple(out.shape) != (x.shape[0], sizes[-1]):
        raise ValueError(f"forward_noalloc: x must be [batch, {sizes[0]}] and out [batch, {sizes[-1]}]")
    wp = (C.c_void_p * n)(*[w.data_ptr() for w in weights])
    bp = None
    if biases is not None:
        bp = (C.c_void_p * n)(*[None if b is None else b.data_ptr() for b in biases])
    nat.check(lib.poz_forward_noalloc(dev.index, n, _arr(C.c_int, sizes),
                                      _arr(C.c_int, [nat.ACT_CODES[a] for a in acts]), wp, bp,
                                      _ptr(x), x.shape[0], _ptr(out), _stream_ptr(dev)))
    return out


class _DevPtr:
    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False),
                                         "version": 2, "strides": None}


def _wrap_device_pointer(ptr: int, n: int, dtype: torch.dtype, device) -> torch.Tensor:
    typestr = "<f4" if dtype == torch.float32 else "<f8"
    with torch.cuda.device(device):
        return torch.as_tensor(_DevPtr(ptr, n, typestr), device=device)


def bind_host_to_gpu_numa_node(device) -> dict:
    """Pin this process (and therefore the pinned buffers it allocates next) to the CPUs of the NUMA
    node the GPU hangs off, so that host->device copies and launches of a feeder do not cross the
    inter-socket link on a multi-socket box.  A no-op where the platform reports no NUMA node for the
    GPU (e.g. single-node VMs).  Best effort: returns what it did, never raises."""
    import os
    info = {"bound": False}
    try:
        idx = device.index if isinstance(device, torch.device) else int(device)
        pr = torch.cuda.get_device_properties(idx)
        bus = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bus}/numa_node") as f:
            node = int(f.read().strip())
        info["pci"] = bus
        info["numa_node"] = node
        if node < 0:
            return info
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            spec = f.read().strip()
        cpus = set()
        for part in spec.split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = os.sched_getaffinity(0) & cpus
        if allowed:
            os.sched_setaffinity(0, allowed)
            info["bound"] = True
            info["cpus"] = len(allowed)
    except Exception as e:          # no sysfs, no permission, not Linux ...
        info["error"] = str(e)
    return info


class HostBatchFeeder:
    """End-to-end stepping from HOST buffers: every step copies its batch from pinned host memory
    to the device (on a copy stream, double-buffered so the copy of step i+1 overlaps the kernels
    of step i) and reads that step's cost back to the host, like the reference's per-epoch
    ``cost.item()`` (neural_net_model.py:503)."""

    def __init__(self, eng: DeviceMLP, batch: int):
        self.eng = eng
        dev = eng.device
        self.batch = batch
        self.x_dev = [torch.empty((batch, eng.sizes[0]), dtype=eng.tdtype, device=dev) for _ in range(2)]
        if eng.spec.loss == "ce":
            self.t_dev = [torch.empty((batch,), dtype=torch.int32, device=dev) for _ in range(2)]
        else:
            self.t_dev = [torch.empty((batch, eng.sizes[-1]), dtype=eng.tdtype, device=dev) for _ in range(2)]
        self.copy_stream = torch.cuda.Stream(device=dev)
        self.read_stream = torch.cuda.Stream(device=dev)     # cost read-back, never queued behind a batch copy
        self.copied = [torch.cuda.Event() for _ in range(2)]
        self.consumed = [torch.cuda.Event() for _ in range(2)]
        self.h2d_bytes = self.x_dev[0].numel() * self.x_dev[0].element_size() + \
            self.t_dev[0].numel() * self.t_dev[0].element_size()
        self.d2h_bytes = 8
        self._n = 0
        self._staged = -1

    def stage(self, x_pinned: torch.Tensor, t_pinned: torch.Tensor):
        """Enqueue the host->device copy of the next batch (returns immediately)."""
        slot = (self._staged + 1) % 2
        with torch.cuda.stream(self.copy_stream):
            if self._staged >= 1:
                self.copy_stream.wait_event(self.consumed[slot])
            self.x_dev[slot].copy_(x_pinned, non_blocking=True)
            self.t_dev[slot].copy_(t_pinned, non_blocking=True)
            self.copied[slot].record(self.copy_stream)
        self._staged += 1

    def step(self, epoch_index: int) -> None:
        """Launch the train step on the oldest staged batch."""
        slot = self._n % 2
        cur = torch.cuda.current_stream(self.eng.device)
        cur.wait_event(self.copied[slot])
        self.eng.train_step(self.x_dev[slot], self.t_dev[slot])
        self.consumed[slot].record(cur)
        self._n += 1

    def cost(self, epoch_index: int, step_number: int | None = None) -> float:
        """Device->host read of one step's cost.  With ``step_number`` (the 0-based count of
        ``step()`` calls that produced it) the read runs on its own stream behind that step's
        completion event only, so a later step already in flight is not waited for; without it the
        compute stream is synchronised."""
        if step_number is None:
            return float(self.eng.costs(epoch_index, 1)[0])
        self.read_stream.wait_event(self.consumed[step_number % 2])
        return float(self.eng.costs(epoch_index, 1, stream=self.read_stream)[0])
