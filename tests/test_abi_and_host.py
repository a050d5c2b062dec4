"""CPU-side checks: the C-ABI library loads and exports every symbol include/pozmlp.h declares
(no compute without a GPU), and the host mirror of the reference class behaves like the reference
where no device is needed (topology, initialisation on a seed, buffering, errors)."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT, golden_names, load_golden
from oracle import mlp_oracle as orc
from penr_oz_neural_network_torch_b200 import _native as nat
from penr_oz_neural_network_torch_b200 import NeuralNetworkModel, StackSpec
from penr_oz_neural_network_torch_b200.neural_net_model import expand_algos


@pytest.fixture(scope="module")
def built():
    return nat.build()


def header_functions():
    text = open(os.path.join(ROOT, "include", "pozmlp.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(poz_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_what_python_binds():
    assert header_functions() == sorted(nat.SIGNATURES)


def test_library_exports_every_declared_symbol(built):
    handle = ctypes.CDLL(built)
    for name in header_functions():
        assert hasattr(handle, name), name
    lib = nat.lib()
    assert lib.poz_version() >= 100
    assert lib.poz_launch_count() >= 0


def test_create_fails_cleanly_without_device(built):
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = nat.lib()
    handle = ctypes.c_void_p()
    sizes = (ctypes.c_int * 2)(3, 3)
    acts = (ctypes.c_int * 1)(1)
    rc = lib.poz_ctx_create(0, 1, sizes, acts, None, None, 1, 0, 0, ctypes.byref(handle))
    assert rc != 0 and not handle.value
    assert lib.poz_last_error()          # a message, not an abort


def test_no_cpu_fallback_for_compute(built):
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    model = NeuralNetworkModel("t", [3, 3], activation_algos=["relu"])
    with pytest.raises(RuntimeError):
        model.compute_output([0.1, 0.2, 0.3])


@pytest.mark.parametrize("name", golden_names("step_") + golden_names("init_"))
def test_expansion_and_dropout_points_match_reference(name):
    _, meta = load_golden(name)
    algos = meta["algos"] or ["relu"] * (len(meta["sizes"]) - 1)
    assert expand_algos(algos) == meta["expanded_algos"]
    if any(a in ("embedding", "flatten", "batchnorm") for a in meta["expanded_algos"]):
        return      # the unit structure of those stacks is exercised against the reference on the GPU
    spec = StackSpec.from_algos(meta["expanded_algos"])
    ref = orc.parse_stack(meta["expanded_algos"])
    assert spec.acts == [u.act for u in ref.linears]
    assert spec.drop_z == [u.drop_z for u in ref.linears]
    assert spec.drop_a == [u.drop_a for u in ref.linears]
    assert spec.loss == ref.loss


@pytest.mark.parametrize("name", golden_names("init_"))
def test_construction_on_a_seed_matches_reference(name):
    arrays, meta = load_golden(name)
    torch.manual_seed(meta["seed"])
    model = NeuralNetworkModel("g", meta["sizes"], meta["weight_algo"], meta["bias"], meta["algos"],
                               meta["optimizer"], confidence=meta["confidence"])
    assert model.algos == meta["expanded_algos"]
    assert [bool(l.hidden) for l in model.layers] == meta["hidden"]
    assert model.num_params == meta["num_params"] == model.training_buffer_size
    assert (model.optimizer is not None) == (meta["optimizer"] == "adam")
    for i, p in enumerate(model.params):
        np.testing.assert_array_equal(p.numpy(), arrays[f"param{i}"])
    assert model.progress == [] and model.avg_cost is None and model.stats is None
    assert model.status == "Created"


@pytest.mark.parametrize("sizes,algos,expected", [
    ([3, 3], None, 12), ([9, 9, 9], None, 180), ([18, 9, 3], ["sigmoid"] * 2, 201),
])
def test_num_params_like_reference_table(sizes, algos, expected):
    # test_neural_net_model.py:112-131 (Linear+activation rows)
    model = NeuralNetworkModel("test", sizes, activation_algos=algos)
    assert model.training_buffer_size == expected == model.num_params


def test_unknown_algo_is_value_error():
    with pytest.raises(ValueError) as e:
        NeuralNetworkModel(model_id="test", layer_sizes=[9, 9, 9], activation_algos=["relu", "unknown_algo"])
    assert str(e.value) == "Unsupported activation algorithm: unknown_algo"


@pytest.mark.parametrize("algos", [["relu", "batchnorm", "tanh"], ["softmax", "relu"], ["tanh", "embedding", "relu"]])
def test_off_path_entries_are_refused_not_emulated(algos):
    """Entry orders the device path has no kernels for (a batchnorm that does not follow a Linear, a
    softmax or an embedding in the middle) are refused at construction -- never run on the CPU."""
    with pytest.raises(ValueError):
        NeuralNetworkModel("test", [9, 2, 6, 9], activation_algos=algos)


@pytest.mark.parametrize("sizes,algos,expected", [
    ([10, 3, 6, 20, 10], ["embedding", "tanh", "softmax"], 380),
    ([10, 3, 6, 20, 10], ["embedding", "linear", "batchnorm", "tanh", "softmax"], 420),
])
def test_embedding_and_batchnorm_stacks_are_built_like_the_reference(sizes, algos, expected):
    # test_neural_net_model.py:116-117
    model = NeuralNetworkModel("test", sizes, activation_algos=algos)
    assert model.num_params == expected == model.training_buffer_size


def test_insufficient_data_is_buffered_and_persisted(workdir):
    # test_neural_net_model.py:238-260 -- needs no device: the gate returns before any compute
    model = NeuralNetworkModel(model_id="test", layer_sizes=[9, 9, 9], activation_algos=["relu"] * 2)
    data = [([0.5] * 9, [1.0] * 9)] * (model.training_buffer_size - 1)
    model.train(training_data=data, epochs=1)
    assert len(model.progress) == 0
    assert len(model.training_data_buffer) >= len(data)
    again = NeuralNetworkModel.deserialize("test")
    assert len(again.training_data_buffer) == len(model.training_data_buffer)
    assert again.algos == model.algos
    for a, b in zip(again.params, model.params):
        np.testing.assert_array_almost_equal(a.tolist(), b.tolist(), 8)
    assert again.optimizer is not None


def test_missing_model_is_key_error_and_delete_is_silent(workdir):
    with pytest.raises(KeyError):
        NeuralNetworkModel.deserialize("nonexistent_model")
    NeuralNetworkModel.delete("nonexistent")
    model = NeuralNetworkModel("test", [3, 3])
    model.serialize()
    NeuralNetworkModel.delete("test")
    with pytest.raises(KeyError):
        NeuralNetworkModel.deserialize("test")


def test_randint_hands_out_the_same_indices_in_one_call_or_many():
    """NeuralNetworkModel.train() draws the row indices of up to 16 epochs in one torch.randint call (the reference draws
    them epoch by epoch from the same global generator, neural_net_model.py:460): the numbers, and the generator's state
    afterwards, must not depend on how the draws are grouped."""
    import torch
    for batch, epochs, n in [(1024, 16, 2048), (5, 7, 9), (201, 3, 201), (180, 16, 1 << 20)]:
        torch.manual_seed(1234)
        one_by_one = torch.cat([torch.randint(0, n, (batch,)) for _ in range(epochs)])
        after_a = torch.randint(0, 1 << 30, (4,))
        torch.manual_seed(1234)
        at_once = torch.randint(0, n, (epochs * batch,))
        after_b = torch.randint(0, 1 << 30, (4,))
        assert torch.equal(one_by_one, at_once)
        assert torch.equal(after_a, after_b)


def test_fast_list_conversion_equals_torch_tensor():
    """compute_output() / train() convert their Python lists with numpy.fromiter where the shape allows; the result must be
    what torch.tensor(data, dtype=float64) gives (neural_net_model.py:379,472), and anything irregular must reach torch.tensor."""
    import numpy as np
    import torch
    from penr_oz_neural_network_torch_b200.neural_net_model import _list_to_f64
    rng = np.random.default_rng(3)
    rows = rng.standard_normal((37, 11)).tolist()
    assert torch.equal(_list_to_f64(rows), torch.tensor(rows, dtype=torch.float64))
    ints = [[1, 2, 3], [4, 5, 6]]                                   # token indices, integers, booleans
    assert torch.equal(_list_to_f64(ints), torch.tensor(ints, dtype=torch.float64))
    assert torch.equal(_list_to_f64([[True, False]]), torch.tensor([[True, False]], dtype=torch.float64))
    vec = [0.5, 1.5, -2.0]                                          # a single vector: torch.tensor's own path
    assert torch.equal(_list_to_f64(vec), torch.tensor(vec, dtype=torch.float64))
    deep = [[[1, 2], [3, 4]], [[5, 6], [7, 8]]]
    assert torch.equal(_list_to_f64(deep), torch.tensor(deep, dtype=torch.float64))
    for bad in ([[1.0, 2.0], [3.0]], [[1.0, "x"]], [[1.0, None]]):  # ragged rows, non-numbers: torch.tensor's exception
        with pytest.raises((ValueError, TypeError)):
            torch.tensor(bad, dtype=torch.float64)
        with pytest.raises((ValueError, TypeError)):
            _list_to_f64(bad)
    assert _list_to_f64([]).shape == torch.tensor([], dtype=torch.float64).shape
