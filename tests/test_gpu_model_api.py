"""The drop-in class through the reference's own list-level API, on the GPU."""
import numpy as np
import pytest
import torch

from conftest import golden_names, load_golden
from oracle.detgen import det_labels, det_params, det_uniform

pytestmark = pytest.mark.gpu


def make_model(meta, dtype):
    from penr_oz_neural_network_torch_b200 import NeuralNetworkModel
    m = NeuralNetworkModel("golden", meta["sizes"], "xavier", meta["bias"], meta["algos"], meta["optimizer"],
                           dtype=dtype)
    params = det_params(meta["sizes"], bool(meta["bias"]), meta["salt"])
    it = iter(params)
    for layer in m.layers:
        if layer.algo == "linear":
            layer.state_dict = {"params": [next(it).tolist()] + ([next(it).tolist()] if meta["bias"] else [])}
    if meta["optimizer"] == "adam":
        m.optimizer = torch.optim.Adam(m.params)
    return m


@pytest.mark.parametrize("dtype", ["float32", "float64"])
@pytest.mark.parametrize("name", golden_names("train_"))
def test_train_list_api_matches_reference_run(name, dtype, workdir):
    """Same seed, same data, dropout 0: the whole train() call (sampling with torch.randint, lr
    decay, L2, optimizer, progress, statistics) reproduces the reference's outputs."""
    arrays, meta = load_golden(name)
    m = make_model(meta, dtype)
    n = meta["n_data"]
    xs = det_uniform((n, meta["sizes"][0]), meta["salt"] + 3).tolist()
    if m.algos[-1] == "softmax":
        ts = [[int(v)] for v in det_labels(n, meta["sizes"][-1], meta["salt"] + 7)]
    else:
        ts = det_uniform((n, meta["sizes"][-1]), meta["salt"] + 9, 0.0, 1.0).tolist()
    torch.manual_seed(meta["seed"])
    m.train(list(zip(xs, ts)), epochs=meta["epochs"], learning_rate=meta["lr"], batch_size=meta["batch"],
            decay_rate=meta["decay"], dropout_rate=0.0, l2_lambda=meta["l2"])
    f32 = dtype == "float32"
    rtol = 2e-4 if f32 else 1e-9          # errors compound over up to 120 dependent steps
    assert m.status == meta["status"] == "Trained"
    np.testing.assert_array_equal([p["epoch"] for p in m.progress], arrays["progress_epoch"])
    np.testing.assert_allclose([p["cost"] for p in m.progress], arrays["progress_cost"], rtol=rtol)
    wur = [[v for v in p["weight_upd_ratio"] if v is not None] for p in m.progress]
    np.testing.assert_allclose(wur, arrays["progress_wur"], rtol=5e-3 if f32 else 1e-7)
    np.testing.assert_allclose(m.avg_cost, arrays["avg_cost"], rtol=rtol)
    for i, p in enumerate(m.params):
        a = arrays[f"param{i}"]
        np.testing.assert_allclose(p.numpy(), a, rtol=rtol, atol=(2e-4 if f32 else 1e-10) * np.abs(a).max())
    st = m.stats
    tol = dict(rtol=1e-3, atol=1e-6) if f32 else dict(rtol=1e-7, atol=1e-12)
    np.testing.assert_allclose([l["activation"]["mean"] for l in st["layers"]], arrays["stats_act_mean"], **tol)
    np.testing.assert_allclose([l["activation"]["std"] for l in st["layers"]], arrays["stats_act_std"], **tol)
    np.testing.assert_allclose([l["activation"]["saturated"] for l in st["layers"]], arrays["stats_act_sat"],
                               atol=0.02)
    got_gm = [np.nan if l["gradient"] is None else l["gradient"]["mean"] for l in st["layers"]]
    assert [np.isnan(v) for v in got_gm] == [bool(np.isnan(v)) for v in arrays["stats_grad_mean"]]
    np.testing.assert_allclose(np.nan_to_num(got_gm), np.nan_to_num(arrays["stats_grad_mean"]), **tol)
    np.testing.assert_allclose([w["gradient"]["std"] for w in st["weights"] if w], arrays["stats_wgrad_std"], **tol)
    assert len(st["layers"][0]["activation"]["histogram"]["y"]) == 100
    if m.optimizer is not None:
        sd = m.optimizer.state_dict()["state"]
        assert float(sd[0]["step"]) == float(arrays["adam_step"])
        np.testing.assert_allclose(sd[0]["exp_avg"].numpy(), arrays["adam_m0"], rtol=rtol,
                                   atol=(1e-4 if f32 else 1e-10) * np.abs(arrays["adam_m0"]).max())


TRAIN_TABLE = [   # the Linear+activation rows of test_neural_net_model.py:161-179
    ([9, 9, 9], ["sigmoid"] * 2, None, 1.0, [0.5] * 9, [1.0] + [0.0] * 8),
    ([9, 9, 9], ["relu", "softmax"], "adam", 1.0, [0.5] * 9, [1]),
    ([9, 9, 9], ["tanh"] * 2, None, 1.0, [0.5] * 9, [1.0] + [0.0] * 8),
    ([18, 9, 3], ["relu", "sigmoid"], "adam", 0.9, [0.5] * 18, [1.0] + [0.0] * 2),
    ([9, 18, 9], ["sigmoid", "linear", "softmax"], None, 0.7, [0.5] * 9, [1]),
    ([4, 8, 16], ["sigmoid"] * 2, None, 1.0, [0.5] * 4, [1.0] + [0.0] * 15),
    ([3, 3, 3, 3], ["relu", "relu", "softmax"], "adam", 0.5, [0.5] * 3, [0]),
    ([18, 9, 3], ["relu"] * 2, None, 1.0, [0.5] * 18, [1.0] + [0.0] * 2),
    ([9, 18, 9], ["relu", "tanh"], "adam", 1.0, [0.5] * 9, [1.0] + [0.0] * 8),
]


@pytest.mark.parametrize("sizes,algos,optimizer,confidence,sample_input,target", TRAIN_TABLE)
def test_reference_test_train_flow(sizes, algos, optimizer, confidence, sample_input, target, workdir):
    """Re-statement of the reference's own test_train (test_neural_net_model.py:187-236)."""
    from penr_oz_neural_network_torch_b200 import NeuralNetworkModel
    NeuralNetworkModel.delete("test")
    model = NeuralNetworkModel("test", sizes, activation_algos=algos, optimizer_algo=optimizer,
                               confidence=confidence)
    initial = [p.tolist() for p in model.params]
    _, initial_cost = model.compute_output(sample_input, target)
    data = [(sample_input, target)] * model.training_buffer_size
    model.avg_cost_history = [1.0] * 100
    model.train(data, epochs=1, dropout_rate=0.001)
    updated = [p.tolist() for p in model.params]
    for a, e in zip(updated, initial):
        np.testing.assert_array_almost_equal(a, e, 1)
    assert len(model.progress) > 0
    assert model.progress[0]["cost"] != initial_cost
    assert sum(p["cost"] for p in model.progress) / len(model.progress) == model.avg_cost
    assert len(model.avg_cost_history) == 100
    assert model.avg_cost_history[0] == 1.0 and model.avg_cost_history[-1] == model.avg_cost
    assert len(model.training_data_buffer) == 0
    assert model.stats is not None and model.status == "Trained"
    again = NeuralNetworkModel.deserialize(model.model_id)
    for a, e in zip([p.tolist() for p in again.params], updated):
        np.testing.assert_array_almost_equal(a, e, 8)
    assert len(again.progress) == len(model.progress)
    assert again.avg_cost == model.avg_cost and again.avg_cost_history == model.avg_cost_history
    assert again.stats == model.stats and again.status == model.status
    if optimizer == "adam":
        assert again.optimizer is not None
        # a second train() call resumes from the persisted Adam state
        again.train(data, epochs=1, dropout_rate=0.0)
        assert float(again.optimizer.state_dict()["state"][0]["step"]) == 2.0


@pytest.mark.parametrize("sizes,algos,sample_input,target", [
    ([9, 9, 9], ["sigmoid"] * 2, [0.5] * 9, None),
    ([9, 9], ["softmax"], [1] + [0] * 8, [4]),
    ([18, 9, 3], ["relu", "softmax"], [1] + [0] * 17, None),
    ([9, 18, 9], ["tanh"] * 2, [0.5] * 9, [0.5] * 9),
    ([4, 8, 16], ["tanh", "softmax"], [0.5] * 4, [13]),
    ([3, 3, 3, 3], ["relu", "relu", "softmax"], [0.5] * 3, None),
    ([4, 8, 16], ["tanh", "softmax"], [[0.5] * 4, [0.1] * 4], [[13], [2]]),
])
def test_reference_test_compute_output_flow(sizes, algos, sample_input, target):
    """test_neural_net_model.py:133-159."""
    from penr_oz_neural_network_torch_b200 import NeuralNetworkModel
    model = NeuralNetworkModel("test", sizes, activation_algos=algos)
    output, cost = model.compute_output(sample_input, target)
    if isinstance(sample_input[0], list):
        assert len(output) == len(sample_input) and len(output[0]) == sizes[-1]
    else:
        assert len(output) == sizes[-1]
    assert (cost is None) == (target is None)


def test_params_edited_on_host_are_picked_up():
    """Callers mutate parameters in place (neural_net_model.py:283) or replace them through
    state_dict (:39-45); the device image must follow."""
    from penr_oz_neural_network_torch_b200 import NeuralNetworkModel
    model = NeuralNetworkModel("test", [4, 3], activation_algos=["sigmoid"])
    x = [0.1, 0.2, 0.3, 0.4]
    a, _ = model.compute_output(x)
    with torch.no_grad():
        model.weights[0] *= 0.0
    b, _ = model.compute_output(x)
    np.testing.assert_allclose(b, [0.5] * 3, atol=1e-6)
    assert not np.allclose(a, b)
    model.layers[0].state_dict = {"params": [[[1.0] * 3] * 4, [0.0] * 3]}
    c, _ = model.compute_output(x)
    np.testing.assert_allclose(c, [1 / (1 + np.exp(-1.0))] * 3, rtol=1e-5)


NET_TRAIN_TABLE = [   # the remaining rows of test_neural_net_model.py:180-185 and the model of test_torch_backward.py
    ([9, 2, 6, 18, 9], ["embedding", "tanh", "softmax"], "adam", 1.0, [0, 5, 7], [5]),
    ([9, 2, 6, 18, 9], ["embedding", "linear", "batchnorm", "tanh", "softmax"], None, 1.0, [0, 5, 7], [5]),
    ([9, 2, 4, 9, 18, 9], ["embedding", "linear", "batchnorm", "tanh", "flatten", "linear", "softmax"], None, 0.1,
     [0, 5, 7, 8], [5]),
    ([27, 10, 30, 64, 27], ["embedding", "linear", "batchnorm", "tanh", "linear", "softmax"], "adam", 1.0, [1, 2, 3], [4]),
]


@pytest.mark.parametrize("sizes,algos,optimizer,confidence,sample_input,target", NET_TRAIN_TABLE)
def test_reference_test_train_flow_with_embedding_and_batchnorm(sizes, algos, optimizer, confidence, sample_input,
                                                                target, workdir):
    test_reference_test_train_flow(sizes, algos, optimizer, confidence, sample_input, target, workdir)


@pytest.mark.parametrize("sizes,algos,sample_input,target", [
    ([9, 2, 6, 18, 9], ["embedding", "tanh", "softmax"], [0, 5, 8], [2]),
    ([9, 2, 6, 18, 9], ["embedding", "tanh", "softmax"], [[0, 5, 8], [1, 3, 7]], [[2], [4]]),
    ([9, 2, 4, 9, 18, 9], ["embedding", "tanh", "flatten", "linear", "softmax"], [[0, 5, 8, 2], [1, 3, 7, 4]], [[2], [4]]),
])
def test_reference_test_compute_output_flow_with_embedding(sizes, algos, sample_input, target):
    test_reference_test_compute_output_flow(sizes, algos, sample_input, target)
