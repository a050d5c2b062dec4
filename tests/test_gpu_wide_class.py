"""The wide bf16 / tcgen05 path and data-parallel training behind the unchanged class API:
`NeuralNetworkModel(..., dtype="bfloat16").train(list, ...)` -- sampling (row gather), dropout, the last epoch's
diagnostics and the statistics all run in the bf16 context -- against the float64 oracle driven with the same
sampled indices; and `train()` as a rank of a torch.distributed group (2 GPUs) against the single-GPU run.

Stated bf16 tolerance (derivation in tests/test_gpu_round2.py): costs rtol 2e-2, outputs rms error <= 2.5e-2 of
the reference's rms.
"""
import os
import socket

import numpy as np
import pytest
import torch

from oracle import mlp_oracle as orc
from oracle.detgen import det_normalish, det_uniform

pytestmark = pytest.mark.gpu

SIZES = [512, 768, 512, 256]
ACTS = ["relu", "tanh", "sigmoid"]
N_DATA, BATCH, EPOCHS = 2048, 1024, 3


def make_data():
    x = torch.tensor(det_normalish((N_DATA, SIZES[0]), 301)).to(torch.bfloat16).double().numpy()     # bf16-representable
    t = torch.tensor(det_uniform((N_DATA, SIZES[-1]), 302, 0.0, 1.0)).to(torch.bfloat16).double().numpy()
    return x, t, [(x[i].tolist(), t[i].tolist()) for i in range(N_DATA)]


def rms(v):
    return float(np.sqrt(np.mean(np.square(v))))


def test_bf16_model_trains_through_the_class_api(workdir):
    from penr_oz_neural_network_torch_b200 import NeuralNetworkModel
    x, t, data = make_data()
    torch.manual_seed(11)
    model = NeuralNetworkModel("wide", SIZES, "he", "random", ACTS, "adam", dtype="bfloat16")
    model.training_buffer_size = N_DATA
    params0 = [p.detach().numpy().copy() for p in model.params]
    torch.manual_seed(12)
    model.train(data, epochs=EPOCHS, learning_rate=1e-3, batch_size=BATCH, decay_rate=0.9, dropout_rate=0.0, l2_lambda=1e-4)
    assert model.status == "Trained" and len(model.progress) == EPOCHS
    # the oracle on the same sampled indices (torch.randint on the same seed, neural_net_model.py:460)
    torch.manual_seed(12)
    stack = orc.parse_stack(model.algos)
    adam = orc.AdamState()
    p, ref_costs = params0, []
    for e in range(EPOCHS):
        idx = torch.randint(0, N_DATA, (BATCH,)).numpy()
        p, c, *_ = orc.train_step(stack, p, x[idx], t[idx], lr=1e-3 * 0.9 ** e, optimizer="adam", adam=adam, l2_lambda=1e-4)
        ref_costs.append(c)
    np.testing.assert_allclose([q["cost"] for q in model.progress], ref_costs, rtol=2e-2)
    out, cost = model.compute_output(x[:256].tolist(), t[:256].tolist())
    ref = orc.forward(stack, p, x[:256])[0][-1]
    assert rms(np.asarray(out) - ref) <= 2.5e-2 * rms(ref)
    np.testing.assert_allclose(cost, np.mean((ref - t[:256]) ** 2), rtol=2e-2)
    # the last epoch's statistics came from the bf16 tensors on the device (neural_net_model.py:542-583)
    assert model.stats is not None and len(model.stats["layers"]) == len(model.layers)
    for entry, layer in zip(model.stats["layers"], model.layers):
        assert np.isfinite(entry["activation"]["mean"]) and np.isfinite(entry["activation"]["std"])
        assert entry["gradient"] is not None and np.isfinite(entry["gradient"]["std"])
        assert len(entry["activation"]["histogram"]["y"]) == 100
    relu_stats = model.stats["layers"][1]["activation"]
    assert 0.2 < relu_stats["saturated"] < 0.8                      # relu: share of outputs <= 0
    for w in model.stats["weights"]:
        if w is not None:
            assert np.isfinite(w["gradient"]["std"]) and w["gradient"]["std"] > 0
    assert all(np.isfinite(r) for r in model.progress[-1]["weight_upd_ratio"] if r is not None)
    # Adam's state went back into the torch optimizer object (.pth interop)
    st = model.optimizer.state_dict()["state"]
    assert int(st[0]["step"]) == EPOCHS


def test_bf16_model_trains_with_the_default_dropout(workdir):
    """dropout_rate 0.2 is train()'s default (neural_net_model.py:411): Philox masks in the tcgen05 epilogues."""
    from penr_oz_neural_network_torch_b200 import NeuralNetworkModel
    x, t, data = make_data()
    torch.manual_seed(21)
    model = NeuralNetworkModel("wide-drop", SIZES, "he", "zeros", ACTS, "adam", dtype="bfloat16")
    model.training_buffer_size = N_DATA
    model.train(data, epochs=4, learning_rate=1e-3, batch_size=BATCH)
    costs = [q["cost"] for q in model.progress]
    assert np.isfinite(costs).all() and model.status == "Trained"
    acts, cost = model._forward(torch.tensor(x[:512]), t[:512].tolist(), 0.5)
    zero = (acts[1] == 0).double().mean().item()                    # relu output: its own zeros, then half of the rest dropped
    plain, _ = model._forward(torch.tensor(x[:512]), t[:512].tolist(), 0.0)
    base = (plain[1] == 0).double().mean().item()
    assert abs(zero - (base + 0.5 * (1 - base))) < 0.05, (zero, base)


# ---- train() as a rank of a process group ------------------------------------------------------------------------
def _dp_worker(rank, world, port, out_dir, dtype):
    import torch.distributed as dist
    from penr_oz_neural_network_torch_b200 import NeuralNetworkModel
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    os.chdir(out_dir)
    os.makedirs(f"rank{rank}", exist_ok=True)
    os.chdir(f"rank{rank}")
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    x, t, data = make_data()
    torch.manual_seed(31)                                            # same model and same samples on every rank
    model = NeuralNetworkModel("dp", SIZES, "he", "random", ACTS, "adam", dtype=dtype, device=rank)
    model.training_buffer_size = N_DATA
    torch.manual_seed(32)
    model.train(data, epochs=EPOCHS, learning_rate=1e-3, batch_size=BATCH, dropout_rate=0.0, l2_lambda=1e-4)
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), costs=np.asarray([q["cost"] for q in model.progress]),
             **{f"p{i}": p.detach().numpy() for i, p in enumerate(model.params)})
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("dtype", ["float32", "bfloat16"])
def test_train_shards_over_the_process_group(dtype, tmp_path, monkeypatch):
    import torch.multiprocessing as mp
    from penr_oz_neural_network_torch_b200 import NeuralNetworkModel
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_dp_worker, args=(2, port, str(tmp_path), dtype), nprocs=2, join=True)
    r0, r1 = np.load(tmp_path / "r0.npz"), np.load(tmp_path / "r1.npz")
    monkeypatch.chdir(tmp_path)
    x, t, data = make_data()
    torch.manual_seed(31)
    model = NeuralNetworkModel("single", SIZES, "he", "random", ACTS, "adam", dtype=dtype)
    model.training_buffer_size = N_DATA
    torch.manual_seed(32)
    model.train(data, epochs=EPOCHS, learning_rate=1e-3, batch_size=BATCH, dropout_rate=0.0, l2_lambda=1e-4)
    single = np.asarray([q["cost"] for q in model.progress])
    np.testing.assert_allclose(r0["costs"], r1["costs"], rtol=1e-12)                 # the reduced cost is the same number
    np.testing.assert_allclose(r0["costs"], single, rtol=1e-4 if dtype == "float32" else 2e-3)
    for i, p in enumerate(model.params):
        np.testing.assert_array_equal(r0[f"p{i}"], r1[f"p{i}"])                      # replicas identical
        ref = p.detach().numpy()
        upd = rms(ref - r0[f"p{i}"])
        assert upd <= 0.1 * max(rms(ref), 1e-12) + 1e-6
