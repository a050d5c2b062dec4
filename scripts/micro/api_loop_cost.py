"""Wall time of NeuralNetworkModel.train()'s epoch loop (train_loop_seconds) against the number of epochs, configs[1] stack."""
import os, sys, time, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
os.chdir(tempfile.mkdtemp())
import numpy as np, torch
from penr_oz_neural_network_torch_b200 import NeuralNetworkModel
rng = np.random.default_rng(0)
n = 2048
rows = rng.standard_normal((n, 784)).astype(np.float32)
data = [(rows[i].tolist(), [int(rng.integers(0, 10))]) for i in range(n)]
for epochs in (200, 1, 10, 100, 200, 400, 1000, 2000, 200, 2000):
    m = NeuralNetworkModel("rate", [784, 256, 128, 10], activation_algos=["relu", "relu", "softmax"])
    m.training_buffer_size = n
    t0 = time.perf_counter()
    m.train(data, epochs=epochs, batch_size=1024, dropout_rate=0.0)
    t1 = time.perf_counter()
    print(f"epochs {epochs:5d}: loop {m.train_loop_seconds * 1e3:8.2f} ms  ({m.train_loop_seconds / epochs * 1e6:8.1f} us/epoch), whole call {1e3 * (t1 - t0):8.1f} ms", flush=True)
