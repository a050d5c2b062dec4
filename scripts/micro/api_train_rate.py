"""Per-epoch time of the drop-in class API: NeuralNetworkModel.train() on the configs[1] stack, batch 1024.
The one-off ingestion of the Python lists is excluded by differencing two epoch counts."""
import os, sys, time, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
os.chdir(tempfile.mkdtemp())
import numpy as np, torch
from penr_oz_neural_network_torch_b200 import NeuralNetworkModel
rng = np.random.default_rng(0)
n = 4096
data = [(rng.standard_normal(784).astype(np.float32).tolist(), [int(rng.integers(0, 10))]) for _ in range(n)]
times = {}
for epochs in (50, 200, 2200):
    m = NeuralNetworkModel("rate", [784, 256, 128, 10], activation_algos=["relu", "relu", "softmax"])
    m.training_buffer_size = n if hasattr(m, "training_buffer_size") else None
    t0 = time.perf_counter()
    m.train(data, epochs=epochs, batch_size=1024, dropout_rate=0.0)
    times[epochs] = time.perf_counter() - t0
    print(epochs, "epochs:", round(times[epochs], 3), "s, last cost", m.progress[-1]["cost"])
print("per epoch: %.1f us" % ((times[2200] - times[200]) / 2000 * 1e6))
