"""cProfile of NeuralNetworkModel.train()'s epoch loop (2000 epochs, configs[1] stack): where the host's time per epoch goes."""
import os, sys, tempfile, cProfile, pstats
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
os.chdir(tempfile.mkdtemp())
import numpy as np, torch
from penr_oz_neural_network_torch_b200 import NeuralNetworkModel
rng = np.random.default_rng(0)
n = 2048
rows = rng.standard_normal((n, 784)).astype(np.float32)
data = [(rows[i].tolist(), [int(rng.integers(0, 10))]) for i in range(n)]
m = NeuralNetworkModel("rate", [784, 256, 128, 10], activation_algos=["relu", "relu", "softmax"])
m.training_buffer_size = n
m.train(data, epochs=200, batch_size=1024, dropout_rate=0.0)
m = NeuralNetworkModel("rate", [784, 256, 128, 10], activation_algos=["relu", "relu", "softmax"])
m.training_buffer_size = n
pr = cProfile.Profile()
pr.enable()
m.train(data, epochs=2000, batch_size=1024, dropout_rate=0.0)
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(18)
