"""cProfile of one NeuralNetworkModel.train() call on the C1 row of the reference's tests ([9, 9, 9], 180 samples, 1 epoch)."""
import os, sys, tempfile, cProfile, pstats, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
os.chdir(tempfile.mkdtemp())
import torch
from penr_oz_neural_network_torch_b200 import NeuralNetworkModel
def one():
    m = NeuralNetworkModel("c1", [9, 9, 9], activation_algos=["relu", "softmax"])
    d1 = [([0.5] * 9, [1])] * m.num_params
    t0 = time.perf_counter()
    m.train(list(d1), epochs=1, dropout_rate=0.0)
    return time.perf_counter() - t0
for _ in range(3):
    print("call: %.2f ms" % (one() * 1e3))
pr = cProfile.Profile(); pr.enable(); one(); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
