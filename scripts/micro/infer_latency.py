"""Latency of one small inference call (launch -> result visible to the host) and device time per call,
for the few-row cluster kernel (default) against the per-layer kernels (POZ_SMALL_INFER=0)."""
import os, sys, time, torch, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
algos = ["linear", "relu", "linear", "relu", "linear", "softmax"]
sizes = [784, 256, 128, 10]
eng = DeviceMLP(sizes, StackSpec.from_algos(algos), True, "float32", 0)
rng = np.random.default_rng(0)
for l in range(3):
    eng.upload_params(l, rng.standard_normal((sizes[l], sizes[l + 1])) * 0.05, np.zeros(sizes[l + 1]))
side = torch.cuda.Stream()
for B in (1, 4, 8):
    x = torch.randn(B, 784, device="cuda:0")
    out = torch.empty(B, 10, device="cuda:0")
    with torch.cuda.stream(side):
        for _ in range(20):
            eng.forward(x, out)
        side.synchronize()
        t0 = time.perf_counter()
        for _ in range(300):
            eng.forward(x, out)
            side.synchronize()
        lat = (time.perf_counter() - t0) / 300 * 1e6
        eng.profile_begin()
        for _ in range(100):
            eng.forward(x, out)
        prof = eng.profile_end()
    dev = sum(v["ms"] for v in prof.values()) / 100 * 1e3
    print(f"POZ_SMALL_INFER={os.environ.get('POZ_SMALL_INFER', '1')} batch {B}: call+sync {lat:.1f} us, "
          f"device (events around eager launches) {dev:.1f} us over {sum(v['launches'] for v in prof.values()) // 100} launches")
eng.close()
