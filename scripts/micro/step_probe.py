"""Device time of the configs[1] train step (784-256-128-10, batch 1024, Adam + L2) on four L2-resident batches,
for A/B runs of the library's environment switches on one box:  POZ_x=... python scripts/micro/step_probe.py
With POZ_KERNEL_TRACE=1 the per-CTA kernel trace is printed when the engine is closed."""
import os, sys, time, torch, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
from penr_oz_neural_network_torch_b200 import _native
algos = ["linear", "relu", "linear", "relu", "linear", "softmax"]
sizes = [784, 256, 128, 10]
eng = DeviceMLP(sizes, StackSpec.from_algos(algos), True, "float32", 0)
rng = np.random.default_rng(0)
for l in range(3):
    eng.upload_params(l, rng.standard_normal((sizes[l], sizes[l+1])) * 0.05, np.zeros(sizes[l+1]))
B = 1024
xs = [torch.randn(B, 784, device="cuda:0") for _ in range(4)]
ys = [torch.randint(0, 10, (B,), device="cuda:0", dtype=torch.int32) for _ in range(4)]
eng.train_begin(optimizer="adam", learning_rate=0.001, l2_lambda=0.001, epochs=100000)
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    for i in range(40):
        eng.train_step(xs[i % 4], ys[i % 4])
    torch.cuda.synchronize()
    lib = _native.lib()
    print("last error:", lib.poz_last_error())
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(side)
    for i in range(400):
        eng.train_step(xs[i % 4], ys[i % 4])
    e1.record(side)
    torch.cuda.synchronize()
    print("us/step", e0.elapsed_time(e1) * 1000 / 400, "launches", lib.poz_launch_count())
eng.close()
