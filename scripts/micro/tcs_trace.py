"""POZ_KERNEL_TRACE=1 python scripts/micro/tcs_trace.py : per-CTA time stamps of the C2 step's kernels (stderr)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import bench

eng = bench.make_engine(0)
xs, ys = bench.synth_pool(8, 1024, torch.device("cuda", 0), 1)
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    eng.train_begin(optimizer="adam", epochs=64, **bench.HP)
    for i in range(24):
        eng.train_step(xs[i % 8], ys[i % 8])
side.synchronize()
eng.close()
