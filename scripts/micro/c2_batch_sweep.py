"""Device time per train step of the configs[1] stack (784-256-128-10, Adam, L2) against the batch size."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import bench

dev = torch.device("cuda", 0)
side = torch.cuda.Stream(device=dev)
torch.cuda.synchronize()
torch.cuda.set_stream(side)
for batch in [int(b) for b in os.environ.get("SWEEP_BATCHES", "64,128,256,512,1024,1536,2048,2560,4096,8192,16384").split(",")]:
    eng = bench.make_engine(0)
    n_pool = max(2, min(64, int(210e6 / (batch * 784 * 4))))
    xs, ys = bench.synth_pool(n_pool, batch, dev, 7)
    steps = 300
    eng.train_begin(optimizer="adam", epochs=2 * n_pool + steps + 40, **bench.HP)
    for i in range(2 * n_pool + 20):
        eng.train_step(xs[i % n_pool], ys[i % n_pool])
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        eng.train_step(xs[i % n_pool], ys[i % n_pool])
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / steps * 1e3
    c = np.asarray(eng.costs(2 * n_pool + 20, steps))
    assert np.isfinite(c).all()
    print(f"batch {batch:6d}: {us:8.1f} us/step  {batch / us:8.2f} M samples/s", flush=True)
    eng.close()
    del xs, ys
    torch.cuda.empty_cache()
