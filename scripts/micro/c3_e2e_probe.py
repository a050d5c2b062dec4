"""Probe of bench.py's inference loops: python c3_e2e_probe.py ROWS MODE(h2d|dev|pool) ITERS SYNC_EVERY [side]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import bench

batch = int(sys.argv[1]); mode = sys.argv[2]; iters = int(sys.argv[3]); every = int(sys.argv[4])
eng = bench.make_engine(0)
dev = torch.device("cuda", 0)
if len(sys.argv) > 5:
    side = torch.cuda.Stream(device=dev)
    torch.cuda.synchronize()
    torch.cuda.set_stream(side)
out = torch.empty((batch, 10), device=dev, dtype=torch.float32)
hx = torch.randn((batch, 784), dtype=torch.float32).pin_memory()
ho = torch.empty((batch, 10), dtype=torch.float32).pin_memory()
xd = torch.empty((batch, 784), device=dev, dtype=torch.float32)
pool = torch.randn((2, batch, 784), device=dev, dtype=torch.float32) if mode == "pool" else None
for i in range(iters):
    if mode == "h2d":
        xd.copy_(hx, non_blocking=True)
        eng.forward(xd, out)
        ho.copy_(out, non_blocking=True)
    elif mode == "pool":
        eng.forward(pool[i % 2], out)
    else:
        xd.normal_()
        eng.forward(xd, out)
    if (i + 1) % every == 0:
        torch.cuda.synchronize()
        print(i, "ok", float(out.sum()), flush=True)
torch.cuda.synchronize()
print("done", flush=True)
eng.close()
