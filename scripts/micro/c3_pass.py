"""python scripts/micro/c3_pass.py [rows] : a few large-batch inference passes of the 784-256-128-10 stack (for ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import bench

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
eng = bench.make_engine(0)
g = torch.Generator(device="cuda:0").manual_seed(1)
x = torch.randn((rows, 784), generator=g, device="cuda:0", dtype=torch.float32)
for _ in range(3):
    out = eng.forward(x)
torch.cuda.synchronize()
print(float(out.sum()))
eng.close()
