"""python scripts/micro/c3_stress.py ROWS CALLS : repeated large-batch inference passes on two inputs; every result must be
bit-identical to the first result for the same input (the kernels are deterministic), no sync between the calls."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import bench

rows = int(sys.argv[1]); calls = int(sys.argv[2])
eng = bench.make_engine(0)
dev = torch.device("cuda", 0)
side = torch.cuda.Stream(device=dev)
torch.cuda.synchronize()
torch.cuda.set_stream(side)
pool = torch.randn((2, rows, 784), device=dev, dtype=torch.float32)
outs = [torch.empty((rows, 10), device=dev) for _ in range(2)]
ref = [eng.forward(pool[i]).clone() for i in range(2)]
torch.cuda.synchronize()
bad = 0
for i in range(calls):
    eng.forward(pool[i % 2], outs[i % 2])
    if i % 10 >= 8:
        if not torch.equal(outs[i % 2], ref[i % 2]):
            bad += 1
            print("call", i, "differs:", int((outs[i % 2] != ref[i % 2]).sum()), "values", flush=True)
torch.cuda.synchronize()
print("calls", calls, "bad", bad, flush=True)
eng.close()
