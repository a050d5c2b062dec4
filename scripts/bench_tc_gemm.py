"""Times the three tcgen05 GEMM variants at the C4 layer shape (batch 65536, 4096x4096)."""
import json
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from penr_oz_neural_network_torch_b200 import _native as nat

lib = nat.lib()
M, N, K = int(os.environ.get("M", 65536)), 4096, 4096
dev = "cuda:0"
x = torch.randn((M, K), device=dev).to(torch.bfloat16)
w = (torch.randn((K, N), device=dev) * 0.02).to(torch.bfloat16)
dz = torch.randn((M, N), device=dev).to(torch.bfloat16)
y = torch.empty((M, N), device=dev, dtype=torch.bfloat16)
g = torch.empty((4, K, N), device=dev, dtype=torch.float32)
bias = torch.zeros(N, device=dev)
s = int(torch.cuda.current_stream().cuda_stream)


def run(name, fn, flops, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    print(json.dumps({"kernel": name, "ms": round(ms, 4), "tflops": round(flops / ms / 1e9, 1)}), flush=True)


fl = 2.0 * M * N * K
run("fwd  (A k-major, B n-major)", lambda: nat.check(lib.poz_tc_gemm(0, 0, x.data_ptr(), w.data_ptr(), M, N, K, y.data_ptr(), bias.data_ptr(), None, 1, 1, s)), fl)
run("bwd_x (A k-major, B k-major)", lambda: nat.check(lib.poz_tc_gemm(0, 1, dz.data_ptr(), w.data_ptr(), M, K, N, y.data_ptr(), None, x.data_ptr(), 1, 1, s)), fl)
for sp in (1, 2, 4):
    run(f"bwd_w splits={sp} (A m-major, B n-major)", lambda: nat.check(lib.poz_tc_gemm(0, 2, x.data_ptr(), dz.data_ptr(), K, N, M, g.data_ptr(), None, None, 0, sp, s)), fl)
run("torch.matmul bf16 (cuBLAS) same shape", lambda: torch.matmul(x, w, out=y), fl)
