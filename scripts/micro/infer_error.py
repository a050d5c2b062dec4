"""Error of the inference paths against the float64 oracle on the 784-256-128-10 stack (one process per path).
usage: python scripts/micro/infer_error.py            -> runs itself once per path and prints one line each"""
import os, subprocess, sys
import numpy as np

PATHS = {
    "simt fp32 FFMA": {"POZ_TC3_MIN_BATCH": "1000000000"},
    "tc3 bf16 x 3 planes in tmem, six MMAs per product": {"POZ_TC3_F16": "0"},
    "tc3 fp16 x 2 planes in tmem, three MMAs per product (default)": {},
    "  the same, inputs scaled by 1e-3": {"INFER_ERROR_SCALE": "1e-3"},
    "  the same, inputs scaled by 1e+3": {"INFER_ERROR_SCALE": "1e3"},
}

def child():
    import torch
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
    from oracle import mlp_oracle as orc
    from oracle.detgen import det_params
    from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
    sizes = [784, 256, 128, 10]
    algos = ["linear", "relu", "linear", "relu", "linear", "softmax"]
    params = det_params(sizes, True, 5)
    eng = DeviceMLP(sizes, StackSpec.from_algos(algos), True, "float32", 0)
    for l in range(3):
        eng.upload_params(l, params[2 * l], params[2 * l + 1])
    g = torch.Generator(device="cuda:0").manual_seed(1)
    xd = torch.randn((32768, 784), generator=g, device="cuda:0", dtype=torch.float32) * float(os.environ.get("INFER_ERROR_SCALE", "1"))
    out = eng.forward(xd).cpu().numpy().astype(np.float64)
    ref = orc.forward(orc.parse_stack(algos), params, xd.cpu().numpy().astype(np.float64))[0][-1]
    rel = np.abs(out - ref) / ref
    print(f"max rel {rel.max():.2e}  mean rel {rel.mean():.2e}  max abs {np.abs(out - ref).max():.2e}  "
          f"argmax agree {(out.argmax(1) == ref.argmax(1)).mean():.5f}")

if __name__ == "__main__":
    if os.environ.get("INFER_ERROR_CHILD"):
        child()
    else:
        for name, env in PATHS.items():
            e = dict(os.environ, INFER_ERROR_CHILD="1", **env)
            r = subprocess.run([sys.executable, os.path.abspath(__file__)], env=e, capture_output=True, text=True, timeout=300)
            print(f"{name:60s} {r.stdout.strip() or r.stderr.strip()[-300:]}")
