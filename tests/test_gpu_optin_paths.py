"""The variants of the narrow float32 train step and of large-batch inference that an environment switch selects
(read once per process / per context).  Each must give the oracle's numbers like the default path: three steps of
BASELINE.json configs[1] (784-256-128-10, softmax + cross-entropy, batch 1024, Adam + L2) are run in a
child process per switch (same seeded inputs as test_config2_full_size_against_live_oracle) and compared
there with oracle/mlp_oracle.py at the float32 tolerance of
tests/test_gpu_parity.py (rtol 1e-5, atol 1e-5 * max|reference|, relu-flip outliers allowed)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r'''
import sys
sys.path.insert(0, sys.argv[1])
sys.path.insert(0, sys.argv[1] + "/tests")
import numpy as np, torch
from oracle import mlp_oracle as orc
from oracle.detgen import det_params, det_normalish, det_labels
from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec
from test_gpu_parity import close, ADAM_PARAM_ATOL

sizes = [784, 256, 128, 10]
algos = ["linear", "relu", "linear", "relu", "linear", "softmax"]
B = 1024
params = det_params(sizes, True, 77)
x = det_normalish((B, 784), 78)
y = det_labels(B, 10, 79)
eng = DeviceMLP(sizes, StackSpec.from_algos(algos), True, "float32", 0)
for l in range(3):
    eng.upload_params(l, params[2 * l], params[2 * l + 1])
xd = torch.tensor(x, dtype=torch.float32, device="cuda:0")
yd = torch.tensor(y, dtype=torch.int32, device="cuda:0")
eng.train_begin(optimizer="adam", learning_rate=0.01, decay_rate=0.9, l2_lambda=0.001, epochs=3)
side = torch.cuda.Stream()
with torch.cuda.stream(side):          # a non-default stream, so that graph replay is exercised too
    for _ in range(3):
        eng.train_step(xd, yd)
torch.cuda.synchronize()
stack = orc.parse_stack(algos)
adam = orc.AdamState()
p, costs = params, []
tgt = [[int(v)] for v in y]
for step in range(3):
    p, c, *_ = orc.train_step(stack, p, x, tgt, lr=0.01 * 0.9 ** step, optimizer="adam", adam=adam, l2_lambda=0.001)
    costs.append(c)
close(eng.costs(0, 3), costs, "float32", "costs")
for l in range(3):
    w, b = eng.download_params(l)
    close(w, p[2 * l], "float32", f"W{l}", atol=ADAM_PARAM_ATOL["float32"], outliers=True, loose_atol=0.02)
    close(b, p[2 * l + 1], "float32", f"b{l}", atol=ADAM_PARAM_ATOL["float32"], outliers=True, loose_atol=0.02)
out = eng.forward(xd).cpu().numpy()
ref = orc.forward(stack, p, x)[0][-1]
close(out, ref, "float32", "probs", rtol=1e-4)
print("CHILD-OK")
'''


def run_child(env_extra):
    env = dict(os.environ)
    env.update(env_extra)
    r = subprocess.run([sys.executable, "-c", CHILD, ROOT], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "CHILD-OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.gpu
@pytest.mark.parametrize("env", [{"POZ_TCS": "0"}, {"POZ_TCS_FUSE_UPDATE": "0"}, {"POZ_TCS_GROUP": "1"},
                                 {"POZ_TCS_GROUP": "1", "POZ_TCS_FUSE_UPDATE": "0"}, {"POZ_TAIL": "0"},
                                 {"POZ_TAIL": "0", "POZ_TCS": "0"}, {"POZ_TAIL_GROUP": "0"},
                                 {"POZ_TAIL_CLUSTER": "1"}, {"POZ_TAIL_CLUSTER": "2"}, {"POZ_PDL": "0"}],
                         ids=["simt-layer0", "gradient-stored", "grouped", "grouped-stored", "separate-tail", "round-1-step",
                              "tail-without-grouped-weight-gradients", "tail-without-cluster", "tail-cluster-of-2",
                              "plain-edges-instead-of-programmatic-launches"])
def test_small_batch_tensor_core_variants_match_the_oracle(env):
    """The default C2 step runs layer 0's forward and weight gradient as cluster split-K tcgen05 launches with the
    optimizer in the weight-gradient epilogue (csrc/poz_tcs_kernel.cuh).  The same three steps with that path off,
    with the gradient stored for the update kernel instead, and with every layer's weight gradient grouped into one
    launch must give the oracle's numbers too."""
    run_child(env)


@pytest.mark.gpu
def test_default_path_in_a_child_process_matches_too():
    run_child({})


# ---- inference variants of the tcgen05 path (csrc/poz_tc3_x32t_kernel.cuh and its predecessors) ----------------
INFER_CHILD = r'''
import sys
sys.path.insert(0, sys.argv[1])
sys.path.insert(0, sys.argv[1] + "/tests")
import numpy as np, torch
from oracle import mlp_oracle as orc
from oracle.detgen import det_params, det_normalish
from penr_oz_neural_network_torch_b200 import DeviceMLP, StackSpec

for sizes, last_act, batch in [([784, 256, 128, 10], "softmax", 4096 + 77), ([100, 200, 72, 5], "sigmoid", 2048 + 300)]:
    nl = len(sizes) - 1
    algos = []
    for l in range(nl):
        algos += ["linear", last_act if l == nl - 1 else "relu"]
    params = det_params(sizes, True, 31)
    eng = DeviceMLP(sizes, StackSpec.from_algos(algos), True, "float32", 0)
    for l in range(nl):
        eng.upload_params(l, params[2 * l], params[2 * l + 1])
    x = det_normalish((batch, sizes[0]), 32)
    out = eng.forward(torch.tensor(x, dtype=torch.float32, device="cuda:0")).cpu().numpy().astype(np.float64)
    ref = orc.forward(orc.parse_stack(algos), params, x)[0][-1].reshape(batch, -1)
    # the contract's float32 tolerance, on every output: |out - ref| <= 1e-5 * |ref| (+ 1e-7 absolute)
    bad = np.abs(out - ref) > 1e-5 * np.abs(ref) + 1e-7
    assert not bad.any(), (sizes, int(bad.sum()), float((np.abs(out - ref) / np.abs(ref)).max()))
    assert (out.argmax(1) == ref.argmax(1)).mean() > 0.9995
    eng.close()
print("CHILD-OK")
'''


def run_infer_child(env_extra):
    env = dict(os.environ)
    env.update(env_extra)
    r = subprocess.run([sys.executable, "-c", INFER_CHILD, ROOT], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "CHILD-OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.gpu
@pytest.mark.parametrize("env", [{}, {"POZ_TC3_F32_CHAIN": "0"}, {"POZ_TC3_FUSE_HEAD": "0"},
                                 {"POZ_TC3_X32": "0"}, {"POZ_TC3_MIN_BATCH": "1000000000"}],
                         ids=["default", "plane-hand-over", "separate-head", "pre-split-input", "simt"])
def test_inference_variants_within_the_float32_contract(env):
    """Every variant of the large-batch inference path (and the SIMT kernels) keeps every output within rtol 1e-5
    of the float64 oracle -- the tolerance BASELINE.json's north_star states for float32."""
    run_infer_child(env)
