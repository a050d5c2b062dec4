"""tcgen05 GEMM variants against a plain PyTorch fp32 reference of the same op (bf16 inputs,
fp32 accumulation).  Tolerances: fp32 output rtol 1e-4 + atol 1e-4*max (different summation order
over K); bf16 output additionally carries one bf16 rounding (2^-9 relative)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

ACT = {"none": 0, "relu": 1, "sigmoid": 2, "tanh": 3}


def _call(variant, a, b, m, n, k, out, bias=None, aux=None, act=0, splits=1):
    from penr_oz_neural_network_torch_b200 import _native as nat
    lib = nat.lib()
    torch.cuda.synchronize()
    nat.check(lib.poz_tc_gemm(0, variant, a.data_ptr(), b.data_ptr(), m, n, k, out.data_ptr(),
                              None if bias is None else bias.data_ptr(),
                              None if aux is None else aux.data_ptr(), act, splits,
                              int(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()


def _rand(shape, seed, scale=1.0):
    g = torch.Generator(device="cuda:0").manual_seed(seed)
    return (torch.randn(shape, generator=g, device="cuda:0") * scale).to(torch.bfloat16)


def _close(got, want, bf16_out):
    got, want = got.float(), want.float()
    scale = want.abs().max().item()
    rtol, atol = (1e-2, 1e-2 * scale) if bf16_out else (1e-4, 1e-4 * scale)
    torch.testing.assert_close(got, want, rtol=rtol, atol=atol)


SHAPES = [(256, 256, 64), (256, 256, 256), (512, 512, 1024), (300, 264, 200), (128, 8, 72), (1000, 4096, 4096)]


@pytest.mark.parametrize("m,n,k", SHAPES)
@pytest.mark.parametrize("act", ["relu", "none"])
def test_forward_variant(m, n, k, act):
    a, b = _rand((m, k), 1), _rand((k, n), 2, 0.1)
    bias = torch.randn(n, device="cuda:0")
    out = torch.full((m, n), float("nan"), device="cuda:0", dtype=torch.bfloat16)
    _call(0, a, b, m, n, k, out, bias=bias, act=ACT[act])
    want = a.float() @ b.float() + bias
    if act == "relu":
        want = want.relu()
    _close(out, want, True)


@pytest.mark.parametrize("m,n,k", SHAPES)
def test_backward_data_variant(m, n, k):
    dz, w = _rand((m, k), 3), _rand((n, k), 4, 0.1)
    a_prev = _rand((m, n), 5).relu()
    out = torch.full((m, n), float("nan"), device="cuda:0", dtype=torch.bfloat16)
    _call(1, dz, w, m, n, k, out, aux=a_prev, act=ACT["relu"])
    want = (dz.float() @ w.float().T) * (a_prev.float() > 0)
    _close(out, want, True)


@pytest.mark.parametrize("m,n,k,splits", [(256, 256, 64, 1), (256, 256, 4096, 4), (264, 520, 1000, 3),
                                          (4096, 512, 8192, 2), (72, 8, 136, 1)])
def test_backward_weight_variant(m, n, k, splits):
    x, dz = _rand((k, m), 6), _rand((k, n), 7, 0.1)
    out = torch.full((splits, m, n), float("nan"), device="cuda:0", dtype=torch.float32)
    _call(2, x, dz, m, n, k, out, splits=splits)
    want = x.float().T @ dz.float()
    _close(out.sum(0), want, False)
