"""Split-bf16 tcgen05 GEMM (float32 in / out) against a float64 PyTorch reference of the same op.
Stated tolerance: rtol 1e-5 + atol 1e-5 * max|reference| -- the float32 contract of the narrow
configurations; the measured error is printed."""
import pytest
import torch

pytestmark = pytest.mark.gpu
ACT = {"none": 0, "relu": 1, "sigmoid": 2, "tanh": 3}


def _call(variant, a, b, m, n, k, out, bias=None, aux=None, act=0, splits=1):
    from penr_oz_neural_network_torch_b200 import _native as nat
    lib = nat.lib()
    torch.cuda.synchronize()
    nat.check(lib.poz_tc3_gemm(0, variant, a.data_ptr(), b.data_ptr(), m, n, k, out.data_ptr(),
                               None if bias is None else bias.data_ptr(), None if aux is None else aux.data_ptr(),
                               act, splits, int(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()


def _rand(shape, seed, scale=1.0):
    g = torch.Generator(device="cuda:0").manual_seed(seed)
    return torch.randn(shape, generator=g, device="cuda:0") * scale


def _close(got, want, what):
    want = want.double()
    scale = want.abs().max().item()
    err = (got.double() - want).abs().max().item() / scale
    print(f"{what}: max error / max|ref| = {err:.2e}")
    torch.testing.assert_close(got.double(), want, rtol=1e-5, atol=1e-5 * scale)


SHAPES = [(128, 64, 64), (256, 256, 256), (1024, 256, 784), (300, 264, 200), (128, 10, 128), (5000, 128, 256),
          (1000, 72, 1000)]


@pytest.mark.parametrize("m,n,k", SHAPES)
def test_forward(m, n, k):
    a, b, bias = _rand((m, k), 1), _rand((k, n), 2, 0.1), _rand((n,), 3)
    out = torch.full((m, n), float("nan"), device="cuda:0")
    _call(0, a, b, m, n, k, out, bias=bias, act=ACT["relu"])
    _close(out, (a.double() @ b.double() + bias.double()).relu(), f"fwd {m}x{n}x{k}")


@pytest.mark.parametrize("m,n,k", SHAPES)
def test_backward_data(m, n, k):
    dz, w, aux = _rand((m, k), 4), _rand((n, k), 5, 0.1), _rand((m, n), 6).relu()
    out = torch.full((m, n), float("nan"), device="cuda:0")
    _call(1, dz, w, m, n, k, out, aux=aux, act=ACT["relu"])
    _close(out, (dz.double() @ w.double().T) * (aux > 0), f"bwd_x {m}x{n}x{k}")


@pytest.mark.parametrize("m,n,k,splits", [(128, 64, 64, 1), (784, 256, 1024, 4), (264, 520, 1000, 3), (128, 10, 1024, 8)])
def test_backward_weight(m, n, k, splits):
    x, dz = _rand((k, m), 7), _rand((k, n), 8, 0.1)
    out = torch.full((splits, m, n), float("nan"), device="cuda:0")
    _call(2, x, dz, m, n, k, out, splits=splits)
    _close(out.sum(0), x.double().T @ dz.double(), f"bwd_w {m}x{n}x{k}/{splits}")
