"""GPU tests of the tcgen05/TMEM/TMA GEMM (K5) against torch.matmul in fp32 on the same bf16
operands.  Tolerance: bf16 output rounding (2^-8 relative) + fp32 accumulation-order noise."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from yolov8_vit_b200 import ops

DEV = "cuda:0"


def _mk(M, N, K, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    a = (torch.randn(M, K, generator=g) * 0.5).to(torch.bfloat16)
    w = (torch.randn(N, K, generator=g) * 0.05).to(torch.bfloat16)
    bias = torch.randn(N, generator=g) * 0.1
    return a.to(DEV), w.to(DEV), bias.to(DEV)


def _close(got, ref, rtol, atol):
    err = (got.float() - ref).abs()
    tol = atol + rtol * ref.abs()
    assert bool((err <= tol).all()), f"max err {float(err.max()):.4g}, worst excess {float((err - tol).max()):.4g}"


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (128, 256, 768), (256, 768, 768), (197 * 3, 2304, 768),
                                   (1000, 768, 3072), (50, 1000, 768), (129, 264, 128)])
def test_gemm_plain_f32_out(M, N, K):
    a, w, _ = _mk(M, N, K, seed=M + N + K)
    ref = a.float() @ w.float().t()
    got = ops.gemm_bf16(a, w, out_f32=True)
    _close(got, ref, 1e-4, 1e-3)


def test_gemm_epilogues():
    M, N, K = 197 * 5, 3072, 768
    a, w, bias = _mk(M, N, K, seed=3)
    base = a.float() @ w.float().t() + bias
    _close(ops.gemm_bf16(a, w, bias=bias), base, 1e-2, 1e-2)
    _close(ops.gemm_bf16(a, w, bias=bias, gelu=True, out_f32=True), torch.nn.functional.gelu(base), 1e-4, 2e-3)
    res = (torch.randn(M, N, device=DEV) * 0.3).to(torch.bfloat16)
    _close(ops.gemm_bf16(a, w, bias=bias, residual=res, out_f32=True), base + res.float(), 1e-4, 2e-3)


def test_gemm_many_tiles_persistent_phases():
    """More tiles than SMs x 2 so every CTA wraps both the smem ring and the TMEM double buffer."""
    M, N, K = 128 * 40, 256 * 9, 256
    a, w, bias = _mk(M, N, K, seed=5)
    ref = a.float() @ w.float().t() + bias
    _close(ops.gemm_bf16(a, w, bias=bias, out_f32=True), ref, 1e-4, 2e-3)
    # determinism
    assert torch.equal(ops.gemm_bf16(a, w, bias=bias), ops.gemm_bf16(a, w, bias=bias))
