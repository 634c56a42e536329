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


@pytest.mark.parametrize("M,N,K", [(197 * 5, 768, 768), (1000, 768, 3072), (129, 264, 128), (31, 40, 64), (256 * 5 + 7, 512, 256)])
def test_gemm_fp16_residual_stream_update(M, N, K):
    """The fused residual form: x16 = x16 + a w^T + b, residual boxes in and sums out by TMA (fp16 stream).  Row /
    column counts that are not multiples of the 32 x 32 box exercise the clipped loads and stores; in place and
    with a separate output.  Tolerance: fp16 output rounding (2^-11 relative) + accumulation-order noise."""
    a, w, bias = _mk(M, N, K, seed=M + N + K + 1)
    x = (torch.randn(M, N, device=DEV) * 1.5).to(torch.float16)
    ref = x.float() + a.float() @ w.float().t() + bias
    out = torch.full((M, N), float("nan"), dtype=torch.float16, device=DEV)
    ops.gemm_bf16(a, w, bias=bias, residual=x, out=out)
    _close(out, ref, 2e-3, 2e-3)
    xi = x.clone()
    ops.gemm_bf16(a, w, bias=bias, residual=xi, out=xi)                 # in place
    assert torch.equal(xi, out)
    # no residual, fp16 output (LSU form)
    o2 = torch.empty(M, N, dtype=torch.float16, device=DEV)
    ops.gemm_bf16(a, w, bias=bias, out=o2)
    _close(o2, a.float() @ w.float().t() + bias, 2e-3, 2e-3)


def test_gemm_fp16_stream_saturates_instead_of_overflowing():
    a = torch.full((64, 64), 8.0, dtype=torch.bfloat16, device=DEV)
    w = torch.full((32, 64), 8.0, dtype=torch.bfloat16, device=DEV)      # 64 * 64 = 4096 per output
    x = torch.full((64, 32), 65000.0, dtype=torch.float16, device=DEV)
    ops.gemm_bf16(a, w, residual=x, out=x)
    assert bool(torch.isfinite(x).all()) and float(x.max()) == 65504.0


@pytest.mark.parametrize("name,N,K,kw", [("qkv", 2304, 768, {}), ("proj", 768, 768, {"stream": True}),
                                         ("fc1", 3072, 768, {"gelu": True}), ("fc2", 768, 3072, {"stream": True})])
def test_gemm_block_shapes_at_bench_m(name, N, K, kw):
    """The four block GEMMs at the BENCHMARKED row count (1072 crops x 197 = 211 184 rows, 825 row tiles, one
    pass) against torch.matmul in fp32 on sampled rows: first / last rows, both sides of CTA-pair tile (256) and
    TMA box (32) boundaries, and 64 random rows."""
    M = 1072 * 197
    g = torch.Generator(device="cpu").manual_seed(N + K)
    a = (torch.randn(M, K, generator=g, dtype=torch.float32) * 0.5).to(torch.bfloat16).to(DEV)
    w = (torch.randn(N, K, generator=g) * 0.05).to(torch.bfloat16).to(DEV)
    bias = (torch.randn(N, generator=g) * 0.1).to(DEV)
    rows = sorted(set([0, 1, 31, 32, 127, 128, 255, 256, 257, M // 2, 824 * 256 - 1, 824 * 256, M - 33, M - 32, M - 2, M - 1]
                      + torch.randint(0, M, (64,), generator=g).tolist()))
    ri = torch.tensor(rows, device=DEV)
    base = a[ri].float() @ w.float().t() + bias
    if kw.get("stream"):
        x = (torch.randn(M, N, device=DEV) * 1.5).to(torch.float16)
        ref = x[ri].float() + base
        ops.gemm_bf16(a, w, bias=bias, residual=x, out=x)
        _close(x[ri], ref, 2e-3, 2e-3)
    elif kw.get("gelu"):
        got = ops.gemm_bf16(a, w, bias=bias, gelu=True)
        _close(got[ri], torch.nn.functional.gelu(base), 1e-2, 1e-2)
    else:
        got = ops.gemm_bf16(a, w, bias=bias)
        _close(got[ri], base, 1e-2, 1e-2)


@pytest.mark.parametrize("M,N,K,gelu", [(128, 256, 256, False), (197 * 5, 2304, 768, False), (197 * 5, 3072, 768, True),
                                        (1000, 768, 1024, False), (333, 272, 512, True), (1, 256, 768, False)])
def test_linear_with_folded_layernorm(M, N, K, gelu):
    """yv_linear_ln_f16: Linear(LayerNorm(x)) with the LayerNorm folded into the GEMM (f16 rows x fp16 g(.)W on the
    tensor cores, rstd (acc - mu colsum) + c in the epilogue) against LayerNorm + matmul in fp32 on the same f16 rows
    and bf16 weights.  Rows carry a large common offset (mean 3, spread 0.5..4) so the mean subtraction in the
    epilogue is exercised; tolerance = bf16 output rounding + the fp16 rounding of the folded weights (2^-11 per
    term, sqrt(K) accumulation)."""
    g = torch.Generator().manual_seed(M + N + K)
    scale = torch.rand(M, 1, generator=g) * 3.5 + 0.5
    x = (torch.randn(M, K, generator=g) * scale + 3.0 * torch.randn(M, 1, generator=g)).to(torch.float16)
    w = (torch.randn(N, K, generator=g) * 0.05).to(torch.bfloat16)
    bias = torch.randn(N, generator=g) * 0.1
    gamma = 1.0 + 0.3 * torch.randn(K, generator=g)
    beta = 0.2 * torch.randn(K, generator=g)
    ref = torch.nn.functional.layer_norm(x.double(), (K,), gamma.double(), beta.double(), 1e-6) @ w.double().t() + bias.double()
    if gelu:
        ref = torch.nn.functional.gelu(ref)
    got = ops.linear_ln_f16(x.to(DEV), w.to(DEV), bias.to(DEV), gamma.to(DEV), beta.to(DEV), 1e-6, gelu).cpu()
    _close(got, ref.float(), rtol=2 ** -7, atol=6e-3)
    # and it is the same operator as the two-kernel form (LayerNorm kernel to bf16, then the bf16 GEMM), whose
    # bf16 rounding of the normalised rows is the larger error of the two
    y = torch.nn.functional.layer_norm(x.float(), (K,), gamma, beta, 1e-6).to(torch.bfloat16)
    two = ops.gemm_bf16(y.to(DEV), w.to(DEV), bias.to(DEV), gelu=gelu).cpu()
    e_fold = float((got.double() - ref).abs().max())
    e_two = float((two.double() - ref).abs().max())
    assert e_fold <= 1.5 * e_two + 1e-3, (e_fold, e_two)


@pytest.mark.parametrize("M,N,K", [(197 * 5, 768, 768), (197 * 5, 768, 3072), (1000, 256, 128), (77, 1024, 192), (1, 768, 64)])
def test_residual_update_with_row_statistics(M, N, K):
    """yv_residual_linear_f16: x += a w^T + b in place (the fused TMA residual epilogue) with the (mean, rstd) of the
    UPDATED rows as a by-product of the epilogue (partial sums per 64 columns + combine kernel).  The update must be
    bit-identical to the same GEMM without statistics, and the statistics must match fp64 statistics of the stored
    fp16 rows: the epilogue sums the fp32 values BEFORE their fp16 rounding, so the mean may differ by the average
    rounding error (<= 2^-12 max|row|) and rstd by 1e-3 relative.  Rows carry a large common offset."""
    g = torch.Generator().manual_seed(M + N + K)
    a, w, bias = _mk(M, N, K, seed=M + K)
    x0 = (torch.randn(M, N, generator=g) * (torch.rand(M, 1, generator=g) * 3 + 0.3) + 4.0 * torch.randn(M, 1, generator=g)).to(torch.float16).to(DEV)
    x_plain = x0.clone()
    ops.gemm_bf16(a, w, bias, residual=x_plain, out=x_plain)
    x = x0.clone()
    stats = ops.residual_linear_f16(a, w, bias, x)
    assert torch.equal(x, x_plain)
    ref = x0.double() + a.double() @ w.double().t() + bias.double()
    _close(x, ref.float(), rtol=2 ** -10, atol=2e-3)
    xd = x.double()
    mean = xd.mean(1)
    rstd = (xd.var(1, unbiased=False) + 1e-6).rsqrt()
    tol_mean = 2.0 ** -12 * xd.abs().amax(1) + 1e-5
    assert bool(((stats[:, 0].double() - mean).abs() <= tol_mean).all()), float((stats[:, 0].double() - mean).abs().max())
    assert bool(((stats[:, 1].double() / rstd - 1).abs() <= 1e-3).all()), float((stats[:, 1].double() / rstd - 1).abs().max())
    # without statistics the entry point is the plain in-place update
    x2 = x0.clone()
    assert ops.residual_linear_f16(a, w, bias, x2, want_stats=False) is None and torch.equal(x2, x_plain)


def test_cluster_of_four_multicast_form_parity():
    """YV_GEMM_CLUSTER=4 / 6: two / three CTA pairs per cluster on M-adjacent tiles with the W operand multicast to all of them (opt-in,
    measured slower than pairs at the power cap).  Same results bit for bit as the default form on the
    three TMA epilogues (plain / GELU bf16, fused fp16 residual with statistics, folded LayerNorm), including an odd
    number of M-tiles (the second pair of the last cluster works on an all-padding tile) and a single-tile M."""
    import os, subprocess, sys, tempfile
    code = r"""
import sys, torch
sys.path.insert(0, %r)
from yolov8_vit_b200 import ops
dev = 'cuda:0'
g = torch.Generator().manual_seed(5)
out = []
for M, N, K in ((197 * 13, 768, 768), (300, 1024, 768), (1290, 512, 256), (256, 256, 256)):
    a = (torch.randn(M, K, generator=g) * 0.5).to(torch.bfloat16).to(dev)
    w = (torch.randn(N, K, generator=g) * 0.05).to(torch.bfloat16).to(dev)
    b = torch.randn(N, generator=g).to(dev)
    out.append(ops.gemm_bf16(a, w, b).cpu())
    out.append(ops.gemm_bf16(a, w, b, gelu=True).cpu())
    x = torch.randn(M, N, generator=g).to(torch.float16).to(dev)
    st = ops.residual_linear_f16(a, w, b, x)
    out += [x.cpu(), st.cpu()]
    xf = (torch.randn(M, K, generator=g) + 2).to(torch.float16).to(dev)
    out.append(ops.linear_ln_f16(xf, w, b, torch.ones(K, device=dev), torch.zeros(K, device=dev), 1e-6, True).cpu())
torch.save(out, sys.argv[1])
print('ok')
""" % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    res = []
    with tempfile.TemporaryDirectory() as d:
        for mode in ("2", "4", "6"):
            f = os.path.join(d, mode + ".pt")
            r = subprocess.run([sys.executable, "-c", code, f], env=dict(os.environ, YV_GEMM_CLUSTER=mode),
                               capture_output=True, text=True, timeout=280)
            assert r.returncode == 0 and "ok" in r.stdout, r.stdout + r.stderr
            res.append(torch.load(f))
    assert len(res[0]) == len(res[1]) == len(res[2]) == 20
    for k, (p, q, r6) in enumerate(zip(*res)):
        assert torch.equal(p, q) and torch.equal(p, r6), k
