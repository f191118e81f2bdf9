"""GPU parity: the tcgen05 GEMM (T5 linears) vs a plain fp32 torch matmul of the same bf16 operands."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def run_gemm(A, B, out, mode=0, relu=0, row_scale=None, cta_group=2):
    from pacednegatives_b200 import _lib
    _lib.require_gpu()
    M, K = A.shape
    N = B.shape[0]
    import ctypes
    vp = lambda t: ctypes.c_void_p(t.data_ptr())  # row-strided views are allowed (lda/ldc)
    _lib.check(_lib.lib().pn_gemm_bf16(vp(A), A.stride(0), vp(B), B.stride(0), vp(out), out.stride(0),
                                       M, N, K, mode, relu, _lib.ptr(row_scale), cta_group,
                                       _lib.stream()))
    return out


def ref_mm(A, B):
    return A.float() @ B.float().t()


@pytest.mark.parametrize("cta_group", [1, 2])
@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (256, 256, 128), (512, 768, 768), (1000, 128, 192),
                                   (77, 384, 512), (4096, 2304, 768), (2048, 768, 3072),
                                   (148 * 128 * 3 + 40, 1536, 768)])
def test_gemm_bf16_out(cta_group, M, N, K):
    torch.manual_seed(M + N + K)
    A = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
    B = (torch.randn(N, K, device="cuda") * 0.5).bfloat16()
    out = torch.full((M, N), float("nan"), device="cuda", dtype=torch.bfloat16)
    run_gemm(A, B, out, cta_group=cta_group)
    torch.cuda.synchronize()
    ref = ref_mm(A, B)
    err = (out.float() - ref).abs().max().item()
    tol = 0.02 * (K ** 0.5) * 0.25 + 0.02
    assert err <= tol, f"max err {err} > {tol}"
    # bf16 rounding of the exact fp32 result is the only error allowed: check tightly in relative terms
    assert torch.allclose(out.float(), ref, rtol=1.6e-2, atol=1e-2 * K ** 0.5 * 0.05)


@pytest.mark.parametrize("cta_group", [1, 2])
def test_gemm_relu_rowscale_and_residual(cta_group):
    torch.manual_seed(5)
    M, N, K = 640, 512, 256
    A = torch.randn(M, K, device="cuda").bfloat16()
    B = (torch.randn(N, K, device="cuda") * 0.1).bfloat16()
    rs = torch.rand(M, device="cuda") + 0.5
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    run_gemm(A, B, out, relu=1, row_scale=rs, cta_group=cta_group)
    ref = torch.relu(ref_mm(A, B) * rs[:, None])
    assert torch.allclose(out.float(), ref, rtol=1e-2, atol=2e-2)
    resid = torch.randn(M, N, device="cuda")
    acc = resid.clone()
    run_gemm(A, B, acc, mode=1, cta_group=cta_group)
    assert torch.allclose(acc, resid + ref_mm(A, B), rtol=1e-4, atol=1e-3)
    # strided views: A is a column slice of a wider buffer, out a slice too
    wide = torch.randn(M, K + 128, device="cuda").bfloat16()
    Av = wide[:, 64:64 + K]
    outw = torch.zeros(M, N + 256, device="cuda", dtype=torch.bfloat16)
    run_gemm(Av, B, outw[:, 128:128 + N], cta_group=cta_group)
    assert torch.allclose(outw[:, 128:128 + N].float(), ref_mm(Av, B), rtol=1.6e-2, atol=2e-2)
    assert (outw[:, :128] == 0).all() and (outw[:, 128 + N:] == 0).all()


@pytest.mark.parametrize("M,N,K", [(100, 256, 64), (148 * 128 * 2 + 77, 768, 768), (4096 + 31, 1024, 1024),
                                   (2048 + 5, 768, 3072)])
def test_gemm_residual_epilogue_ragged_strided(M, N, K):
    """fp32 read-modify-write epilogue: narrow-K launches take the TMA-fed tile ring (several tiles per CTA pair, a
    ragged last slab, a pair whose second CTA lies beyond M), K = 3072 the register path; the output is a column
    slice of a wider buffer whose margins must stay untouched."""
    torch.manual_seed(M + K)
    A = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
    B = (torch.randn(N, K, device="cuda") * 0.1).bfloat16()
    rs = torch.rand(M, device="cuda") + 0.5
    wide = torch.randn(M, N + 64, device="cuda")
    before = wide.clone()
    run_gemm(A, B, wide[:, 32:32 + N], mode=1, row_scale=rs, cta_group=2)
    ref = before[:, 32:32 + N] + ref_mm(A, B) * rs[:, None]
    assert torch.allclose(wide[:, 32:32 + N], ref, rtol=1e-4, atol=2e-3)
    assert torch.equal(wide[:, :32], before[:, :32]) and torch.equal(wide[:, 32 + N:], before[:, 32 + N:])


def test_gemm_rejects_bad_shapes():
    from pacednegatives_b200 import _lib
    A = torch.zeros(64, 100, device="cuda", dtype=torch.bfloat16)
    B = torch.zeros(128, 100, device="cuda", dtype=torch.bfloat16)
    out = torch.zeros(64, 128, device="cuda", dtype=torch.bfloat16)
    with pytest.raises(RuntimeError):
        run_gemm(A, B, out)


@pytest.mark.gpu
@pytest.mark.parametrize("cta_group", [1, 2])
@pytest.mark.parametrize("K,M,N", [(4096, 768, 768), (1000, 2304, 768), (520, 3000, 256), (64, 128, 128),
                                   (20000, 768, 3072)])
def test_wgrad_gemm_vs_torch(cta_group, K, M, N):
    """dW[M,N] += dY[K,M]^T X[K,N] on the MN-major tcgen05 path with split-K fp32 atomics."""
    from pacednegatives_b200 import _lib
    _lib.require_gpu()
    torch.manual_seed(K + M + N)
    dY = (torch.randn(K, M, device="cuda") * 0.5).bfloat16()
    X = (torch.randn(K, N, device="cuda") * 0.5).bfloat16()
    out = torch.randn(M, N, device="cuda")
    ref = out + dY.float().t() @ X.float()
    _lib.check(_lib.lib().pn_gemm_bf16_wgrad(_lib.ptr(dY), M, _lib.ptr(X), N, _lib.ptr(out), N, M, N, K, cta_group,
                                             _lib.stream()))
    err = (out - ref).abs().max().item()
    scale = ref.abs().max().item()
    assert err <= 2e-5 * scale + 1e-3, f"K={K} M={M} N={N}: max err {err} (scale {scale})"


# ---------------------------------------------------------------------------------------------------------------
# Building blocks of the fused RMSNorm backward (include/pn_b200.h: pn_gemm_bf16_relu_bwd / _norm_bwd / pn_t5_norm_coef)
@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (1000, 768, 256), (148 * 256 + 77, 3072, 768)])
def test_relu_backward_epilogue_and_its_row_dots(M, N, K):
    """out = rs * (dY Wt^T) where act > 0; dot partials: sum over each 256-column tile of out * act.
    Tolerance: bf16 output rounding (2^-8 relative) on the product; the dots are fp32 sums of the UNROUNDED products,
    compared with the fp32 torch value at 2e-3 of the row's absolute-sum scale."""
    from pacednegatives_b200 import _lib
    _lib.require_gpu()
    import ctypes
    torch.manual_seed(M + N)
    dY = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
    Wt = (torch.randn(N, K, device="cuda") * 0.1).bfloat16()
    act = torch.relu(torch.randn(M, N, device="cuda")).bfloat16()
    rs = torch.rand(M, device="cuda") + 0.5
    out = torch.full((M, N), float("nan"), device="cuda").bfloat16()
    parts = torch.full((N // 256, M), float("nan"), device="cuda")
    vp = lambda t: ctypes.c_void_p(t.data_ptr())
    _lib.check(_lib.lib().pn_gemm_bf16_relu_bwd(vp(dY), K, vp(Wt), K, vp(act), N, vp(rs), vp(out), N, vp(parts), M, N, K,
                                                _lib.stream()))
    torch.cuda.synchronize()
    ref = ref_mm(dY, Wt) * rs[:, None] * (act > 0)
    assert torch.allclose(out.float(), ref, rtol=1.6e-2, atol=2e-2)
    prod = ref * act.float()
    ref_parts = prod.view(M, N // 256, 256).sum(2).t()
    scale = prod.abs().view(M, N // 256, 256).sum(2).t() + 1e-3
    assert torch.isfinite(parts).all()
    assert ((parts - ref_parts).abs() / scale).max().item() < 2e-3


@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (1000, 768, 2304), (148 * 256 + 77, 768, 3072), (4096 + 31, 1024, 1024)])
def test_norm_backward_residual_epilogue(M, N, K):
    """dh += Ys Wt^T - coef (.) hb and its bf16 copy, on a column slice of a wider buffer (margins untouched)."""
    from pacednegatives_b200 import _lib
    _lib.require_gpu()
    import ctypes
    torch.manual_seed(M + K)
    Ys = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
    Wt = (torch.randn(N, K, device="cuda") * 0.1).bfloat16()
    hb = torch.randn(M, N, device="cuda").bfloat16()
    coef = torch.randn(M, device="cuda") * 0.3
    wide = torch.randn(M, N + 64, device="cuda")
    before = wide.clone()
    dhb = torch.full((M, N), float("nan"), device="cuda").bfloat16()
    vp = lambda t: ctypes.c_void_p(t.data_ptr())
    dh = wide[:, 32:32 + N]
    _lib.check(_lib.lib().pn_gemm_bf16_norm_bwd(vp(Ys), K, vp(Wt), K, vp(hb), N, vp(coef), vp(dh), wide.stride(0),
                                                vp(dhb), N, M, N, K, _lib.stream()))
    torch.cuda.synchronize()
    ref = before[:, 32:32 + N] + ref_mm(Ys, Wt) - coef[:, None] * hb.float()
    assert torch.allclose(dh, ref, rtol=1e-4, atol=2e-3)
    assert torch.equal(wide[:, :32], before[:, :32]) and torch.equal(wide[:, 32 + N:], before[:, 32 + N:])
    assert torch.equal(dhb, dh.bfloat16())


def test_norm_coef_kernel():
    from pacednegatives_b200 import _lib
    _lib.require_gpu()
    rows, parts, d = 1000, 12, 768
    part = torch.randn(parts, rows, device="cuda")
    rs = torch.rand(rows, device="cuda") + 0.5
    coef = torch.zeros(rows, device="cuda")
    _lib.check(_lib.lib().pn_t5_norm_coef(_lib.ptr(part), parts, _lib.ptr(rs), rows, d, _lib.ptr(coef), _lib.stream()))
    torch.cuda.synchronize()
    assert torch.allclose(coef, rs * part.sum(0) / d, rtol=1e-5, atol=1e-6)
