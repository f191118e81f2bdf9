"""GPU parity: backward of the encoder self-attention (flash-style tcgen05 kernel) vs torch autograd of
softmax(QK^T + bias) V in fp32 on the same bf16 inputs.  The kernel recomputes P from the forward's base-2
logsumexp and rounds P / dS to bf16 before the accumulating MMAs, hence the bf16-level tolerances."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def run_case(docs, S, H, with_scale, seed, DC=0):
    from pacednegatives_b200 import _lib
    lib = _lib.lib()
    torch.manual_seed(seed)
    di = H * 64
    rows = docs * S
    qkv = (torch.randn(rows, 3 * di, device="cuda") * 0.5).bfloat16()
    bias = torch.randn(H, 1023, device="cuda")
    if DC:   # T5-like table: one value per side beyond |j - i| >= DC
        bias[:, 511 + DC:] = bias[:, 1022:1023]
        bias[:, :511 - DC + 1] = bias[:, 0:1]
    dctx = (torch.randn(rows, di, device="cuda") * 0.5).bfloat16()
    rs = (torch.rand(rows, device="cuda") + 0.5) if with_scale else None
    ctx = torch.zeros(rows, di, device="cuda", dtype=torch.bfloat16)
    lse = torch.zeros(rows, H, device="cuda")
    _lib.check(lib.pn_t5_self_attention_lse(_lib.ptr(qkv), docs, S, H, _lib.ptr(bias), _lib.ptr(ctx),
                                            _lib.ptr(lse), _lib.stream()))
    dqkv = torch.full((rows, 3 * di), float("nan"), device="cuda").bfloat16()
    diag = torch.zeros(H, 1023, device="cuda")
    dvec = torch.zeros(rows, H, device="cuda")
    _lib.check(lib.pn_t5_self_attention_bwd(_lib.ptr(qkv), docs, S, H, _lib.ptr(bias), _lib.ptr(ctx),
                                            _lib.ptr(dctx), _lib.ptr(lse), _lib.ptr(rs) if with_scale else None,
                                            _lib.ptr(dqkv), _lib.ptr(diag), _lib.ptr(dvec), _lib.stream()))
    torch.cuda.synchronize()

    x = qkv.float().requires_grad_(True)
    bp = bias.clone().requires_grad_(True)
    q, k, v = [t.view(docs, S, H, 64).transpose(1, 2) for t in x.split(di, dim=1)]
    pos = torch.arange(S, device="cuda")
    b = bp[:, (pos[None, :] - pos[:, None]) + 511]
    s = q @ k.transpose(2, 3) + b[None]
    p = torch.softmax(s, dim=-1)
    ref_ctx = (p @ v).transpose(1, 2).reshape(rows, di)
    ref_lse2 = torch.logsumexp(s, dim=-1).transpose(1, 2).reshape(rows, H) * 1.4426950408889634
    ref_ctx.backward(dctx.float())
    gx = x.grad
    if with_scale:
        gx = gx * rs[:, None]
    assert (lse - ref_lse2).abs().max().item() < 2e-2
    got = dqkv.float()
    assert torch.isfinite(got).all(), "unwritten or non-finite gradient rows"
    scale = gx.abs().max().item()
    for name, sl in (("dq", slice(0, di)), ("dk", slice(di, 2 * di)), ("dv", slice(2 * di, 3 * di))):
        err = (got[:, sl] - gx[:, sl]).abs().max().item()
        assert err < 2e-2 * max(1.0, scale), f"{name}: docs={docs} S={S} H={H}: err {err} (scale {scale})"
    gb = bp.grad
    if DC:   # the constant regions are one parameter each: compare their totals, and the near diagonals one by one
        def fold(x):
            return torch.cat([x[:, :511 - DC + 1].sum(1, keepdim=True), x[:, 511 - DC + 1:511 + DC],
                              x[:, 511 + DC:].sum(1, keepdim=True)], dim=1)
        diag, gb = fold(diag), fold(gb)
    err = (diag - gb).abs().max().item()
    assert err < 2e-2 * max(1.0, gb.abs().max().item()), f"dbias: docs={docs} S={S} H={H}: err {err}"


@pytest.mark.parametrize("docs,S,H,with_scale", [
    (1, 128, 1, False), (2, 128, 2, False), (1, 256, 1, False), (3, 512, 2, True), (2, 384, 4, False),
    (1, 200, 2, True), (2, 77, 2, False), (5, 512, 12, True),
])
def test_self_attention_backward_vs_autograd(docs, S, H, with_scale):
    from pacednegatives_b200 import _lib
    _lib.require_gpu()
    run_case(docs, S, H, with_scale, seed=docs * 1000 + S + H)


@pytest.mark.parametrize("docs,S,H,DC", [(3, 512, 2, 91), (2, 384, 4, 128), (1, 200, 2, 60), (5, 512, 12, 91),
                                         (2, 512, 1, 300)])
def test_self_attention_backward_saturated_bias(docs, S, H, DC):
    """T5-shaped bias tables (one value per side beyond the bucket saturation distance, 91 for 32 buckets /
    max_distance 128): the saturated regions are a single parameter each, so their gradient is compared as a sum."""
    from pacednegatives_b200 import _lib
    _lib.require_gpu()
    run_case(docs, S, H, True, seed=docs * 1000 + S + H + DC, DC=DC)


@pytest.mark.parametrize("docs,S,H,with_scale", [(3, 512, 2, True), (1, 200, 2, True), (2, 77, 4, False),
                                                 (5, 512, 12, True)])
def test_self_attention_backward_row_dots(docs, S, H, with_scale):
    """pn_t5_self_attention_bwd_dot: the 4*H planes summed = sum_col dqkv[row, col] * qkv[row, col] (the RMSNorm
    backward's row dot), and dqkv itself is bit-identical to the plain entry point's.  The kernel sums the fp32
    (unrounded) gradient entries, the check the stored bf16 ones: 1e-2 of the row's absolute-sum scale."""
    from pacednegatives_b200 import _lib
    _lib.require_gpu()
    lib = _lib.lib()
    torch.manual_seed(7 * docs + S + H)
    di, rows = H * 64, docs * S
    qkv = (torch.randn(rows, 3 * di, device="cuda") * 0.5).bfloat16()
    bias = torch.randn(H, 1023, device="cuda")
    dctx = (torch.randn(rows, di, device="cuda") * 0.5).bfloat16()
    rs = (torch.rand(rows, device="cuda") + 0.5) if with_scale else None
    ctx = torch.zeros(rows, di, device="cuda", dtype=torch.bfloat16)
    lse = torch.zeros(rows, H, device="cuda")
    _lib.check(lib.pn_t5_self_attention_lse(_lib.ptr(qkv), docs, S, H, _lib.ptr(bias), _lib.ptr(ctx), _lib.ptr(lse),
                                            _lib.stream()))
    outs = []
    for with_dot in (False, True):
        dqkv = torch.full((rows, 3 * di), float("nan"), device="cuda").bfloat16()
        diag = torch.zeros(H, 1023, device="cuda")
        dvec = torch.zeros(rows, H, device="cuda")
        dots = torch.full((4 * H, rows), float("nan"), device="cuda")
        args = [_lib.ptr(qkv), docs, S, H, _lib.ptr(bias), _lib.ptr(ctx), _lib.ptr(dctx), _lib.ptr(lse),
                _lib.ptr(rs) if with_scale else None, _lib.ptr(dqkv), _lib.ptr(diag), _lib.ptr(dvec)]
        if with_dot:
            _lib.check(lib.pn_t5_self_attention_bwd_dot(*args, _lib.ptr(dots), _lib.stream()))
        else:
            _lib.check(lib.pn_t5_self_attention_bwd(*args, _lib.stream()))
        torch.cuda.synchronize()
        outs.append((dqkv, dots))
    assert torch.equal(outs[0][0], outs[1][0])
    dqkv, dots = outs[1]
    assert torch.isfinite(dots).all()
    prod = dqkv.float() * qkv.float()
    ref = prod.sum(1)
    scale = prod.abs().sum(1) + 1e-3
    assert ((dots.sum(0) - ref).abs() / scale).max().item() < 1e-2
