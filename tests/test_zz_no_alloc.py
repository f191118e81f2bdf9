"""pn_t5_self_attention_bwd_static — the allocation-free variant of the attention backward (scratch in a file-scope
device array of libpn_b200.so; no counterpart in the reference) —
against the existing entry point pn_t5_self_attention_bwd, through the C-ABI, on two calls in a row (the second call
finds the first one's contents in the shared array).  Collected last on purpose.

Tolerance: the two entry points run the SAME kernels on the same inputs; dq / dk / dv are deterministic (bit-equal),
the bias-gradient diagonals are flushed with fp32 atomics across CTAs, so they agree to summation order — checked with
the 2e-2 absolute (x scale) bound tests/test_gpu_attn_bwd.py uses, and in practice far tighter."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(docs, S, H, seed):
    from pacednegatives_b200 import _lib
    lib = _lib.lib()
    torch.manual_seed(seed)
    di, rows = H * 64, docs * S
    qkv = (torch.randn(rows, 3 * di, device="cuda") * 0.5).bfloat16()
    bias = torch.randn(H, 1023, device="cuda")
    dctx = (torch.randn(rows, di, device="cuda") * 0.5).bfloat16()
    rs = torch.rand(rows, device="cuda") + 0.5
    ctx = torch.zeros(rows, di, device="cuda", dtype=torch.bfloat16)
    lse = torch.zeros(rows, H, device="cuda")
    _lib.check(lib.pn_t5_self_attention_lse(_lib.ptr(qkv), docs, S, H, _lib.ptr(bias), _lib.ptr(ctx), _lib.ptr(lse),
                                            _lib.stream()))
    return qkv, bias, ctx, dctx, lse, rs


def _run(static, docs, S, H, t):
    from pacednegatives_b200 import _lib
    lib = _lib.lib()
    qkv, bias, ctx, dctx, lse, rs = t
    di, rows = H * 64, docs * S
    dqkv = torch.full((rows, 3 * di), float("nan"), device="cuda").bfloat16()
    diag = torch.zeros(H, 1023, device="cuda")
    if static:
        _lib.check(lib.pn_t5_self_attention_bwd_static(_lib.ptr(qkv), docs, S, H, _lib.ptr(bias), _lib.ptr(ctx),
                                                       _lib.ptr(dctx), _lib.ptr(lse), _lib.ptr(rs), _lib.ptr(dqkv),
                                                       _lib.ptr(diag), _lib.stream()))
    else:
        dvec = torch.zeros(rows, H, device="cuda")
        _lib.check(lib.pn_t5_self_attention_bwd(_lib.ptr(qkv), docs, S, H, _lib.ptr(bias), _lib.ptr(ctx),
                                                _lib.ptr(dctx), _lib.ptr(lse), _lib.ptr(rs), _lib.ptr(dqkv),
                                                _lib.ptr(diag), _lib.ptr(dvec), _lib.stream()))
    torch.cuda.synchronize()
    return dqkv.float(), diag


def test_static_scratch_attention_backward_matches_the_workspace_entry_point():
    from pacednegatives_b200 import _lib
    _lib.require_gpu()
    # two calls in a row on different shapes: the second finds the first's values in the shared scratch array
    for docs, S, H, seed in ((3, 512, 4, 11), (2, 200, 2, 12)):
        t = _inputs(docs, S, H, seed)
        ref_dqkv, ref_diag = _run(False, docs, S, H, t)
        got_dqkv, got_diag = _run(True, docs, S, H, t)
        assert torch.isfinite(got_dqkv).all()
        scale = max(1.0, ref_dqkv.abs().max().item())
        assert (got_dqkv - ref_dqkv).abs().max().item() < 2e-2 * scale
        assert torch.equal(got_dqkv, ref_dqkv)          # same kernels, deterministic outputs
        dscale = max(1.0, ref_diag.abs().max().item())
        assert (got_diag - ref_diag).abs().max().item() < 2e-2 * dscale


def test_static_scratch_rejects_shapes_beyond_its_maximum():
    from pacednegatives_b200 import _lib
    _lib.require_gpu()
    lib = _lib.lib()
    z = torch.zeros(8, device="cuda")
    # 129 documents x 512 tokens x 16 heads = 1 056 768 > 1 048 576 entries: refused before anything is launched
    rc = lib.pn_t5_self_attention_bwd_static(_lib.ptr(z), 129, 512, 16, _lib.ptr(z), _lib.ptr(z), _lib.ptr(z),
                                             _lib.ptr(z), None, _lib.ptr(z), _lib.ptr(z), _lib.stream())
    assert rc == -5
    msg = lib.pn_last_error().decode()
    assert "exceeds the built-in scratch" in msg and "1048576" in msg
