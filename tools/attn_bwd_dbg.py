"""Debug driver for the attention backward kernel: one case per subprocess (a hang only costs its timeout)."""
import subprocess
import sys

CASES = [(1, 128, 1, 0), (2, 128, 2, 0), (1, 256, 1, 0), (1, 512, 1, 0), (3, 512, 2, 1), (1, 200, 2, 1), (5, 512, 12, 1)]


def one(docs, S, H, ws):
    import torch
    from pacednegatives_b200 import _lib
    lib = _lib.lib()
    torch.manual_seed(1)
    di = H * 64
    rows = docs * S
    qkv = (torch.randn(rows, 3 * di, device="cuda") * 0.5).bfloat16()
    bias = torch.randn(H, 1023, device="cuda")
    DC = 0
    dctx = (torch.randn(rows, di, device="cuda") * 0.5).bfloat16()
    rs = (torch.rand(rows, device="cuda") + 0.5) if ws else None
    ctx = torch.zeros(rows, di, device="cuda", dtype=torch.bfloat16)
    lse = torch.zeros(rows, H, device="cuda")
    _lib.check(lib.pn_t5_self_attention_lse(_lib.ptr(qkv), docs, S, H, _lib.ptr(bias), _lib.ptr(ctx),
                                            _lib.ptr(lse), _lib.stream()))
    torch.cuda.synchronize()
    dqkv = torch.full((rows, 3 * di), float("nan"), device="cuda").bfloat16()
    diag = torch.zeros(H, 1023, device="cuda")
    dvec = torch.zeros(rows, H, device="cuda")
    _lib.check(lib.pn_t5_self_attention_bwd(_lib.ptr(qkv), docs, S, H, _lib.ptr(bias), _lib.ptr(ctx),
                                            _lib.ptr(dctx), _lib.ptr(lse), _lib.ptr(rs) if ws else None,
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
    if ws:
        gx = gx * rs[:, None]
    D = (dctx.float() * ctx.float()).view(rows, H, 64).sum(-1)
    print("  lse err", (lse - ref_lse2).abs().max().item(), "dvec err", (dvec - D).abs().max().item())
    got = dqkv.float()
    for name, sl in (("dq", slice(0, di)), ("dk", slice(di, 2 * di)), ("dv", slice(2 * di, 3 * di))):
        g, r = got[:, sl], gx[:, sl]
        nan = (~torch.isfinite(g)).sum().item()
        err = (torch.nan_to_num(g) - r).abs()
        print(f"  {name}: nan {nan} maxerr {err.max().item():.4f} ref max {r.abs().max().item():.3f} "
              f"rows with err>0.05: {(err.max(dim=1).values > 0.05).sum().item()}/{rows}")
    gb = bp.grad
    e = (diag - gb).abs()
    print(f"  dbias: maxerr {e.max().item():.4f} ref max {gb.abs().max().item():.3f} bad {int((e > 0.05).sum())}")
    if e.max().item() > 0.05:
        h0 = 0
        idx = torch.nonzero(e[h0] > 0.05).flatten()[:12].tolist()
        print("   bad diag idx", idx, [round(diag[h0, i].item(), 3) for i in idx], [round(gb[h0, i].item(), 3) for i in idx])


if __name__ == "__main__":
    if len(sys.argv) > 1:
        one(*[int(a) for a in sys.argv[1:5]])
    else:
        for c in CASES:
            print("case", c, flush=True)
            try:
                r = subprocess.run([sys.executable, __file__] + [str(x) for x in c], timeout=60,
                                   capture_output=True, text=True)
                print(r.stdout[-3000:], r.stderr[-1500:] if r.returncode else "", flush=True)
            except subprocess.TimeoutExpired:
                print("  TIMEOUT (hang)", flush=True)
