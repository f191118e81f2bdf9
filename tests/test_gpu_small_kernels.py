"""GPU parity: bandwidth kernels vs the oracle and the reference-generated golden vectors.
Everything goes through the ctypes C-ABI (pacednegatives_b200._lib)."""
import os

import numpy as np
import pytest
import torch

from oracle import weights as ow, loss as ol, select as osel, metrics as om

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pn():
    import pacednegatives_b200.weights as w
    import pacednegatives_b200.loss as l
    import pacednegatives_b200.dataloader as d
    import pacednegatives_b200.metrics as m
    from pacednegatives_b200 import _lib
    _lib.require_gpu()
    return type("PN", (), dict(w=w, l=l, d=d, m=m, lib=_lib))


def load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


# ------------------------------------------------------------------ weights (a4/a5/a6)
def test_weights_golden_bit_exact(pn, golden_dir):
    g = load(golden_dir, "weights_golden.npz")
    loss = torch.tensor(g["loss"]).cuda()
    for ei, eta in enumerate(g["etas"]):
        eta_t = torch.tensor([eta]).cuda()
        w, keep, stats = pn.w.spl_weights(loss, eta_t, 0.0, pn.w.LCE, True, True)
        np.testing.assert_array_equal(w.cpu().numpy(), g[f"lce_w_{ei}"])
        np.testing.assert_array_equal(keep.cpu().numpy().astype(bool), ~(g["loss"] > eta))
        st = stats.cpu().numpy()
        np.testing.assert_allclose(st[2] / len(g["loss"]), g[f"lce_dmean_deta_{ei}"][0], rtol=1e-5)
        # statistics: assembled from fp64 raw sums of l and l^2 (DESIGN: 1e-5 relative), not from rounded w
        np.testing.assert_allclose(st, ow.spl_stats(g["loss"], eta, 0.0, ow.LCE), rtol=1e-5)
        for mask in (0.0, 0.25):
            for name, var in (("eta", pn.w.ETA), ("lce", pn.w.LCE)):
                w, _, _ = pn.w.spl_weights(loss, eta_t, mask, var)
                np.testing.assert_array_equal(w.cpu().numpy(),
                                              g[f"nograd_{name}_{ei}_m{int(mask * 100)}"])


@pytest.mark.parametrize("n", [0, 1, 3, 4, 5, 1023, 1 << 20, (1 << 22) + 7])
@pytest.mark.parametrize("variant", [0, 1])
def test_weights_sizes_and_alignment(pn, n, variant):
    rng = np.random.RandomState(5678 + n % 97)
    loss = rng.exponential(0.7, size=n + 1).astype(np.float32)
    for off in (0, 1):  # off=1 -> 4-byte aligned only: scalar path
        l = torch.tensor(loss).cuda()[off:off + n]
        for eta in (0.3466, 1.0):
            w, keep, stats = pn.w.spl_weights(l, torch.tensor([eta]).cuda(), 0.5, variant, True, True)
            ew, ek = ow.spl_weights(loss[off:off + n], eta, 0.5, variant)
            np.testing.assert_array_equal(w.cpu().numpy(), ew)
            np.testing.assert_array_equal(keep.cpu().numpy().astype(bool), ek)
            if n:
                np.testing.assert_allclose(stats.cpu().numpy(),
                                           ow.spl_stats(loss[off:off + n], eta, 0.5, variant),
                                           rtol=1e-5, atol=1e-6 * max(1, n))


def test_weights_modules_autograd(pn):
    loss = torch.tensor([0.1, 0.5, 0.7, 0.3]).cuda()
    m = pn.w.LCEWeights({"eta": 0.5})
    w = m(loss)
    np.testing.assert_array_equal(w.detach().cpu().numpy(),
                                  np.array([0.2, 1.0, 0.0, 0.6], dtype=np.float32))
    (w * loss).mean().backward()
    assert abs(m.eta.grad.item() - (-0.35)) < 1e-6
    e = pn.w.EtaWeights(0.5, min=1e-10, max=1.0)
    e.set_mask(0.25)
    w2 = e(loss)
    ew, _ = ow.spl_weights(loss.cpu().numpy(), 0.5, 0.25, ow.ETA)
    np.testing.assert_array_equal(w2.detach().cpu().numpy(), ew)
    lreq = loss.clone().requires_grad_(True)
    e(lreq).sum().backward()
    # dw/dl: -1/eta below the threshold, mask*eta above
    np.testing.assert_allclose(lreq.grad.cpu().numpy(), [-2.0, -2.0, 0.125, -2.0], rtol=1e-6)
    e.eta.data.fill_(5.0)
    e.clamp()
    assert e.eta.item() == 1.0
    np.testing.assert_array_equal(e.no_grad(loss, torch.tensor([0.5]).cuda()).cpu().numpy(), ew)


def test_table_update_golden(pn, golden_dir):
    g = load(golden_dir, "table_update_golden.npz")
    B = len(g["init"])
    W = pn.w.Weights((2, B))
    W.weight.data[1] = torch.tensor(g["init"]).cuda()
    W.meta_update(1, torch.tensor(g["ce_n"]).cuda(), float(g["lr"]))
    np.testing.assert_allclose(W.weight[1].cpu().numpy(), g["updated"], rtol=1e-6, atol=1e-7)
    assert (W.weight[0] == 1).all()
    both = pn.w.Weights((1, B))
    both.meta_update(0, torch.tensor(g["ce_n"]).cuda(), 0.05, torch.tensor(g["ce_p"]).cuda())
    exp = ow.table_update(np.ones(B, np.float32), g["ce_p"], g["ce_n"], 1.0 / B, 0.05)
    np.testing.assert_allclose(both.weight[0].cpu().numpy(), exp, rtol=1e-6, atol=1e-7)


# ------------------------------------------------------------------ loss (a3)
def test_pair_loss_golden(pn, golden_dir):
    g = load(golden_dir, "loss_golden.npz")
    for c in range(int(g["n_cases"])):
        T, n, use_mean, Q, V = [int(x) for x in g[f"c{c}_meta"]]
        p = torch.tensor(g[f"c{c}_p"]).cuda()
        nl = torch.tensor(g[f"c{c}_n"]).cuda().view(-1, n, V)
        op = torch.tensor(g[f"c{c}_op"]).cuda()
        on = torch.tensor(g[f"c{c}_on"]).cuda()
        ce = pn.l.pair_loss(p, nl, op, on, bool(use_mean))
        assert ce.shape == (Q * T,)
        np.testing.assert_allclose(ce.cpu().numpy(), g[f"c{c}_pair_loss"], rtol=1e-5, atol=1e-5)
        fn = pn.l.init_LCEcrossentropy(ignore_index=-100, use_mean=bool(use_mean))
        np.testing.assert_allclose(fn(p, nl, op, on).item(), g[f"c{c}_lcece_mean"], rtol=1e-5)


@pytest.mark.parametrize("V,dtype", [(32128, torch.float32), (32128, torch.bfloat16),
                                     (1001, torch.float32), (37, torch.bfloat16)])
def test_token_ce_fwd_bwd(pn, V, dtype):
    rng = np.random.RandomState(V)
    rows = 24
    x = (rng.randn(rows, V) * 4).astype(np.float32)
    t = rng.randint(0, V, size=rows).astype(np.int64)
    t[3] = -100
    xt = torch.tensor(x).cuda().to(dtype).requires_grad_(True)
    xr = xt.detach().float().cpu().numpy()
    ce = pn.l.token_ce(xt, torch.tensor(t).cuda())
    ece, _ = ol.token_ce(xr, t)
    np.testing.assert_allclose(ce.detach().cpu().numpy(), ece, rtol=1e-5, atol=2e-5)
    gvec = rng.rand(rows).astype(np.float32)
    (ce * torch.tensor(gvec).cuda()).sum().backward()
    eg = ol.token_ce_grad(xr, t, gvec)
    tol = 1e-6 if dtype == torch.float32 else 4e-3
    np.testing.assert_allclose(xt.grad.float().cpu().numpy(), eg, atol=tol)


@pytest.mark.parametrize("G", [2, 8, 16, 32, 33])
def test_lce_restricted(pn, G):
    rng = np.random.RandomState(G)
    Q = 37
    z = (rng.randn(Q * G, 2) * 3).astype(np.float32)
    zt = torch.tensor(z).cuda().requires_grad_(True)
    loss, score = pn.l.lce_restricted(zt, G)
    el, es = ol.lce_restricted(z, G)
    np.testing.assert_allclose(loss.detach().cpu().numpy(), el, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(score.cpu().numpy(), es, rtol=1e-5, atol=1e-6)
    gq = rng.rand(Q).astype(np.float32)
    (loss * torch.tensor(gq).cuda()).sum().backward()
    np.testing.assert_allclose(zt.grad.cpu().numpy(), ol.lce_restricted_grad(z, G, gq), atol=2e-6)


def test_group_reduce_grad(pn):
    pce = torch.rand(6).cuda().requires_grad_(True)
    nce = torch.rand(6 * 7).cuda().requires_grad_(True)
    from pacednegatives_b200.loss import _GroupReduce
    for use_mean in (True, False):
        pce.grad = nce.grad = None
        out = _GroupReduce.apply(pce, nce, 7, use_mean)
        exp = ol.group_reduce(pce.detach().cpu().numpy(), nce.detach().cpu().numpy(), 7, use_mean)
        np.testing.assert_allclose(out.detach().cpu().numpy(), exp, rtol=1e-6)
        out.sum().backward()
        assert torch.allclose(pce.grad, torch.ones(6).cuda())
        assert torch.allclose(nce.grad, torch.full((42,), 1 / 7 if use_mean else 1.0).cuda())


# ------------------------------------------------------------------ level selection (a7)
def test_level_select_golden(pn, golden_dir):
    g = load(golden_dir, "select_golden.npz")
    cand = g["level_cand"]
    cm = pn.d.CandidateMatrix(cand)
    cr = cm.rank_major()
    Q = cand.shape[0]
    for use_max in (0, 1):
        for wi, w in enumerate(g["level_w64"]):
            for m in (cm, cr):
                _, doc = m.select_level(float(w), bool(use_max))
                np.testing.assert_array_equal(doc.cpu().numpy(), g[f"level_doc64_max{use_max}"][wi])
        for wi, w in enumerate(g["level_w32"]):
            wt = torch.full((Q,), float(w), dtype=torch.float32).cuda()
            for m in (cm, cr):
                _, doc = m.select_items(wt, bool(use_max))
                np.testing.assert_array_equal(doc.cpu().numpy(), g[f"level_doc32_max{use_max}"][wi])


def test_level_select_large_and_errors(pn):
    Q, C = 4099, 1000
    q = np.arange(Q, dtype=np.int64)[:, None]
    r = np.arange(C, dtype=np.int64)[None, :]
    cand = ((q * 7919 + r * 104729) % 8841823).astype(np.int32)
    cm = pn.d.CandidateMatrix(cand)
    cr = cm.rank_major()
    rng = np.random.RandomState(1)
    w = rng.rand(Q).astype(np.float32)
    w[:5] = [0.0, 1.0, 0.5, 1e-10, 0.999999]
    for use_ceil in (False, True):
        ei, ed = osel.level_select(cand, w.astype(np.float64), use_ceil)
        for m in (cm, cr):
            idx, doc = m.select_items(torch.tensor(w).cuda(), use_ceil)
            np.testing.assert_array_equal(idx.cpu().numpy(), ei)
            np.testing.assert_array_equal(doc.cpu().numpy(), ed)
    for level in range(1, 11):
        ei, ed = osel.level_select(cand, level / 10, False)
        for m in (cm, cr):
            idx, doc = m.select_level(level / 10)
            assert idx.item() == ei[0]
            np.testing.assert_array_equal(doc.cpu().numpy(), ed)
    with pytest.raises(IndexError):
        cm.select_level(1.5)
    idx, doc = cm.select_items(torch.tensor([2.0] * Q).cuda(), check=False)
    assert (idx == -1).all() and (doc == -1).all()


# ------------------------------------------------------------------ binomial selection (a8)
def test_binomial_table_matches_numpy(pn, golden_dir):
    g = load(golden_dir, "select_golden.npz")
    cm = pn.d.CandidateMatrix(g["binom_cand"])
    for mi, mean in enumerate(g["binom_means"]):
        for var in (0.01, 0.1):
            table, status = cm.binomial_table(mean, var, pmf=g[f"binom_pmf_{mi}"])
            p, cdf, nz = osel.binomial_table(mean, 1000, var, pmf=g[f"binom_pmf_{mi}"])
            t = table.cpu().numpy()
            np.testing.assert_array_equal(t[:1000], p)      # bit-exact fp64
            np.testing.assert_array_equal(t[1000:], cdf)
            assert status.cpu().tolist() == [nz, 0]
        # device lgamma pmf: close to scipy's, same support
        table, status = cm.binomial_table(mean, 0.01)
        p, _, nz = osel.binomial_table(mean, 1000, 0.01, pmf=g[f"binom_pmf_{mi}"])
        np.testing.assert_allclose(table.cpu().numpy()[:1000], p, rtol=1e-9, atol=1e-300)


def test_binomial_select_golden(pn, golden_dir):
    g = load(golden_dir, "select_golden.npz")
    cand = g["binom_cand"]
    cm = pn.d.CandidateMatrix(cand)
    seeds = g["binom_seeds"]
    for mi, mean in enumerate(g["binom_means"]):
        pmf = g[f"binom_pmf_{mi}"]
        for ki, k in enumerate(g["binom_ks"]):
            k = int(k)
            for var in (0.01, 0.1):
                table, _ = cm.binomial_table(mean, var, pmf=pmf)
                ref = g[f"binom_idx_m{mi}_k{ki}_v{int(var * 100)}"]
                # per-query substreams: one "query" per seed
                U = 64 * k
                u = np.stack([np.random.RandomState(int(s)).random_sample(U) for s in seeds])
                big = pn.d.CandidateMatrix(np.tile(cand[:1], (len(seeds), 1)))
                idx, docs, consumed, status = big.select_binomial(table, k, torch.tensor(u).cuda(), 0)
                assert status.item() == 0
                np.testing.assert_array_equal(idx.cpu().numpy(), ref)
                np.testing.assert_array_equal(docs.cpu().numpy(), cand[0][ref])
            # one global stream over the queries, as the reference's DataLoader would consume it
            table, _ = cm.binomial_table(mean, 0.01, pmf=pmf)
            u = np.random.RandomState(1234).random_sample(64 * k * len(cand))
            idx, docs, consumed, status = cm.select_binomial(table, k, torch.tensor(u).cuda(), 1)
            assert status.item() == 0
            np.testing.assert_array_equal(docs.cpu().numpy(), g[f"binom_seqdocs_m{mi}_k{ki}"])


def test_binomial_select_many_queries_vs_oracle(pn):
    pytest.importorskip("scipy")
    Q, C, k = 300, 1000, 15
    q = np.arange(Q, dtype=np.int64)[:, None]
    r = np.arange(C, dtype=np.int64)[None, :]
    cand = ((q * 7919 + r * 104729) % 8841823).astype(np.int32)
    cm = pn.d.CandidateMatrix(cand)
    rng = np.random.RandomState(77)
    u = rng.random_sample((Q, 40 * k))
    for mean in (0.05, 0.3466, 0.97):
        p, _, _ = osel.binomial_table(mean, C, 0.01)
        from scipy.stats import binom
        pmf = binom.pmf(np.arange(C), C - 1, mean)
        table, _ = cm.binomial_table(mean, 0.01, pmf=pmf)
        for layout_m in (cm, cm.rank_major()):
            idx, docs, consumed, status = layout_m.select_binomial(table, k, torch.tensor(u).cuda(), 0)
            ei, ed, ec = osel.binomial_select(cand, mean, 0.01, k, u, 0, pmf=pmf)
            np.testing.assert_array_equal(idx.cpu().numpy(), ei)
            np.testing.assert_array_equal(docs.cpu().numpy(), ed)
            np.testing.assert_array_equal(consumed.cpu().numpy(), ec)
    # philox stream: valid, distinct, deterministic, right distribution centre
    table, _ = cm.binomial_table(0.5, 0.01)
    a = cm.select_binomial(table, k, None, 0, seed=42)[0].cpu().numpy()
    b = cm.select_binomial(table, k, None, 0, seed=42)[0].cpu().numpy()
    c = cm.select_binomial(table, k, None, 0, seed=43)[0].cpu().numpy()
    np.testing.assert_array_equal(a, b)
    assert (a != c).any()
    assert all(len(set(row)) == k for row in a.tolist())
    assert abs(a.mean() - 499.5) < 3.0
    # running out of uniforms is reported, not silently wrong
    idx, _, consumed, status = cm.select_binomial(table, k, torch.zeros(Q, k, dtype=torch.float64).cuda(), 0)
    assert status.item() == Q and (idx == -1).all()


def test_lcedataset_dropin_global_rng(pn, golden_dir):
    """LCEDataset.sample consumes np.random exactly like the reference (same indices, same end state)."""
    pytest.importorskip("scipy")
    g = load(golden_dir, "select_golden.npz")
    cand = g["binom_cand"]

    class Corpus:
        def docs_iter(self):
            for x in sorted(set(cand.reshape(-1).tolist())):
                yield {"doc_id": str(x), "text": f"d{x}"}
            for qq in range(len(cand)):
                yield {"doc_id": f"p{qq}", "text": f"pos{qq}"}

        def queries_iter(self):
            for qq in range(len(cand)):
                yield {"query_id": str(qq), "text": f"q{qq}"}

    neg_idx = [[str(x) for x in row] for row in cand]
    pairs = [(str(qq), f"p{qq}") for qq in range(len(cand))]
    for mi, mean in enumerate(g["binom_means"][:5]):
        for ki, k in enumerate(g["binom_ks"]):
            ds = pn.d.LCEDataset(pairs, neg_idx, Corpus(), None, float(mean), 16, var=0.01, n=int(k),
                                 min=1e-10, max=1 - 1e-10)
            ref = g[f"binom_idx_m{mi}_k{ki}_v1"]
            for si, seed in enumerate(g["binom_seeds"]):
                np.random.seed(int(seed))
                np.testing.assert_array_equal(ds.sample(float(mean)), ref[si])
            np.random.seed(1234)
            docs = []
            for qq in range(len(cand)):
                item = ds[qq]
                assert item["pos"] == f"Query: q{qq} Document: pos{qq} Relevant:"
                docs.append([int(s.split(" Document: d")[1].split(" Relevant:")[0])
                             for s in item["neg"]])
            np.testing.assert_array_equal(np.array(docs), g[f"binom_seqdocs_m{mi}_k{ki}"])


def test_triplet_loaders_dropin(pn, golden_dir):
    g = load(golden_dir, "select_golden.npz")
    cand = g["level_cand"]
    Q = len(cand)

    class Corpus:
        def docs_iter(self):
            for i in range(5003):
                yield {"doc_id": str(i), "text": f"d{i}"}

        def queries_iter(self):
            for i in range(Q):
                yield {"query_id": str(i), "text": f"q{i}"}

    neg_idx = [[str(x) for x in row] for row in cand]
    pairs = [(str(q), str(q)) for q in range(Q)]
    for use_max in (False, True):
        ds = pn.d.TripletDataset(pairs, neg_idx, Corpus(), max=use_max)
        assert ds.n_neg == cand.shape[1] - 1 and len(ds) == Q
        for wi in (0, 2, 3, 9, 11):
            w = float(g["level_w64"][wi])
            for q in range(Q):
                qq, p, n = ds.get_items(q, w)
                assert (qq, p, n) == (f"q{q}", f"d{q}", f"d{g[f'level_doc64_max{int(use_max)}'][wi][q]}")
            lv = pn.d.LevelLoader(ds, 3).get_batch(1, w)
            assert lv[1] == [f"Query: q{q} Document: d{g[f'level_doc64_max{int(use_max)}'][wi][q]} Relevant:"
                             for q in range(3, 6)]
            assert lv[2] == ["true"] * 3 and lv[3] == ["false"] * 3
        w32 = torch.tensor(g["level_w32"][[5, 7, 9]])
        pl = pn.d.PairLoader(ds, 3).get_batch(0, w32)
        exp = [g[f"level_doc32_max{int(use_max)}"][wi][q] for q, wi in enumerate([5, 7, 9])]
        assert pl[1] == [f"Query: q{q} Document: d{d} Relevant:" for q, d in enumerate(exp)]


# ------------------------------------------------------------------ top-k + metrics (a12)
def test_topk_metrics_dense_relevance_and_many_queries(pn):
    """The sort-free path (K <= 32) when most documents are judged relevant (several counting passes), with more
    queries than resident CTAs, K = 32, C not a multiple of 4, and grades up to 255."""
    rng = np.random.RandomState(99)
    Q, C, K = 1500, 203, 32
    scores = np.round(rng.randn(Q, C).astype(np.float32), 1)            # heavy ties
    docids = np.stack([rng.permutation(5000)[:C] for _ in range(Q)]).astype(np.int32)
    rels = rng.choice([0, 1, 2, 3, 255], size=(Q, C), p=[.2, .4, .2, .15, .05]).astype(np.uint8)
    rels[7] = 1        # every document relevant
    rels[8] = 0
    out = pn.m.topk_metrics(torch.tensor(scores).cuda(), torch.tensor(docids).cuda(), torch.tensor(rels).cuda(), K)
    eids, esc, em = om.batch_metrics(scores, docids, rels, K)
    np.testing.assert_array_equal(out["top_ids"].cpu().numpy(), eids)
    np.testing.assert_array_equal(out["top_scores"].cpu().numpy(), esc)
    np.testing.assert_allclose(out["metrics"].cpu().numpy(), em, rtol=1e-6, atol=1e-7)
    # both paths agree on the same input (K = 33 takes the bitonic sort)
    out33 = pn.m.topk_metrics(torch.tensor(scores).cuda(), torch.tensor(docids).cuda(), torch.tensor(rels).cuda(), 33)
    np.testing.assert_array_equal(out33["top_ids"].cpu().numpy()[:, :32], eids)
    np.testing.assert_allclose(out33["metrics"].cpu().numpy()[:, [1, 3]], em[:, [1, 3]], rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("Q,C,K", [(200, 1000, 10), (3, 7, 10), (5, 1024, 10), (2, 4096, 100), (1, 1, 1), (4, 4096, 10)])
def test_topk_metrics(pn, Q, C, K):
    rng = np.random.RandomState(3456 + C)
    scores = rng.randn(Q, C).astype(np.float32)
    scores[:, : C // 3] = np.round(scores[:, : C // 3], 1)  # plenty of ties
    docids = np.stack([rng.permutation(8841823 if C > 10 else 50)[:C] for _ in range(Q)]).astype(np.int32)
    rels = rng.choice(4, size=(Q, C), p=[.95, .03, .015, .005]).astype(np.uint8)
    if Q > 1:
        rels[1] = 0  # a query with no relevant document
    out = pn.m.topk_metrics(torch.tensor(scores).cuda(), torch.tensor(docids).cuda(),
                            torch.tensor(rels).cuda(), K)
    eids, esc, em = om.batch_metrics(scores, docids, rels, K)
    np.testing.assert_array_equal(out["top_ids"].cpu().numpy(), eids)   # bit-exact ranking
    np.testing.assert_array_equal(out["top_scores"].cpu().numpy(), esc)
    np.testing.assert_allclose(out["metrics"].cpu().numpy(), em, rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("C", [1000, 999, 1024, 64])
def test_topk_metrics_warp_kernel_paths(pn, C):
    """Every branch of the second-generation warp kernel (csrc/pn_topk_metrics.cu, topk_warp2_kernel): the number of
    relevant documents on both sides of the direct / sorted-list / too-many thresholds (12 | 13, 256 | 257), none and
    all; the tau filter with few survivors (scores falling with the position, either lane layout), 33..64 survivors and
    the fall-back to the K sweeps (the best keys crowded into one lane's slots); vector (C % 4 == 0, aligned) and scalar
    loads (odd C, rows starting off a 16-byte boundary); no docids.  Bit-exact ranking against the oracle."""
    rng = np.random.RandomState(77 + C)
    K = 10
    rows = []
    def lane_slots(vec):      # candidate positions in the order lane 0's slots, lane 1's slots, ...
        if vec:
            p = [4 * (lane + 32 * (j >> 2)) + (j & 3) for lane in range(32) for j in range(32)]
        else:
            p = [lane + 32 * j for lane in range(32) for j in range(32)]
        return [x for x in p if x < C]
    for nrel in [0, 1, 12, 13, 40, 256, 257, C]:
        nrel = min(nrel, C)
        s = rng.randn(C).astype(np.float32)
        s[: C // 4] = np.round(s[: C // 4], 1)
        r = np.zeros(C, np.uint8)
        r[rng.permutation(C)[:nrel]] = rng.randint(1, 4, size=nrel)
        rows.append((s, r))
    for crowd in ["falling", "vec", "scalar"]:
        s = np.empty(C, np.float32)
        order = np.arange(C) if crowd == "falling" else np.array(lane_slots(crowd == "vec"))
        s[order] = np.linspace(5.0, -5.0, C, dtype=np.float32) + (rng.randn(C) * 1e-4).astype(np.float32)
        r = (rng.rand(C) < 0.05).astype(np.uint8) * 2
        rows.append((s, r))
    scores = np.stack([x for x, _ in rows])
    rels = np.stack([x for _, x in rows])
    Q = len(rows)
    docids = np.stack([rng.permutation(8841823)[:C] for _ in range(Q)]).astype(np.int32)
    for ids in (docids, None):
        eids, esc, em = om.batch_metrics(scores, docids if ids is not None else np.tile(np.arange(C), (Q, 1)), rels, K)
        out = pn.m.topk_metrics(torch.tensor(scores).cuda(), None if ids is None else torch.tensor(ids).cuda(),
                                torch.tensor(rels).cuda(), K)
        np.testing.assert_array_equal(out["top_ids"].cpu().numpy(), eids)
        np.testing.assert_array_equal(out["top_scores"].cpu().numpy(), esc)
        np.testing.assert_allclose(out["metrics"].cpu().numpy(), em, rtol=1e-6, atol=1e-7)
    # rows that start off a 16-byte boundary (a view into a larger buffer): the scalar-load path, same results
    flat_s = torch.zeros(Q * C + 1).cuda()
    flat_d = torch.zeros(Q * C + 1, dtype=torch.int32).cuda()
    flat_r = torch.zeros(Q * C + 1, dtype=torch.uint8).cuda()
    flat_s[1:] = torch.tensor(scores).cuda().reshape(-1)
    flat_d[1:] = torch.tensor(docids).cuda().reshape(-1)
    flat_r[1:] = torch.tensor(rels).cuda().reshape(-1)
    eids, esc, em = om.batch_metrics(scores, docids, rels, K)
    out = pn.m.topk_metrics(flat_s[1:].view(Q, C), flat_d[1:].view(Q, C), flat_r[1:].view(Q, C), K)
    np.testing.assert_array_equal(out["top_ids"].cpu().numpy(), eids)
    np.testing.assert_allclose(out["metrics"].cpu().numpy(), em, rtol=1e-6, atol=1e-7)
    # no relevance labels: ranking only
    out = pn.m.topk_metrics(torch.tensor(scores).cuda(), torch.tensor(docids).cuda(), None, 32)
    eids32, _, _ = om.batch_metrics(scores, docids, np.zeros_like(rels), 32)
    np.testing.assert_array_equal(out["top_ids"].cpu().numpy(), eids32)


def test_topk_metrics_with_qrel_totals(pn):
    """AP / nDCG denominators from the queries' full qrels (relevant documents missing from the candidate list):
    hand case of tests/test_oracle_golden.py + a random batch against the oracle, both kernel variants (K <= 32 and
    the sort path)."""
    from pacednegatives_b200.metrics import qrel_totals, docno_order_ids
    sc = torch.tensor([[0.5, 0.9, 0.5, 0.1]]).cuda()
    ids = torch.tensor([[10, 11, 12, 13]], dtype=torch.int32).cuda()
    rels = torch.tensor([[1, 0, 2, 0]], dtype=torch.uint8).cuda()
    num_rel, idcg = qrel_totals([[1, 0, 2, 0, 3]], K=3)
    out = pn.m.topk_metrics(sc, ids, rels, 3, num_rel=num_rel, idcg=idcg)
    dcg = 2 / np.log2(3) + 1 / np.log2(4)
    assert abs(out["ndcg"].item() - dcg / idcg[0]) < 1e-6
    assert abs(out["ap"].item() - (1 / 2 + 2 / 3) / 3) < 1e-6 and abs(out["rr"].item() - 0.5) < 1e-7
    rng = np.random.RandomState(3)
    Q, C = 37, 200
    scores = rng.randn(Q, C).astype(np.float32)
    docids = np.stack([rng.permutation(5000)[:C] for _ in range(Q)]).astype(np.int32)
    rl = rng.choice(4, size=(Q, C), p=[.9, .05, .03, .02]).astype(np.uint8)
    missing = [rng.choice(4, size=rng.randint(0, 6), p=[.3, .3, .2, .2]) for _ in range(Q)]   # judged, not retrieved
    for K in (10, 40):
        nr, ig = qrel_totals([np.concatenate([rl[q], missing[q]]) for q in range(Q)], K)
        enr, eig = om.qrel_totals([np.concatenate([rl[q], missing[q]]) for q in range(Q)], K)
        assert np.array_equal(nr, enr) and np.array_equal(ig, eig)
        out = pn.m.topk_metrics(torch.tensor(scores).cuda(), torch.tensor(docids).cuda(), torch.tensor(rl).cuda(), K,
                                num_rel=nr, idcg=ig)
        eids, esc, em = om.batch_metrics(scores, docids, rl, K, num_rel=enr, idcg=eig)
        np.testing.assert_array_equal(out["top_ids"].cpu().numpy(), eids)
        np.testing.assert_allclose(out["metrics"].cpu().numpy(), em, rtol=1e-6, atol=1e-7)
    # docno strings -> ids with the strings' order (trec_eval's tie-break is on the docno string)
    d = docno_order_ids([["D10", "D9", "D100"], ["a", "c", "b"]])
    assert d[0].tolist() == [0, 2, 1] or (d[0, 1] > d[0, 0] > d[0, 2])     # "D9" > "D10" > "D100" as strings
    assert d[1, 1] > d[1, 2] > d[1, 0]


def test_argument_errors(pn):
    with pytest.raises(ValueError):
        pn.lib.check(pn.lib.lib().pn_weights_fwd(None, -1, None, 0.0, 0, None, None, None, None))
    assert b"bad argument" in pn.lib.lib().pn_last_error()
    with pytest.raises(ValueError):
        pn.m.topk_metrics(torch.zeros(1, 5000).cuda())


def _ragged(rng, n, lo, hi, vocab=32000):
    return [rng.randint(2, vocab, size=rng.randint(lo, hi + 1)).tolist() for _ in range(n)]


@pytest.mark.parametrize("n", [0, 1, 37, 3000])
def test_pack_prompts_vs_oracle(n):
    """pn_pack_prompts: bit-exact against the oracle for ragged stores, pad-to-longest and fixed
    width, document truncation, and out-of-range indices."""
    from oracle import select as osel
    from pacednegatives_b200.dataloader import PromptPacker, TokenStore
    rng = np.random.RandomState(n + 5)
    q_tokens, d_tokens = _ragged(rng, 50, 0, 12), _ragged(rng, 400, 0, 90)
    prefix, middle, suffix = [3, 4], [9], [7, 8, 1]
    qs, ds = TokenStore(q_tokens), TokenStore(d_tokens)
    pk = PromptPacker(prefix=prefix, middle=middle, suffix=suffix, pad_id=0)
    q_idx = rng.randint(0, 50, size=n).astype(np.int32)
    d_idx = rng.randint(0, 400, size=n).astype(np.int32)
    tq, td = torch.tensor(q_idx).cuda(), torch.tensor(d_idx).cuda()
    if n == 0:
        ids, lens = pk.pack(qs, ds, tq, td, S=16)
        assert ids.shape == (0, 16) and lens.numel() == 0
        return
    ids, lens = pk.pack(qs, ds, tq, td)                                   # pad to the longest row
    eids, elens = osel.pack_prompts(prefix, middle, suffix, q_tokens, d_tokens, q_idx, d_idx)
    np.testing.assert_array_equal(ids.cpu().numpy(), eids)
    np.testing.assert_array_equal(lens.cpu().numpy(), elens)
    ids, lens = pk.pack(qs, ds, tq, td, S=128)                            # fixed width, everything fits
    eids, _ = osel.pack_prompts(prefix, middle, suffix, q_tokens, d_tokens, q_idx, d_idx, S=128)
    np.testing.assert_array_equal(ids.cpu().numpy(), eids)
    pk.check()
    ids, lens = pk.pack(qs, ds, tq, td, S=64, max_len=40)                 # truncated documents
    eids, elens = osel.pack_prompts(prefix, middle, suffix, q_tokens, d_tokens, q_idx, d_idx, S=64, max_len=40)
    np.testing.assert_array_equal(ids.cpu().numpy(), eids)
    np.testing.assert_array_equal(lens.cpu().numpy(), elens)
    assert int(lens.max()) <= 40
    if n >= 37:
        ids, lens = pk.pack(qs, ds, tq, td, S=48)                         # rows longer than S are flagged, all pad
        eids, elens = osel.pack_prompts(prefix, middle, suffix, q_tokens, d_tokens, q_idx, d_idx, S=48)
        np.testing.assert_array_equal(ids.cpu().numpy(), eids)
        assert (elens > 48).any()
        with pytest.raises(ValueError):
            pk.check()
        bad = td.clone()
        bad[3] = 400
        with pytest.raises(IndexError):
            pk.pack(qs, ds, tq, bad)


def test_pack_prompts_feeds_the_reference_collate_batches(tmp_path):
    """End to end against the reference: the positives' prompts of tests/golden/datamodule_golden.npz, packed on the
    GPU from token stores, equal the ids LCEDataModule.collate produced (lceT5.py:38-54)."""
    from test_oracle_golden import _datamodule_positive_case
    from pacednegatives_b200.dataloader import PromptPacker, TokenStore
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "datamodule_golden.npz"))
    (prefix, middle, suffix), q_tokens, d_tokens, q_idx, d_idx = _datamodule_positive_case(tmp_path)
    qs, ds = TokenStore(q_tokens), TokenStore(d_tokens)
    pk = PromptPacker(prefix=prefix, middle=middle, suffix=suffix)
    for i in range(int(g["batches"])):
        ids, _ = pk.pack(qs, ds, torch.tensor(q_idx[4 * i:4 * i + 4]), torch.tensor(d_idx[4 * i:4 * i + 4]))
        np.testing.assert_array_equal(ids.cpu().numpy(), g[f"p{i}"])


def test_empty_inputs_are_no_ops(pn):
    """Zero-length batches go through every bandwidth entry point without a launch error and with empty results
    (the reference's python loops simply do not execute on an empty batch)."""
    from pacednegatives_b200.loss import check_success, token_ce, lce_restricted
    dev = "cuda"
    e = torch.zeros(0, device=dev)
    eta = torch.tensor([0.5], device=dev)
    w, keep, stats = pn.w.spl_weights(e, eta, 0.0, 0, want_keep=True, want_stats=True)
    assert w.numel() == 0 and keep.numel() == 0 and float(stats.abs().sum()) == 0.0
    ce = token_ce(torch.zeros(0, 32, device=dev), torch.zeros(0, dtype=torch.int64, device=dev))
    assert ce.numel() == 0
    loss, sc = lce_restricted(torch.zeros(0, 2, device=dev), 4)
    assert loss.numel() == 0 and sc.numel() == 0
    assert check_success(torch.zeros(0, 1, 8, device=dev), torch.zeros(0, 1, 8, device=dev), 1, 2).item() == 0.0
    out = pn.m.topk_metrics(torch.zeros(0, 10, device=dev), None, None, 5)
    assert out["top_ids"].shape == (0, 5)
    cm = pn.d.CandidateMatrix(np.zeros((0, 7), dtype=np.int32), pn.d.ROW, Q=0, C=7)
    idx, doc = cm.select_level(0.5)
    assert idx.numel() == 1 and doc.numel() == 0
