"""CPU: the oracle restatements replayed against vectors produced by the reference itself
(tests/golden/make_golden.py) and against hand-checked anchors from SURVEY §4/§8c."""
import os

import numpy as np
import pytest

from oracle import weights as ow, loss as ol, select as osel, metrics as om, t5 as ot


def load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def test_weights_hand_anchor():
    # eta=.5, loss=[.1,.5,.7,.3] -> [.2,1,0,.6]; d mean(w*l)/d eta = -0.35
    w, keep = ow.spl_weights([0.1, 0.5, 0.7, 0.3], 0.5)
    np.testing.assert_array_equal(w, np.array([0.2, 1.0, 0.0, 0.6], dtype=np.float32))
    assert keep.tolist() == [True, True, False, True]
    st = ow.spl_stats([0.1, 0.5, 0.7, 0.3], 0.5)
    assert abs(st[2] / 4 - (-0.35)) < 1e-6


def test_weights_vs_reference(golden_dir):
    g = load(golden_dir, "weights_golden.npz")
    loss, etas = g["loss"], g["etas"]
    for ei, eta in enumerate(etas):
        w, _ = ow.spl_weights(loss, eta, 0.0, ow.LCE)
        np.testing.assert_array_equal(w, g[f"lce_w_{ei}"])  # bit-exact
        st = ow.spl_stats(loss, eta, 0.0, ow.LCE)
        np.testing.assert_allclose(st[2] / len(loss), g[f"lce_dmean_deta_{ei}"][0], rtol=1e-5)
        for mask in (0.0, 0.25):
            for name, var in (("eta", ow.ETA), ("lce", ow.LCE)):
                w, _ = ow.spl_weights(loss, eta, mask, var)
                np.testing.assert_array_equal(w, g[f"nograd_{name}_{ei}_m{int(mask * 100)}"])


def test_table_update_vs_reference(golden_dir):
    g = load(golden_dir, "table_update_golden.npz")
    B = len(g["init"])
    # NewWrapper differentiates only the negative term (two separate `v` leaves, pairwrapper.py:225-235)
    got = ow.table_update(g["init"], None, g["ce_n"], 1.0 / B, float(g["lr"]))
    np.testing.assert_allclose(got, g["updated"], rtol=1e-6, atol=1e-7)
    ce_p, _ = ol.token_ce(g["logits_p"], g["o_p"])
    np.testing.assert_allclose(ce_p, g["ce_p"], rtol=1e-5, atol=1e-6)


def test_loss_vs_reference(golden_dir):
    g = load(golden_dir, "loss_golden.npz")
    for c in range(int(g["n_cases"])):
        T, n, use_mean, Q, V = g[f"c{c}_meta"]
        nview = g[f"c{c}_n"].reshape(-1, n, V)
        ce = ol.pair_loss(g[f"c{c}_p"], nview, g[f"c{c}_op"], g[f"c{c}_on"], bool(use_mean))
        assert ce.shape == (Q * T,)
        np.testing.assert_allclose(ce, g[f"c{c}_pair_loss"], rtol=1e-5, atol=1e-6)
        m = ol.lce_crossentropy(g[f"c{c}_p"], nview, g[f"c{c}_op"], g[f"c{c}_on"], bool(use_mean))
        np.testing.assert_allclose(m, g[f"c{c}_lcece_mean"], rtol=1e-5)


def test_level_select_vs_reference(golden_dir):
    g = load(golden_dir, "select_golden.npz")
    cand = g["level_cand"]
    for use_max in (0, 1):
        for wi, w in enumerate(g["level_w64"]):
            _, doc = osel.level_select(cand, float(w), bool(use_max))
            np.testing.assert_array_equal(doc, g[f"level_doc64_max{use_max}"][wi])
        for wi, w in enumerate(g["level_w32"]):
            _, doc = osel.level_select(cand, float(np.float32(w)), bool(use_max))
            np.testing.assert_array_equal(doc, g[f"level_doc32_max{use_max}"][wi])


def test_binomial_select_vs_reference(golden_dir):
    g = load(golden_dir, "select_golden.npz")
    cand = g["binom_cand"]
    C = cand.shape[1]
    for mi, mean in enumerate(g["binom_means"]):
        pmf = g[f"binom_pmf_{mi}"]
        for ki, k in enumerate(g["binom_ks"]):
            for var in (0.01, 0.1):
                p, _, nz = osel.binomial_table(mean, C, var, pmf=pmf)
                assert nz >= k
                ref = g[f"binom_idx_m{mi}_k{ki}_v{int(var * 100)}"]
                for si, seed in enumerate(g["binom_seeds"]):
                    u = np.random.RandomState(int(seed)).random_sample(64 * k)
                    got, _ = osel.legacy_choice(p, int(k), u)
                    np.testing.assert_array_equal(got, ref[si])
            # one global stream across queries (the DataLoader order of the reference)
            u = np.random.RandomState(1234).random_sample(64 * k * len(cand))
            _, doc, _ = osel.binomial_select(cand, mean, 0.01, int(k), u, stream_mode=1, pmf=pmf)
            np.testing.assert_array_equal(doc, g[f"binom_seqdocs_m{mi}_k{ki}"])


def test_binomial_anchor():
    # SURVEY §8c anchor: seed 0, mean .5, C=1000, k=7 -> [501,508,504,496,505,497,520]
    pytest.importorskip("scipy")
    p, _, _ = osel.binomial_table(0.5, 1000, 0.01)
    u = np.random.RandomState(0).random_sample(128)
    got, _ = osel.legacy_choice(p, 7, u)
    assert got.tolist() == [501, 508, 504, 496, 505, 497, 520]
    p2, _, _ = osel.binomial_table(0.5, 1000, 0.1)
    got2, _ = osel.legacy_choice(p2, 7, u)
    assert got2.tolist() == got.tolist()  # var is a no-op


def test_pairwise_sum_matches_numpy():
    rng = np.random.RandomState(3)
    for n in (5, 8, 100, 129, 1000, 1001):
        a = rng.rand(n) * rng.choice([1e-9, 1.0, 1e6], n)
        assert osel.np_pairwise_sum(a) == np.sum(a)


def test_metrics_hand_cases():
    # ranking 3 docs: scores tie broken by docid desc
    ids, sc, m = om.batch_metrics(np.array([[0.5, 0.9, 0.5, 0.1]], np.float32),
                                  np.array([[10, 11, 12, 13]]), np.array([[1, 0, 2, 0]]), K=3)
    assert ids[0].tolist() == [11, 12, 10]
    dcg = 0 / np.log2(2) + 2 / np.log2(3) + 1 / np.log2(4)
    idcg = 2 / np.log2(2) + 1 / np.log2(3)
    assert abs(m[0, 0] - dcg / idcg) < 1e-6
    assert abs(m[0, 1] - 0.5) < 1e-7 and abs(m[0, 2] - 0.5) < 1e-7
    assert abs(m[0, 3] - (1 / 2 + 2 / 3) / 2) < 1e-6
    # no relevant docs
    _, _, m = om.batch_metrics(np.array([[0.3, 0.2]], np.float32), np.array([[1, 2]]),
                               np.array([[0, 0]]), K=10)
    assert m[0].tolist() == [0, 0, 0, 0]


def test_metrics_with_a_relevant_document_missing_from_the_candidates():
    """trec_eval judges against ALL judged documents of the query (eval.py:26: pt.Experiment over the qrels): a
    relevant document BM25 did not retrieve lowers AP and nDCG.  Hand case: candidates (grades 1, 0, 2, 0 ranked
    0.9 > 0.5(id 12) > 0.5(id 10) > 0.1), qrels also hold an unretrieved grade-3 document."""
    sc = np.array([[0.5, 0.9, 0.5, 0.1]], np.float32)
    ids = np.array([[10, 11, 12, 13]])
    rels = np.array([[1, 0, 2, 0]])
    num_rel, idcg = om.qrel_totals([[1, 0, 2, 0, 3]], K=3)
    assert num_rel.tolist() == [3] and abs(idcg[0] - (3 / np.log2(2) + 2 / np.log2(3) + 1 / np.log2(4))) < 1e-12
    _, _, m = om.batch_metrics(sc, ids, rels, K=3, num_rel=num_rel, idcg=idcg)
    dcg = 0 / np.log2(2) + 2 / np.log2(3) + 1 / np.log2(4)
    assert abs(m[0, 0] - dcg / idcg[0]) < 1e-6                    # nDCG@3 against the full ideal ranking
    assert abs(m[0, 3] - (1 / 2 + 2 / 3) / 3) < 1e-6              # AP over 3 judged-relevant documents, not 2
    assert abs(m[0, 1] - 0.5) < 1e-7                              # RR does not depend on the totals
    _, _, m0 = om.batch_metrics(sc, ids, rels, K=3)
    assert m0[0, 0] > m[0, 0] and m0[0, 3] > m[0, 3]              # candidate-only totals inflate both


def test_t5_oracle_vs_hf_golden(golden_dir):
    g = load(golden_dir, "t5_tiny_golden.npz")
    v, d, kv, ff, L, Ld, H = [int(x) for x in g["dims"]]
    dims = ot.T5Dims(vocab=v, d_model=d, d_kv=kv, d_ff=ff, n_layers=L, n_dec_layers=Ld, n_heads=H)
    w = ot.init_weights(dims, seed=int(g["seed"]))
    pl = ot.logits(w, dims, g["pids"], g["plab"])
    nl = ot.logits(w, dims, g["nids"], g["nlab"])
    np.testing.assert_allclose(pl, g["plogits"], rtol=1e-4, atol=1e-4)
    np.testing.assert_allclose(nl, g["nlogits"], rtol=1e-4, atol=1e-4)
    k = int(g["k"])
    ce = ol.pair_loss(pl, nl.reshape(-1, k, v), g["plab"], g["nlab"], True)
    np.testing.assert_allclose(ce, g["pair_loss"], rtol=1e-5, atol=1e-5)
    s, z2 = ot.score(w, dims, g["pids"], int(g["rel_ids"][0]), int(g["rel_ids"][1]))
    np.testing.assert_allclose(z2, g["p1_z2"], rtol=1e-4, atol=1e-4)
    np.testing.assert_allclose(s, g["p1_score"], rtol=1e-4, atol=1e-5)


def test_relative_bucket_vs_hf():
    tr = pytest.importorskip("transformers.models.t5.modeling_t5")
    import torch
    rel = torch.arange(-600, 601)
    hb = tr.T5Attention._relative_position_bucket(rel, bidirectional=True).numpy()
    np.testing.assert_array_equal(ot.relative_bucket(rel.numpy(), True), hb)


def _datamodule_positive_case(tmp_path):
    """Pieces of the prompts in tests/golden/datamodule_golden.npz (written by the reference's LCEDataModule)."""
    import json
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    import make_golden as mg
    path, corpus = mg.datamodule_case(str(tmp_path))
    tok = mg.fake_tokenizer()
    enc = lambda s: tok([s])["input_ids"][0].tolist()      # ids + </s>
    docs = list(corpus.docs_iter())
    queries = list(corpus.queries_iter())
    d_tokens = [enc(d["text"])[:-1] for d in docs]
    q_tokens = [enc(q["text"])[:-1] for q in queries]
    with open(path) as f:
        recs = json.load(f)
    q_idx = [int(r["query_id"]) for r in recs]
    d_idx = [int(r["doc_id_a"]) for r in recs]
    pieces = (enc("Query:")[:-1], enc("Document:")[:-1], enc("Relevant:"))
    return pieces, q_tokens, d_tokens, q_idx, d_idx


def test_pack_prompts_oracle_vs_reference_collate(tmp_path):
    """oracle.select.pack_prompts == tokenizer(format(q, d), padding=True) as the reference's LCEDataModule.collate
    produced it for the positives of every batch (lceT5.py:38-54 through dataloader.py:153-162)."""
    from oracle import select as osel
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "datamodule_golden.npz"))
    (prefix, middle, suffix), q_tokens, d_tokens, q_idx, d_idx = _datamodule_positive_case(tmp_path)
    B = 4
    for i in range(int(g["batches"])):
        ids, lens = osel.pack_prompts(prefix, middle, suffix, q_tokens, d_tokens, q_idx[i * B:(i + 1) * B],
                                      d_idx[i * B:(i + 1) * B])
        np.testing.assert_array_equal(ids, g[f"p{i}"])
        assert lens.max() == g[f"p{i}"].shape[1]
    # truncation keeps the suffix and only shortens the document
    ids, lens = osel.pack_prompts(prefix, middle, suffix, q_tokens, d_tokens, q_idx[:4], d_idx[:4], S=16, max_len=12)
    assert (lens <= 12).all() and ids.shape == (4, 16)
    for r, n in zip(ids, lens):
        assert list(r[n - len(suffix):n]) == list(suffix) and (r[n:] == 0).all()


@pytest.mark.parametrize("S", [128, 200, 384])
def test_oracle_attention_mask_vs_hf_golden(golden_dir, S):
    """oracle.t5.score(attention_mask=) against HF run with the tokenizer-style mask (gold_t5_mask)."""
    g = np.load(os.path.join(golden_dir, "t5_mask_golden.npz"))
    v, d, kv, ff, L, Ld, H = [int(x) for x in g["dims"]]
    dims = ot.T5Dims(vocab=v, d_model=d, d_kv=kv, d_ff=ff, n_layers=L, n_dec_layers=Ld, n_heads=H)
    w = ot.init_weights(dims, seed=int(g["seed"]))
    ids, lens = g[f"ids_{S}"], g[f"lens_{S}"]
    mask = (np.arange(S)[None, :] < lens[:, None]).astype(np.int64)
    rel, nrel = [int(x) for x in g["rel_ids"]]
    s, z2 = ot.score(w, dims, ids, rel, nrel, attention_mask=mask)
    np.testing.assert_allclose(z2, g[f"z2_{S}"], atol=2e-5)
    np.testing.assert_allclose(s, g[f"score_{S}"], atol=2e-5)
    _, z0 = ot.score(w, dims, ids, rel, nrel)
    np.testing.assert_allclose(z0, g[f"z2_nomask_{S}"], atol=2e-5)
