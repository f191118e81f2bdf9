"""Oracle: ranking + trec_eval measures (test infrastructure only; see oracle/__init__.py).

The reference evaluates through PyTerrier -> pytrec_eval (eval.py:26,41-43, eval_metrics =
["map", "ndcg_cut_10", "recip_rank"]); both are third-party, unpinned and absent from the reference
tree and from this image (no JVM).  This restates trec_eval's published definitions:
  ranking     score descending, ties broken by docno descending
  ndcg_cut_K  gain = rel, discount log2(rank + 1), ideal = the query's judged rels sorted descending
  recip_rank  1 / rank of the first document with rel >= 1
  map         average precision over documents with rel >= 1
Parity unpinned against trec_eval itself (not installable offline); pinned by hand-computed cases in
tests/test_oracle_golden.py.
"""
import numpy as np


def rank_order(scores, docids):
    """argsort by (score desc, docid desc)."""
    scores = np.asarray(scores, dtype=np.float32)
    docids = np.asarray(docids, dtype=np.int64)
    return np.lexsort((-docids, -scores.astype(np.float64)))


def query_metrics(scores, docids, rels, K=10, num_rel=None, idcg=None):
    """Returns (top_ids[K], top_scores[K], ndcg@K, rr, rr@K, ap) for one query.  ``num_rel`` / ``idcg``: the
    query's totals over ALL judged documents (trec_eval's denominators; a relevant document the first stage missed
    still counts); None = every judged-relevant document is among the candidates."""
    scores = np.asarray(scores, dtype=np.float32)
    docids = np.asarray(docids, dtype=np.int64)
    rels = np.asarray(rels, dtype=np.int64)
    order = rank_order(scores, docids)
    r = rels[order]
    C = len(scores)
    top = order[:K]
    disc = 1.0 / np.log2(np.arange(2, K + 2, dtype=np.float64))
    dcg = float((r[:K] * disc[:min(K, C)]).sum())
    ideal = np.sort(rels[rels > 0])[::-1][:K]
    if idcg is None:
        idcg = float((ideal * disc[:len(ideal)]).sum())
    ndcg = dcg / idcg if idcg > 0 else 0.0
    hits = np.nonzero(r >= 1)[0]
    rr = 1.0 / (hits[0] + 1) if len(hits) else 0.0
    rrk = rr if len(hits) and hits[0] < K else 0.0
    nrel = int((rels >= 1).sum()) if num_rel is None else int(num_rel)
    ap = float(((np.arange(1, len(hits) + 1)) / (hits + 1.0)).sum() / nrel) if nrel else 0.0
    return docids[top], scores[top], ndcg, rr, rrk, ap


def qrel_totals(judged_grades, K=10):
    """(num_rel [Q], idcg [Q]) over each query's full list of judged grades (trec_eval: relevant = grade >= 1)."""
    disc = 1.0 / np.log2(np.arange(2, K + 2, dtype=np.float64))
    num, idcg = [], []
    for g in judged_grades:
        g = np.asarray(g, dtype=np.int64)
        pos = np.sort(g[g > 0])[::-1][:K]
        num.append(int((g >= 1).sum()))
        idcg.append(float((pos * disc[:len(pos)]).sum()))
    return np.asarray(num, dtype=np.int32), np.asarray(idcg, dtype=np.float64)


def batch_metrics(scores, docids, rels, K=10, num_rel=None, idcg=None):
    """scores/docids/rels [Q, C] -> (top_ids [Q,K] int32, top_scores [Q,K] f32, metrics [Q,4] f32)."""
    Q, C = np.asarray(scores).shape
    ids = np.full((Q, K), -1, dtype=np.int32)
    sc = np.zeros((Q, K), dtype=np.float32)
    m = np.zeros((Q, 4), dtype=np.float32)
    for q in range(Q):
        t, s, a, b, c, d = query_metrics(scores[q], docids[q], rels[q], K,
                                         None if num_rel is None else num_rel[q],
                                         None if idcg is None else idcg[q])
        ids[q, :len(t)] = t
        sc[q, :len(s)] = s
        m[q] = (a, b, c, d)
    return ids, sc, m
