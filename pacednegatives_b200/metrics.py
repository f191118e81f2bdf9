"""Batched top-k re-ranking + nDCG@10 / MRR@10 / MAP on the B200 — the measure side of eval.py:26,41-43
(pt.Experiment with eval_metrics=["map", "ndcg_cut_10", "recip_rank"]).  One launch for all queries
(csrc/pn_topk_metrics.cu); ranking rule score desc, docid desc (trec_eval breaks score ties by docno descending:
``docno_order_ids`` gives integer ids with the docno strings' order)."""
import numpy as np
import torch

from . import _lib


def qrel_totals(judged_grades, K=10):
    """Per-query totals over ALL judged documents of the qrels, retrieved or not — what trec_eval divides by.
    judged_grades: one sequence of integer grades per query.  Returns (num_rel int32 [Q], idcg float64 [Q]) for
    ``topk_metrics(..., num_rel=, idcg=)``: number of grades >= 1 and the ideal DCG@K (linear gain, log2(i + 1))."""
    disc = 1.0 / np.log2(np.arange(2, K + 2, dtype=np.float64))
    num, idcg = [], []
    for g in judged_grades:
        g = np.asarray(g, dtype=np.int64)
        pos = np.sort(g[g > 0])[::-1][:K]
        num.append(int((g >= 1).sum()))
        idcg.append(float((pos * disc[:len(pos)]).sum()))
    return np.asarray(num, dtype=np.int32), np.asarray(idcg, dtype=np.float64)


def docno_order_ids(docnos):
    """int32 ids whose numeric order is the docno STRINGS' order, so that the kernel's tie-break (docid descending)
    is trec_eval's (docno descending).  docnos: array-like of str, any shape."""
    arr = np.asarray(docnos, dtype=object)
    uniq, inv = np.unique(arr.reshape(-1).astype(str), return_inverse=True)
    return inv.reshape(arr.shape).astype(np.int32)


def topk_metrics(scores, docids=None, rels=None, K=10, num_rel=None, idcg=None):
    """scores fp32 [Q,C]; docids int32 [Q,C] (None -> candidate position); rels uint8 [Q,C] graded
    relevance (None -> no metrics); num_rel int32 [Q] / idcg float64 [Q]: the queries' qrel totals
    (``qrel_totals``) — without them AP and nDCG assume every judged-relevant document is among the candidates.
    Returns dict(top_ids [Q,K], top_scores [Q,K], ndcg, rr, rr_cut, ap, and their means as device scalars)."""
    _lib.require_gpu()
    s = scores.detach().contiguous().float()
    Q, C = s.shape
    d = docids.contiguous().to(torch.int32) if docids is not None else None
    r = rels.contiguous().to(torch.uint8) if rels is not None else None
    top_ids = torch.empty(Q, K, dtype=torch.int32, device=s.device)
    top_scores = torch.empty(Q, K, dtype=torch.float32, device=s.device)
    m = torch.empty(Q, 4, dtype=torch.float32, device=s.device) if r is not None else None
    nr = torch.as_tensor(num_rel).to(device=s.device, dtype=torch.int32).contiguous() if num_rel is not None else None
    ig = torch.as_tensor(idcg).to(device=s.device, dtype=torch.float64).contiguous() if idcg is not None else None
    for t in (nr, ig):
        if t is not None and t.numel() != Q:
            raise ValueError(f"qrel totals must have one entry per query ({Q})")
    _lib.check(_lib.lib().pn_topk_metrics_qrels(_lib.ptr(s), _lib.ptr(d), _lib.ptr(r), _lib.ptr(nr), _lib.ptr(ig),
                                                Q, C, K, _lib.ptr(top_ids), _lib.ptr(top_scores), _lib.ptr(m),
                                                _lib.stream()))
    out = {"top_ids": top_ids, "top_scores": top_scores}
    if m is not None:
        out.update(ndcg=m[:, 0], rr=m[:, 1], rr_cut=m[:, 2], ap=m[:, 3], metrics=m)
    return out
