"""Batched top-k re-ranking + nDCG@10 / MRR@10 / MAP on the B200 — the measure side of eval.py:26,41-43
(pt.Experiment with eval_metrics=["map", "ndcg_cut_10", "recip_rank"]).  One launch for all queries
(csrc/pn_topk_metrics.cu); ranking rule score desc, docid desc (trec_eval)."""
import torch

from . import _lib


def topk_metrics(scores, docids=None, rels=None, K=10):
    """scores fp32 [Q,C]; docids int32 [Q,C] (None -> candidate position); rels uint8 [Q,C] graded
    relevance (None -> no metrics).  Returns dict(top_ids [Q,K], top_scores [Q,K], ndcg, rr, rr_cut, ap,
    and their means as device scalars)."""
    _lib.require_gpu()
    s = scores.detach().contiguous().float()
    Q, C = s.shape
    d = docids.contiguous().to(torch.int32) if docids is not None else None
    r = rels.contiguous().to(torch.uint8) if rels is not None else None
    top_ids = torch.empty(Q, K, dtype=torch.int32, device=s.device)
    top_scores = torch.empty(Q, K, dtype=torch.float32, device=s.device)
    m = torch.empty(Q, 4, dtype=torch.float32, device=s.device) if r is not None else None
    _lib.check(_lib.lib().pn_topk_metrics(_lib.ptr(s), _lib.ptr(d), _lib.ptr(r), Q, C, K,
                                          _lib.ptr(top_ids), _lib.ptr(top_scores), _lib.ptr(m),
                                          _lib.stream()))
    out = {"top_ids": top_ids, "top_scores": top_scores}
    if m is not None:
        out.update(ndcg=m[:, 0], rr=m[:, 1], rr_cut=m[:, 2], ap=m[:, 3], metrics=m)
    return out
