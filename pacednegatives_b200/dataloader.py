"""Paced negative selection on the B200 — drop-in for ``pacednegatives.dataloader`` (dataloader.py:16-162).

Same classes and call surface (``TripletDataset.get_items``, ``PairLoader/LevelLoader.get_batch``,
``LCEDataset.sample/get/__getitem__``, mutable ``LCEDataset.weight``).  The candidate rank lists live on
the GPU as an int32 matrix (``CandidateMatrix``) and every index decision is made by
csrc/pn_select.cu; the host only maps the returned document ids back to text.  Indices are bit-identical
to the reference's: level mode multiplies in fp64 like the Python float the reference uses, binomial mode
replays numpy's legacy ``choice`` on the very uniforms ``np.random`` would have produced, and leaves the
global numpy RNG in the state the reference would have left it in.
"""
from functools import partial
from math import ceil, floor

import numpy as np
import torch

from . import _lib

OUTPUTS = ["true", "false"]
ROW, RANK = 0, 1


def _field(rec, name):
    return rec[name] if isinstance(rec, dict) else getattr(rec, name)


class CandidateMatrix:
    """int32 candidate ids [Q, C] (row-major: one rank list per query, index 0 = easiest, as written by
    utilities/compute_all_bm25.py:43-44) or rank-major [C, Q] (one contiguous row per difficulty rank)."""

    def __init__(self, ids, layout=ROW, Q=None, C=None):
        _lib.require_gpu()
        t = torch.as_tensor(ids)
        if t.dtype != torch.int32:
            t = t.to(torch.int32)
        self.ids = t.cuda().contiguous()
        self.layout = layout
        if layout == ROW:
            self.Q, self.C = self.ids.shape
        else:
            self.C, self.Q = self.ids.shape

    def rank_major(self):
        if self.layout == RANK:
            return self
        return CandidateMatrix(self.ids.t().contiguous(), RANK)

    # -- level mode ---------------------------------------------------------------------------
    def select_level(self, weight, use_ceil=False, check=True, rows=None):
        """One scalar difficulty for all queries (LevelLoader.get_batch, dataloader.py:69-79).
        Returns (idx int32[1], docs int32[Q])."""
        ids, Q = self.ids, self.Q
        if rows is not None:
            assert self.layout == ROW
            ids = self.ids[rows[0]:rows[1]]
            Q = ids.shape[0]
        docs = torch.empty(Q, dtype=torch.int32, device=ids.device)
        idx = torch.empty(1, dtype=torch.int32, device=ids.device)
        status = torch.empty(1, dtype=torch.int32, device=ids.device)
        _lib.check(_lib.lib().pn_select_level_scalar(
            _lib.ptr(ids), Q, self.C, self.layout, float(weight), int(use_ceil),
            _lib.ptr(idx), _lib.ptr(docs), _lib.ptr(status), _lib.stream()))
        if check and int(status.item()):
            raise IndexError(f"list index out of range: weight {weight} with n_neg {self.C - 1}")
        return idx, docs

    def select_items(self, weights, use_ceil=False, rows=None, check=True):
        """Per-item difficulties (PairLoader.get_batch, dataloader.py:46-56).  ``weights`` fp32 [B];
        ``rows`` = (start, stop) slice of queries (row-major only) or None for all."""
        w = weights.detach().contiguous().float()
        ids, Q = self.ids, self.Q
        if rows is not None:
            assert self.layout == ROW
            ids = self.ids[rows[0]:rows[1]]
            Q = ids.shape[0]
        assert w.numel() == Q
        docs = torch.empty(Q, dtype=torch.int32, device=ids.device)
        idx = torch.empty(Q, dtype=torch.int32, device=ids.device)
        status = torch.empty(1, dtype=torch.int32, device=ids.device)
        _lib.check(_lib.lib().pn_select_level_items(
            _lib.ptr(ids), Q, self.C, self.layout, _lib.ptr(w), int(use_ceil), _lib.ptr(idx),
            _lib.ptr(docs), _lib.ptr(status), _lib.stream()))
        if check and int(status.item()):
            raise IndexError("list index out of range in level selection")
        return idx, docs

    # -- binomial mode ------------------------------------------------------------------------
    def binomial_table(self, mean, var, clip_min=1e-10, clip_max=1.0 - 1e-10, pmf=None):
        """Device table (p, first-round cdf) for one difficulty mean.  ``pmf``: optional float64[C]
        (scipy's binom.pmf) for strict parity; otherwise the device lgamma formula."""
        dev = self.ids.device
        table = torch.empty(2 * self.C, dtype=torch.float64, device=dev)
        status = torch.empty(2, dtype=torch.int32, device=dev)
        pmf_t = None
        if pmf is not None:
            pmf_t = torch.as_tensor(np.asarray(pmf, dtype=np.float64)).to(dev).contiguous()
        _lib.check(_lib.lib().pn_binomial_table(float(mean), float(clip_min), float(clip_max),
                                                float(var), self.C, _lib.ptr(pmf_t), _lib.ptr(table),
                                                _lib.ptr(status), _lib.stream()))
        return table, status

    def select_binomial(self, table, k, uniforms=None, stream_mode=0, seed=0, rows=None,
                        want_docs=True):
        """k distinct ranks per query.  Returns (idx [Q,k], docs [Q,k] | None, consumed [Q], status)."""
        ids, Q = self.ids, self.Q
        if rows is not None:
            assert self.layout == ROW
            ids = self.ids[rows[0]:rows[1]]
            Q = ids.shape[0]
        dev = ids.device
        idx = torch.empty(Q, k, dtype=torch.int32, device=dev)
        docs = torch.empty(Q, k, dtype=torch.int32, device=dev) if want_docs else None
        consumed = torch.empty(Q, dtype=torch.int32, device=dev)
        status = torch.empty(1, dtype=torch.int32, device=dev)
        U = 0
        if uniforms is not None:
            uniforms = uniforms.contiguous()
            assert uniforms.dtype == torch.float64
            U = uniforms.shape[-1] if stream_mode == 0 else uniforms.numel()
        _lib.check(_lib.lib().pn_select_binomial(
            _lib.ptr(ids), Q, self.C, self.layout, _lib.ptr(table), k, _lib.ptr(uniforms), U,
            stream_mode, seed, _lib.ptr(idx), _lib.ptr(docs), _lib.ptr(consumed), _lib.ptr(status),
            _lib.stream()))
        return idx, docs, consumed, status


def _text_maps(corpus):
    docs = {_field(d, "doc_id"): _field(d, "text") for d in corpus.docs_iter()}
    queries = {_field(q, "query_id"): _field(q, "text") for q in corpus.queries_iter()}
    return docs, queries


def _dense_ids(neg_idx):
    """doc-id strings -> dense int32 ids (first-seen order) + the reverse table."""
    table, rev = {}, []
    out = np.empty((len(neg_idx), len(neg_idx[0])), dtype=np.int32)
    for i, row in enumerate(neg_idx):
        for j, d in enumerate(row):
            v = table.get(d)
            if v is None:
                v = table[d] = len(rev)
                rev.append(d)
            out[i, j] = v
    return out, rev


class TripletDataset:
    """dataloader.py:16-33."""

    def __init__(self, pairs, neg_idx, corpus, max=False):
        self.neg_idx = neg_idx
        self.n_neg = len(neg_idx[0]) - 1
        self.docs, self.queries = _text_maps(corpus)
        self.round = ceil if max else floor
        self.use_ceil = bool(max)
        self.data = [(self.queries[q], self.docs[p]) for q, p in pairs]
        dense, self._rev = _dense_ids(neg_idx)
        self.cand = CandidateMatrix(dense, ROW)

    def __len__(self):
        return len(self.data)

    def _doc_text(self, dense_id):
        return self.docs[self._rev[dense_id]]

    def get_items(self, idx, weight=None):
        assert weight is not None, "Weights not set"
        q, p = self.data[idx]
        _, d = self.cand.select_level(float(weight), self.use_ceil, rows=(idx, idx + 1))
        return q, p, self._doc_text(int(d.item()))

    def get_batch_docs(self, start, stop, weights):
        """Negative texts for items [start, stop) — one launch + one D2H copy for the batch."""
        if torch.is_tensor(weights) and weights.numel() > 1:
            _, d = self.cand.select_items(weights.to("cuda"), self.use_ceil, rows=(start, stop))
            d = d.cpu().tolist()
        else:
            w = float(weights.item()) if torch.is_tensor(weights) else float(weights)
            _, d = self.cand.select_level(w, self.use_ceil, rows=(start, stop))
            d = d.cpu().tolist()
        return [self._doc_text(x) for x in d]


class _Loader:
    def __init__(self, dataset, batch_size) -> None:
        self.dataset = dataset
        self.batch_size = batch_size

    def __len__(self):
        return len(self.dataset)

    def format(self, q, d):
        return 'Query: ' + q + ' Document: ' + d + ' Relevant:'

    def _batch(self, idx, weights):
        lo, hi = idx * self.batch_size, (idx + 1) * self.batch_size
        negs = self.dataset.get_batch_docs(lo, hi, weights)
        px, nx, o_p, o_n = [], [], [], []
        for j, n in zip(range(lo, hi), negs):
            q, p = self.dataset.data[j]
            px.append(self.format(q, p))
            nx.append(self.format(q, n))
            o_p.append(OUTPUTS[0])
            o_n.append(OUTPUTS[1])
        return px, nx, o_p, o_n


class PairLoader(_Loader):
    """dataloader.py:35-56: per-item weights."""

    def get_batch(self, idx, weights=None):
        if weights is None:
            weights = torch.ones(self.batch_size)
        return self._batch(idx, weights)


class LevelLoader(_Loader):
    """dataloader.py:58-79: one scalar level for the batch."""

    def get_batch(self, idx, weight=None):
        if weight is None:
            weight = 0.0
        return self._batch(idx, weight)


class LCEDataset(torch.utils.data.Dataset):
    """dataloader.py:97-162: binomial-paced sampling of n negatives without replacement."""

    def __init__(self, pairs, neg_idx, corpus, tokenizer, init_weight, batch_size: int, var: float,
                 n: int, min: float, max: float, use_max=False) -> None:
        self.neg_idx = [np.array(x) for x in neg_idx]
        self.n_neg = len(neg_idx[0])
        self.docs, self.queries = _text_maps(corpus)
        self.tokenizer = tokenizer
        self.data = [(self.queries[q], self.docs[p]) for q, p in pairs]
        self.batch_size = batch_size
        self.var = var
        self.n = n
        self.round = torch.floor
        self.min = min
        self.max = max
        self.weight = init_weight
        dense, self._rev = _dense_ids(neg_idx)
        self.cand = CandidateMatrix(dense, ROW)
        self._table_key, self._table = None, None
        self.strict_pmf = True  # use scipy's pmf table when scipy is importable (bit parity)

    def __len__(self):
        return len(self.data)

    def get(self, idx):
        return self.data[idx[0]], [self.docs[x] for x in self.neg_idx[idx[0]][idx[1]].tolist()]

    def _get_table(self, mean):
        key = (float(mean), float(self.var))
        if key != self._table_key:
            pmf = None
            if self.strict_pmf:
                try:
                    from scipy.stats import binom
                    m = np.clip(mean, self.min, self.max)
                    pmf = binom.pmf(np.arange(self.n_neg), self.n_neg - 1, m)
                except ImportError:
                    pmf = None
            table, status = self.cand.binomial_table(mean, self.var, self.min, self.max, pmf)
            nz, bad = status.cpu().tolist()
            if bad or nz < self.n:
                # the reference prints and exit()s here (dataloader.py:149-151); raise instead
                raise ValueError(f'Failed to sample with mean {mean} and var {self.var}')
            self._table_key, self._table = key, table
        return self._table

    def sample_rows(self, mean, start, stop):
        """Sample for queries [start, stop) from the global numpy RNG exactly as consecutive
        ``sample(mean)`` calls would, in one launch.  Returns int32 [stop-start, n] on the host."""
        table = self._get_table(mean)
        Q = stop - start
        state = np.random.get_state()
        budget = 8 * self.n * Q + 64
        while True:
            u = torch.from_numpy(np.random.random_sample(budget)).cuda()
            idx, _, consumed, status = self.cand.select_binomial(
                table, self.n, u, stream_mode=1, rows=(start, stop), want_docs=False)
            np.random.set_state(state)
            if int(status.item()) == 0:
                break
            budget *= 4
        used = int(consumed.sum().item())
        np.random.random_sample(used)  # leave the global stream where the reference would
        return idx.cpu().numpy()

    def sample(self, mean):
        return self.sample_rows(mean, 0, 1)[0].astype(np.int64)

    def format(self, q, d):
        return 'Query: ' + q + ' Document: ' + d + ' Relevant:'

    def __getitem__(self, idx):
        _idx = self.sample(self.weight)
        qp, n = self.get((idx, _idx))
        q, p = qp
        return {'pos': self.format(q, p), 'neg': list(map(partial(self.format, q), n))}


class TokenStore:
    """Pre-tokenised texts as a CSR pair in HBM: ``tok`` int32 [total], ``off`` int64 [n + 1].  Built once from the
    tokenizer's id lists (no </s>); MS MARCO passage (8.8 M x ~75 tokens) is ~2.7 GB."""

    def __init__(self, id_lists):
        lens = np.fromiter((len(x) for x in id_lists), dtype=np.int64, count=len(id_lists))
        off = np.zeros(len(id_lists) + 1, dtype=np.int64)
        np.cumsum(lens, out=off[1:])
        flat = np.empty(int(off[-1]), dtype=np.int32)
        for i, x in enumerate(id_lists):
            flat[off[i]:off[i + 1]] = x
        _lib.require_gpu()
        self.n = len(id_lists)
        self.tok = torch.from_numpy(flat).cuda()
        self.off = torch.from_numpy(off).cuda()


class PromptPacker:
    """'Query: ' + q + ' Document: ' + d + ' Relevant:' (``format``, dataloader.py:43-44) assembled on the GPU from
    token stores (pn_pack_prompts): rows come out as the int32 [n, S] matrix the scorer reads, padded like
    ``tokenizer(prompts, padding=True)``; nothing passes through the host but the 8-byte status word.

    ``tokenizer`` is only used here, once, for the three literal pieces (any callable ``encode(str) -> ids`` whose
    result ends in </s>, like T5Tokenizer.encode)."""

    def __init__(self, tokenizer=None, prefix=None, middle=None, suffix=None, pad_id=0):
        if tokenizer is not None:
            prefix = list(tokenizer.encode('Query:'))[:-1]
            middle = list(tokenizer.encode('Document:'))[:-1]
            suffix = list(tokenizer.encode('Relevant:'))
        _lib.require_gpu()
        self.pieces = [torch.tensor(list(x), dtype=torch.int32, device="cuda") for x in (prefix, middle, suffix)]
        self.pad_id = int(pad_id)

    def pack(self, queries, docs, q_idx, d_idx, S=None, max_len=0):
        """queries / docs: TokenStore; q_idx / d_idx: int32 device tensors [n] (d_idx straight from the selection
        kernels).  S None -> pad to the longest row of this batch (one extra lengths-only launch and a D2H of the
        status word); a fixed S (e.g. 512) needs no host round trip before the scorer.  Returns (ids, lengths)."""
        q_idx = q_idx.to(device="cuda", dtype=torch.int32).contiguous()
        d_idx = d_idx.to(device="cuda", dtype=torch.int32).contiguous()
        n = q_idx.numel()
        if d_idx.numel() != n:
            raise ValueError("q_idx and d_idx must have the same length")
        lens = torch.empty(n, dtype=torch.int32, device="cuda")
        status = torch.empty(2, dtype=torch.int32, device="cuda")
        lib = _lib.lib()

        def call(ids, width):
            p = self.pieces
            _lib.check(lib.pn_pack_prompts(_lib.ptr(p[0]), p[0].numel(), _lib.ptr(p[1]), p[1].numel(), _lib.ptr(p[2]),
                                           p[2].numel(), _lib.ptr(queries.tok), _lib.ptr(queries.off), queries.n,
                                           _lib.ptr(docs.tok), _lib.ptr(docs.off), docs.n, _lib.ptr(q_idx),
                                           _lib.ptr(d_idx), n, width, int(max_len), self.pad_id, _lib.ptr(ids),
                                           _lib.ptr(lens), _lib.ptr(status), _lib.stream()))
        if S is None:
            call(None, 0)
            longest, bad = status.cpu().tolist()
            if bad:
                raise IndexError(f"pack: {bad} (query, document) indices out of range")
            S = max(int(longest), 1)
        ids = torch.empty(n, int(S), dtype=torch.int32, device="cuda")
        call(ids, int(S))
        self.last_status = status
        return ids, lens

    def check(self):
        """Raise if the last pack() had rows longer than S or indices out of range (reads the status word)."""
        longest, bad = self.last_status.cpu().tolist()
        if bad:
            raise ValueError(f"pack: {bad} rows out of range or longer than the row width (longest {longest})")
