"""Generate golden vectors by running the REFERENCE itself (build container only).

The reference ships no tests or fixtures, so parity is pinned against its own functions executed here:
this script imports /root/reference/pacednegatives (with stubs for the packages that are absent
offline: lightning, ir_datasets, transformers.AdamW), runs the hot-path functions on seeded inputs and
writes inputs + outputs to tests/golden/*.npz.  /root/reference does not exist on the GPU box, so the
tests only ever read the .npz files.

    python tests/golden/make_golden.py
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"


def import_reference():
    """Stub the absent third-party packages, then import the reference modules unmodified."""
    import torch.nn as nn

    class _Hp(dict):
        __getattr__ = dict.__getitem__
        __setattr__ = dict.__setitem__

    class LightningModule(nn.Module):
        def __init__(self):
            super().__init__()
            object.__setattr__(self, "_hp", _Hp())

        @property
        def hparams(self):
            return self._hp

        def save_hyperparameters(self):
            pass

    lightning = types.ModuleType("lightning")
    lightning.LightningModule = LightningModule
    lightning.LightningDataModule = object
    lightning.Callback = object
    sys.modules["lightning"] = lightning
    irds = types.ModuleType("ir_datasets")
    irds.load = lambda name: None
    sys.modules["ir_datasets"] = irds
    import transformers

    def adamw(params, lr):  # transformers.AdamW was removed upstream; same defaults (eps 1e-6, wd 0)
        return torch.optim.AdamW(params, lr=lr, eps=1e-6, weight_decay=0.0)

    def patch():
        for mod in {transformers, sys.modules["transformers"]}:
            mod.__dict__["AdamW"] = adamw

    sys.path.insert(0, REF)
    import pacednegatives.weights as rw
    import pacednegatives.dataloader as rd
    import pacednegatives.utilities.loss as rl
    patch()
    import pacednegatives.lceT5 as rt
    patch()
    import pacednegatives.pairwrapper as rp
    return rw, rd, rl, rt, rp


class FakeCorpus:
    """Duck-type of the ir_datasets corpus the reference iterates (dataloader.py:20-21)."""

    def __init__(self, n_docs, n_queries):
        self.n_docs, self.n_queries = n_docs, n_queries

    def docs_iter(self):
        for i in range(self.n_docs):
            yield {"doc_id": str(i), "text": f"d{i}"}

    def queries_iter(self):
        for i in range(self.n_queries):
            yield {"query_id": str(i), "text": f"q{i}"}


def gold_weights(rw, rt, out):
    rng = np.random.RandomState(5678)
    loss = rng.exponential(0.7, size=257).astype(np.float32)
    loss[:4] = [0.1, 0.5, 0.7, 0.3]
    etas = np.array([0.1, 0.3466, 0.5, 1.0], dtype=np.float32)
    loss[5] = etas[1]  # boundary: loss == eta takes the else branch
    res = {"loss": loss, "etas": etas}
    for ei, eta in enumerate(etas):
        hp = types.SimpleNamespace(eta=float(eta))
        m = rt.LCEWeights(hp)
        lt = torch.tensor(loss)
        w = m(lt)
        (w * lt).mean().backward()
        res[f"lce_w_{ei}"] = w.detach().numpy()
        res[f"lce_dmean_deta_{ei}"] = m.eta.grad.numpy().copy()
        for mask in (0.0, 0.25):
            for name, cls in (("eta", rw.EtaWeights), ("lce", rw.LCEWeights)):
                mod = cls(float(eta), device=torch.device("cpu"))
                mod.set_mask(mask)
                wn = mod.no_grad(lt, torch.tensor([float(eta)]))
                res[f"nograd_{name}_{ei}_m{int(mask * 100)}"] = wn.numpy()
    np.savez_compressed(out, **res)


def gold_loss(rl, rt, out):
    rng = np.random.RandomState(91)
    V = 97
    res = {}
    case = 0
    for T in (1, 2):
        for n in (2, 7, 15):
            for use_mean in (True, False):
                Q = 4
                p = rng.randn(Q, T, V).astype(np.float32) * 3
                nl = rng.randn(Q * n, T, V).astype(np.float32) * 3
                op = rng.randint(0, V, size=(Q, T)).astype(np.int64)
                on = rng.randint(0, V, size=(Q * n, T)).astype(np.int64)
                if case % 3 == 0:
                    on[1, 0] = -100  # ignore_index path
                fn = rl.init_LCEcrossentropy(ignore_index=-100, use_mean=use_mean)
                # the callers re-view the negatives as [-1, n, V] first (lceT5.py:134)
                nview = torch.tensor(nl).view(-1, n, V)
                loss_mean = fn(torch.tensor(p), nview, torch.tensor(op), torch.tensor(on))
                model = types.SimpleNamespace(
                    loss_fn=torch.nn.CrossEntropyLoss(ignore_index=-100, reduction="none"),
                    hparams=types.SimpleNamespace(use_mean=use_mean))
                ce = rt.LCEModel.pair_loss(model, torch.tensor(p), nview, torch.tensor(op),
                                           torch.tensor(on))
                res[f"c{case}_meta"] = np.array([T, n, int(use_mean), Q, V])
                res[f"c{case}_p"] = p
                res[f"c{case}_n"] = nl
                res[f"c{case}_op"] = op
                res[f"c{case}_on"] = on
                res[f"c{case}_pair_loss"] = ce.numpy()
                res[f"c{case}_lcece_mean"] = np.array(loss_mean.item(), dtype=np.float32)
                case += 1
    res["n_cases"] = np.array(case)
    np.savez_compressed(out, **res)


def gold_select(rd, out):
    res = {}
    # level mode: TripletDataset.get_items
    Q, C = 6, 101
    cand = ((np.arange(Q)[:, None] * 7919 + np.arange(C)[None, :] * 104729) % 5003).astype(np.int32)
    neg_idx = [[str(x) for x in row] for row in cand]
    pairs = [(str(q), str(q)) for q in range(Q)]
    corpus = FakeCorpus(5003, Q)
    n_neg = C - 1
    ws = [0.0, 1.0, 0.5, 0.3466, 1e-10, 1 - 1e-10]
    for kk in range(0, n_neg + 1, 7):
        x = kk / n_neg
        ws += [x, float(np.nextafter(x, 0)), float(np.nextafter(x, 1))]
    ws = [w for w in ws if 0.0 <= w <= 1.0]
    w32 = np.array(ws, dtype=np.float32)
    for use_max in (False, True):
        ds = rd.TripletDataset(pairs, neg_idx, corpus, max=use_max)
        got64 = np.zeros((len(ws), Q), dtype=np.int32)
        got32 = np.zeros((len(ws), Q), dtype=np.int32)
        for wi, w in enumerate(ws):
            for q in range(Q):
                got64[wi, q] = int(ds.get_items(q, w)[2][1:])
                got32[wi, q] = int(ds.get_items(q, torch.tensor(w32)[wi].item())[2][1:])
        res[f"level_doc64_max{int(use_max)}"] = got64
        res[f"level_doc32_max{int(use_max)}"] = got32
    res["level_cand"] = cand
    res["level_w64"] = np.array(ws, dtype=np.float64)
    res["level_w32"] = w32
    # binomial mode: LCEDataset.sample with the global numpy RNG
    from scipy.stats import binom
    C = 1000
    Qb = 3
    candb = ((np.arange(Qb)[:, None] * 7919 + np.arange(C)[None, :] * 104729) % 8841823).astype(np.int32)
    neg_idx = [[str(x) for x in row] for row in candb]
    docs = {}

    class Corpus2:
        def docs_iter(self):
            seen = set()
            for row in candb:
                for x in row:
                    if int(x) not in seen:
                        seen.add(int(x))
                        yield {"doc_id": str(int(x)), "text": f"d{int(x)}"}
            for q in range(Qb):
                if -q - 1 not in seen:
                    yield {"doc_id": f"p{q}", "text": f"pos{q}"}

        def queries_iter(self):
            for q in range(Qb):
                yield {"query_id": str(q), "text": f"q{q}"}

    pairs = [(str(q), f"p{q}") for q in range(Qb)]
    means = [1e-10, 0.01, 0.3466, 0.5, 0.9, 1.5]
    ks = [2, 7, 15, 31]
    seeds = list(range(6))
    res["binom_cand"] = candb
    res["binom_means"] = np.array(means)
    res["binom_ks"] = np.array(ks)
    res["binom_seeds"] = np.array(seeds)
    for mi, mean in enumerate(means):
        res[f"binom_pmf_{mi}"] = binom.pmf(np.arange(C), C - 1, np.clip(mean, 1e-10, 1 - 1e-10))
        for ki, k in enumerate(ks):
            for var in (0.01, 0.1):
                ds = rd.LCEDataset(pairs, neg_idx, Corpus2(), None, mean, 16, var=var, n=k,
                                   min=1e-10, max=1 - 1e-10)
                per_seed = np.zeros((len(seeds), k), dtype=np.int64)
                for si, seed in enumerate(seeds):
                    np.random.seed(seed)
                    per_seed[si] = ds.sample(mean)
                res[f"binom_idx_m{mi}_k{ki}_v{int(var * 100)}"] = per_seed
            # one global stream over the Qb queries via __getitem__ (the reference's DataLoader order)
            ds = rd.LCEDataset(pairs, neg_idx, Corpus2(), None, mean, 16, var=0.01, n=k,
                               min=1e-10, max=1 - 1e-10)
            np.random.seed(1234)
            docs_seq = np.zeros((Qb, k), dtype=np.int32)
            for q in range(Qb):
                item = ds[q]
                docs_seq[q] = [int(s.split(" Document: d")[1].split(" Relevant:")[0])
                               for s in item["neg"]]
            res[f"binom_seqdocs_m{mi}_k{ki}"] = docs_seq
    np.savez_compressed(out, **res)


def gold_table_update(rp, rw, out):
    """NewWrapper.meta_loop (pairwrapper.py:219-239) with a fake model/tokenizer/loader."""
    B, V = 8, 23
    rng = np.random.RandomState(7)
    logits_p = torch.tensor(rng.randn(B, 1, V).astype(np.float32) * 2)
    logits_n = torch.tensor(rng.randn(B, 1, V).astype(np.float32) * 2)
    o_p = torch.tensor(rng.randint(0, V, size=(B, 1)))
    o_n = torch.tensor(rng.randint(0, V, size=(B, 1)))
    wrap = object.__new__(rp.NewWrapper)
    wrap.device = torch.device("cpu")
    wrap.loss_fn = torch.nn.CrossEntropyLoss(ignore_index=-100, reduction="none")
    wrap.weights = rw.Weights((2, B), device=torch.device("cpu"))
    init = rng.rand(B).astype(np.float32)
    with torch.no_grad():
        wrap.weights.weight[1] = torch.tensor(init)
    wrap.train_loader = types.SimpleNamespace(get_batch=lambda j, w: ("px", "nx", "op", "on"))
    wrap.prep_batch = lambda batch: ("px", "nx", o_p, o_n)
    wrap.model = lambda input_ids, labels: types.SimpleNamespace(
        logits=logits_p if input_ids == "px" else logits_n)
    lr = 0.05
    wrap.scheduler = types.SimpleNamespace(get_last_lr=lambda: [lr])
    wrap.meta_loop(1)
    ce_p = wrap.loss_fn(logits_p.view(-1, V), o_p.view(-1)).numpy()
    ce_n = wrap.loss_fn(logits_n.view(-1, V), o_n.view(-1)).numpy()
    np.savez_compressed(out, init=init, ce_p=ce_p, ce_n=ce_n, lr=np.float32(lr),
                        updated=wrap.weights.weight[1].detach().numpy(),
                        logits_p=logits_p.numpy(), logits_n=logits_n.numpy(),
                        o_p=o_p.numpy(), o_n=o_n.numpy())


def gold_t5(rt, out):
    """HF T5ForConditionalGeneration itself (the reference's model, lceT5.py:106) on seeded inputs."""
    sys.path.insert(0, ROOT)
    from oracle import t5 as ot
    from transformers import T5Config, T5ForConditionalGeneration
    dims = ot.T5Dims(vocab=384, d_model=128, d_kv=64, d_ff=256, n_layers=2, n_dec_layers=2,
                     n_heads=2)
    w = ot.init_weights(dims, seed=11)
    cfg = T5Config(vocab_size=dims.vocab, d_model=dims.d_model, d_kv=dims.d_kv, d_ff=dims.d_ff,
                   num_layers=dims.n_layers, num_decoder_layers=dims.n_dec_layers,
                   num_heads=dims.n_heads, dropout_rate=0.0, decoder_start_token_id=0,
                   pad_token_id=0, eos_token_id=1, feed_forward_proj="relu",
                   tie_word_embeddings=True)
    m = T5ForConditionalGeneration(cfg).eval()
    sd = dict(w)
    sd["encoder.embed_tokens.weight"] = sd["decoder.embed_tokens.weight"] = w["shared.weight"]
    sd["lm_head.weight"] = w["shared.weight"]
    missing = m.load_state_dict(sd, strict=False)
    assert not missing.missing_keys and not missing.unexpected_keys, missing
    rng = np.random.RandomState(1234)
    Q, k, S = 2, 3, 128
    rel_id, nrel_id = 117, 213
    pids = rng.randint(2, dims.vocab, size=(Q, S)).astype(np.int64)
    nids = rng.randint(2, dims.vocab, size=(Q * k, S)).astype(np.int64)
    pids[:, -1] = 1
    nids[:, -1] = 1
    plab = np.tile([[rel_id, 1]], (Q, 1)).astype(np.int64)
    nlab = np.tile([[nrel_id, 1]], (Q * k, 1)).astype(np.int64)
    with torch.no_grad():
        pl = m(input_ids=torch.tensor(pids), labels=torch.tensor(plab)).logits
        nl = m(input_ids=torch.tensor(nids), labels=torch.tensor(nlab)).logits
        # pairwrapper / re-rank path: one decoder step (pairwrapper.py:81-82, distill/wrapper.py:34-38)
        p1 = m(input_ids=torch.tensor(pids), labels=torch.tensor(plab[:, :1])).logits
        score = torch.log_softmax(p1[:, 0, (rel_id, nrel_id)], dim=1)[:, 0]
    model = types.SimpleNamespace(
        loss_fn=torch.nn.CrossEntropyLoss(ignore_index=-100, reduction="none"),
        hparams=types.SimpleNamespace(use_mean=True))
    ce = rt.LCEModel.pair_loss(model, pl, nl.view(-1, k, dims.vocab), torch.tensor(plab),
                               torch.tensor(nlab))
    np.savez_compressed(
        out, dims=np.array([dims.vocab, dims.d_model, dims.d_kv, dims.d_ff, dims.n_layers,
                            dims.n_dec_layers, dims.n_heads]), seed=np.array(11),
        pids=pids, nids=nids, plab=plab, nlab=nlab, plogits=pl.numpy(), nlogits=nl.numpy(),
        p1_z2=p1[:, 0, (rel_id, nrel_id)].numpy(), p1_score=score.numpy(),
        pair_loss=ce.numpy(), rel_ids=np.array([rel_id, nrel_id]), k=np.array(k))


def gold_t5_mask(out):
    """HF T5ForConditionalGeneration with the tokenizer's attention_mask, as the re-rank path calls it
    (pyterrier_t5.MonoT5ReRanker in eval.py:14,23; distill/wrapper.py:36 `self.model(**enc)`): right-padded
    prompts of different lengths, one decoder step, the two-row score."""
    sys.path.insert(0, ROOT)
    from oracle import t5 as ot
    from transformers import T5Config, T5ForConditionalGeneration
    dims = ot.T5Dims(vocab=384, d_model=128, d_kv=64, d_ff=256, n_layers=2, n_dec_layers=2, n_heads=2)
    w = ot.init_weights(dims, seed=13)
    cfg = T5Config(vocab_size=dims.vocab, d_model=dims.d_model, d_kv=dims.d_kv, d_ff=dims.d_ff,
                   num_layers=dims.n_layers, num_decoder_layers=dims.n_dec_layers,
                   num_heads=dims.n_heads, dropout_rate=0.0, decoder_start_token_id=0,
                   pad_token_id=0, eos_token_id=1, feed_forward_proj="relu", tie_word_embeddings=True)
    m = T5ForConditionalGeneration(cfg).eval()
    sd = dict(w)
    sd["encoder.embed_tokens.weight"] = sd["decoder.embed_tokens.weight"] = w["shared.weight"]
    sd["lm_head.weight"] = w["shared.weight"]
    missing = m.load_state_dict(sd, strict=False)
    assert not missing.missing_keys and not missing.unexpected_keys, missing
    rng = np.random.RandomState(77)
    rel_id, nrel_id = 117, 213
    res = {}
    for S in (128, 200, 384):
        N = 7
        lens = np.array([S, 1, 5, 31, S // 2, S - 1, 130 if S > 130 else S - 3], dtype=np.int32)
        ids = rng.randint(2, dims.vocab, size=(N, S)).astype(np.int64)
        mask = (np.arange(S)[None, :] < lens[:, None]).astype(np.int64)
        ids[mask == 0] = 0
        ids[np.arange(N), lens - 1] = 1
        with torch.no_grad():
            lg = m(input_ids=torch.tensor(ids), attention_mask=torch.tensor(mask),
                   decoder_input_ids=torch.zeros(N, 1, dtype=torch.long)).logits
            sc = torch.log_softmax(lg[:, 0, (rel_id, nrel_id)], dim=1)[:, 0]
            lg_nomask = m(input_ids=torch.tensor(ids), decoder_input_ids=torch.zeros(N, 1, dtype=torch.long)).logits
        res[f"ids_{S}"], res[f"lens_{S}"] = ids, lens
        res[f"z2_{S}"] = lg[:, 0, (rel_id, nrel_id)].numpy()
        res[f"score_{S}"] = sc.numpy()
        res[f"z2_nomask_{S}"] = lg_nomask[:, 0, (rel_id, nrel_id)].numpy()
    np.savez_compressed(out, dims=np.array([dims.vocab, dims.d_model, dims.d_kv, dims.d_ff, dims.n_layers,
                                            dims.n_dec_layers, dims.n_heads]), seed=np.array(13),
                        rel_ids=np.array([rel_id, nrel_id]), **res)


class LegacyAdamW(torch.optim.Optimizer):
    """Restatement of transformers.AdamW.step (transformers <= 4.x optimization.py, the class lceT5.py:5,171-172
    constructs; removed from the transformers 5.x in this image): defaults betas (0.9, 0.999), eps 1e-6,
    weight_decay 0, correct_bias True."""

    def __init__(self, params, lr, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.0):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))

    @torch.no_grad()
    def step(self):
        for g in self.param_groups:
            for p in g["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if not st:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p)
                    st["exp_avg_sq"] = torch.zeros_like(p)
                b1, b2 = g["betas"]
                st["step"] += 1
                st["exp_avg"].mul_(b1).add_(p.grad, alpha=1.0 - b1)
                st["exp_avg_sq"].mul_(b2).addcmul_(p.grad, p.grad, value=1.0 - b2)
                denom = st["exp_avg_sq"].sqrt().add_(g["eps"])
                step_size = g["lr"] * (1.0 - b2 ** st["step"]) ** 0.5 / (1.0 - b1 ** st["step"])
                p.addcdiv_(st["exp_avg"], denom, value=-step_size)
                if g["weight_decay"] > 0.0:
                    p.add_(p, alpha=-g["lr"] * g["weight_decay"])


def gold_train(rt, out):
    """The reference's own LCEModel.training_step (lceT5.py:124-167) executed unmodified on a tiny random-init
    HF T5 (fp32, CPU) for two steps: losses, eta, every parameter's gradient (norm + leading slice) of step 1."""
    sys.path.insert(0, ROOT)
    from oracle import t5 as ot
    from transformers import T5Config, T5ForConditionalGeneration, get_linear_schedule_with_warmup
    dims = ot.T5Dims(vocab=384, d_model=128, d_kv=64, d_ff=256, n_layers=2, n_dec_layers=2, n_heads=2)
    w = ot.init_weights(dims, seed=23)
    cfg = T5Config(vocab_size=dims.vocab, d_model=dims.d_model, d_kv=dims.d_kv, d_ff=dims.d_ff,
                   num_layers=dims.n_layers, num_decoder_layers=dims.n_dec_layers,
                   num_heads=dims.n_heads, dropout_rate=0.0, decoder_start_token_id=0,
                   pad_token_id=0, eos_token_id=1, feed_forward_proj="relu", tie_word_embeddings=True)
    m = T5ForConditionalGeneration(cfg).train()
    sd = {k_: torch.tensor(v) for k_, v in w.items()}
    sd["encoder.embed_tokens.weight"] = sd["decoder.embed_tokens.weight"] = sd["shared.weight"]
    sd["lm_head.weight"] = sd["shared.weight"]
    missing = m.load_state_dict(sd, strict=False)
    assert not missing.missing_keys and not missing.unexpected_keys, missing
    hp = types.SimpleNamespace(n=3, use_mean=True, ignore_index=-100, lr=1e-3, meta_lr=1e-2, warmup_steps=0,
                               total_steps=10, eta=20.0)
    rng = np.random.RandomState(4321)
    Q, k, S = 2, hp.n, 128
    rel_id, nrel_id = 117, 213
    pids = rng.randint(2, dims.vocab, size=(Q, S)).astype(np.int64)
    nids = rng.randint(2, dims.vocab, size=(Q * k, S)).astype(np.int64)
    pids[:, -1] = 1
    nids[:, -1] = 1
    plab = np.tile([[rel_id, 1]], (Q, 1)).astype(np.int64)
    nlab = np.tile([[nrel_id, 1]], (Q * k, 1)).astype(np.int64)
    batch = ({"input_ids": torch.tensor(pids), "labels": torch.tensor(plab)},
             {"input_ids": torch.tensor(nids), "labels": torch.tensor(nlab)})

    self = types.SimpleNamespace()
    self.hparams = hp
    self.model = m
    self.weights = rt.LCEWeights(hp)
    self.loss_fn = torch.nn.CrossEntropyLoss(ignore_index=hp.ignore_index, reduction="none")
    self.pair_loss = lambda p, n, op, on: rt.LCEModel.pair_loss(self, p, n, op, on)
    meta_opt = LegacyAdamW(self.weights.parameters(), lr=hp.meta_lr)
    opt = LegacyAdamW(m.parameters(), lr=hp.lr)
    scheds = [get_linear_schedule_with_warmup(o, num_warmup_steps=hp.warmup_steps,
                                              num_training_steps=hp.total_steps) for o in (meta_opt, opt)]
    self.optimizers = lambda: (meta_opt, opt)
    self.lr_schedulers = lambda: scheds
    self.manual_backward = lambda loss: loss.backward()
    logged = {}
    self.log = lambda name, v: logged.__setitem__(name, float(v))
    names = [n_ for n_, _ in m.named_parameters()]
    res = {}
    for step in range(2):
        before = {n_: p.detach().clone() for n_, p in m.named_parameters()}
        o = rt.LCEModel.training_step(self, batch, step)
        res[f"main_loss{step}"] = float(o["main_loss"])
        res[f"meta_loss{step}"] = float(o["meta_loss"])
        res[f"eta{step}"] = float(self.weights.eta.item())
        res[f"avg_weight{step}"] = logged["avg_weight"]
        if step == 0:
            for n_, p in m.named_parameters():
                g = p.grad.detach().reshape(-1)
                res["gnorm/" + n_] = float(g.norm())
                res["ghead/" + n_] = g[:64].numpy().copy()
                res["dnorm/" + n_] = float((p.detach() - before[n_]).norm())
    np.savez_compressed(
        out, dims=np.array([dims.vocab, dims.d_model, dims.d_kv, dims.d_ff, dims.n_layers, dims.n_dec_layers,
                            dims.n_heads]), seed=np.array(23), pids=pids, nids=nids, plab=plab, nlab=nlab,
        hp=np.array([hp.n, int(hp.use_mean), hp.lr, hp.meta_lr, hp.warmup_steps, hp.total_steps, hp.eta]),
        names=np.array(names), **{k_: np.asarray(v) for k_, v in res.items()})


def _sample_index(name, numel, count=8192):
    """Seeded element sample of a gradient tensor (the wide goldens cannot commit 130 MB of full gradients): the
    relative L2 error over the sample estimates the tensor's; tests re-derive the same indices from the name."""
    import zlib
    if numel <= 65536:
        return np.arange(numel, dtype=np.int64)
    rng = np.random.RandomState(zlib.crc32(name.encode()) & 0x7FFFFFFF)
    return np.sort(rng.choice(numel, size=count, replace=False)).astype(np.int64)


def gold_train_wide(rt, out, dims, Q, k, S, seed, eta=None, ids_seed=4321, rel_id=1176, nrel_id=6136, steps=1,
                    lr=1e-3, meta_lr=1e-2, dropout=None, S_neg=None):
    """The reference's own LCEModel.training_step (lceT5.py:124-167), unmodified, on HF T5 fp32 CPU at the widths
    the bench runs: per parameter the gradient norm, a seeded element sample (all elements for
    tensors <= 65536) and, for matrices, the row / column sums and the norm of every 64 x 64 block (a wrong head /
    block shows up there even when the sample misses it).  eta=None puts the threshold midway between the two middle losses of the step so both weight branches run."""
    sys.path.insert(0, ROOT)
    from oracle import t5 as ot
    from transformers import T5Config, T5ForConditionalGeneration, get_linear_schedule_with_warmup
    w = ot.init_weights(dims, seed=seed)
    # dropout = (p, base_seed): HF in train mode WITH dropout_rate p, as Lightning runs the reference; the masks of
    # every dropout call are the counter-hash masks of csrc/pn_dropout.cuh, replayed inside HF (oracle/dropout.py)
    from oracle import dropout as od
    drop_p, drop_base = dropout if dropout else (0.0, 0)
    cfg = T5Config(vocab_size=dims.vocab, d_model=dims.d_model, d_kv=dims.d_kv, d_ff=dims.d_ff,
                   num_layers=dims.n_layers, num_decoder_layers=dims.n_dec_layers,
                   num_heads=dims.n_heads, dropout_rate=drop_p, decoder_start_token_id=0,
                   pad_token_id=0, eos_token_id=1, feed_forward_proj="relu", tie_word_embeddings=True)
    m = T5ForConditionalGeneration(cfg).train()
    sd = {k_: torch.tensor(v) for k_, v in w.items()}
    sd["encoder.embed_tokens.weight"] = sd["decoder.embed_tokens.weight"] = sd["shared.weight"]
    sd["lm_head.weight"] = sd["shared.weight"]
    missing = m.load_state_dict(sd, strict=False)
    assert not missing.missing_keys and not missing.unexpected_keys, missing

    def replay(step, forwards=(0, 1)):
        """masks of the step's forwards: the no-grad pair (p, n) then the gradient pair, seeds (base << 32) + 2 *
        step + which, positives numbered from document 0 and negatives from Q"""
        sched = []
        for which in forwards:
            sd_ = (drop_base << 32) + 2 * step + which
            sched += [(sd_, 0), (sd_, Q)]
        return od.HFDropoutReplay(drop_p, dims.n_layers, dims.n_dec_layers, sched)
    rng = np.random.RandomState(ids_seed)
    hi = min(dims.vocab, 32100)
    pids = rng.randint(2, hi, size=(Q, S)).astype(np.int64)
    nids = rng.randint(2, hi, size=(Q * k, S_neg or S)).astype(np.int64)
    pids[:, -1] = 1
    nids[:, -1] = 1
    plab = np.tile([[rel_id, 1]], (Q, 1)).astype(np.int64)
    nlab = np.tile([[nrel_id, 1]], (Q * k, 1)).astype(np.int64)
    batch = ({"input_ids": torch.tensor(pids), "labels": torch.tensor(plab)},
             {"input_ids": torch.tensor(nids), "labels": torch.tensor(nlab)})
    loss_fn = torch.nn.CrossEntropyLoss(ignore_index=-100, reduction="none")
    if eta is None:
        probe = types.SimpleNamespace(loss_fn=loss_fn, hparams=types.SimpleNamespace(use_mean=True))
        with torch.no_grad(), replay(0, (0,)):      # the losses eta is compared with: step 0's no-grad forward
            pl = m(**batch[0]).logits
            nl = m(**batch[1]).logits.view(-1, k, dims.vocab)
            ls = np.sort(rt.LCEModel.pair_loss(probe, pl, nl, batch[0]["labels"], batch[1]["labels"]).numpy())
            # midway between the two middle losses: no element sits on the threshold, so bf16 noise cannot flip one
            eta = float(np.float32(0.5 * (ls[len(ls) // 2 - 1] + ls[len(ls) // 2])))
    hp = types.SimpleNamespace(n=k, use_mean=True, ignore_index=-100, lr=lr, meta_lr=meta_lr, warmup_steps=0,
                               total_steps=10, eta=eta)
    self = types.SimpleNamespace()
    self.hparams = hp
    self.model = m
    self.weights = rt.LCEWeights(hp)
    self.loss_fn = loss_fn
    self.pair_loss = lambda p, n, op, on: rt.LCEModel.pair_loss(self, p, n, op, on)
    meta_opt = LegacyAdamW(self.weights.parameters(), lr=hp.meta_lr)
    opt = LegacyAdamW(m.parameters(), lr=hp.lr)
    scheds = [get_linear_schedule_with_warmup(o, num_warmup_steps=hp.warmup_steps,
                                              num_training_steps=hp.total_steps) for o in (meta_opt, opt)]
    self.optimizers = lambda: (meta_opt, opt)
    self.lr_schedulers = lambda: scheds
    self.manual_backward = lambda loss: loss.backward()
    logged = {}
    self.log = lambda name, v: logged.__setitem__(name, float(v))
    names = [n_ for n_, _ in m.named_parameters()]
    res = {}
    for step in range(steps):
        with replay(step) as rp_:
            o = rt.LCEModel.training_step(self, batch, step)
        if drop_p > 0.0:
            assert rp_.calls == 4 * len(rp_.sites), (rp_.calls, len(rp_.sites))
        res[f"main_loss{step}"] = float(o["main_loss"])
        res[f"meta_loss{step}"] = float(o["meta_loss"])
        res[f"eta{step}"] = float(self.weights.eta.item())
        res[f"avg_weight{step}"] = logged["avg_weight"]
        if step == 0:
            for n_, p in m.named_parameters():
                g = p.grad.detach()
                flat = g.reshape(-1)
                res["gnorm/" + n_] = float(flat.double().norm())
                res["gsample/" + n_] = flat[torch.from_numpy(_sample_index(n_, flat.numel()))].numpy().copy()
                if g.dim() == 2 and flat.numel() > 65536:
                    res["growsum/" + n_] = g.double().sum(1).float().numpy()
                    res["gcolsum/" + n_] = g.double().sum(0).float().numpy()
                    # Frobenius norm of every 64 x 64 block (one head's slice of a projection): no cancellation, so a
                    # wrong head / block / diagonal shows up here even when the element sample misses it
                    R, C = g.shape
                    gb = g.double()[:R - R % 64, :C - C % 64].reshape(R // 64, 64, C // 64, 64)
                    res["gblock/" + n_] = gb.pow(2).sum((1, 3)).sqrt().float().numpy()
    np.savez_compressed(
        out, dims=np.array([dims.vocab, dims.d_model, dims.d_kv, dims.d_ff, dims.n_layers, dims.n_dec_layers,
                            dims.n_heads]), seed=np.array(seed), pids=pids, nids=nids, plab=plab, nlab=nlab,
        hp=np.array([hp.n, int(hp.use_mean), hp.lr, hp.meta_lr, hp.warmup_steps, hp.total_steps, hp.eta],
                    dtype=np.float64),
        steps=np.array(steps), dropout=np.array([drop_p, drop_base], dtype=np.float64), names=np.array(names),
        **{k_: np.asarray(v) for k_, v in res.items()})


def gold_train_dropout(rt):
    """a10 with the reference's train-mode dropout (0.1): tiny model for two steps (the seeds advance), and the
    t5-base width (12 heads, S 512) for one."""
    sys.path.insert(0, ROOT)
    from oracle import t5 as ot
    gold_train_wide(rt, os.path.join(HERE, "train_golden_dropout_tiny.npz"),
                    ot.T5Dims(vocab=384, d_model=128, d_kv=64, d_ff=256, n_layers=2, n_dec_layers=2, n_heads=2),
                    Q=2, k=3, S=128, seed=23, steps=2, rel_id=117, nrel_id=213, dropout=(0.1, 7))
    gold_train_wide(rt, os.path.join(HERE, "train_golden_dropout_base.npz"),
                    ot.T5Dims(vocab=32128, d_model=768, d_kv=64, d_ff=3072, n_layers=2, n_dec_layers=2, n_heads=12),
                    Q=2, k=3, S=512, seed=31, steps=1, dropout=(0.1, 11))
    # ragged S (odd key pitch is padded to even) and the un-concatenated path (positives and negatives of different
    # lengths run as two model calls): positives 77 tokens, negatives 90
    gold_train_wide(rt, os.path.join(HERE, "train_golden_dropout_ragged.npz"),
                    ot.T5Dims(vocab=384, d_model=128, d_kv=64, d_ff=256, n_layers=2, n_dec_layers=2, n_heads=2),
                    Q=2, k=3, S=77, seed=29, steps=1, rel_id=117, nrel_id=213, dropout=(0.1, 13), S_neg=90)


def gold_train_widths(rt):
    """VERDICT r1 'next' item 1: (i) t5-base width, (ii) T5-large width (2+2 layers, S 512, vocab 32128), (iii)
    SURVEY 8d cfg1 exactly (full t5-small, Q 8, k 7, S 128, eta 0.3466, one optimiser step)."""
    sys.path.insert(0, ROOT)
    from oracle import t5 as ot
    gold_train_wide(rt, os.path.join(HERE, "train_golden_base.npz"),
                    ot.T5Dims(vocab=32128, d_model=768, d_kv=64, d_ff=3072, n_layers=2, n_dec_layers=2, n_heads=12),
                    Q=2, k=3, S=512, seed=31, steps=2)
    gold_train_wide(rt, os.path.join(HERE, "train_golden_large.npz"),
                    ot.T5Dims(vocab=32128, d_model=1024, d_kv=64, d_ff=4096, n_layers=2, n_dec_layers=2, n_heads=16),
                    Q=2, k=3, S=512, seed=37, steps=1)
    gold_train_wide(rt, os.path.join(HERE, "train_golden_cfg1.npz"), ot.T5_SMALL, Q=8, k=7, S=128, seed=0,
                    eta=0.3466, ids_seed=1234, steps=1, meta_lr=1e-3)


def gold_newwrapper(rp, rw, out):
    """The reference's NewWrapper.meta_loop + main_loop (pairwrapper.py:219-258) executed unmodified on a tiny
    random-init HF T5 (fp32, CPU) for three batches: table rows, losses and parameter movement."""
    sys.path.insert(0, ROOT)
    from oracle import t5 as ot
    from transformers import T5Config, T5ForConditionalGeneration, get_linear_schedule_with_warmup
    dims = ot.T5Dims(vocab=384, d_model=128, d_kv=64, d_ff=256, n_layers=2, n_dec_layers=2, n_heads=2)
    w = ot.init_weights(dims, seed=29)
    cfg = T5Config(vocab_size=dims.vocab, d_model=dims.d_model, d_kv=dims.d_kv, d_ff=dims.d_ff,
                   num_layers=dims.n_layers, num_decoder_layers=dims.n_dec_layers,
                   num_heads=dims.n_heads, dropout_rate=0.0, decoder_start_token_id=0,
                   pad_token_id=0, eos_token_id=1, feed_forward_proj="relu", tie_word_embeddings=True)
    m = T5ForConditionalGeneration(cfg).train()
    sd = {k_: torch.tensor(v) for k_, v in w.items()}
    sd["encoder.embed_tokens.weight"] = sd["decoder.embed_tokens.weight"] = sd["shared.weight"]
    sd["lm_head.weight"] = sd["shared.weight"]
    missing = m.load_state_dict(sd, strict=False)
    assert not missing.missing_keys and not missing.unexpected_keys, missing
    B, S, nb = 4, 128, 3
    rel_id, nrel_id = 117, 213
    rng = np.random.RandomState(977)
    pids = rng.randint(2, dims.vocab, size=(nb, B, S)).astype(np.int64)
    nids = rng.randint(2, dims.vocab, size=(nb, B, S)).astype(np.int64)
    pids[:, :, -1] = 1
    nids[:, :, -1] = 1
    lr, warm, total = 2e-3, 2, 12

    class Tok:   # batches arrive as ("p", j) / ("n", j) / label words; prep_batch's own slicing is exercised
        def __call__(self, x, padding=True, return_tensors="pt"):
            if isinstance(x, tuple):
                ids = torch.tensor((pids if x[0] == "p" else nids)[x[1]])
            else:
                ids = torch.tensor([[rel_id if t == "true" else nrel_id, 1] for t in x])
            return types.SimpleNamespace(input_ids=ids)

    seen = []

    class Loader:
        def get_batch(self, j, weights):
            seen.append(weights.detach().numpy().copy())
            return ("p", j), ("n", j), ["true"] * B, ["false"] * B

        def __len__(self):
            return nb * B

    wrap = object.__new__(rp.NewWrapper)
    wrap.device = torch.device("cpu")
    wrap.loss_fn = torch.nn.CrossEntropyLoss(ignore_index=-100, reduction="none")
    wrap.model = m
    wrap.tokenizer = Tok()
    wrap.optimizer = LegacyAdamW(m.parameters(), lr=lr)
    wrap.batch_size = B
    wrap.weights = rw.Weights((nb, B), device=torch.device("cpu"))
    wrap.train_loader = Loader()
    wrap.scheduler = get_linear_schedule_with_warmup(wrap.optimizer, num_warmup_steps=warm,
                                                     num_training_steps=total)
    res = {}
    before = {n_: p.detach().clone() for n_, p in m.named_parameters()}
    for j in range(nb):
        wrap.meta_loop(j)
        res[f"row{j}"] = wrap.weights.weight[j].detach().numpy().copy()
        res[f"loss{j}"] = np.float32(wrap.main_loop(j))
        res[f"lr{j}"] = np.float32(wrap.scheduler.get_last_lr()[0])
    for n_, p in m.named_parameters():
        res["dnorm/" + n_] = float((p.detach() - before[n_]).norm())
    res["seen"] = np.stack(seen)
    np.savez_compressed(
        out, dims=np.array([dims.vocab, dims.d_model, dims.d_kv, dims.d_ff, dims.n_layers, dims.n_dec_layers,
                            dims.n_heads]), seed=np.array(29), pids=pids, nids=nids,
        hp=np.array([B, lr, warm, total, rel_id, nrel_id]), **{k_: np.asarray(v) for k_, v in res.items()})


def gold_stdwrapper(rp, rw, out):
    """The reference's StdWrapper.meta_loop + main_loop (pairwrapper.py:102-179) executed unmodified on a tiny
    random-init HF T5 (fp32, CPU) for three batches.  The harness supplies the one attribute the reference's
    constructor leaves commented out (pairwrapper.py:51: `self.meta_model`), freshly per batch: with transformers
    5.x the reference's second meta_loop call on the SAME meta model raises (update_params replaced `shared.weight`
    with a Parameter no forward path reads, and `grad` refuses unused inputs), so upstream runs exactly one batch."""
    sys.path.insert(0, ROOT)
    from oracle import t5 as ot
    from transformers import T5Config, T5ForConditionalGeneration, get_linear_schedule_with_warmup
    dims = ot.T5Dims(vocab=384, d_model=128, d_kv=64, d_ff=256, n_layers=2, n_dec_layers=2, n_heads=2)
    w = ot.init_weights(dims, seed=43)
    cfg = T5Config(vocab_size=dims.vocab, d_model=dims.d_model, d_kv=dims.d_kv, d_ff=dims.d_ff,
                   num_layers=dims.n_layers, num_decoder_layers=dims.n_dec_layers,
                   num_heads=dims.n_heads, dropout_rate=0.0, decoder_start_token_id=0,
                   pad_token_id=0, eos_token_id=1, feed_forward_proj="relu", tie_word_embeddings=True)

    def model_init():
        m_ = T5ForConditionalGeneration(cfg).train()
        sd = {k_: torch.tensor(v) for k_, v in w.items()}
        sd["encoder.embed_tokens.weight"] = sd["decoder.embed_tokens.weight"] = sd["shared.weight"]
        sd["lm_head.weight"] = sd["shared.weight"]
        missing = m_.load_state_dict(sd, strict=False)
        assert not missing.missing_keys and not missing.unexpected_keys, missing
        return m_

    m = model_init()
    B, S, nb = 4, 128, 3
    rel_id, nrel_id = 117, 213
    rng = np.random.RandomState(1977)
    pids = rng.randint(2, dims.vocab, size=(nb, B, S)).astype(np.int64)
    nids = rng.randint(2, dims.vocab, size=(nb, B, S)).astype(np.int64)
    pids[:, :, -1] = 1
    nids[:, :, -1] = 1
    lr, warm, total = 2e-2, 0, 12     # a look-ahead step large enough to move the losses well past bf16 noise

    class Tok:
        def __call__(self, x, padding=True, return_tensors="pt"):
            if isinstance(x, tuple):
                ids = torch.tensor((pids if x[0] == "p" else nids)[x[1]])
            else:
                ids = torch.tensor([[rel_id if t == "true" else nrel_id, 1] for t in x])
            return types.SimpleNamespace(input_ids=ids)

    seen = []

    class Loader:
        def get_batch(self, j, weights):
            seen.append(weights.detach().numpy().copy())
            return ("p", j), ("n", j), ["true"] * B, ["false"] * B

        def __len__(self):
            return nb * B

    ces = []
    ce_fn = torch.nn.CrossEntropyLoss(ignore_index=-100, reduction="none")

    def loss_fn(logits, target):
        o = ce_fn(logits, target)
        ces.append(o.detach().numpy().copy())
        return o

    wrap = object.__new__(rp.StdWrapper)
    wrap.device = torch.device("cpu")
    wrap.loss_fn = loss_fn
    wrap.model = m
    wrap.tokenizer = Tok()
    wrap.optimizer = LegacyAdamW(m.parameters(), lr=lr)
    wrap.batch_size = B
    wrap.weights = rw.Weights((nb, B), device=torch.device("cpu"))
    wrap.weights.weight.data[:] = torch.tensor(rng.uniform(0.3, 1.0, size=(nb, B)).astype(np.float32))
    wrap.train_loader = Loader()
    wrap.scheduler = get_linear_schedule_with_warmup(wrap.optimizer, num_warmup_steps=warm,
                                                     num_training_steps=total)
    res = {"w0": wrap.weights.weight.detach().numpy().copy()}
    before = {n_: p.detach().clone() for n_, p in m.named_parameters()}
    for j in range(nb):
        wrap.meta_model = model_init()       # pairwrapper.py:51, commented out upstream
        del ces[:]
        wrap.meta_loop(j)
        # loss_fn calls of meta_loop: ce_p(theta), ce_n(theta), ce_p(theta'), ce_n(theta')
        res[f"meta_ce{j}"] = np.stack(ces)
        res[f"row{j}"] = wrap.weights.weight[j].detach().numpy().copy()
        del ces[:]
        res[f"loss{j}"] = np.float32(wrap.main_loop(j))
        res[f"main_ce{j}"] = np.stack(ces)   # ce_p(model), ce_n(meta model = theta')
        res[f"lr{j}"] = np.float32(wrap.scheduler.get_last_lr()[0])
    for n_, p in m.named_parameters():
        res["dnorm/" + n_] = float((p.detach() - before[n_]).norm())
    res["seen"] = np.stack(seen)
    np.savez_compressed(
        out, dims=np.array([dims.vocab, dims.d_model, dims.d_kv, dims.d_ff, dims.n_layers, dims.n_dec_layers,
                            dims.n_heads]), seed=np.array(43), pids=pids, nids=nids,
        hp=np.array([B, lr, warm, total, rel_id, nrel_id]), **{k_: np.asarray(v) for k_, v in res.items()})


def _load_old(name):
    """old/<name>.py by path (the directory is not a package); wandb and ir_datasets.log are stubbed."""
    import importlib.util
    import contextlib
    wandb = types.ModuleType("wandb")
    wandb.run = None
    sys.modules["wandb"] = wandb
    irds = sys.modules["ir_datasets"]

    class _Bar:
        def update(self, n):
            pass

        def set_postfix(self, d):
            pass

    class _Log:
        @contextlib.contextmanager
        def pbar_raw(self, **kw):
            yield _Bar()
    irds.log = types.SimpleNamespace(easy=lambda: _Log())
    spec = importlib.util.spec_from_file_location("ref_old_" + name, os.path.join(REF, "old", name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def gold_scalar_wrappers(rp, rw, out):
    """The reference's LevelWrapper / InterpWrapper / EtaWrapper train() loops (old/levels.py, old/interp.py,
    old/eta_bound.py) executed unmodified on a tiny random-init HF T5 (fp32, CPU): per-step difficulty handed to
    the loader, losses, learning rates, success rates, eta.  Two accommodations, both stated in the test:
    the optimiser is the transformers.AdamW rule (LegacyAdamW; absent from this image's transformers), and
    weights.gen_var returns a non-leaf tensor, which is what `.to(cuda)` makes of it on the reference's own
    device — on CPU the in-place writes of EtaWeights.forward raise on the leaf (weights.py:34-43)."""
    sys.path.insert(0, ROOT)
    from oracle import t5 as ot
    from torch.autograd import Variable
    from transformers import T5Config, T5ForConditionalGeneration
    rw.gen_var = lambda x, y: Variable(x, requires_grad=y).clone()
    dims = ot.T5Dims(vocab=384, d_model=128, d_kv=64, d_ff=256, n_layers=2, n_dec_layers=2, n_heads=2)
    seed = 31
    w = ot.init_weights(dims, seed=seed)
    cfg = T5Config(vocab_size=dims.vocab, d_model=dims.d_model, d_kv=dims.d_kv, d_ff=dims.d_ff,
                   num_layers=dims.n_layers, num_decoder_layers=dims.n_dec_layers,
                   num_heads=dims.n_heads, dropout_rate=0.0, decoder_start_token_id=0,
                   pad_token_id=0, eos_token_id=1, feed_forward_proj="relu", tie_word_embeddings=True)

    def model_init():
        m = T5ForConditionalGeneration(cfg).train()
        sd = {k_: torch.tensor(v) for k_, v in w.items()}
        sd["encoder.embed_tokens.weight"] = sd["decoder.embed_tokens.weight"] = sd["shared.weight"]
        sd["lm_head.weight"] = sd["shared.weight"]
        missing = m.load_state_dict(sd, strict=False)
        assert not missing.missing_keys and not missing.unexpected_keys, missing
        return m

    B, S, nb, n_neg = 4, 128, 10, 4
    rel_id, nrel_id = 117, 213
    rng = np.random.RandomState(1203)
    pids = rng.randint(2, dims.vocab, size=(nb, B, S)).astype(np.int64)
    # one negative prompt per (batch, query, rank): the loader hands out the rank the difficulty selects
    nids = rng.randint(2, dims.vocab, size=(nb, B, n_neg + 1, S)).astype(np.int64)
    pids[..., -1] = 1
    nids[..., -1] = 1
    lr, meta_lr, total = 3e-3, 5e-2, nb * B

    class Tok:
        def __call__(self, x, padding=True, return_tensors="pt"):
            if isinstance(x, tuple):
                ids = torch.tensor(pids[x[1]] if x[0] == "p" else nids[x[1], :, x[2]])
            else:
                ids = torch.tensor([[rel_id if t == "true" else nrel_id, 1] for t in x])
            return types.SimpleNamespace(input_ids=ids)

        def encode(self, word):
            return [rel_id if word == "true" else nrel_id, 1]

    class Loader:
        batch_size = B
        dataset = types.SimpleNamespace(n_neg=n_neg)

        def __init__(self):
            self.seen = []

        def get_batch(self, j, weight):
            self.seen.append(float(weight))
            r = int(np.floor(float(weight) * n_neg))
            return ("p", j), ("n", j, r), ["true"] * B, ["false"] * B

        def __len__(self):
            return nb * B

    res = {}

    def finish(tag, wrap, loader, before):
        res[tag + "/seen"] = np.array(loader.seen, dtype=np.float64)
        res[tag + "/dnorm"] = np.array([float((p_.detach() - before[n_]).norm())
                                        for n_, p_ in wrap.model.named_parameters()])
        res[tag + "/names"] = np.array([n_ for n_, _ in wrap.model.named_parameters()])

    def start(wrap):
        wrap.optimizer = LegacyAdamW(wrap.model.parameters(), lr=lr)
        return {n_: p_.detach().clone() for n_, p_ in wrap.model.named_parameters()}

    # ---- LevelWrapper: threshold 0.3, check every 2 steps
    lv = _load_old("levels")
    wrap = lv.LevelWrapper("synthetic", "tiny", B, model_init, Tok(), lr, -100, 0.45, 2)
    before = start(wrap)
    rates = []
    orig = wrap.check_success

    def spy(pos, neg):
        pp = pos[:, 0, (rel_id, nrel_id)].softmax(dim=-1)[:, 0]
        pn = neg[:, 0, (rel_id, nrel_id)].softmax(dim=-1)[:, 0]
        rates.append((pp - pn).abs().min().item())
        orig(pos, neg)
        spy.last.append(wrap.success_rate[-1])
    spy.last = []
    wrap.check_success = spy
    loader = Loader()
    logs = wrap.train(loader, total, 2)
    res["level/loss"] = np.array(logs["loss"]["train"], dtype=np.float64)
    res["level/lr"] = np.array(logs["lr"], dtype=np.float64)
    res["level/difficulty"] = np.array(logs["difficulty"]["train"], dtype=np.float64)
    res["level/success"] = np.array(spy.last, dtype=np.float64)
    res["level/margin"] = np.array(rates, dtype=np.float64)
    finish("level", wrap, loader, before)

    # ---- InterpWrapper: 0 -> 1 over 24 samples
    ip = _load_old("interp")
    wrap = ip.InterpWrapper("synthetic", "tiny", B, model_init, Tok(), lr, -100, 0.0, 1.0, 24)
    before = start(wrap)
    loader = Loader()
    logs = wrap.train(loader, total, 2 * B)
    res["interp/loss"] = np.array(logs["loss"]["train"], dtype=np.float64)
    res["interp/lr"] = np.array(logs["lr"], dtype=np.float64)
    res["interp/difficulty"] = np.array(logs["difficulty"]["train"], dtype=np.float64)
    finish("interp", wrap, loader, before)

    # ---- EtaWrapper: eta 0.5, Adam(meta_lr) on the threshold, clamp to [1e-10, 1]
    eb = _load_old("eta_bound")
    eb.EtaWeights = rw.EtaWeights
    wrap = eb.EtaWrapper(0.9, "synthetic", "tiny", B, model_init, Tok(), lr, meta_lr, -100)
    before = start(wrap)
    loader = Loader()
    pm, rm = [], []
    orig_probs, orig_rate = wrap.check_probs, wrap.check_success_rate

    def spy_probs(pos, neg):
        pp = pos[:, 0, (rel_id, nrel_id)].softmax(dim=-1)[:, 0]
        pn = neg[:, 0, (rel_id, nrel_id)].softmax(dim=-1)[:, 0]
        pm.append((pp - pn).abs().min().item())
        orig_probs(pos, neg)

    def spy_rate(loss):
        rm.append((loss - wrap.weights.eta.item()).abs().min().item())
        orig_rate(loss)
    wrap.check_probs, wrap.check_success_rate = spy_probs, spy_rate
    logs = wrap.train(loader, total, 2 * B)
    res["eta/probs_margin"] = np.array(pm, dtype=np.float64)
    res["eta/rate_margin"] = np.array(rm, dtype=np.float64)
    res["eta/loss_main"] = np.array(logs["loss"]["main"], dtype=np.float64)
    res["eta/loss_meta"] = np.array(logs["loss"]["meta"], dtype=np.float64)
    res["eta/lr_main"] = np.array(logs["lr"]["main"], dtype=np.float64)
    res["eta/lr_meta"] = np.array(logs["lr"]["meta"], dtype=np.float64)
    res["eta/eta"] = np.array(logs["eta"], dtype=np.float64)
    res["eta/difficulty"] = np.array(logs["difficulty"], dtype=np.float64)
    res["eta/probs"] = np.array(logs["probs"], dtype=np.float64)
    res["eta/running_rate"] = np.array(wrap.running_rate, dtype=np.float64)
    finish("eta", wrap, loader, before)
    for k_, v in res.items():
        if not k_.endswith("names") and not k_.endswith("dnorm"):
            print(k_, np.asarray(v))
    np.savez_compressed(
        out, dims=np.array([dims.vocab, dims.d_model, dims.d_kv, dims.d_ff, dims.n_layers, dims.n_dec_layers,
                            dims.n_heads]), seed=np.array(seed), pids=pids, nids=nids,
        hp=np.array([B, lr, meta_lr, total, rel_id, nrel_id, n_neg]), **{k_: np.asarray(v) for k_, v in res.items()})


def fake_tokenizer():
    """Deterministic stand-in for T5Tokenizer (SentencePiece is out of scope): word -> 2 + crc32 % 30000, </s> = 1,
    pad-to-longest with 0 — the call surface LCEDataModule.collate uses (lceT5.py:38-54).  Shared by the test."""
    import zlib

    class Enc(dict):
        __getattr__ = dict.__getitem__

    class Tok:
        def __call__(self, texts, padding=True, return_tensors="pt"):
            rows = [[2 + zlib.crc32(w.encode()) % 30000 for w in t.split()] + [1] for t in texts]
            L = max(len(r) for r in rows)
            ids = torch.tensor([r + [0] * (L - len(r)) for r in rows], dtype=torch.int64)
            return Enc(input_ids=ids, attention_mask=(ids != 0).long())
    return Tok()


def datamodule_case(tmpdir):
    """Synthetic candidate file in the format utilities/compute_all_bm25.py:34-49 writes."""
    import json
    rng = np.random.RandomState(77)
    Q, C, D = 12, 50, 400
    recs = []
    for q in range(Q):
        docs = rng.permutation(D)[:C + 1]
        recs.append({"query_id": str(q), "doc_id_a": str(docs[0]), "doc_id_b": [str(x) for x in docs[1:]]})
    path = os.path.join(tmpdir, "cands.json")
    with open(path, "w") as f:
        json.dump(recs, f)

    class Corpus:
        def docs_iter(self):
            for i in range(D):
                yield {"doc_id": str(i), "text": " ".join(f"w{(i * 7 + j) % 97}" for j in range(3 + i % 5))}

        def queries_iter(self):
            for i in range(Q):
                yield {"query_id": str(i), "text": f"what is q{i} about"}
    return path, Corpus()


def gold_datamodule(rt, out):
    """LCEDataModule.setup / collate / train_dataloader + ChangeDifficulty (lceT5.py:13-80) executed unmodified
    (num_workers=0 so the global numpy stream is the one consumed): every batch's token ids and labels."""
    import tempfile
    with tempfile.TemporaryDirectory() as tmp:
        path, corpus = datamodule_case(tmp)
        sys.modules["ir_datasets"].load = lambda name: corpus
        dm = rt.LCEDataModule(data_dir=path, corpus="synthetic", tokenizer=fake_tokenizer(), batch_size=4,
                              shuffle=False, use_max=False, var=0.01, n=3, init_weight=0.3, num_workers=0)
        dm.setup()
        np.random.seed(2024)
        dl = dm.train_dataloader()
        cb = rt.ChangeDifficulty()
        trainer = types.SimpleNamespace(train_dataloader=dl)
        res = {}
        for i, (p, n) in enumerate(dl):
            res[f"p{i}"], res[f"n{i}"] = p["input_ids"].numpy(), n["input_ids"].numpy()
            res[f"pl{i}"], res[f"nl{i}"] = p["labels"].numpy(), n["labels"].numpy()
            cb.on_train_batch_start(trainer, types.SimpleNamespace(difficulty=0.2 + 0.35 * i), (p, n), i)
            res[f"w{i}"] = np.float64(dl.dataset.weight)
        res["after"] = np.float64(np.random.random_sample())
        res["batches"] = np.array(i + 1)
        np.savez_compressed(out, **res)


def main():
    rw, rd, rl, rt, rp = import_reference()
    gold_weights(rw, rt, os.path.join(HERE, "weights_golden.npz"))
    gold_loss(rl, rt, os.path.join(HERE, "loss_golden.npz"))
    gold_select(rd, os.path.join(HERE, "select_golden.npz"))
    gold_table_update(rp, rw, os.path.join(HERE, "table_update_golden.npz"))
    gold_t5(rt, os.path.join(HERE, "t5_tiny_golden.npz"))
    gold_t5_mask(os.path.join(HERE, "t5_mask_golden.npz"))
    gold_train(rt, os.path.join(HERE, "train_golden.npz"))
    if "--wide" in sys.argv or "--all" in sys.argv:   # minutes of CPU: only on request
        gold_train_widths(rt)
        gold_train_dropout(rt)
    gold_newwrapper(rp, rw, os.path.join(HERE, "newwrapper_golden.npz"))
    gold_stdwrapper(rp, rw, os.path.join(HERE, "stdwrapper_golden.npz"))
    gold_scalar_wrappers(rp, rw, os.path.join(HERE, "scalar_wrappers_golden.npz"))
    gold_datamodule(rt, os.path.join(HERE, "datamodule_golden.npz"))
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
