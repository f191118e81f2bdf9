"""Per-sample paced training loops — the host side of pacednegatives/pairwrapper.py on the B200 kernels.

``NewWrapper`` is the live trainer of the reference (pairwrapper.py:208-283): every batch first updates that
batch's row of the per-sample weight table from the no-grad losses (``meta_loop``, :219-239), the row then *is* the
difficulty that picks each query's negative (``PairLoader.get_batch(j, weights[j])``, dataloader.py:46-56), and
``main_loop`` (:241-258) takes one AdamW step on ``mean(ce_p) + mean(ce_n)``.

Same constructor arguments, attributes (``logs``, ``weights``, ``scheduler``) and method names.  ``model_init``
must return a ``pacednegatives_b200.lceT5.MonoT5Model`` (the reference passes a callable returning
``T5ForConditionalGeneration``); everything numeric runs in libpn_b200.so:

* forward / backward: pn_t5_logits, pn_t5_train_forward / pn_t5_train_backward
* per-token cross-entropy: pn_token_ce_fwd / _bwd
* table row update (closed form of the autograd graph at :226-238, negative term only): pn_weights_table_update
* optimiser: pn_t5_adamw_step (transformers.AdamW defaults, :53) under ``get_linear_schedule_with_warmup``

``StdWrapper`` (:91-206) dereferences ``self.meta_model``, which the reference's constructor leaves commented out
(:51), so it cannot run upstream either; it is declared here and raises.
"""
import time
from collections import defaultdict

import torch

from .lceT5 import linear_schedule_with_warmup
from .loss import token_ce
from .weights import Weights

RND = 42


def adjust_lr(optimizer, init_lr, num_warmup_steps, num_training_steps):
    """pairwrapper.py:16-25: returns ``update(step)`` that writes the warmup/linear-decay rate into
    ``optimizer.param_groups`` (any object with that attribute, e.g. ``_Schedule`` below)."""
    def update(step):
        if step < num_warmup_steps:
            lr = init_lr * float(step) / float(max(1, num_warmup_steps))
        else:
            lr = init_lr * max(0.0, float(num_training_steps - step)
                               / float(max(1, num_training_steps - num_warmup_steps)))
        for param_group in optimizer.param_groups:
            param_group['lr'] = lr
    return update


class _Schedule:
    """get_linear_schedule_with_warmup over one learning rate: ``get_last_lr()`` / ``step()``."""

    def __init__(self, init_lr, num_warmup_steps, num_training_steps):
        self.init_lr, self.warmup, self.total = float(init_lr), int(num_warmup_steps), int(num_training_steps)
        self.steps = 0

    def get_last_lr(self):
        return [self.init_lr * linear_schedule_with_warmup(self.steps, self.warmup, self.total)]

    def step(self):
        self.steps += 1


class PacedWrapper:
    """pairwrapper.py:27-89."""

    def __init__(self, dataset, model_name, batch_size, model_init, tokenizer, lr, ignore_index) -> None:
        self.logs = {
            'dataset': dataset,
            'model_name': model_name,
            'batch_size': batch_size,
            'lr': lr,
            'loss': defaultdict(list),
            'avg_weight': defaultdict(list),
        }
        self.device = torch.device('cuda')
        self.ignore_index = ignore_index
        self.loss_fn = lambda logits, target: token_ce(logits, target, self.ignore_index)
        self.model = model_init()
        if not hasattr(self.model, "step") or not hasattr(self.model, "zero_grad"):
            raise TypeError("model_init must return a pacednegatives_b200.lceT5.MonoT5Model")
        self.tokenizer = tokenizer
        self.lr = lr
        self.batch_size = batch_size

    def prep_batch(self, batch):
        """pairwrapper.py:76-89: prompts padded to the batch's longest (mask dropped), labels = the first
        sentencepiece id of 'true' / 'false' as a [B, 1] column."""
        px, nx, o_p, o_n = batch
        px = self.tokenizer(px, padding=True, return_tensors='pt').input_ids.to(self.device)
        nx = self.tokenizer(nx, padding=True, return_tensors='pt').input_ids.to(self.device)
        o_p = self.tokenizer(o_p, padding=True, return_tensors='pt').input_ids[:, 0].view(-1, 1).to(self.device)
        o_n = self.tokenizer(o_n, padding=True, return_tensors='pt').input_ids[:, 0].view(-1, 1).to(self.device)
        return px, nx, o_p, o_n

    def _logits_no_grad(self, x, o):
        logits, _ = self.model.logits(x, o)
        return logits

    def _ce_no_grad(self, x, o):
        logits = self._logits_no_grad(x, o)
        return self.loss_fn(logits.reshape(-1, logits.size(-1)), o.view(-1))


class StdWrapper(PacedWrapper):
    """pairwrapper.py:91-206 — look-ahead meta model; not runnable in the reference (``meta_model`` is never
    created, :51) and needs a double backward through the T5 kernels."""

    def __init__(self, *args, **kw) -> None:
        raise NotImplementedError("StdWrapper needs self.meta_model, which pairwrapper.py:51 never creates; "
                                  "use NewWrapper")


class NewWrapper(PacedWrapper):
    """pairwrapper.py:208-283."""

    def meta_loop(self, j):
        """``weights.forward(idx=j)`` hands out a fresh clone each call (weights.py:87-88), so the reference's
        ``grad(weighted_ce, (v,))`` (:236) only sees the second clone: d/dv = ce_n / B - 1.  The positive
        term never reaches the table, so its no-grad forward (:222-224) is not run here."""
        px, nx, o_p, o_n = self.prep_batch(self.train_loader.get_batch(j, self.weights[j]))
        with torch.no_grad():
            ce_n = self._ce_no_grad(nx, o_n)
            self.weights.meta_update(j, ce_n, self.scheduler.get_last_lr()[0])   # W[j] <- sigmoid(v - lr d/dv)

    def main_loop(self, j):
        px, nx, o_p, o_n = self.prep_batch(self.train_loader.get_batch(j, self.weights[j]))
        if px.shape == nx.shape and o_p.shape == o_n.shape:    # one model pass over both halves of the batch
            logits = self.model(torch.cat([px, nx], 0), torch.cat([o_p, o_n], 0))
            ce = self.loss_fn(logits.reshape(-1, logits.size(-1)), torch.cat([o_p, o_n], 0).view(-1))
            rows = o_p.numel()
            loss = torch.mean(ce[:rows]) + torch.mean(ce[rows:])
        else:
            logits = self.model(px, o_p)
            ce_p = self.loss_fn(logits.reshape(-1, logits.size(-1)), o_p.view(-1))
            logits = self.model(nx, o_n)
            ce_n = self.loss_fn(logits.reshape(-1, logits.size(-1)), o_n.view(-1))
            loss = torch.mean(ce_p) + torch.mean(ce_n)

        self.model.zero_grad()
        loss.backward()
        gscale = self.model.all_reduce_grads()
        self.model.step(self.scheduler.get_last_lr()[0], gscale=gscale)
        self.scheduler.step()
        return loss.item()

    def train(self, train_loader, epochs, warmup_steps, progress=None):
        torch.manual_seed(RND)
        self.weights = Weights((len(train_loader) // self.batch_size, self.batch_size), device=self.device)
        self.train_loader = train_loader
        self.epochs = epochs
        self.scheduler = _Schedule(self.lr, warmup_steps if warmup_steps else len(train_loader) // 4,
                                   epochs * len(train_loader))
        start = time.time()
        for epoch in range(epochs):
            for j in range(len(train_loader) // self.batch_size):
                self.meta_loop(j)
                loss = self.main_loop(j)
                self.logs['loss']['train'].append(loss)
                self.logs['avg_weight'][epoch].append(self.weights[j].mean().item())
                if progress is not None:
                    progress(epoch, j, loss)
        self.logs['time'] = time.time() - start
