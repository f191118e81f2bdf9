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

``StdWrapper`` (:91-206) adds the one-step look-ahead copy (``meta_model``; the reference's constructor leaves it
commented out at :51, so upstream the class raises AttributeError — here it is built): pn_t5_sgd_step.
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
    """pairwrapper.py:91-206 — per-sample weights updated from a one-step LOOK-AHEAD copy of the model.

    The reference never constructs ``self.meta_model`` (:51 is commented out), so upstream this class raises
    AttributeError; here the constructor builds it with ``model_init()``.  What the reference's code then computes
    (pinned by ``tests/golden/stdwrapper_golden.npz`` = the reference's own meta_loop / main_loop, run with that one
    attribute supplied):

    * meta_loop (:102-151): meta_model <- model; theta' = theta - lr_t * d/dtheta [sum(ce_p v)/B + sum(ce_n v)/B]
      (``update_params``; ``nn.Parameter(param - lr * grad)`` makes theta' a LEAF, so despite ``create_graph=True`` no
      second-order term survives, and the later losses run under ``no_grad`` anyway); then
      ``grad(weighted_ce, (v,))`` sees only the LAST clone ``weights.forward`` handed out (the negatives' one):
      d/dv = ce_n(theta') / B - 1, and W[j] <- sigmoid(v - lr_t * d/dv).  With transformers 5.x ``set_param`` on
      'shared.weight' does not reach the tied embedding / head modules, so theta' keeps theta there.
    * main_loop (:153-179): positives through ``self.model``, negatives through ``self.meta_model`` (theta'), loss
      = sum(ce_p v)/B + sum(ce_n v)/B; only the positive term reaches the optimised parameters.
    Upstream the second meta_loop call fails under transformers 5.x (``grad`` over a replaced, now unused
    ``shared.weight``); the golden re-creates the meta model per batch, which is what this class does in effect
    (``load_parameters_from`` overwrites every master value)."""

    def __init__(self, dataset, model_name, batch_size, model_init, tokenizer, lr, ignore_index,
                 lookahead_updates_embedding=False) -> None:
        super().__init__(dataset, model_name, batch_size, model_init, tokenizer, lr, ignore_index)
        self.meta_model = model_init()
        self.lookahead_updates_embedding = bool(lookahead_updates_embedding)

    def _ce(self, model, x, o, grad):
        if grad:
            logits = model(x, o)
        else:
            with torch.no_grad():
                logits, _ = model.logits(x, o)
        return self.loss_fn(logits.reshape(-1, logits.size(-1)), o.view(-1)), logits.shape[-2]

    def meta_loop(self, j):
        px, nx, o_p, o_n = self.prep_batch(self.train_loader.get_batch(j, self.weights[j]))
        self.meta_model.load_parameters_from(self.model)
        v = self.weights.weight.data[j]
        ce, T = self._ce(self.meta_model, px, o_p, True)
        weighted_ce_p = torch.sum(ce * v.view(-1, 1).repeat(1, T).view(-1)) / len(ce)
        ce, T = self._ce(self.meta_model, nx, o_n, True)
        weighted_ce_n = torch.sum(ce * v.view(-1, 1).repeat(1, T).view(-1)) / len(ce)
        self.meta_model.zero_grad()
        (weighted_ce_p + weighted_ce_n).backward()
        lr = self.scheduler.get_last_lr()[0]
        self.meta_model.sgd_step(lr, include_embedding=self.lookahead_updates_embedding)      # theta'
        ce_n, _ = self._ce(self.meta_model, nx, o_n, False)
        self.weights.meta_update(j, ce_n, lr)          # W[j] <- sigmoid(v - lr (ce_n(theta') / B - 1))

    def main_loop(self, j):
        px, nx, o_p, o_n = self.prep_batch(self.train_loader.get_batch(j, self.weights[j]))
        v = self.weights.weight.data[j]
        ce, T = self._ce(self.model, px, o_p, True)
        weighted_ce_p = torch.sum(ce * v.view(-1, 1).repeat(1, T).view(-1)) / len(ce)
        ce, T = self._ce(self.meta_model, nx, o_n, False)
        weighted_ce_n = torch.sum(ce * v.view(-1, 1).repeat(1, T).view(-1)) / len(ce)
        loss = torch.mean(weighted_ce_p + weighted_ce_n)
        self.model.zero_grad()
        loss.backward()
        gscale = self.model.all_reduce_grads()
        self.model.step(self.scheduler.get_last_lr()[0], gscale=gscale)
        self.scheduler.step()
        return loss.item()

    def train(self, train_loader, epochs, warmup_steps, progress=None):
        return NewWrapper.train(self, train_loader, epochs, warmup_steps, progress)


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
