"""The scalar-difficulty paced trainers — host side of the reference's ``old/levels.py``, ``old/interp.py`` and
``old/eta_bound.py`` on the B200 kernels.

All three drive ``LevelLoader.get_batch(j, difficulty)`` (dataloader.py:69-79): ONE difficulty for the whole
batch picks every query's negative at rank ``floor|ceil(difficulty * n_neg)`` (pn_select_level_scalar), and each
step is one AdamW update on ``mean(ce_p) + mean(ce_n)``.  They differ in who moves the difficulty:

* ``LevelWrapper``  (old/levels.py:11-98): every ``heuristic_step_check`` steps, if the mean success rate
  (``check_success``: P(true | positive) > P(true | negative), pn_check_success) since the last check exceeds
  ``success_threshold``, difficulty += 1 / n_neg (capped at 1).
* ``InterpWrapper`` (old/interp.py:20-98): difficulty = interpolate_scalar(start, max, interpolate_steps)(i * B).
* ``EtaWrapper``    (old/eta_bound.py:21-173): difficulty = eta, the self-paced threshold of ``EtaWeights``
  (w = 1 - l/eta below eta), learnt by Adam on ``mean(pce v) + mean(nce v) - sum(v)`` and clamped to [1e-10, 1].

Same constructor arguments, ``logs`` keys and method names as the reference; ``model_init`` must return a
``pacednegatives_b200.lceT5.MonoT5Model``.  wandb / ir_datasets progress bars are not reproduced (control plane).
"""
import time

import torch

from .lceT5 import _AdamWTensor
from .loss import check_success as _check_success
from .pairwrapper import PacedWrapper, RND, _Schedule
from .weights import EtaWeights, spl_weights, ETA


def interpolate_scalar(start_value, end_value, num_steps):
    """old/interp.py:10-18 (Python float arithmetic, same expression order)."""
    step_size = (end_value - start_value) / num_steps

    def get_interpolated_value(step):
        if step > num_steps:
            return end_value
        return start_value + step * step_size

    return get_interpolated_value


class _ScalarPacedWrapper(PacedWrapper):
    """What the three loops share: the gradient step on one (positives, negatives) batch."""

    def __init__(self, dataset, model_name, batch_size, model_init, tokenizer, lr, ignore_index) -> None:
        super().__init__(dataset, model_name, batch_size, model_init, tokenizer, lr, ignore_index)
        self.REL = self.tokenizer.encode('true')[0]
        self.NREL = self.tokenizer.encode('false')[0]

    def _step(self, px, nx, o_p, o_n):
        """logits_p, logits_n, ce_p, ce_n, loss after one optimiser step on mean(ce_p) + mean(ce_n)
        (old/levels.py:44-56, old/interp.py:50-61, old/eta_bound.py:107-124).  Both halves go through the model
        as one batch when their shapes agree (documents are independent)."""
        if px.shape == nx.shape and o_p.shape == o_n.shape:
            rows = px.shape[0]
            o = torch.cat([o_p, o_n], 0)
            logits = self.model(torch.cat([px, nx], 0), o)
            ce = self.loss_fn(logits.reshape(-1, logits.size(-1)), o.view(-1))
            split = o_p.numel()
            ce_p, ce_n = ce[:split], ce[split:]
            logits_p, logits_n = logits[:rows], logits[rows:]
        else:
            logits_p = self.model(px, o_p)
            ce_p = self.loss_fn(logits_p.reshape(-1, logits_p.size(-1)), o_p.view(-1))
            logits_n = self.model(nx, o_n)
            ce_n = self.loss_fn(logits_n.reshape(-1, logits_n.size(-1)), o_n.view(-1))
        loss = torch.mean(ce_p) + torch.mean(ce_n)
        self.model.zero_grad()
        loss.backward()
        gscale = self.model.all_reduce_grads()
        self.model.step(self.scheduler.get_last_lr()[0], gscale=gscale)
        self.scheduler.step()
        return logits_p, logits_n, ce_p.detach(), ce_n.detach(), loss

    def _success(self, pos, neg):
        return _check_success(pos, neg, self.REL, self.NREL).item()


class LevelWrapper(_ScalarPacedWrapper):
    """old/levels.py:11-98."""

    def __init__(self, dataset, model_name, batch_size, model_init, tokenizer, lr, ignore_index,
                 success_threshold, heuristic_step_check) -> None:
        super().__init__(dataset, model_name, batch_size, model_init, tokenizer, lr, ignore_index)
        self.difficulty = 0.0
        self.success_threshold = success_threshold
        self.heuristic_step_check = heuristic_step_check
        self.success_rate = []
        self.logs['difficulty'] = {'train': []}
        self.logs['lr'] = []

    def check_success(self, pos, neg):
        self.success_rate.append(self._success(pos, neg))

    def main_loop(self, j):
        px, nx, o_p, o_n = self.prep_batch(self.train_loader.get_batch(j, self.difficulty))
        logits_p, logits_n, _, _, loss = self._step(px, nx, o_p, o_n)
        self.check_success(logits_p, logits_n)
        return loss.item()

    def train(self, train_loader, total_steps, warmup_steps, progress=None):
        torch.manual_seed(RND)
        self.train_loader = train_loader
        self.total_steps = total_steps
        self.scheduler = _Schedule(self.lr, warmup_steps if warmup_steps else total_steps // 100, total_steps)
        start = time.time()
        for i in range(total_steps // self.train_loader.batch_size):
            if i % self.heuristic_step_check == 0 and len(self.success_rate) > 0:
                agg_success_rate = sum(self.success_rate) / len(self.success_rate)
                if agg_success_rate > self.success_threshold:
                    self.difficulty = min(1.0, self.difficulty + 1 / train_loader.dataset.n_neg)
                self.success_rate = []
            loss = self.main_loop(i)
            self.logs['loss']['train'].append(loss)
            self.logs['success_rate'] = self.success_rate[-1]
            self.logs['lr'].append(self.scheduler.get_last_lr()[0])
            self.logs['difficulty']['train'].append(self.difficulty)
            if progress is not None:
                progress(i, loss)
        self.logs['time'] = time.time() - start
        return self.logs


class InterpWrapper(_ScalarPacedWrapper):
    """old/interp.py:20-98."""

    def __init__(self, dataset, model_name, batch_size, model_init, tokenizer, lr, ignore_index,
                 start_difficulty, max_difficulty, interpolate_steps) -> None:
        super().__init__(dataset, model_name, batch_size, model_init, tokenizer, lr, ignore_index)
        self.difficulty = start_difficulty
        self.start_difficulty = start_difficulty
        self.max_difficulty = max_difficulty
        self.interpolate_steps = interpolate_steps
        self.interpolate_difficulty = interpolate_scalar(start_difficulty, max_difficulty, interpolate_steps)
        self.logs['difficulty'] = {'train': []}
        self.logs['lr'] = []

    def main_loop(self, j):
        px, nx, o_p, o_n = self.prep_batch(self.train_loader.get_batch(j, self.difficulty))
        return self._step(px, nx, o_p, o_n)[4].item()

    def train(self, train_loader, total_steps, warmup_steps, progress=None):
        torch.manual_seed(RND)
        self.train_loader = train_loader
        self.total_steps = total_steps
        B = self.train_loader.batch_size
        self.scheduler = _Schedule(self.lr, warmup_steps // B if warmup_steps else (total_steps // 100), total_steps)
        start = time.time()
        for i in range(total_steps // B):
            loss = self.main_loop(i)
            self.difficulty = self.interpolate_difficulty(i * B)
            self.logs['loss']['train'].append(loss)
            self.logs['lr'].append(self.scheduler.get_last_lr()[0])
            self.logs['difficulty']['train'].append(self.difficulty)
            if progress is not None:
                progress(i, loss)
        self.logs['time'] = time.time() - start
        return self.logs


class EtaWrapper(_ScalarPacedWrapper):
    """old/eta_bound.py:21-173.  The meta objective's gradient flows to eta through pn_weights_bwd; its
    optimiser is torch.optim.Adam in the reference (eps 1e-8, no weight decay) = the AdamW kernel at wd 0."""

    def __init__(self, eta, dataset, model_name, batch_size, model_init, tokenizer, lr, meta_lr,
                 ignore_index) -> None:
        super().__init__(dataset, model_name, batch_size, model_init, tokenizer, lr, ignore_index)
        self.weights = EtaWeights(eta, device=self.device, min=0. + 1e-10, max=1.)
        self.logs['eta'] = []
        self.logs['loss'] = {'main': [], 'meta': []}
        self.logs['difficulty'] = []
        self.logs['lr'] = {'main': [], 'meta': []}
        self.logs['success_rate'] = []
        self.logs['probs'] = []
        self.running_rate = []
        self.meta_lr = meta_lr
        self.meta_optimizer = _AdamWTensor(self.weights.eta, lr=self.meta_lr, eps=1e-8, weight_decay=0.0)

    def check_probs(self, pos, neg):
        self.logs['probs'].append(self._success(pos, neg))

    def check_success_rate(self, loss):
        # mean(loss < eta): count from the weighting kernel's statistics (stats[3] = #(l < eta))
        _, _, stats = spl_weights(loss, self.weights.eta, 0.0, ETA, False, True)
        self.running_rate.append(float(stats[3].item()) / max(loss.numel(), 1))

    def meta_loop(self, j):
        px, nx, o_p, o_n = self.prep_batch(self.train_loader.get_batch(j, self.difficulty))
        with torch.no_grad():
            if px.shape == nx.shape and o_p.shape == o_n.shape:
                ce = self._ce_no_grad(torch.cat([px, nx], 0), torch.cat([o_p, o_n], 0))
                pce, nce = ce[:o_p.numel()], ce[o_p.numel():]
            else:
                pce, nce = self._ce_no_grad(px, o_p), self._ce_no_grad(nx, o_n)
        ce = torch.div(pce + nce, 2)
        v = self.weights.forward(loss=ce)
        weighted_ce = torch.mean(pce * v) + torch.mean(nce * v) - torch.sum(v)
        self.meta_optimizer.zero_grad()
        weighted_ce.backward()
        self.meta_optimizer.step(lr=self.meta_scheduler.get_last_lr()[0])
        self.weights.clamp()
        self.meta_scheduler.step()
        self.logs['loss']['meta'].append(weighted_ce.item())
        return weighted_ce.item()

    def main_loop(self, j):
        px, nx, o_p, o_n = self.prep_batch(self.train_loader.get_batch(j, self.difficulty))
        plogits, nlogits, pce, nce, loss = self._step(px, nx, o_p, o_n)
        self.check_probs(plogits, nlogits)
        self.check_success_rate(torch.div(pce + nce, 2))
        return loss.item()

    def train(self, train_loader, total_steps, warmup_steps, progress=None):
        torch.manual_seed(RND)
        self.train_loader = train_loader
        self.total_steps = total_steps
        self.difficulty = 0.0
        B = self.train_loader.batch_size
        warm = warmup_steps // B if warmup_steps else (total_steps // 100)
        self.scheduler = _Schedule(self.lr, warm, total_steps)
        self.meta_scheduler = _Schedule(self.meta_lr, warm, total_steps)
        start = time.time()
        for i in range(total_steps // B):
            meta_loss = self.meta_loop(i)
            loss = self.main_loop(i)
            self.logs['loss']['main'].append(loss)
            self.logs['lr']['main'].append(self.scheduler.get_last_lr()[0])
            self.logs['lr']['meta'].append(self.meta_scheduler.get_last_lr()[0])
            self.logs['difficulty'].append(self.difficulty)
            self.logs['eta'].append(self.weights.eta.item())
            self.difficulty = self.weights.eta.item()
            if progress is not None:
                progress(i, loss, meta_loss)
        self.logs['time'] = time.time() - start
        return self.logs
