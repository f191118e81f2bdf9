"""The LCE training step on the B200 — host-side mirror of ``pacednegatives/lceT5.py``.

``LCEModel`` keeps the reference's interface for the hot path (``pair_loss`` lceT5.py:116-123,
``training_step`` lceT5.py:124-167, ``configure_optimizers`` lceT5.py:170-183) with the same hparams names;
the model is ``MonoT5Model`` (libpn_b200.so: tcgen05 forward + backward, flat fp32 parameters, AdamW kernel)
instead of ``transformers.T5ForConditionalGeneration`` + autograd + ``transformers.AdamW``.  A batch is the
``(p, n)`` pair of ``{'input_ids', 'labels'}`` dicts that ``LCEDataModule.collate`` returns (lceT5.py:38-53);
``LCEDataModule`` / ``ChangeDifficulty`` (lceT5.py:13-80) are mirrored as plain classes (Lightning itself, the
SentencePiece tokenizer and ir_datasets are out of scope: pass any tokenizer callable and a corpus object).

Data parallel (Lightning DDP in the reference): every rank runs the same step on its own queries; the flat
gradient buffer and eta's gradient are all-reduced (mean) once per step over ``torch.distributed``.
"""
from collections import OrderedDict
from types import SimpleNamespace

import torch
import torch.distributed as dist

from . import _lib
from .loss import pair_loss as _pair_loss
from .scorer import MonoT5Scorer
from .weights import LCEWeights


class LCEDataModule:
    """lceT5.py:13-67.  Same constructor arguments and methods (``setup``, ``collate``, ``train_dataloader``).
    ``corpus`` is an object with ``docs_iter()`` / ``queries_iter()`` (what ``ir_datasets.load(name)`` returns); a
    string is resolved through ir_datasets when that package is importable.  The candidate file is the
    ``[{query_id, doc_id_a, doc_id_b[C]}]`` json utilities/compute_all_bm25.py writes; it becomes the dense int32
    [Q, C] candidate matrix in HBM that pn_select_binomial samples from (dataloader.LCEDataset).  The sampler runs
    on this process's GPU, so batches are produced in the main process: ``num_workers`` is accepted and ignored
    (forked DataLoader workers cannot use the CUDA context)."""

    def __init__(self, data_dir: str = "path/to/dir", corpus="msmarco-passage", tokenizer=None,
                 batch_size: int = 32, shuffle=False, use_max=False, var=0.01, n=2, init_weight=0. + 1e-10,
                 num_workers=4):
        self.data_dir = data_dir
        if isinstance(corpus, str):
            try:
                import ir_datasets as irds
            except ImportError as e:
                raise RuntimeError("LCEDataModule: pass a corpus object with docs_iter()/queries_iter(); "
                                   f"ir_datasets is not installed to resolve {corpus!r}") from e
            corpus = irds.load(corpus)
        self.corpus = corpus
        self.tokenizer = tokenizer
        self.batch_size = batch_size
        self.shuffle = shuffle
        self.use_max = use_max
        self.var = var
        self.n = n
        self.num_workers = num_workers
        self.weight = init_weight

    def collate(self):
        tokenizer = self.tokenizer

        def collate_fn(batch):
            pos = [x['pos'] for x in batch]
            neg = [d for x in batch for d in x['neg']]      # query-major, lceT5.py:43
            pos = tokenizer(pos, padding=True, return_tensors='pt')
            neg = tokenizer(neg, padding=True, return_tensors='pt')
            plab = tokenizer(['true'] * len(pos['input_ids']), padding=True, return_tensors='pt')['input_ids']
            nlab = tokenizer(['false'] * len(neg['input_ids']), padding=True, return_tensors='pt')['input_ids']
            return ({'input_ids': pos['input_ids'], 'labels': plab}, {'input_ids': neg['input_ids'], 'labels': nlab})
        return collate_fn

    def setup(self, stage: str = None):
        import pandas as pd
        from .dataloader import LCEDataset
        with open(self.data_dir, 'r') as f:
            dataset = pd.read_json(f, orient='records', dtype={'query_id': str, 'doc_id_a': str})
        if self.shuffle:
            dataset = dataset.sample(frac=1).reset_index(drop=True)
        self.pairs = dataset[['query_id', 'doc_id_a']].values.tolist()
        self.neg_idx = [[str(d) for d in row] for row in dataset['doc_id_b'].values]
        self.dataset = LCEDataset(self.pairs, self.neg_idx, self.corpus, self.tokenizer, self.weight,
                                  self.batch_size, var=self.var, n=self.n, min=0. + 1e-10, max=1.0 - 1e-10,
                                  use_max=self.use_max)

    def train_dataloader(self):
        from torch.utils.data import DataLoader
        return DataLoader(self.dataset, batch_size=self.batch_size, num_workers=0, collate_fn=self.collate(),
                          pin_memory=True)


class ChangeDifficulty:
    """lceT5.py:77-80: before each batch the dataset's sampling mean follows the model's eta."""

    def on_train_batch_start(self, trainer, pl_module, batch, batch_idx):
        trainer.train_dataloader.dataset.weight = min(1 - 1e-10, pl_module.difficulty)


def linear_schedule_with_warmup(step, warmup_steps, total_steps):
    """lr multiplier of transformers.get_linear_schedule_with_warmup after `step` scheduler steps."""
    if step < warmup_steps:
        return float(step) / float(max(1, warmup_steps))
    return max(0.0, float(total_steps - step) / float(max(1, total_steps - warmup_steps)))


def all_reduce_mean_scale(t, group=None):
    """DDP's gradient averaging on a flat buffer: SUM all-reduce in place, returns the 1/world factor the
    optimiser kernel still has to apply (so the mean costs no extra pass)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        return 1.0 / dist.get_world_size(group)
    return 1.0


class _ModelFn(torch.autograd.Function):
    """logits = model(input_ids=, labels=).logits with the backward pass wired to pn_t5_train_backward."""

    @staticmethod
    def forward(ctx, anchor, model, ids, labels):
        N, S = ids.shape
        T = labels.shape[1]
        lib = _lib.lib()
        model._apply_dropout_state()
        need = int(lib.pn_t5_train_workspace_bytes(model._h, N, S, T))
        ws = torch.empty(need, dtype=torch.uint8, device="cuda")
        logits = torch.empty(N * T, model.vocab_pad, dtype=torch.float32, device="cuda")
        _lib.check(lib.pn_t5_train_forward(model._h, _lib.ptr(ids), _lib.ptr(labels), N, S, T, _lib.ptr(logits),
                                           _lib.ptr(ws), need, _lib.stream()))
        if ctx.needs_input_grad[0]:      # under torch.no_grad() (the train-mode meta forward) nothing is kept
            ctx.model, ctx.ws, ctx.ids, ctx.shape = model, ws, ids, (N, S, T)
            model._pending += 1
        return logits.view(N, T, model.vocab_pad)

    @staticmethod
    def backward(ctx, dlogits):
        N, S, T = ctx.shape
        model = ctx.model
        d = dlogits.contiguous().float()
        # gradient ranges may leave for their all-reduce as they complete only if this backward is the step's last
        model._hook_live = model._pending == 1 and not model.accumulating and model._overlap_ready()
        model._pending -= 1
        try:
            _lib.check(_lib.lib().pn_t5_train_backward(model._h, _lib.ptr(ctx.ids), _lib.ptr(d), N, S, T,
                                                       _lib.ptr(ctx.ws), ctx.ws.numel(), _lib.stream()))
        finally:
            model._hook_live = False
        if model._hook_error is not None:
            err, model._hook_error = model._hook_error, None
            raise err
        ctx.ws = None
        return torch.zeros_like(model._anchor), None, None, None


class _ModelZ2Fn(torch.autograd.Function):
    """z2 = model(input_ids=, labels=).logits[..., (rel, nrel)] through the two-row head (pn_t5_train_forward_z2), the
    backward wired to pn_t5_train_backward_z2: the [N*T, vocab] logits and their gradient never exist."""

    @staticmethod
    def forward(ctx, anchor, model, ids, labels):
        N, S = ids.shape
        T = labels.shape[1]
        lib = _lib.lib()
        model._apply_dropout_state()
        need = int(lib.pn_t5_train_workspace_bytes(model._h, N, S, T))
        ws = torch.empty(need, dtype=torch.uint8, device="cuda")
        z2 = torch.empty(N * T, 2, dtype=torch.float32, device="cuda")
        _lib.check(lib.pn_t5_train_forward_z2(model._h, _lib.ptr(ids), _lib.ptr(labels), N, S, T, model.rel_id,
                                              model.nrel_id, _lib.ptr(z2), _lib.ptr(ws), need, _lib.stream()))
        if ctx.needs_input_grad[0]:
            ctx.model, ctx.ws, ctx.ids, ctx.shape = model, ws, ids, (N, S, T)
            model._pending += 1
        return z2.view(N, T, 2)

    @staticmethod
    def backward(ctx, dz2):
        N, S, T = ctx.shape
        model = ctx.model
        d = dz2.contiguous().float()
        model._hook_live = model._pending == 1 and not model.accumulating and model._overlap_ready()
        model._pending -= 1
        try:
            _lib.check(_lib.lib().pn_t5_train_backward_z2(model._h, _lib.ptr(ctx.ids), _lib.ptr(d), model.rel_id,
                                                          model.nrel_id, N, S, T, _lib.ptr(ctx.ws), ctx.ws.numel(),
                                                          _lib.stream()))
        finally:
            model._hook_live = False
        if model._hook_error is not None:
            err, model._hook_error = model._hook_error, None
            raise err
        ctx.ws = None
        return torch.zeros_like(model._anchor), None, None, None


class MonoT5Model(MonoT5Scorer):
    """Trainable monoT5: ``model(input_ids, labels)`` returns logits [N, T, vocab] that backpropagate into the
    flat gradient buffer ``grads`` (HF state_dict order, see ``named_offsets``)."""

    def __init__(self, dims, state_dict, **kw):
        super().__init__(dims, state_dict, train=True, **kw)
        lib = _lib.lib()
        n = int(lib.pn_t5_num_params(self._h))
        self.grads = torch.zeros(n, dtype=torch.float32, device="cuda")
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device="cuda")
        self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device="cuda")
        _lib.check(lib.pn_t5_set_train_buffers(self._h, _lib.ptr(self.grads), _lib.ptr(self.exp_avg),
                                               _lib.ptr(self.exp_avg_sq)))
        self._anchor = torch.zeros(1, device="cuda", requires_grad=True)
        self.opt_steps = 0
        # train-mode dropout of the next forward(s): rate, seed of the forward call, index of the call's first
        # document inside the step's batch (csrc/pn_dropout.cuh).  0.0 = eval-mode arithmetic.
        self.dropout_rate, self.dropout_seed, self.dropout_doc_offset = 0.0, 0, 0
        # gradient all-reduce overlapped with the backward pass (DDP's bucketing): libpn_b200 calls _on_grad_range
        # while it enqueues the backward; ranges go out on a side stream behind an event
        self.overlap_grad_reduce = True
        self.accumulating = False    # True while more micro-batches of this step will still add to `grads`
        self.reduce_group = None
        self.skip_grad_reduce = False   # measurement only (bench.py): run the step without the gradient exchange
        self._pending, self._hook_live, self._hook_error = 0, False, None
        self._reduced, self._comm_stream, self._comm_done = [], None, None
        import ctypes as C
        import weakref
        me = weakref.ref(self)   # the C callback must not keep the model alive

        def hook(_user, off, numel):
            obj = me()
            if obj is not None:
                obj._on_grad_range(off, numel)
        self._hook_c = C.CFUNCTYPE(None, C.c_void_p, C.c_int64, C.c_int64)(hook)
        _lib.check(lib.pn_t5_set_grad_hook(self._h, C.cast(self._hook_c, C.c_void_p), None))

    def _apply_dropout_state(self):
        _lib.check(_lib.lib().pn_t5_set_dropout(self._h, float(self.dropout_rate), int(self.dropout_seed) & (2 ** 64 - 1),
                                                int(self.dropout_doc_offset)))

    def __call__(self, input_ids, labels, attention_mask=None):
        if attention_mask is not None and not bool(torch.as_tensor(attention_mask).all()):
            raise NotImplementedError("padded batches are not supported: every prompt must fill its S tokens")
        ids = self._ids(input_ids)
        lab = torch.as_tensor(labels).to(device="cuda", dtype=torch.int64).contiguous()
        out = _ModelFn.apply(self._anchor, self, ids, lab)
        return out[:, :, :self.dims.vocab]

    def z2(self, input_ids, labels=None):
        """Differentiable two-row head: fp32 [N, T, 2] = (z_true, z_false) at each label position (T = 1 when
        ``labels`` is None: the single decoder step of the re-rank score)."""
        ids = self._ids(input_ids)
        if labels is None:
            lab = torch.zeros(ids.shape[0], 1, dtype=torch.int64, device="cuda")
        else:
            lab = torch.as_tensor(labels).to(device="cuda", dtype=torch.int64).contiguous()
        return _ModelZ2Fn.apply(self._anchor, self, ids, lab)

    def named_grads(self):
        return {n: self.grads[o:o + k] for n, o, k in self.named_offsets()}

    def load_parameters_from(self, other):
        """``self.load_state_dict(other.state_dict())`` (pairwrapper.py:105): one device copy of the flat master
        buffer, then the packed bf16 operands are rebuilt."""
        self.parameters_flat().copy_(other.parameters_flat())
        _lib.check(_lib.lib().pn_t5_finalize(self._h, _lib.stream()))

    def sgd_step(self, lr, include_embedding=False):
        """theta <- theta - lr * grad on this model's own gradient buffer (StdWrapper.update_params,
        pairwrapper.py:68-73).  ``include_embedding=False`` leaves shared.weight alone: with transformers 5.x the
        reference's ``set_param`` replaces ``model.shared.weight`` only, which no forward path reads (the embedding
        and head modules keep the tied original), so its look-ahead model differs from theta everywhere BUT there."""
        first = 0
        if not include_embedding:
            offs = self.named_offsets()
            first = offs[0][2] if offs[0][0] == "shared.weight" else 0
        _lib.check(_lib.lib().pn_t5_sgd_step(self._h, float(lr), first, _lib.stream()))

    def zero_grad(self):
        _lib.check(_lib.lib().pn_t5_zero_grad(self._h, _lib.stream()))

    def _overlap_ready(self):
        return (self.overlap_grad_reduce and not self.skip_grad_reduce and dist.is_available() and dist.is_initialized()
                and dist.get_world_size(self.reduce_group) > 1 and not self._reduced)

    def _on_grad_range(self, off, numel):
        """Host callback from pn_t5_train_backward: gradients [off, off+numel) are final once the work enqueued so
        far has run.  SUM all-reduce of that slice on the side stream, behind an event on the compute stream."""
        if not self._hook_live or numel <= 0:
            return
        try:
            if self._comm_stream is None:
                self._comm_stream = torch.cuda.Stream()
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream())
            self._comm_stream.wait_event(ev)
            with torch.cuda.stream(self._comm_stream):
                dist.all_reduce(self.grads[off:off + numel], op=dist.ReduceOp.SUM, group=self.reduce_group)
                self._comm_done = torch.cuda.Event()
                self._comm_done.record(self._comm_stream)
            self._reduced.append((int(off), int(numel)))
        except Exception as e:   # never unwind through the C frame
            self._hook_error = e

    def all_reduce_grads(self, group=None):
        """DDP's gradient mean; returns the scale still to apply (folded into the AdamW kernel).  Ranges already
        sent from inside the backward pass are only waited for; anything else is reduced here in one call."""
        if self._reduced:
            done = sorted(self._reduced)
            self._reduced = []
            covered = 0
            for off, n in done:
                if off != covered:
                    break
                covered = off + n
            torch.cuda.current_stream().wait_event(self._comm_done)
            if covered != self.grads.numel():
                raise RuntimeError(f"overlapped gradient all-reduce covered {covered} of {self.grads.numel()} elements")
            return 1.0 / dist.get_world_size(self.reduce_group)
        if self.skip_grad_reduce:
            return 1.0
        return all_reduce_mean_scale(self.grads, group if group is not None else self.reduce_group)

    def step(self, lr, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.0, gscale=1.0):
        """transformers.AdamW defaults (the optimiser of lceT5.py:172)."""
        self.opt_steps += 1
        if self._reduced:   # ranges sent from inside the backward pass that nobody waited for
            torch.cuda.current_stream().wait_event(self._comm_done)
            self._reduced = []
        self._pending = 0
        _lib.check(_lib.lib().pn_t5_adamw_step(self._h, float(lr), betas[0], betas[1], eps, weight_decay,
                                               self.opt_steps, float(gscale), _lib.stream()))


class _HParams(dict):
    """Lightning's AttributeDict for the plain-class build: hparams.n and hparams['n'] both work."""
    __getattr__ = dict.__getitem__
    __setattr__ = dict.__setitem__


class _AdamWTensor(torch.optim.Optimizer):
    """transformers.AdamW (eps 1e-6, no weight decay, bias-corrected: lceT5.py:171) on small device tensors — the
    eta parameter — through the pn_adamw kernel.  A real torch Optimizer, so transformers'
    get_linear_schedule_with_warmup (a LambdaLR) and Lightning's optimizer plumbing accept it."""

    def __init__(self, params, lr, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.0):
        if torch.is_tensor(params):
            params = [params]
        super().__init__(params, dict(lr=float(lr), betas=betas, eps=eps, weight_decay=weight_decay))

    @torch.no_grad()
    def step(self, closure=None, lr=None):
        loss = closure() if closure is not None else None
        for g in self.param_groups:
            for p in g["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if not st:
                    st["step"], st["m"], st["v"] = 0, torch.zeros_like(p.data), torch.zeros_like(p.data)
                st["step"] += 1
                grad = p.grad.contiguous().float()
                _lib.check(_lib.lib().pn_adamw(_lib.ptr(p.data), _lib.ptr(grad), _lib.ptr(st["m"]), _lib.ptr(st["v"]),
                                               p.numel(), g["lr"] if lr is None else float(lr), g["betas"][0],
                                               g["betas"][1], g["eps"], g["weight_decay"], st["step"], 1.0,
                                               _lib.stream()))
        return loss


class _ModelAdamW(torch.optim.Optimizer):
    """transformers.AdamW over the model's flat fp32 parameter buffer (lceT5.py:172) = one pn_t5_adamw_step launch
    (+ the bf16 operand repack).  The optimiser's single parameter is a tensor VIEW of the library's master buffer
    whose .grad is the flat gradient buffer, so torch-side tools (LambdaLR, LearningRateMonitor, gradient
    inspection) see real tensors; the data-parallel gradient mean is finished here (``all_reduce_grads``)."""

    def __init__(self, model, lr, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.0):
        self.model = model
        flat = model.parameters_flat().requires_grad_(True)
        flat.grad = model.grads
        super().__init__([flat], dict(lr=float(lr), betas=betas, eps=eps, weight_decay=weight_decay))

    def zero_grad(self, set_to_none=False):
        self.model.zero_grad()     # the buffer stays: the library accumulates into it

    def all_reduce_grads(self, group=None):
        return self.model.all_reduce_grads(group)

    @torch.no_grad()
    def step(self, closure=None, lr=None, gscale=None):
        loss = closure() if closure is not None else None
        g = self.param_groups[0]
        if gscale is None:
            gscale = self.model.all_reduce_grads()
        self.model.step(g["lr"] if lr is None else float(lr), g["betas"], g["eps"], g["weight_decay"], gscale)
        return loss


def _linear_schedule(optimizer, num_warmup_steps, num_training_steps):
    """transformers.get_linear_schedule_with_warmup (lceT5.py:174-179); the same LambdaLR built here when
    transformers is not importable."""
    try:
        from transformers import get_linear_schedule_with_warmup
        return get_linear_schedule_with_warmup(optimizer, num_warmup_steps=num_warmup_steps,
                                               num_training_steps=num_training_steps)
    except ImportError:
        from torch.optim.lr_scheduler import LambdaLR
        return LambdaLR(optimizer, lambda s: linear_schedule_with_warmup(s, num_warmup_steps, num_training_steps))


class TokenCELoss:
    """``nn.CrossEntropyLoss(ignore_index=..., reduction='none')`` (lceT5.py:110) as one streaming kernel per row
    (pn_token_ce_fwd / _bwd): ``loss_fn(logits[rows, V], target[rows]) -> [rows]``."""

    def __init__(self, ignore_index=-100, reduction="none"):
        if reduction != "none":
            raise ValueError("the LCE path uses reduction='none' (lceT5.py:110)")
        self.ignore_index, self.reduction = int(ignore_index), reduction

    def __call__(self, logits, target):
        from .loss import token_ce
        return token_ce(logits, target, self.ignore_index)


try:     # the reference's base class, when the package exists (it is not in this image: plain class then)
    import lightning as _pl
    _LightningBase = _pl.LightningModule
except ImportError:
    _pl, _LightningBase = None, object


class LCEModel(_LightningBase):
    """Drop-in for ``pacednegatives.lceT5.LCEModel`` (lceT5.py:101-182): ``LCEModel(hparams)`` with the reference's
    keys — model_name, total_steps, eta, batch_size, lr, meta_lr, n, warmup_steps, use_mean, ignore_index.

    ``model_name`` is a LOCAL HuggingFace checkpoint directory (no hub access): the model comes from
    ``MonoT5Model.from_pretrained`` and, when the directory holds SentencePiece files, ``self.tokenizer`` from
    ``T5Tokenizer.from_pretrained``.  Without ``model_name`` pass ``dims=`` + ``state_dict=`` (or ``model=``).
    Attributes as in the reference: ``.model``, ``.tokenizer``, ``.weights`` (``.eta``), ``.loss_fn``,
    ``.difficulty``, ``.hparams``, ``automatic_optimization = False``.  Subclasses ``lightning.LightningModule`` when
    that package is importable (manual optimisation: ``self.optimizers()``, ``self.lr_schedulers()``,
    ``self.manual_backward``), a plain class with the same methods otherwise.

    Extra hparams (all optional): ``micro_queries`` (walk the batch in query groups), ``dropout_rate``.
    ``elide_no_grad_forward`` (default: True when dropout is 0) reuses the gradient-tracking forward's logits for the
    meta step: without dropout the reference's no-grad forward (lceT5.py:131-133) computes exactly the same numbers."""

    def __init__(self, hparams, dims=None, state_dict=None, model=None, elide_no_grad_forward=None, tokenizer=None,
                 **model_kw):
        if _pl is not None:
            super().__init__()
        if not isinstance(hparams, dict):
            hparams = dict(vars(hparams))
        if _pl is not None:
            for key in hparams.keys():
                self.hparams[key] = hparams[key]
        else:
            self.hparams = _HParams(hparams)
        self.hparams.setdefault("ignore_index", -100)
        self.hparams.setdefault("use_mean", False)
        self.hparams.setdefault("warmup_steps", 0)
        self.hparams.setdefault("meta_lr", self.hparams.get("lr"))
        name = self.hparams.get("model_name")
        if model is not None:
            self.model = model
        elif dims is not None:
            self.model = MonoT5Model(dims, state_dict, **model_kw)
        elif name is not None:
            self.model = MonoT5Model.from_pretrained(name, **model_kw)
        else:
            raise ValueError("LCEModel: give hparams['model_name'] (a local checkpoint directory) or dims= + state_dict=")
        self.tokenizer = tokenizer
        if tokenizer is None and name is not None and model is None and dims is None:
            self.tokenizer = _load_tokenizer(name)
        self.weights = LCEWeights(self.hparams)
        self.loss_fn = TokenCELoss(ignore_index=int(self.hparams.ignore_index), reduction="none")
        self.automatic_optimization = False
        if _pl is not None:
            self.save_hyperparameters()
        # dropout: hparams['dropout_rate'], else the checkpoint's config.json value when the model came from
        # model_name (t5-* / monot5-*: 0.1, what the reference trains with under Lightning's train mode), else 0
        cfg_base = getattr(self.model, "hf_config_base", None) or {}
        self.dropout_rate = float(self.hparams.get("dropout_rate", cfg_base.get("dropout_rate", 0.0) if name else 0.0))
        self.dropout_seed = int(self.hparams.get("seed", 0))
        self.model.dropout_rate = self.dropout_rate
        if self.dropout_rate > 0.0 and elide_no_grad_forward:
            raise ValueError("elide_no_grad_forward needs dropout_rate 0: with dropout the two forwards of "
                             "lceT5.py:131-133 / 149-150 see different masks")
        self.elide = (bool(elide_no_grad_forward) if elide_no_grad_forward is not None else True) \
            and self.dropout_rate == 0.0
        self.difficulty = self.weights.eta.item()
        self._pn_step = 0
        self.logged = {}
        self._opts = self._scheds = None

    # ---- the slice of the LightningModule interface training_step uses (plain-class build)
    if _pl is None:
        @property
        def global_step(self):
            return self._pn_step

        def log(self, name, value, *a, **kw):
            self.logged[name] = value

        def optimizers(self):
            if self._opts is None:
                self._opts, self._scheds = self.configure_optimizers()
            return self._opts

        def lr_schedulers(self):
            self.optimizers()
            return self._scheds

        def manual_backward(self, loss):
            loss.backward()
    else:
        def log(self, name, value, *a, **kw):
            self.logged[name] = value
            if getattr(self, "_trainer", None) is not None:
                super().log(name, value, *a, **kw)

    def pair_loss(self, p, n, op, on):
        return _pair_loss(p, n, op, on, use_mean=bool(self.hparams.use_mean),
                          ignore_index=int(self.hparams.ignore_index))

    def configure_optimizers(self):
        meta_opt = _AdamWTensor(self.weights.parameters(), lr=self.hparams.meta_lr)
        opt = _ModelAdamW(self.model, lr=self.hparams.lr)
        meta_sched = _linear_schedule(meta_opt, int(self.hparams.warmup_steps), int(self.hparams.total_steps))
        sched = _linear_schedule(opt, int(self.hparams.warmup_steps), int(self.hparams.total_steps))
        return [meta_opt, opt], [meta_sched, sched]

    def _forward_pn(self, p, n, k, V, doc_offset=0):
        """`self.model(**p).logits`, `self.model(**n).logits` (lceT5.py:149-150).  Documents are independent, so
        when both batches have the same sequence and label lengths they go through the model as ONE batch (half
        the kernel launches, fuller GEMM tiles); the logits are the same numbers either way."""
        pi, ni = torch.as_tensor(p["input_ids"]), torch.as_tensor(n["input_ids"])
        pl, nl = torch.as_tensor(p["labels"]), torch.as_tensor(n["labels"])
        Q = pi.shape[0]
        # dropout masks are a function of the document's index in the step's batch: positives doc_offset ..,
        # then the negatives (the same numbering whether the two halves run as one batch or two)
        self.model.dropout_doc_offset = doc_offset
        if pi.shape[1] == ni.shape[1] and pl.shape[1] == nl.shape[1]:
            dev = "cuda"
            logits = self.model(torch.cat([pi.to(dev), ni.to(dev)]), torch.cat([pl.to(dev), nl.to(dev)]))
            return logits[:Q], logits[Q:].reshape(-1, k, V)
        plog = self.model(pi, pl)
        self.model.dropout_doc_offset = doc_offset + Q
        nlog = self.model(ni, nl).view(-1, k, V)
        self.model.dropout_doc_offset = doc_offset
        return plog, nlog

    def _forward_seed(self, which):
        """seed of forward `which` (0 = the no-grad meta forward, 1 = the gradient forward) of this step"""
        self.model.dropout_seed = (self.dropout_seed << 32) + 2 * self._pn_step + which

    def _no_grad_logits(self, p, n, k, V, doc_offset):
        """`with torch.no_grad(): self.model(**p) ... self.model(**n)` (lceT5.py:131-133).  Lightning keeps the module
        in train mode there, so with dropout this is a train-mode forward with its own masks (nothing saved for a
        backward); without dropout it is the scorer's forward."""
        if self.dropout_rate > 0.0:
            return self._forward_pn(p, n, k, V, doc_offset)
        p0, _ = self.model.logits(p["input_ids"], p["labels"])
        n0, _ = self.model.logits(n["input_ids"], n["labels"])
        return p0, n0.reshape(-1, k, V)

    @staticmethod
    def _raw(opt):
        """the optimiser behind Lightning's LightningOptimizer wrapper (or the object itself)"""
        return getattr(opt, "optimizer", opt)

    def _meta_step(self, loss, meta_opt, meta_scheduler):
        """lceT5.py:135-147: eta's AdamW step on mean(w * loss), its scheduler step and the three logged scalars."""
        weights = self.weights(loss)
        meta_loss = weights * loss
        meta_opt.zero_grad()
        self.manual_backward(meta_loss.mean())
        if self.weights.eta.grad is not None:     # DDP's mean over ranks of eta's gradient
            self.weights.eta.grad *= all_reduce_mean_scale(self.weights.eta.grad)
        meta_opt.step()
        meta_scheduler.step()
        self.log("avg_weight", weights.mean())
        self.log("meta_loss", meta_loss.mean())
        self.log("eta", self.weights.eta.item())
        return weights, meta_loss

    def _training_step_micro(self, p, n, mq, meta_opt, opt, meta_scheduler, scheduler):
        """The same step with the batch walked in groups of `micro_queries` queries (T5-large at 1+31 x 512 does
        not keep 512 documents of activations in HBM): every query's loss element, weight and gradient
        contribution is independent of the other queries, so the gradients of mean(main_loss) are
        the sum over micro-batches of d(sum(main_loss_mb) / (Q*T)), accumulated in the flat gradient buffer; the
        all-reduce ranges leave during the LAST micro-batch's backward.  eta only enters the logged weights, so
        running its step after the forwards (the reference runs it before, lceT5.py:135-147) changes nothing."""
        k, V = int(self.hparams.n), self.model.dims.vocab
        pi, ni = torch.as_tensor(p["input_ids"]), torch.as_tensor(n["input_ids"])
        pl, nl = torch.as_tensor(p["labels"]), torch.as_tensor(n["labels"])
        Q, T = pi.shape[0], pl.shape[1]
        if not self.elide:
            with torch.no_grad():
                loss0 = []
                self._forward_seed(0)
                for a in range(0, Q, mq):
                    b = min(Q, a + mq)
                    pc = {"input_ids": pi[a:b], "labels": pl[a:b]}
                    nc = {"input_ids": ni[a * k:b * k], "labels": nl[a * k:b * k]}
                    p0, n0 = self._no_grad_logits(pc, nc, k, V, a * (1 + k))
                    loss0.append(self.pair_loss(p0.contiguous(), n0.contiguous(), pc["labels"].cuda(),
                                                nc["labels"].cuda()))
                    del p0, n0
                loss0 = torch.cat(loss0)
        opt.zero_grad()
        parts = []
        self._forward_seed(1)
        try:
            for a in range(0, Q, mq):
                b = min(Q, a + mq)
                pc = {"input_ids": pi[a:b], "labels": pl[a:b]}
                nc = {"input_ids": ni[a * k:b * k], "labels": nl[a * k:b * k]}
                plogits, nlogits = self._forward_pn(pc, nc, k, V, a * (1 + k))
                main = self.pair_loss(plogits.contiguous(), nlogits.contiguous(), pc["labels"].cuda(),
                                      nc["labels"].cuda())
                parts.append(main.detach())
                self.model.accumulating = b < Q
                self.manual_backward(main.sum() / float(Q * T))
                del plogits, nlogits, main
        finally:
            self.model.accumulating = False
        main_loss = torch.cat(parts)
        loss = main_loss if self.elide else loss0
        weights, meta_loss = self._meta_step(loss, meta_opt, meta_scheduler)
        opt.step()
        scheduler.step()
        self._pn_step += 1
        return self._step_output(loss, weights, main_loss, meta_loss)

    def _step_output(self, loss, weights, main_loss, meta_loss):
        tqdm_dict = {"meta_loss": loss.mean(), "avg_weight": weights.mean(), "main_loss": loss.mean()}
        output = OrderedDict({"main_loss": main_loss.mean(), "meta_loss": meta_loss.mean(),
                              "progress_bar": tqdm_dict, "log": tqdm_dict})
        self.log("main_loss", loss.mean())
        self.difficulty = self.weights.eta.item()
        return output

    def _training_step_restricted(self, p, n, meta_opt, opt, meta_scheduler, scheduler):
        """hparams['mode'] = 'restricted' — the fast mode of north_star items (1) + (2), not a reference formula:
        ONE decoder step, the LM head restricted to the 'true' / 'false' rows (the re-rank score of
        distill/wrapper.py:34-38, s = log_softmax(z_true, z_false)[0]) and the canonical localized contrastive
        estimation over each query's group, loss[q] = logsumexp_j s[q, j] - s[q, 0] (pn_lce_restricted_fwd / _bwd,
        positive first).  Same step structure as lceT5.py:125-168 (eta meta step on the detached losses, unweighted
        main step); the [N*T, vocab] logits, their log-softmax and their gradient never exist."""
        from .loss import lce_restricted
        k = int(self.hparams.n)
        pi, ni = torch.as_tensor(p["input_ids"]).cuda(), torch.as_tensor(n["input_ids"]).cuda()
        Q, G = pi.shape[0], 1 + k
        if pi.shape[1] != ni.shape[1]:
            S = max(pi.shape[1], ni.shape[1])      # pad tokens are attended on this path (no mask, lceT5.py:53)
            pi = torch.nn.functional.pad(pi, (0, S - pi.shape[1]), value=self.model.dims.pad_id)
            ni = torch.nn.functional.pad(ni, (0, S - ni.shape[1]), value=self.model.dims.pad_id)
        self._forward_seed(1)
        self.model.dropout_doc_offset = 0
        z2 = self.model.z2(torch.cat([pi, ni]))[:, 0]                      # [Q + Q*k, 2]
        grouped = torch.cat([z2[:Q].unsqueeze(1), z2[Q:].view(Q, k, 2)], dim=1)   # [Q, G, 2], positive first
        main_loss, _ = lce_restricted(grouped.reshape(-1, 2), G)
        loss = main_loss.detach()
        weights, meta_loss = self._meta_step(loss, meta_opt, meta_scheduler)
        opt.zero_grad()
        self.manual_backward(main_loss.mean())
        opt.step()
        scheduler.step()
        self._pn_step += 1
        return self._step_output(loss, weights, main_loss, meta_loss)

    def training_step(self, batch, batch_nb=0):
        """lceT5.py:125-168, statement for statement; the model calls go to libpn_b200."""
        p, n = batch

        meta_opt, opt = self.optimizers()
        meta_scheduler, scheduler = self.lr_schedulers()
        if self.hparams.get("mode", "parity") == "restricted":
            return self._training_step_restricted(p, n, meta_opt, opt, meta_scheduler, scheduler)
        k, V = int(self.hparams.n), self.model.dims.vocab
        mq = int(self.hparams.get("micro_queries", 0) or 0)
        if 0 < mq < len(p["input_ids"]):
            return self._training_step_micro(p, n, mq, meta_opt, opt, meta_scheduler, scheduler)

        if self.elide:
            self._forward_seed(1)
            plogits, nlogits = self._forward_pn(p, n, k, V)
            p0, n0 = plogits.detach(), nlogits.detach()
        else:
            with torch.no_grad():
                self._forward_seed(0)
                p0, n0 = self._no_grad_logits(p, n, k, V, 0)
        op = torch.as_tensor(p["labels"]).cuda()
        on = torch.as_tensor(n["labels"]).cuda()
        with torch.no_grad():
            loss = self.pair_loss(p0.contiguous(), n0.contiguous(), op, on)

        weights, meta_loss = self._meta_step(loss, meta_opt, meta_scheduler)

        if not self.elide:
            self._forward_seed(1)
            plogits, nlogits = self._forward_pn(p, n, k, V)
        main_loss = self.pair_loss(plogits.contiguous(), nlogits.contiguous(), op, on)

        opt.zero_grad()
        self.manual_backward(main_loss.mean())
        opt.step()          # finishes the data-parallel gradient mean (ranges sent during the backward), then AdamW
        scheduler.step()
        self._pn_step += 1
        return self._step_output(loss, weights, main_loss, meta_loss)


def _load_tokenizer(path):
    """T5Tokenizer.from_pretrained(model_name) (lceT5.py:107) from the local directory; None when it holds no
    SentencePiece / tokenizer files (random-init test models): using `.tokenizer` then fails where it is used."""
    import os
    if not any(os.path.exists(os.path.join(path, f)) for f in ("spiece.model", "tokenizer.json")):
        return None
    from transformers import T5Tokenizer
    return T5Tokenizer.from_pretrained(path)
