"""A stand-in for the `lightning` package (absent from this image; the reference's trainer shell is control plane
and out of scope): `Trainer.fit` is the manual loop Lightning runs for a module with automatic_optimization = False —
callbacks' on_train_batch_start, then training_step — over one epoch of the data module's train_dataloader."""
import sys
import types


class Trainer:
    def __init__(self, **kw):
        self.kw = kw
        self.callbacks = kw.get("callbacks", [])
        self.max_steps = kw.get("max_steps")
        self.train_dataloader = None
        self.outputs = []

    def fit(self, model, datamodule):
        self.train_dataloader = datamodule.train_dataloader()
        for i, batch in enumerate(self.train_dataloader):
            if self.max_steps is not None and i >= self.max_steps:
                break
            for cb in self.callbacks:
                if hasattr(cb, "on_train_batch_start"):
                    cb.on_train_batch_start(self, model, batch, i)
            self.outputs.append(model.training_step(batch, i))


class _Anything:
    def __init__(self, *a, **kw):
        self.args, self.kw = a, kw


def install():
    """lightning, lightning.pytorch (+ .loggers, .callbacks) in sys.modules; returns the modules."""
    L = types.ModuleType("lightning")
    pl = types.ModuleType("lightning.pytorch")
    L.Trainer = Trainer
    L.pytorch = pl
    pl.Trainer = Trainer
    pl.loggers = types.SimpleNamespace(WandbLogger=_Anything)
    pl.callbacks = types.SimpleNamespace(RichProgressBar=_Anything, LearningRateMonitor=_Anything,
                                         ProgressBar=_Anything)
    sys.modules["lightning"] = L
    sys.modules["lightning.pytorch"] = pl
    return L, pl
