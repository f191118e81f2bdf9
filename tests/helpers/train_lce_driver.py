"""The reference's training driver, restated call for call (reference: train/train_lce.py:20-103) so the drop-in
test can run it on the GPU box, where /root/reference does not exist.  `tests/test_dropin_cpu.py` compares the call
sequence of `main` below with the reference file's whenever that file is present (the build container).

Module aliases the caller must have installed in sys.modules before importing this file (the test does):
  pacednegatives / pacednegatives.lceT5  -> pacednegatives_b200 / pacednegatives_b200.lceT5
  lightning, lightning.pytorch           -> the reference's trainer shell (a stub with Trainer.fit = manual loop when
                                            the package is absent: control plane, out of scope)
"""
import os

from pacednegatives.lceT5 import LCEModel, LCEDataModule, ChangeDifficulty
import lightning.pytorch as pl
import lightning as L


def main(data: str,
         dataset_name: str,
         out_dir: str,
         total_steps: int = 100000,
         eta: float = 0.0,
         batch_size: int = 16,
         lr: float = 0.001,
         meta_lr=None,
         var: float = 0.01,
         n: int = 2,
         use_max: bool = False,
         warmup_steps: int = 10000,
         sample: bool = False,
         use_mean: bool = False,
         num_gpus: int = 1,
         wandb_project: str = None,
         num_workers: int = 8,
         max_steps: int = None,
         model_name: str = 't5-base'):       # the reference hard-codes 't5-base' (hub); here a local directory
    os.makedirs(out_dir, exist_ok=True)

    args = {
        'model_name': model_name,
        'total_steps': total_steps // batch_size,
        'eta': eta,
        'batch_size': batch_size,
        'lr': lr,
        'meta_lr': lr if not meta_lr else meta_lr,
        'n': n,
        'warmup_steps': warmup_steps // batch_size,
        'use_mean': use_mean,
        'ignore_index': -100
    }

    model = LCEModel(args)

    data_module = LCEDataModule(data, dataset_name, model.tokenizer, batch_size, sample, use_max, var=var, n=n,
                                init_weight=model.weights.eta.item(), num_workers=num_workers)
    data_module.setup()

    logger = pl.loggers.WandbLogger(project=wandb_project)

    trainer_args = {
        'devices': num_gpus,
        'callbacks': [pl.callbacks.RichProgressBar(), ChangeDifficulty(),
                      pl.callbacks.LearningRateMonitor(logging_interval='step')],
        'logger': logger,
        'max_epochs': 1,
        'default_root_dir': out_dir,
        'profiler': 'simple',
    }
    if max_steps:
        trainer_args['max_steps'] = max_steps

    if num_gpus > 1:
        trainer_args['accelerator'] = 'gpu'
        trainer_args['strategy'] = 'ddp'

    trainer = L.Trainer(**trainer_args)

    trainer.fit(model, data_module)

    model.model.save_pretrained(os.path.join(out_dir, 'finalmodel'))
    return model, trainer
