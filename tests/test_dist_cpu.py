"""Host-side logic of the data-parallel training step on CPU: two gloo ranks averaging a flat gradient buffer the
way LCEModel.training_step does (one SUM all-reduce + the 1/world factor handed to the optimiser), and the
learning-rate schedule against transformers.get_linear_schedule_with_warmup (lceT5.py:174-179)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from pacednegatives_b200.lceT5 import all_reduce_mean_scale
    g = torch.arange(1000, dtype=torch.float32) * (rank + 1)        # rank r holds (r+1) * base
    scale = all_reduce_mean_scale(g)
    eta_grad = torch.tensor([float(rank)])
    eta_grad *= all_reduce_mean_scale(eta_grad)
    if rank == 0:
        torch.save({"g": g * scale, "eta": eta_grad}, out)
    dist.destroy_process_group()


def test_gradient_mean_over_two_ranks(tmp_path):
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    r = torch.load(out)
    base = torch.arange(1000, dtype=torch.float32)
    assert torch.allclose(r["g"], base * 1.5)
    assert torch.allclose(r["eta"], torch.tensor([0.5]))


def test_single_process_is_identity():
    from pacednegatives_b200.lceT5 import all_reduce_mean_scale
    g = torch.ones(8)
    assert all_reduce_mean_scale(g) == 1.0 and torch.equal(g, torch.ones(8))


def test_linear_schedule_matches_transformers():
    from pacednegatives_b200.lceT5 import linear_schedule_with_warmup
    tr = pytest.importorskip("transformers")
    for warm, total in [(0, 10), (3, 10), (5, 5), (100, 1000)]:
        p = torch.nn.Parameter(torch.zeros(1))
        opt = torch.optim.SGD([p], lr=1.0)
        sched = tr.get_linear_schedule_with_warmup(opt, num_warmup_steps=warm, num_training_steps=total)
        for step in range(min(total + 2, 30)):
            assert abs(opt.param_groups[0]["lr"] - linear_schedule_with_warmup(step, warm, total)) < 1e-12
            opt.step()
            sched.step()
