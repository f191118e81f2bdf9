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


def _gather_worker(rank, world, port, out, equal):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from pacednegatives_b200.scorer import gather_rows, reduce_metric_means
    q = 3 if equal else 3 + rank                     # ragged shards: rank 1 holds one more query
    g = torch.Generator().manual_seed(100 + rank)
    scores = torch.randn(q, 7, generator=g)
    metrics = torch.rand(q, 4, generator=g)
    table = gather_rows(scores, equal_shards=equal)
    means = reduce_metric_means(metrics)
    if rank == 0:
        torch.save({"table": table, "means": means}, out)
    dist.destroy_process_group()


@pytest.mark.parametrize("equal", [True, False])
def test_score_gather_equals_single_rank_table(tmp_path, equal):
    """SURVEY 8e collective 3 (eval.py:23-26 evaluates one table over all queries): the gathered [sum Q, C] score
    table is bit-identical to the rank shards stacked in rank order, and the 5-float all-reduce gives the global
    means."""
    out = str(tmp_path / "g.pt")
    mp.spawn(_gather_worker, args=(2, _free_port(), out, equal), nprocs=2, join=True)
    r = torch.load(out)
    shards, mets = [], []
    for rank in range(2):
        q = 3 if equal else 3 + rank
        g = torch.Generator().manual_seed(100 + rank)
        shards.append(torch.randn(q, 7, generator=g))
        mets.append(torch.rand(q, 4, generator=g))
    assert torch.equal(r["table"], torch.cat(shards))
    assert torch.allclose(r["means"], torch.cat(mets).double().mean(0).float(), atol=1e-7)


def test_score_gather_single_process_is_identity():
    from pacednegatives_b200.scorer import gather_rows, reduce_metric_means
    t = torch.arange(12.).view(3, 4)
    assert gather_rows(t) is t or torch.equal(gather_rows(t), t)
    assert torch.allclose(reduce_metric_means(t), t.mean(0))
