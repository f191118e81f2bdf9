"""Run under torchrun with 2 ranks: the gradient all-reduce started from inside the backward pass (ranges handed
out by pn_t5_set_grad_hook) must give bit-identical gradients and parameters to the single all-reduce after it.
With fewer GPUs than ranks both ranks share cuda:0 and the collective runs on gloo (the hook logic is the same)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    ngpu = torch.cuda.device_count()
    torch.cuda.set_device(rank % ngpu)
    if ngpu < world and os.environ.get("PN_REQUIRE_NCCL") == "1":
        raise SystemExit(f"PN_REQUIRE_NCCL=1: {world} ranks need {world} GPUs for NCCL, {ngpu} visible (no gloo fallback)")
    backend = "nccl" if ngpu >= world else "gloo"
    dist.init_process_group(backend, rank=rank, world_size=world)
    from oracle import t5 as ot
    from pacednegatives_b200.lceT5 import LCEModel
    from pacednegatives_b200.scorer import T5Dims
    dims = ot.T5Dims(vocab=384, d_model=128, d_kv=64, d_ff=256, n_layers=4, n_dec_layers=2, n_heads=2)
    w = {k: torch.as_tensor(v) for k, v in ot.init_weights(dims, seed=23).items()}
    hp = dict(n=3, use_mean=True, ignore_index=-100, lr=1e-3, meta_lr=1e-2, warmup_steps=0, total_steps=10, eta=20.0)
    rng = np.random.RandomState(100 + rank)     # every rank trains on its own queries
    Q, k, S = 2, 3, 128
    pids = rng.randint(2, dims.vocab, size=(Q, S)); pids[:, -1] = 1
    nids = rng.randint(2, dims.vocab, size=(Q * k, S)); nids[:, -1] = 1
    batch = ({"input_ids": torch.tensor(pids), "labels": torch.tensor(np.tile([[117, 1]], (Q, 1)))},
             {"input_ids": torch.tensor(nids), "labels": torch.tensor(np.tile([[213, 1]], (Q * k, 1)))})
    out = {}
    for overlap in (True, False):
        lm = LCEModel(hp, dims=T5Dims(**dims.__dict__), state_dict=w)
        lm.model.overlap_grad_reduce = overlap
        for step in range(2):
            lm.training_step(batch, step)
        torch.cuda.synchronize()
        assert (lm.model._comm_done is not None) == overlap, "overlap path not taken as configured"
        out[overlap] = (lm.model.grads.clone(), lm.model.parameters_flat().clone())
    # the backward accumulates with fp32 atomics (split-K weight gradients, embedding rows), so two runs agree to
    # rounding, not bit for bit; a range that left for its all-reduce too early would be off by whole contributions
    g, ref = out[True][0], out[False][0]
    for name, off, numel in lm.model.named_offsets():
        a, b = g[off:off + numel], ref[off:off + numel]
        scale = float(b.abs().max()) + 1e-12
        err = float((a - b).abs().max())
        assert err <= 2e-3 * scale + 1e-7, f"{name}: overlapped vs single all-reduce differ by {err} (scale {scale})"
    dp = float((out[True][1] - out[False][1]).abs().max())
    assert dp <= 2.1e-3, f"parameters differ by {dp} after two steps (lr 1e-3)"
    both = [torch.empty_like(g) for _ in range(world)]
    dist.all_gather(both, g)
    assert all(torch.equal(both[0], b) for b in both), "ranks disagree on the reduced gradient"
    # first step only (afterwards data-parallel and solo parameters diverge): reduced gradient = sum over ranks
    one = LCEModel(hp, dims=T5Dims(**dims.__dict__), state_dict=w)
    one.training_step(batch, 0)
    solo = LCEModel(hp, dims=T5Dims(**dims.__dict__), state_dict=w)
    solo.model.overlap_grad_reduce = False
    solo.model.all_reduce_grads = lambda group=None: 1.0      # this rank's own gradient, no exchange
    solo.training_step(batch, 0)
    torch.cuda.synchronize()
    mine = [torch.empty_like(g) for _ in range(world)]
    dist.all_gather(mine, solo.model.grads)
    total = sum(mine)
    rel = float((total - one.model.grads).abs().max() / total.abs().max())
    if rel >= 2e-3 and rank == 0:
        for name, off, numel in one.model.named_offsets():
            a, b, c = one.model.grads[off:off + numel], total[off:off + numel], solo.model.grads[off:off + numel]
            print(name, "reduced", float(a.norm()), "sum", float(b.norm()), "own", float(c.norm()),
                  "diff", float((a - b).norm()), flush=True)
    assert rel < 2e-3, f"reduced gradient is not the sum of the ranks' gradients (rel {rel})"
    own = float((solo.model.grads - one.model.grads).abs().max() / total.abs().max())
    assert own > 1e-2, "ranks were supposed to train on different queries"
    # the reduced gradient is the SUM over ranks (1/world is applied inside AdamW): differs from a local one
    print(f"rank {rank}: overlap check ok ({g.numel()} gradient elements, backend {backend})", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
