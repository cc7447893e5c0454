"""world_size-2 data-parallel path on CPU with gloo: the Trainer's DDP wiring, per-rank synthetic shards, and
gradient averaging.  The CUDA mixers have no CPU path, so the slot holds the reference's HEAD occupant
(MHPermutMix, pure torch) here -- the thing under test is the host-side sharding/all-reduce logic."""
import os
import socket
from types import SimpleNamespace

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

CFG = dict(img_size=16, patch_size=4, in_channels=3, num_classes=10, embed_dim=16, num_encoders=2, num_heads=2,
           hidden_dim=24, dropout=0.0, activation="gelu", mixer="permut", precision="fp32", random_seed=42,
           learning_rate=1e-3, adam_betas=(0.9, 0.999), adam_weight_decay=0.01)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker_flat(rank, world, port, out):
    """The graph-mode gradient exchange (one flat all-reduce, train.Trainer._allreduce_grads) instead of DDP buckets."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), LOCAL_RANK=str(rank),
                      WORLD_SIZE=str(world))
    torch.set_num_threads(1)
    from vit_spectre_experiments_b200.train import Trainer, init_distributed

    init_distributed("gloo")
    tr = Trainer(SimpleNamespace(**CFG), "cpu", ddp=False)
    tr.world = world
    img, lab = tr.synthetic_batch(4, rank)
    for _ in range(3):
        loss = tr.forward_loss(img, lab)
        tr.opt.zero_grad(set_to_none=True)
        loss.backward()
        tr._allreduce_grads()
        tr.opt.step()
    flat = torch.cat([p.detach().flatten() for p in tr.model.parameters()])
    gathered = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    if rank == 0:
        torch.save({"params": gathered}, out)
    dist.barrier()
    dist.destroy_process_group()


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), LOCAL_RANK=str(rank),
                      WORLD_SIZE=str(world))
    torch.set_num_threads(1)
    from vit_spectre_experiments_b200.train import Trainer, init_distributed

    r, lr, w = init_distributed("gloo")
    assert (r, w) == (rank, world)
    tr = Trainer(SimpleNamespace(**CFG), "cpu")
    assert isinstance(tr.net, torch.nn.parallel.DistributedDataParallel)
    img, lab = tr.synthetic_batch(4, rank)          # each rank: its own shard (seed 42 + rank)
    losses = [tr.step(img, lab).item() for _ in range(3)]
    flat = torch.cat([p.detach().flatten() for p in tr.model.parameters()])
    gathered = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    if rank == 0:
        torch.save({"params": gathered, "losses": losses, "img_sum": img.sum().item()}, out)
    dist.barrier()
    dist.destroy_process_group()


def test_ddp_two_ranks_match_single_process_on_global_batch(tmp_path):
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    res = torch.load(out)
    # replicas stay bit-identical after 3 all-reduced steps
    assert torch.equal(res["params"][0], res["params"][1])

    # a single process stepping on the concatenated global batch gives the same parameters
    from vit_spectre_experiments_b200.train import Trainer

    tr = Trainer(SimpleNamespace(**CFG), "cpu", ddp=False)
    a, la = tr.synthetic_batch(4, 0)
    b, lb = tr.synthetic_batch(4, 1)
    assert abs(a.sum().item() - res["img_sum"]) < 1e-4
    img, lab = torch.cat([a, b]), torch.cat([la, lb])
    for _ in range(3):
        tr.step(img, lab)
    flat = torch.cat([p.detach().flatten() for p in tr.model.parameters()])
    assert (flat - res["params"][0]).abs().max().item() < 2e-5


def test_flat_allreduce_matches_ddp(tmp_path):
    """Graph mode replaces DDP's buckets by one flat all-reduce of every gradient inside the captured step: same
    parameters after 3 steps as the DDP run (gloo has no AVG: the helper must still average)."""
    a, b = str(tmp_path / "ddp.pt"), str(tmp_path / "flat.pt")
    mp.spawn(_worker, args=(2, _free_port(), a), nprocs=2, join=True)
    mp.spawn(_worker_flat, args=(2, _free_port(), b), nprocs=2, join=True)
    pa, pb = torch.load(a)["params"], torch.load(b)["params"]
    assert torch.equal(pb[0], pb[1])
    assert (pa[0] - pb[0]).abs().max().item() < 2e-6
