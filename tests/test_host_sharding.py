"""CPU: the data-parallel host logic -- per-rank index streams (BatchDistributedSampler semantics,
reference src/scvi/dataloaders/_samplers.py:6-89) and the flat-arena gradient mean -- including a
world_size-2 gloo run."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import scvi_oracle as O
from scvi_tools_b200.data import epoch_indices


@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("tail", [False, True])
def test_epoch_indices_equal_oracle_stream(world, tail):
    n = 1003
    seen = []
    for rank in range(world):
        mine = epoch_indices(n, rank, world, epoch=3, seed=7, drop_dataset_tail=tail)
        ref = np.concatenate(O.distributed_batches(n, 128, rank, world, epoch=3, seed=7, drop_dataset_tail=tail))
        assert mine.numpy().tolist() == ref.tolist()
        seen.append(mine)
    lens = {len(s) for s in seen}
    assert len(lens) == 1  # equal length on every rank (reference test_samplers.py)
    allidx = torch.cat(seen)
    if tail:
        assert allidx.unique().numel() == allidx.numel()  # disjoint
    else:
        assert allidx.unique().numel() == n  # every cell covered


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = 501
    mine = epoch_indices(n, rank, world, epoch=1, seed=0)
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine)
    cover = torch.cat(gathered).unique().numel()
    # flat gradient arena: sum-allreduce then the 1/world scale Adam applies == DDP's mean
    g = torch.full((1000,), float(rank + 1))
    dist.all_reduce(g)
    mean = g * (1.0 / world)
    if rank == 0:
        q.put((cover, float(mean[0]), len(mine)))
    dist.destroy_process_group()


def test_gloo_world2_sharding_and_gradient_mean():
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    cover, mean, n_mine = q.get()
    assert cover == 501 and mean == 1.5 and n_mine == 251
