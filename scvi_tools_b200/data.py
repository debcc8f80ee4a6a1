"""HBM-resident count matrix + on-device minibatch index stream.

Replaces the reference's per-step host path -- `AnnDataLoader` / `AnnTorchDataset.__getitem__`
(src/scvi/dataloaders/_ann_dataloader.py:21-138, src/scvi/data/_anntorchdataset.py:115-169) and
the Lightning host->device copy -- the way `DeviceBackedDataSplitter`
(src/scvi/dataloaders/_data_splitting.py:589-732) does, but with the index stream of
`BatchDistributedSampler` (src/scvi/dataloaders/_samplers.py:6-89): every epoch a global
permutation seeded `seed + epoch`, rank-strided shard, padded (or tail-dropped) to equal
length, chunks of the per-replica batch size.  The matrix is uploaded once; a step moves no
count data over PCIe.
"""

from __future__ import annotations

import math

import numpy as np
import torch

from . import ops


def epoch_indices(n_obs: int, rank: int, world: int, epoch: int, seed: int = 0, shuffle: bool = True,
                  drop_dataset_tail: bool = False) -> torch.Tensor:
    """This rank's sample order for `epoch` (host int64 tensor) -- torch `DistributedSampler.__iter__`
    semantics, which `BatchDistributedSampler` inherits."""
    if shuffle:
        g = torch.Generator()
        g.manual_seed(seed + epoch)
        idx = torch.randperm(n_obs, generator=g)
    else:
        idx = torch.arange(n_obs)
    if drop_dataset_tail and n_obs % world != 0:
        num = math.ceil((n_obs - world) / world)
    else:
        num = math.ceil(n_obs / world)
    total = num * world
    if not drop_dataset_tail:
        pad = total - n_obs
        if pad > 0:
            reps = math.ceil(pad / n_obs)
            idx = torch.cat([idx, idx.repeat(reps)[:pad]])
    else:
        idx = idx[:total]
    return idx[rank:total:world].contiguous()


class DeviceResidentCounts:
    """Count matrix [N, G] (fp32, or uint16-compressed) + batch / label columns resident in HBM."""

    def __init__(self, X, batch=None, labels=None, device="cuda", compress_u16: bool = False):
        dev = torch.device(device)
        if isinstance(X, np.ndarray):
            X = torch.from_numpy(X)
        if hasattr(X, "toarray"):  # scipy sparse: densify once on the host (reference :152 does it per batch)
            X = torch.from_numpy(np.asarray(X.toarray()))
        if compress_u16:
            if float(X.max()) >= 65536 or float(X.min()) < 0:
                raise ValueError("uint16 residency needs counts in [0, 65535]")
            X = X.to(torch.uint16)
        elif X.dtype != torch.float32:
            X = X.to(torch.float32)
        self.X = X.to(dev).contiguous()
        n = self.X.shape[0]

        def col(v):
            if v is None:
                return torch.zeros(n, dtype=torch.int64, device=dev)
            v = torch.as_tensor(np.asarray(v) if not isinstance(v, torch.Tensor) else v)
            return v.reshape(-1).to(torch.int64).to(dev).contiguous()

        self.batch = col(batch)
        self.labels = col(labels)

    @property
    def n_obs(self) -> int:
        return self.X.shape[0]

    @property
    def n_vars(self) -> int:
        return self.X.shape[1]

    def minibatch(self, idx: torch.Tensor) -> dict:
        """The dict `AnnTorchDataset.__getitem__` + collate produce, built on the device:
        X float32 (n, G), batch / labels int64 (n, 1)."""
        idx = idx.to(self.X.device, dtype=torch.int64)
        return {
            "X": ops.gather_rows(self.X, idx),
            "batch": ops.gather_i64(self.batch, idx).view(-1, 1),
            "labels": ops.gather_i64(self.labels, idx).view(-1, 1),
        }

    def iter_epoch(self, batch_size: int, epoch: int = 0, seed: int = 0, rank: int = 0, world: int = 1,
                   shuffle: bool = True, drop_last: bool = False, drop_dataset_tail: bool = False):
        """Yield device index tensors, one per minibatch (the sampler's batches)."""
        order = epoch_indices(self.n_obs, rank, world, epoch, seed, shuffle, drop_dataset_tail).to(self.X.device)
        for s in range(0, order.numel(), batch_size):
            chunk = order[s : s + batch_size]
            if chunk.numel() < batch_size and drop_last:
                break
            yield chunk
