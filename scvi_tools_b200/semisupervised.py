"""Semi-supervised training around `SCANVAE`: data split, the labelled / unlabelled minibatch streams and the
training plan -- the callers on either side of the SCANVI hot path.

Mirrors, on device-resident counts and without Lightning / AnnData:

* `validate_data_split`             reference src/scvi/dataloaders/_data_splitting.py:24-92
* `SemiSupervisedDataSplitter`      reference :342-586 (labelled / unlabelled ratio preserved in train / val / test)
* `SemiSupervisedDataLoader`        reference src/scvi/dataloaders/_semi_dataloader.py:11-118 over
                                    `ConcatDataLoader` (_concat_dataloader.py:11-107): the larger of the two streams
                                    drives an epoch, the other one is cycled; optional per-label subsampling that
                                    is redrawn every epoch (`SubSampleLabels` callback, src/scvi/train/_callbacks.py)
* `SemiSupervisedTrainingPlan`      reference src/scvi/train/_trainingplans.py:780-1000: every step feeds the full
                                    minibatch through `SCANVAE.loss` together with a labelled minibatch whose
                                    cross-entropy enters with `classification_ratio` (default 50), KL warm-up as
                                    `TrainingPlan`, Adam(lr 1e-3, eps 0.01, weight decay 1e-6) over the parameters
                                    that receive gradient.

The module step runs on this library's kernels through autograd (`SCANVAE.forward`); the optimizer is the
library's multi-tensor Adam over one flat arena (`scvi_adam_step`), the minibatch gathers are on-device.
"""

from __future__ import annotations

import warnings
from math import ceil, floor

import numpy as np
import torch

from .data import DeviceResidentCounts
from .graphed import WholeStepGraph
from .optim import ArenaAdam, broadcast_module_state
from .train import compute_kl_weight


def validate_data_split(n_samples: int, train_size: float, validation_size: float | None = None,
                        batch_size: int | None = None, drop_last: bool | int = False,
                        train_size_is_none: bool | int = True):
    """(n_train, n_val) for a data set of `n_samples` (reference _data_splitting.py:24-92): train = ceil(train_size
    * n), validation = the rest (or floor(validation_size * n)); when the default train size leaves a last
    minibatch of 1-2 cells, those cells move to the validation set (a BatchNorm layer cannot train on them)."""
    if train_size > 1.0 or train_size <= 0.0:
        raise ValueError("Invalid train_size. Must be: 0 < train_size <= 1")
    n_train = ceil(train_size * n_samples)
    if validation_size is None:
        n_val = n_samples - n_train
    elif validation_size >= 1.0 or validation_size < 0.0:
        raise ValueError("Invalid validation_size. Must be 0 <= validation_size < 1")
    elif (train_size + validation_size) > 1:
        raise ValueError("train_size + validation_size must be between 0 and 1")
    else:
        n_val = floor(n_samples * validation_size)
    if n_train == 0:
        raise ValueError(
            f"With n_samples={n_samples}, train_size={train_size} and validation_size={validation_size}, the "
            "resulting train set will be empty. Adjust any of the aforementioned parameters.")
    if batch_size is not None:
        tail = n_train % batch_size
        if 0 < tail < 3 and drop_last is False:
            if not train_size_is_none:
                warnings.warn(f"Last batch will have a small size of {tail} samples. Consider changing batch_size "
                              f"from currently {batch_size} to avoid errors during model training.", UserWarning,
                              stacklevel=2)
            else:
                n_train -= tail
                if n_val > 0:
                    n_val += tail
    return n_train, n_val


class SemiSupervisedDataSplitter:
    """Train / validation / test indices with the labelled : unlabelled ratio preserved in each
    (reference _data_splitting.py:342-586).  `labels` are integer codes, `unlabeled_category` the code of the cells
    without a label.  The shuffles are `numpy.random.RandomState(seed).choice(idx, len(idx), replace=False)` on the
    labelled and on the unlabelled index set separately, as the reference draws them."""

    def __init__(self, labels, unlabeled_category, train_size: float | None = None,
                 validation_size: float | None = None, shuffle_set_split: bool = True,
                 n_samples_per_label: int | None = None, batch_size: int = 128, drop_last: bool = False, seed: int = 0):
        labels = np.asarray(labels).ravel()
        self.labels = labels
        self.unlabeled_category = unlabeled_category
        self.train_size_is_none = not bool(train_size)
        self.train_size = 0.9 if self.train_size_is_none else float(train_size)
        self.validation_size = validation_size
        self.shuffle_set_split = shuffle_set_split
        self.n_samples_per_label = n_samples_per_label
        self.batch_size, self.drop_last, self.seed = batch_size, drop_last, seed
        self._unlabeled_indices = np.argwhere(labels == unlabeled_category).ravel()
        self._labeled_indices = np.argwhere(labels != unlabeled_category).ravel()
        self.n_train, self.n_val = validate_data_split(labels.shape[0], self.train_size, validation_size, batch_size,
                                                       drop_last, self.train_size_is_none)
        self.setup()

    def _split(self, idx):
        if len(idx) == 0:
            return [], [], []
        n_train, n_val = validate_data_split(len(idx), self.train_size, self.validation_size, self.batch_size,
                                             self.drop_last, self.train_size_is_none)
        perm = idx
        if self.shuffle_set_split:
            perm = np.random.RandomState(seed=self.seed).choice(idx, len(idx), replace=False)
        return perm[n_val : n_val + n_train], perm[:n_val], perm[n_val + n_train :]

    def setup(self):
        lt, lv, lte = self._split(self._labeled_indices)
        ut, uv, ute = self._split(self._unlabeled_indices)
        self.train_idx = np.concatenate((lt, ut)).astype(int)
        self.val_idx = np.concatenate((lv, uv)).astype(int)
        self.test_idx = np.concatenate((lte, ute)).astype(int)

    def _loader(self, data, idx, shuffle, seed):
        if len(idx) == 0:
            return None
        return SemiSupervisedDataLoader(data, self.labels, self.unlabeled_category, indices=idx, shuffle=shuffle,
                                        batch_size=self.batch_size, drop_last=self.drop_last,
                                        n_samples_per_label=self.n_samples_per_label, seed=seed)

    def train_dataloader(self, data):
        return self._loader(data, self.train_idx, True, self.seed)

    def val_dataloader(self, data):
        return self._loader(data, self.val_idx, False, self.seed + 1)

    def test_dataloader(self, data):
        return self._loader(data, self.test_idx, False, self.seed + 2)


class SemiSupervisedDataLoader:
    """Two index streams over device-resident counts: all cells of `indices`, and their labelled subset.

    Iterating yields `(full_minibatch, labelled_minibatch)` dict pairs (`{"X", "batch", "labels"}` built by on-device
    gathers).  The stream with more minibatches runs once per epoch; the other one is cycled -- and, as
    `itertools.cycle` does, repeats the minibatches of its first pass in the same order.  With `n_samples_per_label`
    at most that many cells per label take part, redrawn by `resample_labels()` (called by the plan every epoch).
    When no cell of `indices` is labelled the pairs are `(full_minibatch, None)`."""

    def __init__(self, data: DeviceResidentCounts, labels, unlabeled_category, indices=None, shuffle: bool = False,
                 batch_size: int = 128, drop_last: bool = False, n_samples_per_label: int | None = None,
                 seed: int = 0):
        self.data = data
        labels = np.asarray(labels).ravel()
        self.indices = np.arange(labels.shape[0]) if indices is None else np.asarray(indices)
        self.shuffle, self.batch_size, self.drop_last = shuffle, int(batch_size), drop_last
        self.n_samples_per_label = n_samples_per_label
        self._rng = np.random.RandomState(seed)
        self._gen = torch.Generator().manual_seed(seed)
        # indices of every labelled category, in the order of np.unique (reference _semi_dataloader.py:67-73)
        sub = labels[self.indices]
        self.labeled_locs = [self.indices[np.where(sub == lab)[0]] for lab in np.unique(labels)
                             if lab != unlabeled_category]
        self.labelled_idx = self.subsample_labels()

    def subsample_labels(self):
        if not self.labeled_locs:
            return np.zeros(0, dtype=np.int64)
        if self.n_samples_per_label is None:
            return np.concatenate(self.labeled_locs)
        pick = [loc if len(loc) < self.n_samples_per_label
                else self._rng.choice(loc, self.n_samples_per_label, replace=False) for loc in self.labeled_locs]
        return np.concatenate(pick)

    def resample_labels(self):
        self.labelled_idx = self.subsample_labels()

    def _n_batches(self, n):
        return n // self.batch_size if self.drop_last else -(-n // self.batch_size)

    def __len__(self):
        return max(self._n_batches(len(self.indices)), self._n_batches(len(self.labelled_idx)))

    def _batches(self, idx):
        """Index minibatches of one pass over `idx` (host int64 tensors)."""
        t = torch.as_tensor(np.asarray(idx), dtype=torch.int64)
        if self.shuffle:
            t = t[torch.randperm(t.numel(), generator=self._gen)]
        out = [t[s : s + self.batch_size] for s in range(0, t.numel(), self.batch_size)]
        if self.drop_last and out and out[-1].numel() < self.batch_size:
            out.pop()
        return out

    def index_pairs(self):
        """One epoch as (full index tensor, labelled index tensor or None) pairs."""
        full = self._batches(self.indices)
        lab = self._batches(self.labelled_idx) if len(self.labelled_idx) else []
        if not lab:
            return [(f, None) for f in full]
        n = max(len(full), len(lab))
        return [(full[i % len(full)], lab[i % len(lab)]) for i in range(n)]

    def __iter__(self):
        dev = self.data.X.device
        for f, l in self.index_pairs():
            yield (self.data.minibatch(f.to(dev)), None if l is None else self.data.minibatch(l.to(dev)))


class SemiSupervisedTrainingPlan:
    """`SemiSupervisedTrainingPlan` without Lightning (reference _trainingplans.py:780-1000).

    `training_step(batch)` accepts what the reference's `_split_semisupervised_batch` accepts: a `(full, labelled)`
    pair, or one dict (used for both roles when it carries labels, i.e. has the "labels" key and
    `labelled_from_single=True`).  The optimizer is `ArenaAdam` (optim.py)."""

    def __init__(self, module, n_classes: int, *, classification_ratio: float = 50, lr: float = 1e-3,
                 weight_decay: float = 1e-6, eps: float = 0.01, betas=(0.9, 0.999),
                 n_steps_kl_warmup: int | None = None, n_epochs_kl_warmup: int | None = 400,
                 max_kl_weight: float = 1.0, min_kl_weight: float = 0.0, use_cuda_graph: bool = False,
                 process_group=None, **loss_kwargs):
        self.module = module
        self.n_classes = n_classes
        # whole step (zero_grad, forward, backward, Adam) as one CUDA graph per minibatch shape: graphed.py
        self._graph = WholeStepGraph(self._step) if use_cuda_graph else None
        self._klw = None
        self.lr, self.wd, self.eps, self.betas = lr, weight_decay, eps, betas
        self.n_steps_kl_warmup, self.n_epochs_kl_warmup = n_steps_kl_warmup, n_epochs_kl_warmup
        self.max_kl_weight, self.min_kl_weight = max_kl_weight, min_kl_weight
        self.loss_kwargs = dict(loss_kwargs)
        self.loss_kwargs["classification_ratio"] = classification_ratio
        self.current_epoch = 0
        self.global_step = 0
        broadcast_module_state(module, process_group)
        self.optimizer = ArenaAdam(module.named_parameters, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay,
                                   capturable=use_cuda_graph, process_group=process_group)

    @property
    def kl_weight(self):
        return compute_kl_weight(self.current_epoch, self.global_step, self.n_epochs_kl_warmup, self.n_steps_kl_warmup,
                                 self.max_kl_weight, self.min_kl_weight)

    @staticmethod
    def _split_semisupervised_batch(batch):
        if isinstance(batch, (list, tuple)) and len(batch) == 2:
            return batch[0], batch[1]
        return batch, None

    def zero_grad(self):
        self.optimizer.zero_grad()

    # ------------------------------------------------------------------ steps
    def _step(self, inputs):
        full, labelled, klw = inputs
        kw = dict(self.loss_kwargs)
        kw["kl_weight"] = klw
        kw["labelled_tensors"] = labelled
        self.zero_grad()
        _, _, lo = self.module(full, loss_kwargs=kw)
        lo.loss.backward()
        self.optimizer.step()
        return lo.detach_()

    def training_step(self, batch):
        full, labelled = self._split_semisupervised_batch(batch)
        self.module.train()
        if self._graph is None:
            lo = self._step((full, labelled, self.kl_weight))
        else:
            if self._klw is None:
                self._klw = torch.zeros((), dtype=torch.float32, device=full["X"].device)
            self._klw.fill_(self.kl_weight)  # the KL warm-up weight is graph INPUT, not a captured constant
            lo = self._graph((full, labelled, self._klw))
        self.global_step += 1
        return lo

    @torch.no_grad()
    def validation_step(self, batch):
        full, labelled = self._split_semisupervised_batch(batch)
        kw = dict(self.loss_kwargs)
        kw["labelled_tensors"] = labelled
        self.module.eval()
        _, _, lo = self.module(full, loss_kwargs=kw)
        return lo

    @staticmethod
    def classification_metrics(lo):
        """Accuracy of the labelled minibatch (the reference logs accuracy / F1 / calibration error through
        torchmetrics; micro-averaged accuracy and F1 coincide for single-label multiclass prediction)."""
        if lo.classification_loss is None:
            return {}
        pred = lo.logits.argmax(-1)
        true = lo.true_labels.view(-1)
        acc = float((pred == true).float().mean())
        return {"classification_loss": float(lo.classification_loss), "accuracy": acc, "f1_score": acc}

    def fit(self, loader: SemiSupervisedDataLoader, max_epochs: int, val_loader: SemiSupervisedDataLoader | None = None):
        history = []
        first = self.current_epoch
        for epoch in range(first, first + max_epochs):
            self.current_epoch = epoch
            loader.resample_labels()  # SubSampleLabels callback: on_train_epoch_start
            tot = None  # [loss * n, reconstruction sum, kl sum] accumulated on the device: no per-step host read
            n_cells = n_steps = 0
            last = None
            for batch in loader:
                lo = self.training_step(batch)
                n = lo.n_obs_minibatch
                vec = torch.stack([lo.loss.detach() * n, lo.reconstruction_loss_sum.detach(), lo.kl_local_sum.detach()]).double()
                tot = vec if tot is None else tot + vec
                n_cells += n
                n_steps += 1
                last = lo
            tot = torch.zeros(3, dtype=torch.float64) if tot is None else tot.cpu()
            rec = {"epoch": epoch, "train_loss": float(tot[0]) / max(n_cells, 1),
                   "reconstruction_loss": float(tot[1]) / max(n_cells, 1), "kl_local": float(tot[2]) / max(n_cells, 1),
                   "n_cells": n_cells, "n_steps": n_steps, "kl_weight": self.kl_weight}
            if last is not None:
                rec.update({"train_" + k: v for k, v in self.classification_metrics(last).items()})
            if val_loader is not None:
                accs, losses = [], []
                for batch in val_loader:
                    lo = self.validation_step(batch)
                    m = self.classification_metrics(lo)
                    if m:
                        accs.append(m["accuracy"])
                    losses.append(float(lo.loss))
                rec["validation_loss"] = float(np.mean(losses)) if losses else None
                rec["validation_accuracy"] = float(np.mean(accs)) if accs else None
            history.append(rec)
        self.current_epoch = first + max_epochs
        return history
