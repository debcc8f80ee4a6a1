"""`AdversarialTrainingPlan` without Lightning -- the caller of the TOTALVI hot path (SURVEY.md section 8f rank 2;
reference src/scvi/train/_trainingplans.py:535-777; TOTALVI's defaults -- lr 4e-3, batch size 256, adversarial
classifier on, reduce-lr-on-plateau -- are src/scvi/model/_totalvi.py:240-345).

One step, in the reference's order (manual optimisation, two optimizers):

1. `module(batch, kl_weight)` -> ELBO;  kappa = 1 - kl_weight (`scale_adversarial_loss="auto"`) or the fixed scale;
2. "fool" term: cross-entropy of the adversarial classifier's log-softmax on z against the UNIFORM distribution over
   the wrong batches, added to the ELBO with weight kappa; backward; optimizer 1 (module parameters) steps;
3. the classifier's own loss: cross-entropy against the true batch on z.detach(), times kappa; its gradients from
   step 2 are discarded (optimizer 2 zeroes before this backward); optimizer 2 (Adam lr 1e-3, eps 0.01) steps.

Both optimizers are `ArenaAdam` (one multi-tensor launch each); the module and the classifier run on the library's
kernels through autograd.  `ReduceLROnPlateau` (factor 0.6, patience 30, threshold 0 "abs") is applied by `fit` to
optimizer 1 on the monitored epoch metric, as `on_train_epoch_end` / `on_validation_epoch_end` do.
"""

from __future__ import annotations

import warnings

import numpy as np
import torch
import torch.nn.functional as F

from .graphed import WholeStepGraph
from .optim import ArenaAdam, broadcast_module_state
from .scanvi import Classifier
from .train import compute_kl_weight


class _Plateau:
    """torch.optim.lr_scheduler.ReduceLROnPlateau(mode="min", threshold_mode="abs") on one learning rate."""

    def __init__(self, opt: ArenaAdam, factor, patience, threshold, min_lr):
        self.opt, self.factor, self.patience, self.threshold, self.min_lr = opt, factor, patience, threshold, min_lr
        self.best = float("inf")
        self.bad = 0

    def step(self, metric: float):
        if metric < self.best - self.threshold:
            self.best, self.bad = metric, 0
        else:
            self.bad += 1
        if self.bad > self.patience:
            new = max(self.opt.lr * self.factor, self.min_lr)
            if self.opt.lr - new > 1e-8:
                self.opt.lr = new
            self.bad = 0


class AdversarialTrainingPlan:
    def __init__(self, module, *, lr: float = 1e-3, weight_decay: float = 1e-6, eps: float = 0.01,
                 n_steps_kl_warmup: int | None = None, n_epochs_kl_warmup: int | None = 400,
                 reduce_lr_on_plateau: bool = False, lr_factor: float = 0.6, lr_patience: int = 30,
                 lr_threshold: float = 0.0, lr_scheduler_metric: str = "elbo_validation", lr_min: float = 0.0,
                 adversarial_classifier: bool | Classifier = False, scale_adversarial_loss: float | str = "auto",
                 max_kl_weight: float = 1.0, min_kl_weight: float = 0.0, n_batch: int | None = None,
                 use_cuda_graph: bool = False, process_group=None, **loss_kwargs):
        self.module = module
        self.n_steps_kl_warmup, self.n_epochs_kl_warmup = n_steps_kl_warmup, n_epochs_kl_warmup
        self.max_kl_weight, self.min_kl_weight = max_kl_weight, min_kl_weight
        self.loss_kwargs = dict(loss_kwargs)
        self.scale_adversarial_loss = scale_adversarial_loss
        self.current_epoch = 0
        self.global_step = 0
        n_batch = int(n_batch if n_batch is not None else getattr(module, "n_batch"))
        if adversarial_classifier is True:
            if n_batch == 1:
                warnings.warn("Disabling adversarial classifier.", UserWarning, stacklevel=2)
                self.adversarial_classifier = False
            else:
                self.n_output_classifier = n_batch
                dev = next(module.parameters()).device
                self.adversarial_classifier = Classifier(n_input=module.n_latent, n_hidden=32, n_labels=n_batch,
                                                         n_layers=2, logits=True).to(dev)
        else:
            self.adversarial_classifier = adversarial_classifier
            if adversarial_classifier is not False:
                self.n_output_classifier = n_batch
        self.process_group = process_group
        broadcast_module_state(module, process_group)
        if self.adversarial_classifier is not False:
            broadcast_module_state(self.adversarial_classifier, process_group)
        self.opt1 = ArenaAdam(module.named_parameters, lr=lr, eps=eps, weight_decay=weight_decay, capturable=use_cuda_graph,
                              process_group=process_group)
        self.opt2 = None
        if self.adversarial_classifier is not False:
            self.opt2 = ArenaAdam(self.adversarial_classifier.named_parameters, lr=1e-3, eps=0.01,
                                  weight_decay=weight_decay, capturable=use_cuda_graph, process_group=process_group)
        # whole step (both backward passes, both optimizers) as one CUDA graph per minibatch shape: graphed.py
        self._graph = WholeStepGraph(self._step) if use_cuda_graph else None
        self._klw = None
        self.reduce_lr_on_plateau = reduce_lr_on_plateau
        self.lr_scheduler_metric = lr_scheduler_metric
        self.scheduler = _Plateau(self.opt1, lr_factor, lr_patience, lr_threshold, lr_min) if reduce_lr_on_plateau else None

    @property
    def kl_weight(self):
        return compute_kl_weight(self.current_epoch, self.global_step, self.n_epochs_kl_warmup, self.n_steps_kl_warmup,
                                 self.max_kl_weight, self.min_kl_weight)

    def loss_adversarial_classifier(self, z, batch_index, predict_true_class: bool = True):
        """-mean_b sum_c target[b, c] * log_softmax(classifier(z))[b, c]; target = one-hot(batch) or, to fool the
        classifier, uniform over the other batches (reference :652-668)."""
        n = self.n_output_classifier
        logp = F.log_softmax(self.adversarial_classifier(z), dim=1)
        one_hot = F.one_hot(batch_index.view(-1).long(), n)
        target = one_hot.to(logp.dtype) if predict_true_class else (1 - one_hot).to(logp.dtype) / (n - 1)
        return -(logp * target).sum(dim=1).mean()

    def _step(self, inputs):
        batch, klw = inputs
        kw = dict(self.loss_kwargs)
        kw["kl_weight"] = klw
        kappa = 1 - klw if self.scale_adversarial_loss == "auto" else self.scale_adversarial_loss
        batch_tensor = batch["batch"]
        inf, _, lo = self.module(batch, loss_kwargs=kw)
        z = inf["z"]
        loss = lo.loss
        fool = None
        # a device-resident kappa (graph mode) cannot steer a host branch: the fool term is then always evaluated
        # and contributes exactly kappa * fool = 0 once the warm-up is over
        if self.adversarial_classifier is not False and (isinstance(kappa, torch.Tensor) or kappa > 0):
            self.adversarial_classifier.train()
            fool = self.loss_adversarial_classifier(z, batch_tensor, False)
            loss = loss + fool * kappa
        self.opt1.zero_grad()
        loss.backward()
        self.opt1.step()
        adv = None
        if self.opt2 is not None:
            adv = self.loss_adversarial_classifier(z.detach(), batch_tensor, True) * kappa
            self.opt2.zero_grad()
            adv.backward()
            self.opt2.step()
        last = {"train_loss": loss.detach(), "fool_loss": None if fool is None else fool.detach(),
                "classifier_loss": None if adv is None else adv.detach(), "kappa": kappa}
        return lo.detach_(), last

    def training_step(self, batch):
        self.module.train()
        if self._graph is None:
            lo, self.last = self._step((batch, self.kl_weight))
        else:
            if self._klw is None:
                self._klw = torch.zeros((), dtype=torch.float32, device=batch["X"].device)
            self._klw.fill_(self.kl_weight)
            lo, self.last = self._graph((batch, self._klw))
        self.global_step += 1
        return lo

    @torch.no_grad()
    def sync_buffers(self):
        """Rank 0's BatchNorm buffers to every rank (torch DDP's `broadcast_buffers`; statistics are per-rank during
        the step, as under the reference's DDP)."""
        if self.process_group is None:
            return
        import torch.distributed as dist

        src = dist.get_global_rank(self.process_group, 0)
        mods = [self.module] + ([self.adversarial_classifier] if self.adversarial_classifier is not False else [])
        for mod in mods:
            for b in mod.buffers():
                dist.broadcast(b, src=src, group=self.process_group)

    @torch.no_grad()
    def validation_step(self, batch):
        self.module.eval()
        _, _, lo = self.module(batch, loss_kwargs=dict(self.loss_kwargs))
        return lo

    def fit(self, batches, max_epochs: int, val_batches=None):
        """`batches` / `val_batches`: callables returning an iterable of minibatch dicts for one epoch."""
        history = []
        first = self.current_epoch
        for epoch in range(first, first + max_epochs):
            self.current_epoch = epoch
            acc = None  # accumulated on the device: no per-step host read
            n_cells = 0
            for b in batches():
                lo = self.training_step(b)
                n = lo.n_obs_minibatch
                vec = torch.stack([self.last["train_loss"] * n, lo.reconstruction_loss_sum.detach(),
                                   lo.kl_local_sum.detach()]).double()
                acc = vec if acc is None else acc + vec
                n_cells += n
            acc = torch.zeros(3, dtype=torch.float64) if acc is None else acc.cpu()
            rec = {"epoch": epoch, "train_loss": float(acc[0]) / max(n_cells, 1),
                   "reconstruction_loss_train": float(acc[1]) / max(n_cells, 1),
                   "kl_local_train": float(acc[2]) / max(n_cells, 1), "kl_weight": self.kl_weight, "lr": self.opt1.lr}
            rec["elbo_train"] = rec["reconstruction_loss_train"] + rec["kl_local_train"]
            if val_batches is not None:
                v = torch.zeros(2, dtype=torch.float64)
                nv = 0
                for b in val_batches():
                    lo = self.validation_step(b)
                    v += torch.tensor([float(lo.reconstruction_loss_sum), float(lo.kl_local_sum)], dtype=torch.float64)
                    nv += lo.n_obs_minibatch
                rec["reconstruction_loss_validation"] = float(v[0]) / max(nv, 1)
                rec["kl_local_validation"] = float(v[1]) / max(nv, 1)
                rec["elbo_validation"] = rec["reconstruction_loss_validation"] + rec["kl_local_validation"]
            if self.scheduler is not None and self.lr_scheduler_metric in rec and np.isfinite(rec[self.lr_scheduler_metric]):
                lr_before = self.opt1.lr
                self.scheduler.step(rec[self.lr_scheduler_metric])
                if self._graph is not None and self.opt1.lr != lr_before:
                    self._graph.invalidate()  # the rate is a launch argument of the captured Adam
            self.sync_buffers()
            history.append(rec)
        self.current_epoch = first + max_epochs
        return history
