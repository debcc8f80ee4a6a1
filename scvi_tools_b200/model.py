"""`SCVI`-shaped driver over the B200 kernels: construct -> `train()` -> the post-training getters.

Mirrors the calls a user of `scvi.model.SCVI` makes right around the hot path:
`train` (reference src/scvi/model/base/_training_mixin.py + src/scvi/train/_trainingplans.py:427-532),
`get_latent_representation` / `get_elbo` / `get_reconstruction_error`
(src/scvi/model/base/_vaemixin.py:32-101,196-268,270-383; src/scvi/model/base/_log_likelihood.py:20-136) and
`get_normalized_expression` (src/scvi/model/base/_rnamixin.py:155-330, the `library_size` / `transform_batch`
semantics).  The counts live in HBM (`DeviceResidentCounts`); there is no AnnData registry here -- the caller
hands in the count matrix and the integer batch codes that `AnnDataManager` would have produced.
"""

from __future__ import annotations

import numpy as np
import torch

from .data import DeviceResidentCounts, DeviceResidentCSR, HostStreamedCounts
from .module import MODULE_KEYS, VAE
from .train import FusedTrainingPlan


def get_max_epochs_heuristic(n_obs: int, epochs_cap: int = 400, decay_at_n_obs: int = 20_000) -> int:
    """Reference src/scvi/model/_utils.py `get_max_epochs_heuristic`: min(round(20000 / n_obs * 400), 400)."""
    return max(1, int(min(round((decay_at_n_obs / n_obs) * epochs_cap), epochs_cap)))


class SCVI:
    """single-cell Variational Inference on a device-resident count matrix.

    Parameters follow `scvi.model.SCVI.__init__` (src/scvi/model/_scvi.py): n_hidden, n_latent, n_layers,
    dropout_rate, dispersion, gene_likelihood, latent_distribution and further `VAE` keyword arguments.
    `X`: [n_obs, n_vars] counts (numpy / torch, any device); `batch`: [n_obs] integer batch codes or None.
    `X` may also be a `DeviceResidentCounts` (counts in HBM, the default built from an array) or a
    `HostStreamedCounts` (counts stay in pinned host memory and stream over PCIe in compact integer form).
    `precision` (a `VAE` keyword, not recorded in `init_params_`): None = tcgen05 kernels when n_hidden == 128
    (what `bench.py` times), FFMA fp32 kernels otherwise; 0 / 1 force the FFMA / tensor-core path.

    Differences from `scvi.model.SCVI.train` kept deliberately: every cell is a training cell (the reference
    holds out 10 % by default, `train_size=0.9`, for a validation loop this driver does not run), and there is
    no early stopping / checkpoint callback.
    """

    def __init__(self, X, batch=None, n_hidden: int = 128, n_latent: int = 10, n_layers: int = 1,
                 dropout_rate: float = 0.1, dispersion: str = "gene", gene_likelihood: str = "zinb",
                 latent_distribution: str = "normal", device="cuda:0", var_names=None, precision: int | None = None,
                 **model_kwargs):
        self.device = torch.device(device)
        # the reference's `init_params_` record (BaseModelClass._get_init_params, _base_model.py:747-774)
        self.init_params_ = {
            "non_kwargs": {"n_hidden": n_hidden, "n_latent": n_latent, "n_layers": n_layers,
                           "dropout_rate": dropout_rate, "dispersion": dispersion,
                           "gene_likelihood": gene_likelihood, "latent_distribution": latent_distribution},
            "kwargs": {"kwargs": dict(model_kwargs)},
        }
        if isinstance(X, (DeviceResidentCounts, DeviceResidentCSR, HostStreamedCounts)):
            self.data = X
        elif hasattr(X, "tocsr"):  # scipy sparse stays sparse in HBM
            self.data = DeviceResidentCSR(X, batch=batch, device=self.device)
        else:
            self.data = DeviceResidentCounts(X, batch=batch, device=self.device)
        self.streamed = isinstance(self.data, HostStreamedCounts)
        self.sparse = isinstance(self.data, DeviceResidentCSR)
        codes = self.data.host_batch if self.streamed else self.data.batch
        n_batch = int(codes.max().item()) + 1 if self.data.n_obs else 1
        self.summary = {"n_obs": self.data.n_obs, "n_vars": self.data.n_vars, "n_batch": n_batch}
        self.module = VAE(n_input=self.data.n_vars, n_batch=n_batch, n_hidden=n_hidden, n_latent=n_latent,
                          n_layers=n_layers, dropout_rate=dropout_rate, dispersion=dispersion,
                          gene_likelihood=gene_likelihood, latent_distribution=latent_distribution,
                          precision=precision, **model_kwargs).to(self.device)
        self.var_names = (np.asarray(var_names) if var_names is not None
                          else np.array([str(i) for i in range(self.data.n_vars)]))
        if self.var_names.shape[0] != self.data.n_vars:
            raise ValueError("var_names length differs from the number of genes")
        self.is_trained_ = False
        self.history_: list[dict] = []
        self._plan: FusedTrainingPlan | None = None

    # ------------------------------------------------------------------ training
    def train(self, max_epochs: int | None = None, batch_size: int = 128, lr: float = 1e-3, eps: float = 0.01,
              weight_decay: float = 1e-6, n_epochs_kl_warmup: int | None = 400, n_steps_kl_warmup: int | None = None,
              max_kl_weight: float = 1.0, min_kl_weight: float = 0.0, shuffle: bool = True, seed: int = 0,
              process_group=None, use_cuda_graph: bool = True):
        """`SCVI.train`: `max_epochs` defaults to the reference heuristic; optimizer and KL warm-up defaults are
        `TrainingPlan`'s (Adam lr 1e-3, eps 0.01, weight decay 1e-6, 400 warm-up epochs).  Returns the per-epoch
        history (train loss, reconstruction loss, KL per cell)."""
        if max_epochs is None:
            max_epochs = get_max_epochs_heuristic(self.data.n_obs)
        self.module.train()
        hyper = (batch_size, lr, eps, weight_decay, n_epochs_kl_warmup, n_steps_kl_warmup, max_kl_weight, min_kl_weight,
                 seed, use_cuda_graph)
        # a changed optimizer / warm-up / seed argument takes effect: the plan (and its Adam state) is rebuilt
        new_plan = (self._plan is None or self._plan.pg is not process_group
                    or getattr(self, "_plan_hyper", None) != hyper)
        self._plan_hyper = hyper
        if self.streamed and self.data.B != batch_size:
            raise ValueError(f"HostStreamedCounts was built for batch_size={self.data.B}, train() got {batch_size}")
        if new_plan:
            self._plan = FusedTrainingPlan(
                self.module, self._plan_counts(batch_size), batch_size=batch_size, lr=lr, eps=eps,
                weight_decay=weight_decay,
                n_epochs_kl_warmup=n_epochs_kl_warmup, n_steps_kl_warmup=n_steps_kl_warmup,
                max_kl_weight=max_kl_weight, min_kl_weight=min_kl_weight, seed=seed, use_cuda_graph=use_cuda_graph,
                process_group=process_group)
        if self.streamed:
            hist = self._plan.fit_streamed(self.data, max_epochs=max_epochs, shuffle=shuffle)
        elif self.sparse:
            hist = self._plan.fit_csr(self.data, max_epochs=max_epochs, shuffle=shuffle)
        else:
            hist = self._plan.fit(max_epochs=max_epochs, shuffle=shuffle)
        self.history_.extend(hist)
        self.module.eval()
        self.is_trained_ = True
        return hist

    def _plan_counts(self, batch_size):
        """The dense fp32 rows the plan's kernels read: the resident matrix itself, or a staging matrix."""
        if self.streamed:
            return self.data.stage
        if self.sparse:
            return DeviceResidentCounts(torch.zeros(batch_size, self.data.n_vars),
                                        batch=torch.zeros(batch_size, dtype=torch.int64), device=self.device)
        return self.data

    # ------------------------------------------------------------------ getters
    def _batches(self, indices, batch_size):
        n = self.data.n_obs
        idev = "cpu" if self.streamed else self.device
        idx = torch.arange(n, device=idev) if indices is None else torch.as_tensor(
            np.asarray(indices), dtype=torch.int64, device=idev)
        for s in range(0, idx.numel(), batch_size):
            yield self.data.minibatch(idx[s : s + batch_size])

    @torch.inference_mode()
    def get_latent_representation(self, indices=None, give_mean: bool = True, batch_size: int = 8192,
                                  return_dist: bool = False):
        """Mean of q(z | x) per cell (`give_mean`), a sample otherwise, or (mean, variance) with `return_dist`."""
        self.module.eval()
        zs, means, variances = [], [], []
        for tensors in self._batches(indices, batch_size):
            out = self.module.inference(**self.module._get_inference_input(tensors))
            qz = out[MODULE_KEYS.QZ_KEY]
            if return_dist:
                means.append(qz.loc)
                variances.append(qz.scale.square())
            else:
                zs.append(qz.loc if give_mean else out[MODULE_KEYS.Z_KEY])
        if return_dist:
            return torch.cat(means).cpu().numpy(), torch.cat(variances).cpu().numpy()
        return torch.cat(zs).cpu().numpy()

    @torch.inference_mode()
    def get_elbo(self, indices=None, batch_size: int = 8192, return_mean: bool = True):
        """`-(reconstruction loss + KL)` averaged over the cells (the reference's sign: `-compute_elbo`)."""
        self.module.eval()
        vals = []
        for tensors in self._batches(indices, batch_size):
            _, _, lo = self.module(tensors, get_inference_input_kwargs={"full_forward_pass": True})
            vals.append(lo.reconstruction_loss["reconstruction_loss"] + lo.kl_local["kl_divergence_z"]
                        + lo.kl_local.get("kl_divergence_l", 0.0))
        v = torch.cat(vals)
        return -float(v.mean()) if return_mean else -v.cpu().numpy()

    @torch.inference_mode()
    def get_reconstruction_error(self, indices=None, batch_size: int = 8192, return_mean: bool = True):
        """`{"reconstruction_loss": -log p(x | z)}` averaged over the cells (or per cell)."""
        self.module.eval()
        vals = []
        for tensors in self._batches(indices, batch_size):
            _, _, lo = self.module(tensors, loss_kwargs={"kl_weight": 1},
                                   get_inference_input_kwargs={"full_forward_pass": True})
            vals.append(lo.reconstruction_loss["reconstruction_loss"])
        v = torch.cat(vals)
        return {"reconstruction_loss": float(v.mean()) if return_mean else v.cpu().numpy()}

    @torch.inference_mode()
    def get_marginal_ll(self, indices=None, n_mc_samples: int = 1000, n_mc_samples_per_pass: int = 50,
                        batch_size: int = 4096, return_mean: bool = True):
        """Importance-sampled marginal log-likelihood log p(x) (reference `VAEMixin.get_marginal_ll`,
        src/scvi/model/base/_vaemixin.py:102-194, over `VAE.marginal_ll`): mean over cells or per cell."""
        self.module.eval()
        vals = [self.module.marginal_ll(tensors, n_mc_samples, return_mean=False,
                                        n_mc_samples_per_pass=min(n_mc_samples_per_pass, n_mc_samples))
                for tensors in self._batches(indices, batch_size)]
        v = torch.cat(vals)
        return float(v.mean()) if return_mean else v.numpy()

    @torch.inference_mode()
    def get_normalized_expression(self, indices=None, transform_batch=None, library_size=1, n_samples: int = 1,
                                  batch_size: int = 4096, gene_indices=None):
        """Decoded expression `px_scale * library_size` (`library_size="latent"`: `px_rate`), averaged over the
        batches in `transform_batch` (None = each cell's own batch) and over `n_samples` posterior draws."""
        self.module.eval()
        key = "mu" if library_size == "latent" else "scale"
        scaling = 1.0 if library_size == "latent" else float(library_size)
        tb = [None] if transform_batch is None else (list(transform_batch) if np.ndim(transform_batch) else [transform_batch])
        gsel = None if gene_indices is None else torch.as_tensor(np.asarray(gene_indices), device=self.device)
        out = []
        for tensors in self._batches(indices, batch_size):
            per_batch = []
            for b in tb:
                _, gen = self.module(tensors, inference_kwargs={"n_samples": n_samples},
                                     generative_kwargs={"transform_batch": b}, compute_loss=False)
                e = gen["px"].get_normalized(key) * scaling
                if n_samples > 1:
                    e = e.mean(0)
                per_batch.append(e if gsel is None else e[..., gsel])
            out.append(torch.stack(per_batch).mean(0))
        return torch.cat(out).cpu().numpy()

    # ------------------------------------------------------------------ persistence
    def save(self, dir_path: str, prefix: str | None = None, overwrite: bool = False, **torch_save_kwargs):
        """Write `<dir_path>/<prefix>model.pt` in the reference's layout (`BaseModelClass.save`,
        src/scvi/model/base/_base_model.py:792-905; keys from src/scvi/model/base/_constants.py): a `torch.save`d
        dict {"model_state_dict", "var_names", "attr_dict"} where the state dict has the reference module's keys
        (plus its empty "pyro_param_store" entry) and `attr_dict` holds the trailing-underscore attributes
        (`init_params_`, `is_trained_`, `history_`, `summary_stats_`).  No AnnData is written: there is no AnnData
        registry in this driver, so `attr_dict` carries no `registry_`."""
        import os

        if os.path.exists(dir_path) and not overwrite:
            raise ValueError(f"{dir_path} already exists. Please provide another directory for saving.")
        os.makedirs(dir_path, exist_ok=True)
        sd = {k: v.detach().cpu() for k, v in self.module.state_dict().items()}
        sd["pyro_param_store"] = {"params": {}, "constraints": {}}
        attr = {"init_params_": self.init_params_, "is_trained_": self.is_trained_, "history_": self.history_,
                "summary_stats_": dict(self.summary)}
        torch.save({"model_state_dict": sd, "var_names": self.var_names, "attr_dict": attr},
                   os.path.join(dir_path, f"{prefix or ''}model.pt"), **torch_save_kwargs)

    @classmethod
    def load(cls, dir_path: str, X, batch=None, prefix: str | None = None, device="cuda:0", var_names=None):
        """Rebuild a model from `<dir_path>/<prefix>model.pt` written by `save` -- or by the reference's
        `SCVI.save` (same keys; `registry_` and unknown attributes are ignored) -- on the counts `X`
        (`BaseModelClass.load`, _base_model.py:907-1040; `_load_saved_files` / `_initialize_model`,
        src/scvi/model/base/_save_load.py:69-165).  `var_names`, when given, must equal the saved ones
        (`_validate_var_names`)."""
        import os

        path = os.path.join(dir_path, f"{prefix or ''}model.pt")
        try:
            blob = torch.load(path, map_location="cpu", weights_only=False)
        except FileNotFoundError as exc:
            raise ValueError(f"Failed to load model file at {path}.") from exc
        sd = dict(blob["model_state_dict"])
        sd.pop("pyro_param_store", None)
        attr = dict(blob["attr_dict"])
        if "init_params_" not in attr:
            raise ValueError("No init_params_ were saved by the model.")
        init = attr.pop("init_params_")
        if "non_kwargs" in init:
            non_kwargs = dict(init["non_kwargs"])
            kwargs = {k: v for j in init["kwargs"].values() for k, v in j.items()}
        else:  # pre-0.15 flat layout
            non_kwargs = {k: v for k, v in init.items() if not isinstance(v, dict)}
            kwargs = {k: v for j in init.values() if isinstance(j, dict) for k, v in j.items()}
            non_kwargs.pop("use_cuda", None)
        for k in ("datamodule", "registry", "adata"):
            non_kwargs.pop(k, None)
        saved_names = blob.get("var_names")
        if var_names is not None and saved_names is not None and not np.array_equal(
                np.asarray(var_names).astype(str), np.asarray(saved_names).astype(str)):
            raise ValueError("var_names for the passed data do not match those of the saved model")
        model = cls(X, batch=batch, device=device, var_names=saved_names if var_names is None else var_names,
                    **non_kwargs, **kwargs)
        model.module.load_state_dict(sd)
        model.module.eval()
        model.is_trained_ = bool(attr.get("is_trained_", True))
        model.history_ = attr.get("history_", []) if isinstance(attr.get("history_", []), list) else []
        return model

    def state_dict(self):
        return self.module.state_dict()

    def load_state_dict(self, sd):
        """Load a state dict with the reference's keys (`scvi.module.VAE.state_dict()`); marks the model trained."""
        self.module.load_state_dict(sd)
        self._plan = None  # the parameter arena is rebuilt on the next train()
        self.is_trained_ = True
        self.module.eval()
