"""`SCANVAE` -- the scANVI module over the same B200 kernels as `VAE`.

Mirrors the reference's src/scvi/module/_scanvae.py:26-278 (`SCANVAE.__init__`, `.loss`),
src/scvi/module/_classifier.py:7-79 (`Classifier`), src/scvi/nn/_base_components.py:501-574 (`Decoder`),
src/scvi/module/_utils.py:4-21 (`broadcast_labels`) and the supervised mixin
src/scvi/module/base/_base_module.py:491-560 (`classify`, `classification_loss`): same constructor keyword
arguments, same `state_dict` keys, same `LossOutput`.

What runs where: the [B, G] work -- encoder input layer, decoder heads + gene softmax + (ZI)NB likelihood --
is the fused VAE path unchanged; every `FCLayers` of the label-space networks (classifier trunk,
`encoder_z2_z1`, `decoder_z1_z2`) runs on the library's linear / BatchNorm-ReLU-dropout / latent-head kernels,
with the label one-hot executed as a weight-column gather (labels are passed as indices, never materialised
as the reference's [n_labels * B, n_labels] one-hot); the plain `nn.Linear` heads go through the same linear
kernel.  Only the O(B * n_labels) epilogue (Gaussian log-density of z1, categorical KL, cross-entropy) is
composed from torch elementwise ops on the device.
"""

from __future__ import annotations

from collections.abc import Iterable, Sequence
from typing import Literal

import numpy as np
import torch
from torch import nn
from torch.distributions import Normal
from torch.distributions import kl_divergence as kl
from torch.nn import functional as F

from . import _lib, ops
from .module import MODULE_KEYS, REGISTRY_KEYS, VAE, FusedCountDistribution, LossOutput, auto_move_data
from .nn import Encoder, FCLayers, _require_cuda


def _linear(x: torch.Tensor, lin: nn.Linear) -> torch.Tensor:
    """`nn.Linear` through the library's linear kernel (2-D CUDA fp32 input), else torch."""
    if x.dim() == 2 and x.is_cuda and x.dtype == torch.float32:
        return ops.FusedLinear.apply(x, lin.weight, lin.bias, None, lin.in_features, 0, None, False)[0]
    return lin(x)


def broadcast_labels(o: torch.Tensor, n_broadcast: int):
    """Every cell paired with every label (reference _utils.py:4-21), labels as an index column
    [n_broadcast * B, 1] (label-major, as the reference's `repeat_interleave`d one-hot rows)."""
    n = o.size(-2)
    ys = torch.arange(n_broadcast, device=o.device, dtype=torch.long).repeat_interleave(n).view(-1, 1)
    new_o = o.repeat(n_broadcast, 1) if o.dim() == 2 else o.repeat(1, n_broadcast, 1)
    return ys, new_o


class Classifier(nn.Module):
    """Fully-connected classifier (reference _classifier.py:7-79): `FCLayers` trunk -> Linear (-> Softmax)."""

    def __init__(self, n_input: int, n_hidden: int = 128, n_labels: int = 5, n_layers: int = 1,
                 dropout_rate: float = 0.1, logits: bool = False, use_batch_norm: bool = True,
                 use_layer_norm: bool = False, activation_fn: nn.Module = nn.ReLU, **kwargs):
        super().__init__()
        self.logits = logits
        layers = []
        if n_hidden > 0 and n_layers > 0:
            layers.append(FCLayers(n_in=n_input, n_out=n_hidden, n_layers=n_layers, n_hidden=n_hidden,
                                   dropout_rate=dropout_rate, use_batch_norm=use_batch_norm,
                                   use_layer_norm=use_layer_norm, activation_fn=activation_fn, **kwargs))
        else:
            n_hidden = n_input
        layers.append(nn.Linear(n_hidden, n_labels))
        if not logits:
            layers.append(nn.Softmax(dim=-1))
        self.classifier = nn.Sequential(*layers)

    def forward(self, x):
        _require_cuda(x, "Classifier")
        for layer in self.classifier:
            x = _linear(x, layer) if isinstance(layer, nn.Linear) else layer(x)
        return x


class Decoder(nn.Module):
    """Gaussian decoder: `FCLayers` -> (mean, exp(.) variance) (reference _base_components.py:501-574)."""

    def __init__(self, n_input: int, n_output: int, n_cat_list: Iterable[int] = None, n_layers: int = 1,
                 n_hidden: int = 128, **kwargs):
        super().__init__()
        self.decoder = FCLayers(n_in=n_input, n_out=n_hidden, n_cat_list=n_cat_list, n_layers=n_layers,
                                n_hidden=n_hidden, dropout_rate=0, **kwargs)
        self.mean_decoder = nn.Linear(n_hidden, n_output)
        self.var_decoder = nn.Linear(n_hidden, n_output)

    def forward(self, x: torch.Tensor, *cat_list: int):
        p = self.decoder(x, *cat_list)
        return _linear(p, self.mean_decoder), torch.exp(_linear(p, self.var_decoder))


class SCANVAE(VAE):
    """Single-cell annotation using variational inference (reference _scanvae.py:26-278)."""

    def __init__(
        self,
        n_input: int,
        n_batch: int = 0,
        n_labels: int = 0,
        n_hidden: int = 128,
        n_latent: int = 10,
        n_layers: int = 1,
        n_continuous_cov: int = 0,
        n_cats_per_cov: Iterable[int] | None = None,
        dropout_rate: float = 0.1,
        dispersion: Literal["gene", "gene-batch", "gene-label", "gene-cell"] = "gene",
        log_variational: bool = True,
        gene_likelihood: Literal["zinb", "nb"] = "zinb",
        use_observed_lib_size: bool = True,
        y_prior: torch.Tensor | None = None,
        labels_groups: Sequence[int] = None,
        linear_classifier: bool = False,
        classifier_parameters: dict | None = None,
        use_batch_norm: Literal["encoder", "decoder", "none", "both"] = "both",
        use_layer_norm: Literal["encoder", "decoder", "none", "both"] = "none",
        **vae_kwargs,
    ):
        super().__init__(
            n_input, n_hidden=n_hidden, n_latent=n_latent, n_layers=n_layers, n_continuous_cov=n_continuous_cov,
            n_cats_per_cov=n_cats_per_cov, dropout_rate=dropout_rate, n_batch=n_batch, dispersion=dispersion,
            log_variational=log_variational, gene_likelihood=gene_likelihood,
            use_observed_lib_size=use_observed_lib_size, use_batch_norm=use_batch_norm,
            use_layer_norm=use_layer_norm, **vae_kwargs,
        )
        classifier_parameters = classifier_parameters or {}
        bn_enc = use_batch_norm in ("encoder", "both")
        bn_dec = use_batch_norm in ("decoder", "both")
        ln_enc = use_layer_norm in ("encoder", "both")
        ln_dec = use_layer_norm in ("decoder", "both")
        self.n_labels = n_labels
        cls_parameters = {
            "n_layers": 0 if linear_classifier else n_layers,
            "n_hidden": 0 if linear_classifier else n_hidden,
            "dropout_rate": dropout_rate,
            "logits": True,
        }
        cls_parameters.update(classifier_parameters)
        self.classifier = Classifier(n_latent, n_labels=n_labels, use_batch_norm=bn_enc, use_layer_norm=ln_enc,
                                     **cls_parameters)
        self.encoder_z2_z1 = Encoder(n_latent, n_latent, n_cat_list=[self.n_labels], n_layers=n_layers,
                                     n_hidden=n_hidden, dropout_rate=dropout_rate, use_batch_norm=bn_enc,
                                     use_layer_norm=ln_enc, return_dist=True)
        self.decoder_z1_z2 = Decoder(n_latent, n_latent, n_cat_list=[self.n_labels], n_layers=n_layers,
                                     n_hidden=n_hidden, use_batch_norm=bn_dec, use_layer_norm=ln_dec)
        self.y_prior = nn.Parameter(
            y_prior if y_prior is not None else (1 / n_labels) * torch.ones(1, n_labels), requires_grad=False)
        self.labels_groups = np.array(labels_groups) if labels_groups is not None else None

    # ------------------------------------------------------------------ supervised mixin
    @auto_move_data
    def classify_helper(self, z):
        return self.classifier(z)

    @auto_move_data
    def classify(self, x, batch_index=None, cont_covs=None, cat_covs=None, use_posterior_mean: bool = True):
        """Encoder + classifier: logits [n_obs, n_labels] (reference _base_module.py:500-543).  `log1p` is
        fused into the encoder's input layer when no dense covariate is concatenated to x."""
        cats = torch.split(cat_covs, 1, dim=1) if (cat_covs is not None and self.encode_covariates) else ()
        dense_cov = cont_covs is not None and self.encode_covariates
        if self.log_variational and not dense_cov and x.dim() == 2 and self.z_encoder.encoder.is_fusable():
            qz, z = self.z_encoder(x, batch_index, *cats, _first_flags=_lib.LIN_LOG1P)
        else:
            x_ = torch.log1p(x) if self.log_variational else x
            enc_in = torch.cat((x_, cont_covs), dim=-1) if dense_cov else x_
            qz, z = self.z_encoder(enc_in, batch_index, *cats)
        z = qz.loc if use_posterior_mean else z
        return self.classify_helper(z)

    @auto_move_data
    def classification_loss(self, labelled_dataset):
        """Cross-entropy of `classify` on the labelled minibatch (reference _base_module.py:545-574)."""
        inference_inputs = self._get_inference_input(labelled_dataset)
        data_inputs = {k: v for k, v in inference_inputs.items()
                       if k not in ("batch_index", "cont_covs", "cat_covs", "panel_index")}
        y = labelled_dataset[REGISTRY_KEYS.LABELS_KEY]
        logits = self.classify(
            **data_inputs, batch_index=labelled_dataset[REGISTRY_KEYS.BATCH_KEY],
            cat_covs=labelled_dataset.get(REGISTRY_KEYS.CAT_COVS_KEY, None),
            cont_covs=labelled_dataset.get(REGISTRY_KEYS.CONT_COVS_KEY, None),
        )
        ce_loss = F.cross_entropy(logits, y.view(-1).long())
        return ce_loss, y, logits

    # ------------------------------------------------------------------ loss
    def loss(self, tensors, inference_outputs, generative_ouputs, kl_weight: float = 1.0,
             labelled_tensors=None, classification_ratio: float | None = None):
        """ELBO of the M1+M2 model (reference _scanvae.py:178-278)."""
        px = generative_ouputs[MODULE_KEYS.PX_KEY]
        qz1 = inference_outputs[MODULE_KEYS.QZ_KEY]
        z1 = inference_outputs[MODULE_KEYS.Z_KEY]
        x = tensors[REGISTRY_KEYS.X_KEY]
        batch_index = tensors[REGISTRY_KEYS.BATCH_KEY]

        ys, z1s = broadcast_labels(z1, n_broadcast=self.n_labels)
        if z1.dim() != 2:  # n_samples > 1: the reference's pre-encoded one-hot through the composed FCLayers
            ys = F.one_hot(ys.squeeze(-1), self.n_labels)
        qz2, z2 = self.encoder_z2_z1(z1s, ys)
        pz1_m, pz1_v = self.decoder_z1_z2(z2, ys)
        reconst_loss = px.nll(x) if isinstance(px, FusedCountDistribution) else -px.log_prob(x).sum(-1)

        kl_divergence_z2 = getattr(qz2, "_b200_kl", None)  # KL(q || N(0, I)) from the latent-head kernel
        if kl_divergence_z2 is None:
            kl_divergence_z2 = kl(qz2, Normal(torch.zeros_like(qz2.loc), torch.ones_like(qz2.scale))).sum(dim=-1)
        # Normal(.).log_prob(.).sum(-1): one launch each way instead of torch.distributions' element-wise chain
        loss_z1_unweight = -ops.normal_log_prob_sum(z1s, pz1_m, var=pz1_v)
        loss_z1_weight = ops.normal_log_prob_sum(z1, qz1.loc, scale=qz1.scale)

        probs = self.classifier(z1)
        if self.classifier.logits:
            probs = F.softmax(probs, dim=-1)

        if z1.ndim == 2:
            loss_z1_unweight_ = loss_z1_unweight.view(self.n_labels, -1).t()
            kl_divergence_z2_ = kl_divergence_z2.view(self.n_labels, -1).t()
        else:
            loss_z1_unweight_ = torch.transpose(loss_z1_unweight.view(z1.shape[0], self.n_labels, -1), -1, -2)
            kl_divergence_z2_ = torch.transpose(kl_divergence_z2.view(z1.shape[0], self.n_labels, -1), -1, -2)
        reconst_loss = reconst_loss + loss_z1_weight + (loss_z1_unweight_ * probs).sum(dim=-1)
        kl_divergence = (kl_divergence_z2_ * probs).sum(dim=-1)
        # kl(Categorical(probs), Categorical(y_prior repeated over the cells)), torch's construction, fused
        kl_divergence = kl_divergence + ops.categorical_kl(probs, self.y_prior)

        if not self.use_observed_lib_size:
            ql = inference_outputs[MODULE_KEYS.QL_KEY]
            m, v = self._compute_local_library_params(batch_index)
            kl_divergence_l = kl(ql, Normal(m, torch.sqrt(v))).sum(dim=1)
        else:
            kl_divergence_l = torch.zeros_like(kl_divergence)
        kl_divergence = kl_divergence + kl_divergence_l

        loss = torch.mean(reconst_loss + kl_divergence * kl_weight)
        extra = {}
        if self.extra_payload_autotune:
            extra = {"z": inference_outputs["z"], "batch": tensors[REGISTRY_KEYS.BATCH_KEY],
                     "labels": tensors[REGISTRY_KEYS.LABELS_KEY]}
        if labelled_tensors is not None:
            ce_loss, true_labels, logits = self.classification_loss(labelled_tensors)
            loss = loss + ce_loss * classification_ratio
            return LossOutput(loss=loss, reconstruction_loss=reconst_loss, kl_local=kl_divergence,
                              classification_loss=ce_loss, true_labels=true_labels, logits=logits,
                              extra_metrics=extra)
        return LossOutput(loss=loss, reconstruction_loss=reconst_loss, kl_local=kl_divergence, extra_metrics=extra)
