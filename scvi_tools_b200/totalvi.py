"""`TOTALVAE` -- the totalVI module over the B200 kernels.

Mirrors the reference's src/scvi/module/_totalvae.py:33-761 (`TOTALVAE.__init__`, `_regular_inference`,
`generative`, `get_reconstruction_loss`, `loss`) and src/scvi/nn/_base_components.py:721-1080
(`DecoderTOTALVI`, `EncoderTOTALVI`): same constructor keyword arguments, same `state_dict` keys, same
dictionaries in and out, same `LossOutput`.

What runs where: every `FCLayers` (encoder trunks, the four decoder trunks and the seven one-layer heads with
their `[hidden ‖ z]` skip input and one-hot batch injection as a weight-column gather) runs on the library's
linear / BatchNorm-ReLU-dropout kernels; the latent heads (mean, exp-variance, rsample, KL) on the latent-head
kernel; the gene likelihood on the NB / ZINB log-prob kernels and the protein likelihood on the NB-mixture
kernel (`scvi_mixture_nb_logprob_fwd/_bwd`).  The gene softmax and the O(B x proteins) prior / regulariser terms
are composed from torch elementwise ops on the device.  The fully fused decoder + likelihood kernel of `VAE`
does not apply here: totalVI's gene heads read `[px ‖ z ‖ one-hot(batch)]` (n_hidden + n_latent + n_batch
inputs), which that kernel's operand layout (n_hidden in {64, 128, 256}) does not cover yet.
"""

from __future__ import annotations

from collections.abc import Iterable
from typing import Literal

import numpy as np
import torch
import torch.nn.functional as F
from torch import nn
from torch.distributions import Normal
from torch.distributions import kl_divergence as kl
from torch.nn.functional import one_hot

from . import ops
from .distributions import NegativeBinomial, NegativeBinomialMixture, ZeroInflatedNegativeBinomial
from .module import MODULE_KEYS, REGISTRY_KEYS, LossOutput, auto_move_data
from .nn import FCLayers, _identity, _next_noise, _require_cuda

PROTEIN_EXP_KEY = "proteins"  # REGISTRY_KEYS.PROTEIN_EXP_KEY (reference src/scvi/_constants.py)


class ExpActivation(nn.Module):
    def forward(self, input):
        return torch.exp(input)


def _latent(q, mean_lin: nn.Linear, var_lin: nn.Linear, var_eps: float = 1e-4):
    """(Normal(q_m, sqrt(q_v)), rsample, KL(q || N(0,1)) per row) from the latent-head kernel; composed torch ops
    for 3-D inputs."""
    if q.dim() == 2 and mean_lin.out_features <= 32:
        noise = _next_noise((q.shape[0], mean_lin.out_features), q.device)
        q_m, q_v, sample, kl_row = ops.LatentHeads.apply(q, mean_lin.weight, mean_lin.bias, var_lin.weight,
                                                         var_lin.bias, noise, var_eps)
        dist = Normal(q_m, q_v.sqrt(), validate_args=False)
        dist._b200_kl = kl_row
        return dist, sample
    q_m = mean_lin(q)
    q_v = torch.exp(var_lin(q)) + var_eps
    dist = Normal(q_m, q_v.sqrt(), validate_args=False)
    dist._b200_kl = None
    return dist, q_m + dist.scale * _next_noise(q_m.shape, q.device)


class EncoderTOTALVI(nn.Module):
    """Encodes genes + proteins into q(z | x, y) and q(l | x, y) (reference _base_components.py:938-1080)."""

    def __init__(self, n_input: int, n_output: int, n_cat_list: Iterable[int] = None, n_layers: int = 2,
                 n_hidden: int = 256, dropout_rate: float = 0.1, distribution: str = "ln",
                 use_batch_norm: bool = True, use_layer_norm: bool = False):
        super().__init__()
        self.encoder = FCLayers(n_in=n_input, n_out=n_hidden, n_cat_list=n_cat_list, n_layers=n_layers,
                                n_hidden=n_hidden, dropout_rate=dropout_rate, use_batch_norm=use_batch_norm,
                                use_layer_norm=use_layer_norm)
        self.z_mean_encoder = nn.Linear(n_hidden, n_output)
        self.z_var_encoder = nn.Linear(n_hidden, n_output)
        self.l_gene_encoder = FCLayers(n_in=n_input, n_out=n_hidden, n_cat_list=n_cat_list, n_layers=1,
                                       n_hidden=n_hidden, dropout_rate=dropout_rate, use_batch_norm=use_batch_norm,
                                       use_layer_norm=use_layer_norm)
        self.l_gene_mean_encoder = nn.Linear(n_hidden, 1)
        self.l_gene_var_encoder = nn.Linear(n_hidden, 1)
        self.distribution = distribution
        self.z_transformation = nn.Softmax(dim=-1) if distribution == "ln" else _identity
        self.l_transformation = torch.exp

    def reparameterize_transformation(self, mu, var):
        untran_z = Normal(mu, var.sqrt()).rsample()
        return self.z_transformation(untran_z), untran_z

    def forward(self, data: torch.Tensor, *cat_list: int):
        _require_cuda(data, "EncoderTOTALVI")
        q = self.encoder(data, *cat_list)
        q_z, untran_z = _latent(q, self.z_mean_encoder, self.z_var_encoder)
        z = self.z_transformation(untran_z)
        ql_gene = self.l_gene_encoder(data, *cat_list)
        q_l, log_library_gene = _latent(ql_gene, self.l_gene_mean_encoder, self.l_gene_var_encoder)
        log_library_gene = torch.clamp(log_library_gene, max=15)
        library_gene = self.l_transformation(log_library_gene)
        return q_z, q_l, {"z": z, "l": library_gene}, {"z": untran_z, "l": log_library_gene}


class DecoderTOTALVI(nn.Module):
    """Decodes z into the (ZI)NB parameters of the genes and the NB-mixture parameters of the proteins
    (reference _base_components.py:721-935)."""

    def __init__(self, n_input: int, n_output_genes: int, n_output_proteins: int, n_cat_list: Iterable[int] = None,
                 n_layers: int = 1, n_hidden: int = 256, dropout_rate: float = 0, use_batch_norm: float = True,
                 use_layer_norm: float = False, scale_activation: Literal["softmax", "softplus"] = "softmax",
                 activation_function_bg: Literal["exp", "softplus"] = "softplus"):
        super().__init__()
        self.n_output_genes = n_output_genes
        self.n_output_proteins = n_output_proteins
        self.activation_function_bg = ExpActivation() if activation_function_bg == "exp" else nn.Softplus()
        linear_args = {"n_layers": 1, "use_activation": False, "use_batch_norm": False, "use_layer_norm": False,
                       "dropout_rate": 0}
        trunk = {"n_in": n_input, "n_out": n_hidden, "n_cat_list": n_cat_list, "n_layers": n_layers,
                 "n_hidden": n_hidden, "dropout_rate": dropout_rate, "use_batch_norm": use_batch_norm,
                 "use_layer_norm": use_layer_norm}
        self.px_decoder = FCLayers(**trunk)
        self.px_scale_decoder = FCLayers(n_in=n_hidden + n_input, n_out=n_output_genes, n_cat_list=n_cat_list, **linear_args)
        self.px_scale_activation = nn.Softmax(dim=-1) if scale_activation == "softmax" else nn.Softplus()
        self.py_back_decoder = FCLayers(**trunk)
        self.py_back_mean_log_alpha = FCLayers(n_in=n_hidden + n_input, n_out=n_output_proteins, n_cat_list=n_cat_list,
                                               **linear_args)
        self.py_back_mean_log_beta = FCLayers(n_in=n_hidden + n_input, n_out=n_output_proteins, n_cat_list=n_cat_list,
                                              **linear_args)
        self.py_fore_decoder = FCLayers(**trunk)
        self.py_fore_scale_decoder = FCLayers(n_in=n_hidden + n_input, n_out=n_output_proteins, n_cat_list=n_cat_list,
                                              n_layers=1, use_activation=True, use_batch_norm=False,
                                              use_layer_norm=False, dropout_rate=0, activation_fn=nn.ReLU)
        self.sigmoid_decoder = FCLayers(**trunk)
        self.px_dropout_decoder_gene = FCLayers(n_in=n_hidden + n_input, n_out=n_output_genes, n_cat_list=n_cat_list,
                                                **linear_args)
        self.py_background_decoder = FCLayers(n_in=n_hidden + n_input, n_out=n_output_proteins, n_cat_list=n_cat_list,
                                              **linear_args)

    def forward(self, z: torch.Tensor, library_gene: torch.Tensor, *cat_list: int):
        _require_cuda(z, "DecoderTOTALVI")
        px_, py_ = {}, {}
        px = self.px_decoder(z, *cat_list)
        px_cat_z = torch.cat([px, z], dim=-1)
        px_["scale"] = self.px_scale_activation(self.px_scale_decoder(px_cat_z, *cat_list))
        px_["rate"] = library_gene * px_["scale"]

        py_back = self.py_back_decoder(z, *cat_list)
        py_back_cat_z = torch.cat([py_back, z], dim=-1)
        py_["back_alpha"] = self.py_back_mean_log_alpha(py_back_cat_z, *cat_list)
        py_["back_beta"] = self.activation_function_bg(self.py_back_mean_log_beta(py_back_cat_z, *cat_list)) + 1e-8
        # Normal(back_alpha, back_beta).rsample()
        log_pro_back_mean = py_["back_alpha"] + py_["back_beta"] * _next_noise(py_["back_alpha"].shape, z.device)
        py_["rate_back"] = torch.exp(log_pro_back_mean)

        py_fore = self.py_fore_decoder(z, *cat_list)
        py_fore_cat_z = torch.cat([py_fore, z], dim=-1)
        py_["fore_scale"] = self.py_fore_scale_decoder(py_fore_cat_z, *cat_list) + 1 + 1e-8
        py_["rate_fore"] = py_["rate_back"] * py_["fore_scale"]

        p_mixing = self.sigmoid_decoder(z, *cat_list)
        p_mixing_cat_z = torch.cat([p_mixing, z], dim=-1)
        px_["dropout"] = self.px_dropout_decoder_gene(p_mixing_cat_z, *cat_list)
        py_["mixing"] = self.py_background_decoder(p_mixing_cat_z, *cat_list)
        protein_mixing = 1 / (1 + torch.exp(-py_["mixing"]))
        py_["scale"] = F.normalize((1 - protein_mixing) * py_["rate_fore"], p=1, dim=-1)
        return px_, py_, log_pro_back_mean


class TOTALVAE(nn.Module):
    """Total variational inference for CITE-seq data (reference _totalvae.py:33-761)."""

    def __init__(
        self,
        n_input_genes: int,
        n_input_proteins: int,
        n_batch: int = 0,
        n_labels: int = 0,
        n_hidden: int = 256,
        n_latent: int = 20,
        n_layers_encoder: int = 2,
        n_layers_decoder: int = 1,
        n_continuous_cov: int = 0,
        n_cats_per_cov: Iterable[int] | None = None,
        dropout_rate_decoder: float = 0.2,
        dropout_rate_encoder: float = 0.2,
        gene_dispersion: Literal["gene", "gene-batch", "gene-label"] = "gene",
        protein_dispersion: Literal["protein", "protein-batch", "protein-label"] = "protein",
        log_variational: bool = True,
        gene_likelihood: Literal["zinb", "nb"] = "nb",
        latent_distribution: Literal["normal", "ln"] = "normal",
        protein_batch_mask: dict[str | int, np.ndarray] = None,
        encode_covariates: bool = True,
        protein_background_prior_mean: np.ndarray | None = None,
        protein_background_prior_scale: np.ndarray | None = None,
        use_size_factor_key: bool = False,
        use_observed_lib_size: bool = True,
        extra_payload_autotune: bool = False,
        library_log_means: np.ndarray | None = None,
        library_log_vars: np.ndarray | None = None,
        n_panel: int | None = None,
        panel_key: str = REGISTRY_KEYS.BATCH_KEY,
        use_batch_norm: Literal["encoder", "decoder", "none", "both"] = "both",
        use_layer_norm: Literal["encoder", "decoder", "none", "both"] = "none",
        extra_encoder_kwargs: dict | None = None,
        extra_decoder_kwargs: dict | None = None,
    ):
        super().__init__()
        self.gene_dispersion = gene_dispersion
        self.n_latent = n_latent
        self.log_variational = log_variational
        self.gene_likelihood = gene_likelihood
        self.n_batch = n_batch
        self.n_labels = n_labels
        self.n_panel = n_panel if n_panel is not None else n_batch
        n_panel = self.n_panel
        self.panel_key = panel_key
        self.n_input_genes = n_input_genes
        self.n_input_proteins = n_input_proteins
        self.protein_dispersion = protein_dispersion
        self.latent_distribution = latent_distribution
        self.protein_batch_mask = protein_batch_mask
        self.encode_covariates = encode_covariates
        self.use_size_factor_key = use_size_factor_key
        self.extra_payload_autotune = extra_payload_autotune
        self.use_observed_lib_size = use_size_factor_key or use_observed_lib_size
        self._minified_data_type = None
        if not self.use_observed_lib_size:
            if library_log_means is None or library_log_vars is None:
                raise ValueError("If not using observed_lib_size, must provide library_log_means and library_log_vars.")
            self.register_buffer("library_log_means", torch.from_numpy(library_log_means).float())
            self.register_buffer("library_log_vars", torch.from_numpy(library_log_vars).float())

        # prior on the background protein mean (reference :137-166)
        if protein_background_prior_mean is None:
            if n_panel > 0:
                self.background_pro_alpha = nn.Parameter(3.0 * torch.ones(n_input_proteins, n_panel))
                self.background_pro_log_beta = nn.Parameter(torch.zeros(n_input_proteins, n_panel))
            else:
                self.background_pro_alpha = nn.Parameter(torch.ones(n_input_proteins))
                self.background_pro_log_beta = nn.Parameter(torch.zeros(n_input_proteins))
        else:
            if protein_background_prior_mean.shape[1] == 1 and n_panel != 1:
                init_mean = protein_background_prior_mean.ravel()
                init_scale = protein_background_prior_scale.ravel()
            else:
                init_mean, init_scale = protein_background_prior_mean, protein_background_prior_scale
            self.background_pro_alpha = nn.Parameter(torch.from_numpy(init_mean.astype(np.float32)))
            self.background_pro_log_beta = nn.Parameter(torch.log(torch.from_numpy(init_scale.astype(np.float32))))
        self.log_per_batch_efficiency = nn.Parameter(torch.zeros([n_input_proteins, n_batch]))

        if gene_dispersion == "gene":
            self.px_r = nn.Parameter(torch.randn(n_input_genes))
        elif gene_dispersion == "gene-batch":
            self.px_r = nn.Parameter(torch.randn(n_input_genes, n_batch))
        elif gene_dispersion == "gene-label":
            self.px_r = nn.Parameter(torch.randn(n_input_genes, n_labels))
        if protein_dispersion == "protein":
            self.py_r = nn.Parameter(2 * torch.rand(n_input_proteins))
        elif protein_dispersion == "protein-batch":
            self.py_r = nn.Parameter(2 * torch.rand(n_input_proteins, n_batch))
        elif protein_dispersion == "protein-label":
            self.py_r = nn.Parameter(2 * torch.rand(n_input_proteins, n_labels))

        bn_enc = use_batch_norm in ("encoder", "both")
        bn_dec = use_batch_norm in ("decoder", "both")
        ln_enc = use_layer_norm in ("encoder", "both")
        ln_dec = use_layer_norm in ("decoder", "both")
        n_input = n_input_genes + n_input_proteins
        n_input_encoder = n_input + n_continuous_cov * encode_covariates
        cat_list = [n_batch] + list([] if n_cats_per_cov is None else n_cats_per_cov)
        encoder_cat_list = cat_list if encode_covariates else None
        self.encoder = EncoderTOTALVI(
            n_input_encoder, n_latent, n_layers=n_layers_encoder, n_cat_list=encoder_cat_list, n_hidden=n_hidden,
            dropout_rate=dropout_rate_encoder, distribution=latent_distribution, use_batch_norm=bn_enc,
            use_layer_norm=ln_enc, **(extra_encoder_kwargs or {}),
        )
        self.decoder = DecoderTOTALVI(
            n_latent + n_continuous_cov, n_input_genes, n_input_proteins, n_layers=n_layers_decoder,
            n_cat_list=cat_list, n_hidden=n_hidden, dropout_rate=dropout_rate_decoder, use_batch_norm=bn_dec,
            use_layer_norm=ln_dec, scale_activation="softplus" if use_size_factor_key else "softmax",
            **(extra_decoder_kwargs or {}),
        )

    # ------------------------------------------------------------------ plumbing
    @property
    def device(self):
        devs = list({p.device for p in self.parameters()})
        if len(devs) > 1:
            raise RuntimeError("Module tensors on multiple devices.")
        return devs[0]

    @property
    def minified_data_type(self):
        return self._minified_data_type

    @minified_data_type.setter
    def minified_data_type(self, value):
        self._minified_data_type = value

    def _dispersions(self, batch_index, label):
        if self.gene_dispersion == "gene-label":
            px_r = F.linear(one_hot(label.squeeze(-1), self.n_labels).float(), self.px_r)
        elif self.gene_dispersion == "gene-batch":
            px_r = F.linear(one_hot(batch_index.squeeze(-1), self.n_batch).float(), self.px_r)
        else:
            px_r = self.px_r
        if self.protein_dispersion == "protein-label":
            py_r = F.linear(one_hot(label.squeeze(-1), self.n_labels).float(), self.py_r)
        elif self.protein_dispersion == "protein-batch":
            py_r = F.linear(one_hot(batch_index.squeeze(-1), self.n_batch).float(), self.py_r)
        else:
            py_r = self.py_r
        return torch.exp(px_r), torch.exp(py_r)

    def _get_inference_input(self, tensors, full_forward_pass: bool = False):
        if not (full_forward_pass or self.minified_data_type is None):
            raise NotImplementedError("minified-data inference is not part of the B200 totalVI path")
        return {
            MODULE_KEYS.X_KEY: tensors[REGISTRY_KEYS.X_KEY],
            MODULE_KEYS.Y_KEY: tensors[PROTEIN_EXP_KEY],
            MODULE_KEYS.BATCH_INDEX_KEY: tensors[REGISTRY_KEYS.BATCH_KEY],
            "panel_index": tensors[self.panel_key].long(),
            MODULE_KEYS.CONT_COVS_KEY: tensors.get(REGISTRY_KEYS.CONT_COVS_KEY, None),
            MODULE_KEYS.CAT_COVS_KEY: tensors.get(REGISTRY_KEYS.CAT_COVS_KEY, None),
        }

    def _get_generative_input(self, tensors, inference_outputs):
        return {
            "z": inference_outputs["z"],
            "library_gene": inference_outputs["library_gene"],
            "batch_index": tensors[REGISTRY_KEYS.BATCH_KEY],
            "label": tensors[REGISTRY_KEYS.LABELS_KEY],
            "cat_covs": tensors.get(REGISTRY_KEYS.CAT_COVS_KEY, None),
            "cont_covs": tensors.get(REGISTRY_KEYS.CONT_COVS_KEY, None),
            "size_factor": tensors.get(REGISTRY_KEYS.SIZE_FACTOR_KEY, None),
        }

    # ------------------------------------------------------------------ inference
    @auto_move_data
    def inference(self, x, y, batch_index=None, panel_index=None, label=None, n_samples=1, cont_covs=None,
                  cat_covs=None):
        """q(z | x, y), q(l | x, y) and the background prior of the minibatch (reference :418-509)."""
        x_ = x / (1 + x.mean(1, keepdim=True))
        y_ = y / (1 + y.mean(1, keepdim=True))
        if self.use_observed_lib_size:
            library_gene = x.sum(1).unsqueeze(1)
        if self.log_variational:
            x_ = torch.log(1 + x_)
            y_ = torch.log(1 + y_)
        if cont_covs is not None and self.encode_covariates is True:
            encoder_input = torch.cat((x_, y_, cont_covs), dim=-1)
        else:
            encoder_input = torch.cat((x_, y_), dim=-1)
        cats = torch.split(cat_covs, 1, dim=1) if (cat_covs is not None and self.encode_covariates is True) else ()
        qz, ql, latent, untran_latent = self.encoder(encoder_input, batch_index, *cats)
        z, untran_z, untran_l = latent["z"], untran_latent["z"], untran_latent["l"]
        if not self.use_observed_lib_size:
            library_gene = latent["l"]
        if n_samples > 1:
            with torch.no_grad():
                untran_z = qz.loc + qz.scale * _next_noise((n_samples, *qz.loc.shape), qz.loc.device)
                untran_l = ql.loc + ql.scale * _next_noise((n_samples, *ql.loc.shape), ql.loc.device)
            z = self.encoder.z_transformation(untran_z)
            if self.use_observed_lib_size:
                library_gene = library_gene.unsqueeze(0).expand((n_samples, library_gene.size(0), library_gene.size(1)))
            else:
                library_gene = self.encoder.l_transformation(untran_l)
        if self.n_panel > 0:
            oh = one_hot(panel_index.squeeze(-1), self.n_panel).float()
            py_back_alpha_prior = F.linear(oh, self.background_pro_alpha)
            py_back_beta_prior = F.linear(oh, torch.exp(self.background_pro_log_beta))
        else:
            py_back_alpha_prior = self.background_pro_alpha
            py_back_beta_prior = torch.exp(self.background_pro_log_beta)
        self.back_mean_prior = Normal(py_back_alpha_prior, py_back_beta_prior)
        return {MODULE_KEYS.Z_KEY: z, MODULE_KEYS.QZ_KEY: qz, MODULE_KEYS.QL_KEY: ql, "library_gene": library_gene,
                "untran_z": untran_z, "untran_l": untran_l, "back_mean_prior": self.back_mean_prior}

    # ------------------------------------------------------------------ generative
    @auto_move_data
    def generative(self, z, library_gene, batch_index, label, cont_covs=None, cat_covs=None, size_factor=None,
                   transform_batch: int | None = None):
        """Decoder pass (reference :318-385)."""
        if cont_covs is None:
            decoder_input = z
        elif z.dim() != cont_covs.dim():
            decoder_input = torch.cat([z, cont_covs.unsqueeze(0).expand(z.size(0), -1, -1)], dim=-1)
        else:
            decoder_input = torch.cat([z, cont_covs], dim=-1)
        cats = torch.split(cat_covs, 1, dim=1) if cat_covs is not None else ()
        if transform_batch is not None:
            batch_index = torch.ones_like(batch_index) * transform_batch
        per_batch_efficiency = torch.exp(
            F.linear(one_hot(batch_index.squeeze(-1), self.n_batch).float(), self.log_per_batch_efficiency))
        if not self.use_size_factor_key:
            size_factor = library_gene
        px_, py_, log_pro_back_mean = self.decoder(decoder_input, size_factor, batch_index, *cats)
        px_["r"], py_["r"] = self._dispersions(batch_index, label)
        py_norm_ = py_.copy()
        py_norm_["rate_back"] = per_batch_efficiency * py_["rate_back"]
        py_norm_["rate_fore"] = per_batch_efficiency * py_["rate_fore"]
        return {"px_": px_, "py_": py_, "py_norm_": py_norm_, "per_batch_efficiency": per_batch_efficiency,
                "log_pro_back_mean": log_pro_back_mean}

    # ------------------------------------------------------------------ loss
    def get_reconstruction_loss(self, x, y, px_dict, py_dict, pro_batch_mask_minibatch=None,
                                per_batch_efficiency=None):
        """-log p(x | z, l) per cell for genes and proteins (reference :233-280)."""
        px_, py_ = px_dict, py_dict
        if self.gene_likelihood == "zinb":
            gene = ZeroInflatedNegativeBinomial(mu=px_["rate"], theta=px_["r"], zi_logits=px_["dropout"])
        else:
            gene = NegativeBinomial(mu=px_["rate"], theta=px_["r"])
        reconst_loss_gene = -gene.log_prob(x).sum(dim=-1)
        if per_batch_efficiency is not None:
            mu1, mu2 = per_batch_efficiency * py_["rate_back"], per_batch_efficiency * py_["rate_fore"]
        else:
            mu1, mu2 = py_["rate_back"], py_["rate_fore"]
        py_conditional = NegativeBinomialMixture(mu1=mu1, mu2=mu2, theta1=py_["r"], mixture_logits=py_["mixing"])
        reconst_loss_protein_full = -py_conditional.log_prob(y)
        if pro_batch_mask_minibatch is not None:
            reconst_loss_protein = (pro_batch_mask_minibatch.bool() * reconst_loss_protein_full).sum(dim=-1)
        else:
            reconst_loss_protein = reconst_loss_protein_full.sum(dim=-1)
        return reconst_loss_gene, reconst_loss_protein

    def loss(self, tensors, inference_outputs, generative_outputs, pro_recons_weight=1.0, kl_weight=1.0):
        """ELBO of totalVI (reference :511-616)."""
        qz, ql = inference_outputs["qz"], inference_outputs["ql"]
        px_, py_ = generative_outputs["px_"], generative_outputs["py_"]
        per_batch_efficiency = generative_outputs["per_batch_efficiency"]
        x = tensors[REGISTRY_KEYS.X_KEY]
        batch_index = tensors[REGISTRY_KEYS.BATCH_KEY]
        panel_index = tensors[self.panel_key]
        y = tensors[PROTEIN_EXP_KEY]

        if self.protein_batch_mask is not None:
            pro_batch_mask_minibatch = torch.zeros_like(y)
            for b in torch.unique(panel_index):
                b_indices = (panel_index == b).reshape(-1)
                pro_batch_mask_minibatch[b_indices] = torch.tensor(
                    self.protein_batch_mask[str(int(b.item()))].astype(np.float32), device=y.device)
        else:
            pro_batch_mask_minibatch = None

        reconst_loss_gene, reconst_loss_protein = self.get_reconstruction_loss(
            x, y, px_, py_, pro_batch_mask_minibatch, per_batch_efficiency)

        kl_div_z = getattr(qz, "_b200_kl", None)
        if kl_div_z is None:
            kl_div_z = kl(qz, Normal(0, 1)).sum(dim=1)
        if not self.use_observed_lib_size:
            n_batch = self.library_log_means.shape[1]
            oh = one_hot(batch_index.squeeze(-1), n_batch).float()
            kl_div_l_gene = kl(ql, Normal(F.linear(oh, self.library_log_means),
                                          torch.sqrt(F.linear(oh, self.library_log_vars)))).sum(dim=1)
        else:
            kl_div_l_gene = torch.zeros_like(kl_div_z)

        kl_div_back_pro_full = kl(Normal(py_["back_alpha"], py_["back_beta"]), inference_outputs["back_mean_prior"])
        lkl_back_pro_full = -torch.distributions.LogNormal(
            x.new_zeros(1), x.new_ones(1)).log_prob(per_batch_efficiency)
        lkl_protein_expressed = -1e-3 * torch.distributions.Bernoulli(logits=py_["mixing"]).log_prob(
            torch.ones_like(py_["mixing"]))
        if pro_batch_mask_minibatch is not None:
            kl_div_back_pro = (pro_batch_mask_minibatch.bool() * kl_div_back_pro_full).sum(dim=1)
        else:
            kl_div_back_pro = kl_div_back_pro_full.sum(dim=1)
        kl_div_back_pro = kl_div_back_pro + lkl_back_pro_full.sum(dim=1) + lkl_protein_expressed.sum(dim=1)

        loss = torch.mean(reconst_loss_gene + kl_weight * pro_recons_weight * reconst_loss_protein
                          + kl_weight * kl_div_z + kl_div_l_gene + kl_weight * kl_div_back_pro)
        extra = {}
        if self.extra_payload_autotune:
            extra = {"z": inference_outputs["z"], "batch": tensors[REGISTRY_KEYS.BATCH_KEY],
                     "labels": tensors[REGISTRY_KEYS.LABELS_KEY]}
        return LossOutput(
            loss=loss,
            reconstruction_loss={"reconst_loss_gene": reconst_loss_gene, "reconst_loss_protein": reconst_loss_protein},
            kl_local={"kl_div_z": kl_div_z, "kl_div_l_gene": kl_div_l_gene, "kl_div_back_pro": kl_div_back_pro},
            extra_metrics=extra,
        )

    # ------------------------------------------------------------------ forward
    @auto_move_data
    def forward(self, tensors, get_inference_input_kwargs=None, get_generative_input_kwargs=None,
                inference_kwargs=None, generative_kwargs=None, loss_kwargs=None, compute_loss=True):
        """inference -> generative -> loss (reference `_generic_forward`, _base_module.py:458-488)."""
        d = lambda v: v if isinstance(v, dict) else {}  # noqa: E731
        gi = dict(d(get_inference_input_kwargs))
        gi.pop("full_forward_pass", None)
        inference_outputs = self.inference(**self._get_inference_input(tensors, **gi), **d(inference_kwargs))
        generative_inputs = self._get_generative_input(tensors, inference_outputs, **d(get_generative_input_kwargs))
        generative_outputs = self.generative(**generative_inputs, **d(generative_kwargs))
        if compute_loss:
            return inference_outputs, generative_outputs, self.loss(
                tensors, inference_outputs, generative_outputs, **d(loss_kwargs))
        return inference_outputs, generative_outputs
