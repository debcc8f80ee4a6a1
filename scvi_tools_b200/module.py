"""`VAE` -- the SCVI module surface (`inference` / `generative` / `loss`) over the B200 kernels.

Mirrors the reference's src/scvi/module/_vae.py:31-600 and the module contract of
src/scvi/module/base/_base_module.py:27-209,458-488 (`LossOutput`, `forward` ->
`_generic_forward`), with the same constructor kwargs (so `SCVI.__init__` can build it,
src/scvi/model/_scvi.py:195-214), the same `state_dict` keys and the same dictionary keys
(`MODULE_KEYS`, `REGISTRY_KEYS`).
"""

from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Literal, NamedTuple

import numpy as np
import torch
from torch import nn
from torch.distributions import Normal, kl_divergence
from torch.nn.functional import one_hot

from . import _lib
from .distributions import NegativeBinomial, ZeroInflatedNegativeBinomial
from .nn import DecoderSCVI, Embedding, Encoder, _next_noise


class _REGISTRY_KEYS_NT(NamedTuple):  # reference src/scvi/_constants.py:4-21
    X_KEY: str = "X"
    BATCH_KEY: str = "batch"
    LABELS_KEY: str = "labels"
    CAT_COVS_KEY: str = "extra_categorical_covs"
    CONT_COVS_KEY: str = "extra_continuous_covs"
    INDICES_KEY: str = "ind_x"
    SIZE_FACTOR_KEY: str = "size_factor"
    LATENT_QZM_KEY: str = "latent_qzm"
    LATENT_QZV_KEY: str = "latent_qzv"
    OBSERVED_LIB_SIZE: str = "observed_lib_size"


class _MODULE_KEYS_NT(NamedTuple):  # reference src/scvi/module/_constants.py:4-27
    X_KEY: str = "x"
    Z_KEY: str = "z"
    QZ_KEY: str = "qz"
    QZM_KEY: str = "qzm"
    QZV_KEY: str = "qzv"
    LIBRARY_KEY: str = "library"
    QL_KEY: str = "ql"
    BATCH_INDEX_KEY: str = "batch_index"
    Y_KEY: str = "y"
    CONT_COVS_KEY: str = "cont_covs"
    CAT_COVS_KEY: str = "cat_covs"
    SIZE_FACTOR_KEY: str = "size_factor"
    PX_KEY: str = "px"
    PL_KEY: str = "pl"
    PZ_KEY: str = "pz"
    KL_L_KEY: str = "kl_divergence_l"
    KL_Z_KEY: str = "kl_divergence_z"


REGISTRY_KEYS = _REGISTRY_KEYS_NT()
MODULE_KEYS = _MODULE_KEYS_NT()


def _dict_sum(d):
    return sum(d.values()) if isinstance(d, dict) else d


@dataclass
class LossOutput:
    """Loss record of a minibatch (reference src/scvi/module/base/_base_module.py:27-141):
    `loss` is back-propagated; the per-cell pieces and their sums feed logging."""

    loss: Any
    reconstruction_loss: Any = None
    kl_local: Any = None
    kl_global: Any = None
    classification_loss: Any = None
    logits: Any = None
    true_labels: Any = None
    extra_metrics: dict | None = field(default_factory=dict)
    n_obs_minibatch: int | None = None
    reconstruction_loss_sum: Any = field(default=None)
    kl_local_sum: Any = field(default=None)
    kl_global_sum: Any = field(default=None)

    def __post_init__(self):
        self.loss = _dict_sum(self.loss)
        if self.n_obs_minibatch is None and self.reconstruction_loss is None:
            raise ValueError("Must provide either n_obs_minibatch or reconstruction_loss")
        zero = 0 * self.loss
        for name in ("reconstruction_loss", "kl_local", "kl_global"):
            v = getattr(self, name)
            v = zero if v is None else v
            setattr(self, name, v if isinstance(v, dict) else {name: v})
        self.reconstruction_loss_sum = _dict_sum(self.reconstruction_loss).sum()
        self.kl_local_sum = _dict_sum(self.kl_local).sum()
        self.kl_global_sum = _dict_sum(self.kl_global)
        if self.n_obs_minibatch is None:
            self.n_obs_minibatch = list(self.reconstruction_loss.values())[0].shape[0]
        if self.classification_loss is not None and (self.logits is None or self.true_labels is None):
            raise ValueError("Must provide `logits` and `true_labels` if `classification_loss` is provided.")

    @staticmethod
    def dict_sum(dictionary):
        return _dict_sum(dictionary)

    def detach_(self):
        """Drop every reference to the autograd graph (after `backward`): the record then only carries values."""
        def d(v):
            if isinstance(v, torch.Tensor):
                return v.detach()
            if isinstance(v, dict):
                return {k: d(x) for k, x in v.items()}
            return v

        for f in ("loss", "reconstruction_loss", "kl_local", "kl_global", "classification_loss", "logits", "true_labels",
                  "extra_metrics", "reconstruction_loss_sum", "kl_local_sum", "kl_global_sum"):
            setattr(self, f, d(getattr(self, f)))
        return self

    @property
    def extra_metrics_keys(self):
        return self.extra_metrics.keys()


class FusedCountDistribution:
    """What `generative` returns as ``px``: the (ZI)NB over genes, kept *lazy*.

    `nll(x)` runs the fused decoder-heads + softmax + likelihood kernel, so the cells x genes
    tensors are never written to HBM.  Touching `.mu/.theta/.scale/.zi_logits/.log_prob/.sample/
    .get_normalized` materialises the reference's distribution object (NegativeBinomial /
    ZeroInflatedNegativeBinomial) for the inference-time getters.
    """

    def __init__(self, decoder: DecoderSCVI, px_hidden, library, px_r_log, gene_likelihood, dispersion, precision):
        self._decoder, self._h, self._library = decoder, px_hidden, library
        self._px_r_log, self._lik, self._dispersion, self._precision = px_r_log, gene_likelihood, dispersion, precision
        self._dist = None

    def nll(self, x, row_idx=None):
        return self._decoder.nll(self._h, self._library, x, self._px_r_log, self._lik, row_idx, self._precision)

    def _materialise(self):
        if self._dist is None:
            dec = self._decoder
            h = self._h
            needs_grad = torch.is_grad_enabled() and h.requires_grad
            if needs_grad:
                px_scale = dec.px_scale_decoder(h)
                px_dropout = dec.px_dropout_decoder(h)
                px_rate = torch.exp(self._library) * px_scale
            else:
                from . import ops

                sc, dr = dec.px_scale_decoder[0], dec.px_dropout_decoder
                px_scale, px_rate, px_dropout = ops.decoder_params(
                    h.detach(), sc.weight.detach(), sc.bias.detach(), dr.weight.detach(), dr.bias.detach(),
                    self._library.detach(),
                )
            theta = torch.exp(self._px_r_log)
            if self._lik == "zinb":
                self._dist = ZeroInflatedNegativeBinomial(mu=px_rate, theta=theta, zi_logits=px_dropout, scale=px_scale)
            else:
                self._dist = NegativeBinomial(mu=px_rate, theta=theta, scale=px_scale)
        return self._dist

    def __getattr__(self, name):  # only called for attributes not found normally
        if name.startswith("_"):
            raise AttributeError(name)
        return getattr(self._materialise(), name)

    def __repr__(self):
        return f"FusedCountDistribution({self._lik}, lazy={self._dist is None})"


def auto_move_data(fn):
    """Move tensor arguments to the module's device in eval mode only (reference
    src/scvi/module/base/_decorators.py:9-43)."""
    from functools import wraps

    def _move(obj, device):
        if isinstance(obj, torch.Tensor):
            return obj.to(device)
        if isinstance(obj, dict):
            return {k: _move(v, device) for k, v in obj.items()}
        if isinstance(obj, (list, tuple)):
            return type(obj)(_move(v, device) for v in obj)
        return obj

    @wraps(fn)
    def wrapper(self, *args, **kwargs):
        if not isinstance(self, nn.Module) or self.training:
            return fn(self, *args, **kwargs)
        device = self.device
        return fn(self, *_move(args, device), **_move(kwargs, device))

    return wrapper


class VAE(nn.Module):
    """Variational auto-encoder of scVI (reference src/scvi/module/_vae.py:31-600).

    The default configuration (n_layers=1.., one-hot batch, observed library, dispersion
    "gene"/"gene-batch"/"gene-label", likelihood "zinb"/"nb") runs fused on the B200 kernels;
    see DESIGN.md for what is out of scope.
    """

    def __init__(
        self,
        n_input: int,
        n_batch: int = 0,
        n_labels: int = 0,
        n_hidden: int = 128,
        n_latent: int = 10,
        n_layers: int = 1,
        n_continuous_cov: int = 0,
        n_cats_per_cov: list[int] | None = None,
        dropout_rate: float = 0.1,
        dispersion: Literal["gene", "gene-batch", "gene-label", "gene-cell"] = "gene",
        log_variational: bool = True,
        gene_likelihood: Literal["zinb", "nb", "poisson", "normal"] = "zinb",
        latent_distribution: Literal["normal", "ln"] = "normal",
        encode_covariates: bool = False,
        deeply_inject_covariates: bool = True,
        batch_representation: Literal["one-hot", "embedding"] = "one-hot",
        use_batch_norm: Literal["encoder", "decoder", "none", "both"] = "both",
        use_layer_norm: Literal["encoder", "decoder", "none", "both"] = "none",
        use_size_factor_key: bool = False,
        use_observed_lib_size: bool = True,
        extra_payload_autotune: bool = False,
        library_log_means: np.ndarray | None = None,
        library_log_vars: np.ndarray | None = None,
        var_activation=None,
        extra_encoder_kwargs: dict | None = None,
        extra_decoder_kwargs: dict | None = None,
        batch_embedding_kwargs: dict | None = None,
        precision: int | None = None,
    ):
        super().__init__()
        if gene_likelihood not in ("zinb", "nb"):
            raise NotImplementedError(
                f"gene_likelihood={gene_likelihood!r}: the B200 hot path implements 'zinb' and 'nb' (DESIGN.md)"
            )
        if batch_representation not in ("one-hot", "embedding"):
            raise ValueError("`batch_representation` must be one of 'one-hot', 'embedding'.")
        self.dispersion = dispersion
        self.n_latent = n_latent
        self.log_variational = log_variational
        self.gene_likelihood = gene_likelihood
        self.n_batch = n_batch
        self.n_input = n_input
        self.n_labels = n_labels
        self.n_hidden = n_hidden
        self.n_layers = n_layers
        self.latent_distribution = latent_distribution
        self.encode_covariates = encode_covariates
        self.use_size_factor_key = use_size_factor_key
        self.use_observed_lib_size = use_size_factor_key or use_observed_lib_size
        self.extra_payload_autotune = extra_payload_autotune
        self.batch_representation = batch_representation
        # GEMM arithmetic of the fused kernels: None = the tcgen05 path (error-compensated fp16 split, fp32
        # accumulate, held to the same parity tolerances) whenever its shape constraint n_hidden == 128 holds,
        # FFMA fp32 otherwise; _lib.PREC_FP32 / _lib.PREC_TC force one
        # The tensor-core kernels scale their fp16 operands per cell by exp(library), which bounds every
        # per-gene derivative when the library IS the observed log-count sum; an encoded library
        # (use_observed_lib_size=False) gives no such bound, so that option runs on the FFMA kernels.
        observed = use_size_factor_key or use_observed_lib_size
        if precision is None:
            precision = _lib.PREC_TC if (n_hidden == 128 and observed) else _lib.PREC_FP32
        if precision == _lib.PREC_TC and n_hidden != 128:
            raise ValueError("precision=PREC_TC needs n_hidden == 128")
        if precision == _lib.PREC_TC and not observed:
            raise NotImplementedError("precision=PREC_TC needs the observed library size (use_observed_lib_size=True)")
        self.precision = precision
        self._minified_data_type = None

        if not self.use_observed_lib_size:
            if library_log_means is None or library_log_vars is None:
                raise ValueError(
                    "If not using observed_lib_size, must provide library_log_means and library_log_vars."
                )
            self.register_buffer("library_log_means", torch.from_numpy(library_log_means).float())
            self.register_buffer("library_log_vars", torch.from_numpy(library_log_vars).float())

        if dispersion == "gene":
            self.px_r = nn.Parameter(torch.randn(n_input))
        elif dispersion == "gene-batch":
            self.px_r = nn.Parameter(torch.randn(n_input, n_batch))
        elif dispersion == "gene-label":
            self.px_r = nn.Parameter(torch.randn(n_input, n_labels))
        elif dispersion != "gene-cell":
            raise ValueError("`dispersion` must be one of 'gene', 'gene-batch', 'gene-label', 'gene-cell'.")

        bn_enc = use_batch_norm in ("encoder", "both")
        bn_dec = use_batch_norm in ("decoder", "both")
        ln_enc = use_layer_norm in ("encoder", "both")
        ln_dec = use_layer_norm in ("decoder", "both")
        n_input_encoder = n_input + n_continuous_cov * encode_covariates
        batch_dim = 0
        if batch_representation == "embedding":
            # the batch enters as a learned vector concatenated to the layer inputs instead of one-hot columns
            # (reference _vae.py:218-236)
            self.init_embedding(REGISTRY_KEYS.BATCH_KEY, n_batch, **(batch_embedding_kwargs or {}))
            batch_dim = self.get_embedding(REGISTRY_KEYS.BATCH_KEY).embedding_dim
            n_input_encoder += batch_dim * encode_covariates
            cat_list = list([] if n_cats_per_cov is None else n_cats_per_cov)
        else:
            cat_list = [n_batch] + list([] if n_cats_per_cov is None else n_cats_per_cov)
        encoder_cat_list = cat_list if encode_covariates else None
        enc_kw = extra_encoder_kwargs or {}
        self.z_encoder = Encoder(
            n_input_encoder, n_latent, n_cat_list=encoder_cat_list, n_layers=n_layers, n_hidden=n_hidden,
            dropout_rate=dropout_rate, distribution=latent_distribution, inject_covariates=deeply_inject_covariates,
            use_batch_norm=bn_enc, use_layer_norm=ln_enc, var_activation=var_activation, return_dist=True, **enc_kw,
        )
        self.z_encoder.encoder.tc_linear = precision == _lib.PREC_TC  # tcgen05 input layer (needs n_hidden == 128)
        # library encoder: present (and in the state_dict) even when the library is observed (_vae.py:255-268)
        self.l_encoder = Encoder(
            n_input_encoder, 1, n_layers=1, n_cat_list=encoder_cat_list, n_hidden=n_hidden, dropout_rate=dropout_rate,
            inject_covariates=deeply_inject_covariates, use_batch_norm=bn_enc, use_layer_norm=ln_enc,
            var_activation=var_activation, return_dist=True, **enc_kw,
        )
        dec_kw = extra_decoder_kwargs or {}
        self.decoder = DecoderSCVI(
            n_latent + n_continuous_cov + batch_dim, n_input, n_cat_list=cat_list, n_layers=n_layers, n_hidden=n_hidden,
            inject_covariates=deeply_inject_covariates, use_batch_norm=bn_dec, use_layer_norm=ln_dec,
            scale_activation="softplus" if use_size_factor_key else "softmax", **dec_kw,
        )

    # ------------------------------------------------------------------ embeddings
    # `EmbeddingModuleMixin` (reference src/scvi/module/base/_embedding_mixin.py:8-50): tables live in a ModuleDict
    # registered as `_embeddings_dict`, which fixes the state_dict key `_embeddings_dict.<key>.weight`.
    @property
    def embeddings_dict(self) -> nn.ModuleDict:
        if not hasattr(self, "_embeddings_dict"):
            self._embeddings_dict = nn.ModuleDict()
        return self._embeddings_dict

    def add_embedding(self, key: str, embedding: Embedding, overwrite: bool = False) -> None:
        if key in self.embeddings_dict and not overwrite:
            raise KeyError(f"Embedding {key} already exists.")
        self.embeddings_dict[key] = embedding

    def remove_embedding(self, key: str) -> None:
        if key not in self.embeddings_dict:
            raise KeyError(f"Embedding {key} not found.")
        del self.embeddings_dict[key]

    def get_embedding(self, key: str) -> Embedding:
        if key not in self.embeddings_dict:
            raise KeyError(f"Embedding {key} not found.")
        return self.embeddings_dict[key]

    def init_embedding(self, key: str, num_embeddings: int, embedding_dim: int = 5, **kwargs) -> None:
        self.add_embedding(key, Embedding(num_embeddings, embedding_dim, **kwargs))

    def compute_embedding(self, key: str, indices: torch.Tensor) -> torch.Tensor:
        indices = indices.flatten() if indices.ndim > 1 else indices
        emb = self.get_embedding(key)
        return emb(indices.to(emb.weight.device))

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

    def on_load(self, model, **kwargs):
        """Callback run by `BaseModelClass.load` (reference _base_module.py:166-167)."""

    def _get_inference_input(self, tensors, full_forward_pass: bool = False):
        if full_forward_pass or self.minified_data_type is None:
            return {
                MODULE_KEYS.X_KEY: tensors[REGISTRY_KEYS.X_KEY],
                MODULE_KEYS.BATCH_INDEX_KEY: tensors[REGISTRY_KEYS.BATCH_KEY],
                MODULE_KEYS.CONT_COVS_KEY: tensors.get(REGISTRY_KEYS.CONT_COVS_KEY, None),
                MODULE_KEYS.CAT_COVS_KEY: tensors.get(REGISTRY_KEYS.CAT_COVS_KEY, None),
            }
        if self.minified_data_type in ("latent_posterior_parameters", "latent_posterior_parameters_with_counts"):
            return {
                MODULE_KEYS.QZM_KEY: tensors[REGISTRY_KEYS.LATENT_QZM_KEY],
                MODULE_KEYS.QZV_KEY: tensors[REGISTRY_KEYS.LATENT_QZV_KEY],
                REGISTRY_KEYS.OBSERVED_LIB_SIZE: tensors[REGISTRY_KEYS.OBSERVED_LIB_SIZE],
            }
        raise NotImplementedError(f"Unknown minified-data type: {self.minified_data_type}")

    def _get_generative_input(self, tensors, inference_outputs):
        size_factor = tensors.get(REGISTRY_KEYS.SIZE_FACTOR_KEY, None)
        if size_factor is not None:
            size_factor = torch.log(size_factor)
        return {
            MODULE_KEYS.Z_KEY: inference_outputs[MODULE_KEYS.Z_KEY],
            MODULE_KEYS.LIBRARY_KEY: inference_outputs[MODULE_KEYS.LIBRARY_KEY],
            MODULE_KEYS.BATCH_INDEX_KEY: tensors[REGISTRY_KEYS.BATCH_KEY],
            MODULE_KEYS.Y_KEY: tensors.get(REGISTRY_KEYS.LABELS_KEY, None),
            MODULE_KEYS.CONT_COVS_KEY: tensors.get(REGISTRY_KEYS.CONT_COVS_KEY, None),
            MODULE_KEYS.CAT_COVS_KEY: tensors.get(REGISTRY_KEYS.CAT_COVS_KEY, None),
            MODULE_KEYS.SIZE_FACTOR_KEY: size_factor,
        }

    def _compute_local_library_params(self, batch_index):
        n_batch = self.library_log_means.shape[1]
        oh = one_hot(batch_index.squeeze(-1), n_batch).float()
        return nn.functional.linear(oh, self.library_log_means), nn.functional.linear(oh, self.library_log_vars)

    # ------------------------------------------------------------------ inference
    @auto_move_data
    def inference(self, *args, **kwargs):
        if "qzm" in kwargs and "qzv" in kwargs:
            return self._cached_inference(*args, **kwargs)
        return self._regular_inference(*args, **kwargs)

    @auto_move_data
    def _regular_inference(self, x, batch_index, cont_covs=None, cat_covs=None, n_samples: int = 1):
        """Encoder pass (reference _vae.py:360-417).  `log1p(x)` and `library = log(sum x)` are
        fused into the first encoder GEMM when no dense covariate is concatenated to x."""
        cats = torch.split(cat_covs, 1, dim=1) if (cat_covs is not None and self.encode_covariates) else ()
        embed = self.batch_representation == "embedding"
        # with the embedding representation the encoders take no batch category (reference _vae.py:385-396)
        bcat = () if embed else (batch_index,)
        dense_cov = (cont_covs is not None or embed) and self.encode_covariates
        library = None
        if dense_cov or x.dim() != 2 or not self.z_encoder.encoder.is_fusable():
            if self.use_observed_lib_size:
                library = torch.log(x.sum(1)).unsqueeze(1)
            x_ = torch.log1p(x) if self.log_variational else x
            enc_in = torch.cat((x_, cont_covs), dim=-1) if (cont_covs is not None and self.encode_covariates) else x_
            if embed and self.encode_covariates:
                enc_in = torch.cat([enc_in, self.compute_embedding(REGISTRY_KEYS.BATCH_KEY, batch_index)], dim=-1)
            qz, z = self.z_encoder(enc_in, *bcat, *cats)
        else:
            flags = (_lib.LIN_LOG1P if self.log_variational else 0) | (
                _lib.LIN_LIBRARY if self.use_observed_lib_size else 0
            )
            enc_in = x  # the transform happens inside the kernel
            out = self.z_encoder(x, *bcat, *cats, _first_flags=flags)
            if self.use_observed_lib_size:
                qz, z, library = out
            else:
                qz, z = out
        ql = None
        if not self.use_observed_lib_size:
            x_in = torch.log1p(x) if (self.log_variational and enc_in is x) else enc_in
            ql, library = self.l_encoder(x_in, *bcat, *cats)
        if n_samples > 1:  # `qz.sample((n_samples,))` (:402-411), drawn on the device (or from `rng_feed`)
            with torch.no_grad():
                untran_z = qz.loc + qz.scale * _next_noise((n_samples, *qz.loc.shape), qz.loc.device)
            z = self.z_encoder.z_transformation(untran_z)
            if self.use_observed_lib_size:
                library = library.unsqueeze(0).expand((n_samples, library.size(0), library.size(1)))
            else:
                with torch.no_grad():
                    library = ql.loc + ql.scale * _next_noise((n_samples, *ql.loc.shape), ql.loc.device)
        return {MODULE_KEYS.Z_KEY: z, MODULE_KEYS.QZ_KEY: qz, MODULE_KEYS.QL_KEY: ql, MODULE_KEYS.LIBRARY_KEY: library}

    @auto_move_data
    def _cached_inference(self, qzm, qzv, observed_lib_size, n_samples: int = 1):
        qz = Normal(qzm, qzv.sqrt())
        untran = qz.sample() if n_samples == 1 else qz.sample((n_samples,))
        z = self.z_encoder.z_transformation(untran)
        library = torch.log(observed_lib_size)
        if n_samples > 1:
            library = library.unsqueeze(0).expand((n_samples, library.size(0), library.size(1)))
        return {MODULE_KEYS.Z_KEY: z, MODULE_KEYS.QZ_KEY: qz, MODULE_KEYS.QL_KEY: None, MODULE_KEYS.LIBRARY_KEY: library}

    # ------------------------------------------------------------------ generative
    @auto_move_data
    def generative(self, z, library, batch_index, cont_covs=None, cat_covs=None, size_factor=None, y=None,
                   transform_batch=None):
        """Decoder pass (reference _vae.py:445-549).  Returns the lazy fused ``px`` when the
        configuration allows (2-D z, softmax scale, dispersion != gene-cell)."""
        if cont_covs is None:
            dec_in = z
        elif z.dim() != cont_covs.dim():
            dec_in = torch.cat([z, cont_covs.unsqueeze(0).expand(z.size(0), -1, -1)], dim=-1)
        else:
            dec_in = torch.cat([z, cont_covs], dim=-1)
        cats = torch.split(cat_covs, 1, dim=1) if cat_covs is not None else ()
        if transform_batch is not None:
            batch_index = torch.ones_like(batch_index) * transform_batch
        if not self.use_size_factor_key:
            size_factor = library

        if self.dispersion == "gene-label":
            px_r = nn.functional.linear(one_hot(y.squeeze(-1), self.n_labels).float(), self.px_r)
        elif self.dispersion == "gene-batch":
            px_r = nn.functional.linear(one_hot(batch_index.squeeze(-1), self.n_batch).float(), self.px_r)
        elif self.dispersion == "gene":
            px_r = self.px_r
        else:
            px_r = None

        if self.batch_representation == "embedding":  # reference _vae.py:489-498
            dec_in = torch.cat([dec_in, self.compute_embedding(REGISTRY_KEYS.BATCH_KEY, batch_index)], dim=-1)
            h = self.decoder.hidden(dec_in, *cats)
        else:
            h = self.decoder.hidden(dec_in, batch_index, *cats)
        fused = self.dispersion == "gene" and self.decoder.supports_fused_nll(h) and size_factor.dim() == 2
        if fused:
            px = FusedCountDistribution(self.decoder, h, size_factor, px_r, self.gene_likelihood, self.dispersion,
                                        self.precision)
        else:
            dec = self.decoder
            px_scale = dec.px_scale_decoder(h)
            px_dropout = dec.px_dropout_decoder(h)
            px_rate = torch.exp(size_factor) * px_scale
            if self.dispersion == "gene-cell":
                px_r = dec.px_r_decoder(h)
            theta = torch.exp(px_r)
            if self.gene_likelihood == "zinb":
                px = ZeroInflatedNegativeBinomial(mu=px_rate, theta=theta, zi_logits=px_dropout, scale=px_scale)
            else:
                px = NegativeBinomial(mu=px_rate, theta=theta, scale=px_scale)

        if self.use_observed_lib_size:
            pl = None
        else:
            m, v = self._compute_local_library_params(batch_index)
            pl = Normal(m, v.sqrt())
        pz = Normal(torch.zeros_like(z), torch.ones_like(z), validate_args=False)
        return {MODULE_KEYS.PX_KEY: px, MODULE_KEYS.PL_KEY: pl, MODULE_KEYS.PZ_KEY: pz}

    # ------------------------------------------------------------------ loss
    def loss(self, tensors, inference_outputs, generative_outputs, kl_weight: torch.Tensor | float = 1.0):
        """ELBO pieces (reference _vae.py:551-600; signature pinned by the reference's
        tests/train/test_trainingplans.py:54-69)."""
        if self.minified_data_type == "latent_posterior_parameters":
            raise ValueError("The VAE.loss function currently does not support minified data.")
        x = tensors[REGISTRY_KEYS.X_KEY]
        qz = inference_outputs[MODULE_KEYS.QZ_KEY]
        kl_z = getattr(qz, "_b200_kl", None)
        if kl_z is None:
            kl_z = kl_divergence(qz, generative_outputs[MODULE_KEYS.PZ_KEY]).sum(dim=-1)
        if not self.use_observed_lib_size:
            kl_l = kl_divergence(inference_outputs[MODULE_KEYS.QL_KEY], generative_outputs[MODULE_KEYS.PL_KEY]).sum(dim=1)
        else:
            kl_l = torch.zeros_like(kl_z)
        px = generative_outputs[MODULE_KEYS.PX_KEY]
        reconst = px.nll(x) if isinstance(px, FusedCountDistribution) else -px.log_prob(x).sum(-1)
        loss = torch.mean(reconst + kl_weight * kl_z + kl_l)
        extra = {}
        if self.extra_payload_autotune:
            extra = {"z": inference_outputs["z"], "batch": tensors[REGISTRY_KEYS.BATCH_KEY],
                     "labels": tensors[REGISTRY_KEYS.LABELS_KEY]}
        return LossOutput(
            loss=loss,
            reconstruction_loss=reconst,
            kl_local={MODULE_KEYS.KL_L_KEY: kl_l, MODULE_KEYS.KL_Z_KEY: kl_z},
            extra_metrics=extra,
        )

    # ------------------------------------------------------------------ marginal likelihood
    @torch.inference_mode()
    def marginal_ll(self, tensors, n_mc_samples: int, return_mean: bool = False, n_mc_samples_per_pass: int = 1):
        """Importance-sampled log p(x) per cell (reference _vae.py:665-747): logsumexp over `n_mc_samples` draws
        z ~ q(z|x) of `log p(z) + log p(x|z) - log q(z|x)` (+ the library terms when it is encoded) - log n.

        In eval mode with the fused decoder the draws of one pass are flattened into the row dimension and the
        fused decoder + likelihood kernel reads each cell's counts once per draw through its row index -- no
        [n_samples, B, G] tensor (the reference materialises four) is ever formed."""
        import warnings

        n_per = int(n_mc_samples_per_pass)
        if n_per > n_mc_samples:
            warnings.warn("Number of chunks is larger than the total number of samples, setting it to the number "
                          "of samples", RuntimeWarning, stacklevel=2)
            n_per = n_mc_samples
        dev = self.device
        tensors = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in tensors.items()}
        x = tensors[REGISTRY_KEYS.X_KEY]
        batch_index = tensors[REGISTRY_KEYS.BATCH_KEY]
        B = x.shape[0]
        to_sum = []
        for _ in range(int(np.ceil(n_mc_samples / n_per))):
            inf = self.inference(**self._get_inference_input(tensors, full_forward_pass=True), n_samples=n_per)
            qz, ql = inf[MODULE_KEYS.QZ_KEY], inf[MODULE_KEYS.QL_KEY]
            z, library = inf[MODULE_KEYS.Z_KEY], inf[MODULE_KEYS.LIBRARY_KEY]
            cont = tensors.get(REGISTRY_KEYS.CONT_COVS_KEY, None)
            cat = tensors.get(REGISTRY_KEYS.CAT_COVS_KEY, None)
            flat = (not self.training and self.dispersion == "gene" and not self.use_size_factor_key
                    and cont is None and cat is None and self.decoder.scale_activation == "softmax"
                    and self.decoder.px_decoder.is_fusable() and self.n_hidden in (64, 128, 256))
            if flat:
                S = n_per if z.dim() == 3 else 1
                zf = z.reshape(S * B, -1)
                h = self.decoder.hidden(zf, batch_index.repeat(S, 1))
                rows = torch.arange(B, device=dev).repeat(S)
                reconst = self.decoder.nll(h, library.reshape(S * B, 1).contiguous(), x, self.px_r, self.gene_likelihood,
                                           rows, self.precision)
                reconst = reconst.view(S, B) if z.dim() == 3 else reconst
            else:
                gen = self.generative(**self._get_generative_input(tensors, inf))
                px = gen[MODULE_KEYS.PX_KEY]
                reconst = px.nll(x) if isinstance(px, FusedCountDistribution) else -px.log_prob(x).sum(-1)
            p_z = Normal(torch.zeros_like(qz.loc), torch.ones_like(qz.scale)).log_prob(z).sum(dim=-1)
            q_z_x = qz.log_prob(z).sum(dim=-1)
            log_prob_sum = p_z - reconst - q_z_x
            if not self.use_observed_lib_size:
                m, v = self._compute_local_library_params(batch_index)
                log_prob_sum = log_prob_sum + Normal(m, v.sqrt()).log_prob(library).sum(dim=-1) \
                    - ql.log_prob(library).sum(dim=-1)
            if n_per == 1:
                log_prob_sum = log_prob_sum.unsqueeze(0)
            to_sum.append(log_prob_sum)
        batch_log_lkl = torch.logsumexp(torch.cat(to_sum, dim=0), dim=0) - np.log(n_mc_samples)
        return torch.mean(batch_log_lkl).item() if return_mean else batch_log_lkl.cpu()

    # ------------------------------------------------------------------ forward
    @auto_move_data
    def forward(self, tensors, get_inference_input_kwargs=None, get_generative_input_kwargs=None,
                inference_kwargs=None, generative_kwargs=None, loss_kwargs=None, compute_loss=True):
        """inference -> generative -> loss (reference `_generic_forward`, _base_module.py:458-488)."""
        d = lambda v: v if isinstance(v, dict) else {}  # noqa: E731
        gi = dict(d(get_inference_input_kwargs))
        if not ("latent_qzm" in tensors and "latent_qzv" in tensors):
            gi.pop("full_forward_pass", None)
        inference_outputs = self.inference(**self._get_inference_input(tensors, **gi), **d(inference_kwargs))
        generative_inputs = self._get_generative_input(tensors, inference_outputs, **d(get_generative_input_kwargs))
        generative_outputs = self.generative(**generative_inputs, **d(generative_kwargs))
        if compute_loss:
            return inference_outputs, generative_outputs, self.loss(
                tensors, inference_outputs, generative_outputs, **d(loss_kwargs)
            )
        return inference_outputs, generative_outputs
