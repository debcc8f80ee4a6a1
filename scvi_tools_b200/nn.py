"""`FCLayers`, `Encoder`, `DecoderSCVI` -- the reference's nn surface over the B200 kernels.

Constructor signatures, attribute names and ``state_dict`` layout follow the reference
(src/scvi/nn/_base_components.py:17-442) so checkpoints are interchangeable; the forward
work is done by ``libscvi_b200.so`` through :mod:`scvi_tools_b200.ops`.  Options the kernels
do not cover (3-D inputs, LayerNorm, non-ReLU activations, pre-encoded one-hots) run as
composed torch ops on the same GPU tensors of the same module; CPU tensors are rejected.
"""

from __future__ import annotations

import collections
import contextlib
from collections.abc import Callable, Iterable
from typing import Literal

import torch
from torch import nn
from torch.distributions import Normal

from . import _lib, ops

# --------------------------------------------------------------------------------------
# RNG feed: parity tests hand the kernels the same dropout masks / noise as the oracle
# --------------------------------------------------------------------------------------
_FEED = {"drop_masks": None, "noise": None}


@contextlib.contextmanager
def rng_feed(drop_masks=None, noise=None):
    """Within the context, dropout layers consume ``drop_masks`` (list of uint8 keep-masks, in
    call order) and ``Encoder`` consumes ``noise`` (list of standard-normal draws) instead of
    drawing from torch's generator."""
    old = dict(_FEED)
    _FEED["drop_masks"] = list(drop_masks) if drop_masks is not None else None
    _FEED["noise"] = list(noise) if noise is not None else None
    try:
        yield
    finally:
        _FEED.update(old)


def _next_mask(shape, p, device):
    if _FEED["drop_masks"]:
        m = _FEED["drop_masks"].pop(0)
        return m.to(device=device, dtype=torch.uint8)
    # keep with probability 1 - p: one Philox launch (graph-safe), no uniform -> compare -> cast chain
    return torch.empty(shape, dtype=torch.uint8, device=device).bernoulli_(1.0 - p)


def _next_noise(shape, device):
    if _FEED["noise"]:
        return _FEED["noise"].pop(0).to(device=device, dtype=torch.float32)
    return torch.randn(shape, device=device)


def _require_cuda(x: torch.Tensor, who: str):
    if not x.is_cuda:
        raise RuntimeError(
            f"{who}: scvi_tools_b200 runs on CUDA (sm_100a) only; got a {x.device.type} tensor. "
            "There is no CPU fallback."
        )


def _identity(x):
    return x


class Embedding(nn.Embedding):
    """`scvi.nn.Embedding` (reference src/scvi/nn/_embedding.py:29-98): an embedding table that can be grown
    (`extend`) and that accepts a checkpoint whose table has a different number of rows.  On CUDA the lookup is
    this library's row gather and its gradient a fixed-order segmented sum."""

    @classmethod
    def extend(cls, embedding: "Embedding", init, freeze_prev: bool = True) -> "Embedding":
        """New table = the old rows followed by `init` fresh N(0,1) rows (int) or by copies of the rows listed in
        `init` (list of int).  With `freeze_prev` the old rows receive no gradient."""
        old = embedding.weight
        if isinstance(init, int):
            if init <= 0:
                raise ValueError(f"`init` must be greater than 0, got {init}")
            extra = torch.empty((init, embedding.embedding_dim), device=old.device)
            nn.init.normal_(extra)
        elif isinstance(init, list):
            extra = old[init].clone()
        else:
            raise TypeError(f"`init` must be an `int` or a `list[int]`, got {type(init)}")
        table = torch.cat([old.clone(), extra], dim=0)
        grown = cls(num_embeddings=table.shape[0], embedding_dim=table.shape[1], _weight=table,
                    padding_idx=embedding.padding_idx, max_norm=embedding.max_norm, norm_type=embedding.norm_type,
                    scale_grad_by_freq=embedding.scale_grad_by_freq, sparse=embedding.sparse)
        if freeze_prev:
            n_old = embedding.num_embeddings

            def _zero_old_rows(grad):
                grad = grad.clone()
                grad[:n_old] = 0.0
                return grad

            grown.weight.register_hook(_zero_old_rows)
        return grown

    def _load_from_state_dict(self, state_dict, *args, **kwargs):
        key = [k for k in state_dict.keys() if "weight" in k][0]
        self.weight = nn.Parameter(state_dict[key])
        self.num_embeddings, self.embedding_dim = self.weight.shape[0], self.weight.shape[1]
        return super()._load_from_state_dict(state_dict, *args, **kwargs)

    def forward(self, indices: torch.Tensor) -> torch.Tensor:
        plain = self.padding_idx is None and self.max_norm is None and not self.scale_grad_by_freq and not self.sparse
        if self.weight.is_cuda and plain and self.weight.dtype == torch.float32:
            flat = indices.reshape(-1).to(torch.int64)
            return ops.EmbeddingRows.apply(self.weight, flat).reshape(*indices.shape, self.embedding_dim)
        return super().forward(indices)


class FCLayers(nn.Module):
    """Fully-connected stack: Linear (+ covariate injection) -> BatchNorm -> LayerNorm -> activation -> Dropout.

    Same parameters as the reference (src/scvi/nn/_base_components.py:52-67).  The one-hot
    covariate concat is executed as a weight-column gather inside the linear kernel.
    """

    def __init__(
        self,
        n_in: int,
        n_out: int,
        n_cat_list: Iterable[int] = None,
        n_cont: int = 0,
        n_layers: int = 1,
        n_hidden: int = 128,
        dropout_rate: float = 0.1,
        use_batch_norm: bool = True,
        use_layer_norm: bool = False,
        use_activation: bool = True,
        bias: bool = True,
        inject_covariates: bool = True,
        activation_fn: nn.Module = nn.ReLU,
    ):
        super().__init__()
        self.bias = bias
        self.use_batch_norm = use_batch_norm
        self.use_layer_norm = use_layer_norm
        self.use_activation = use_activation
        self.activation_fn = activation_fn
        self.dropout_rate = dropout_rate
        self.inject_covariates = inject_covariates
        self.n_cont = n_cont
        dims = [n_in] + (n_layers - 1) * [n_hidden] + [n_out]
        # a covariate with a single category carries no information and is dropped (:80)
        self.n_cat_list = [c if c > 1 else 0 for c in n_cat_list] if n_cat_list is not None else []
        self.n_cov = n_cont + sum(self.n_cat_list)
        self.fc_layers = nn.Sequential(
            collections.OrderedDict(
                (f"Layer {i}", self._build_layer(d_in, d_out, i))
                for i, (d_in, d_out) in enumerate(zip(dims[:-1], dims[1:], strict=True))
            )
        )

    def inject_into_layer(self, layer_num) -> bool:
        return layer_num == 0 or (layer_num > 0 and self.inject_covariates)

    def _build_layer(self, n_in: int, n_out: int, layer_num: int) -> nn.Sequential:
        """Positional layout (indices are part of the state_dict keys): 0 Linear, 1 BatchNorm1d,
        2 LayerNorm, 3 activation, 4 Dropout; absent pieces are ``None``."""
        return nn.Sequential(
            nn.Linear(n_in + self.n_cov * self.inject_into_layer(layer_num), n_out, bias=self.bias),
            nn.BatchNorm1d(n_out, momentum=0.01, eps=0.001) if self.use_batch_norm else None,
            nn.LayerNorm(n_out, elementwise_affine=False) if self.use_layer_norm else None,
            self.activation_fn() if self.use_activation else None,
            nn.Dropout(p=self.dropout_rate) if self.dropout_rate > 0 else None,
        )

    def _is_linear_layer(self, layer: nn.Module) -> bool:
        return isinstance(layer, nn.Linear)

    def set_online_update_hooks(self, hook_first_layer=True):
        """scArches hooks: only the covariate columns of injected layers keep gradient
        (reference :121-146)."""
        self.hooks = []

        def keep_cov_columns(grad):
            n = sum(self.n_cat_list)
            out = torch.zeros_like(grad)
            if n > 0:
                out[..., -n:] = grad[..., -n:]
            return out

        def zero(grad):
            return grad * 0

        for i, layers in enumerate(self.fc_layers):
            for layer in layers:
                if i == 0 and not hook_first_layer:
                    continue
                if self._is_linear_layer(layer):
                    fn = keep_cov_columns if self.inject_into_layer(i) else zero
                    self.hooks.append(layer.weight.register_hook(fn))
                    self.hooks.append(layer.bias.register_hook(zero))

    # ------------------------------------------------------------------ forward
    def _layer_is_fusable(self, layers) -> bool:
        lin, bn, ln, act, drop = (layers[j] if j < len(layers) else None for j in range(5))
        return (
            len(layers) == 5
            and type(lin) is nn.Linear
            and (bn is None or type(bn) is nn.BatchNorm1d)
            # LayerNorm (no affine) has its own kernel; stacked on a BatchNorm it runs composed
            and (ln is None or (bn is None and type(ln) is nn.LayerNorm and not ln.elementwise_affine))
            and (act is None or type(act) is nn.ReLU)
            and (drop is None or type(drop) is nn.Dropout)
        )

    def is_fusable(self) -> bool:
        """True when every layer is Linear -> (BatchNorm) -> (ReLU) -> (Dropout), the form the kernels cover."""
        return all(self._layer_is_fusable(layers) for layers in self.fc_layers)

    def forward(self, x: torch.Tensor, *cat_list: int, cont: torch.Tensor | None = None, _first_flags: int = 0,
                _row_idx: torch.Tensor | None = None):
        """``x`` [B, n_in]; ``cat_list`` one (B,1) integer tensor per categorical covariate.

        Private kwargs (used by `VAE`): ``_first_flags`` fuses log1p / library into layer 0;
        ``_row_idx`` makes layer 0 read rows ``x[_row_idx]`` of a resident matrix directly.
        Returns the activations (and the library if requested through ``_first_flags``).
        """
        _require_cuda(x, "FCLayers")
        cat_list = cat_list or []
        if len(self.n_cat_list) > len(cat_list):
            raise ValueError("nb. categorical args provided doesn't match init. params.")
        for n_cat, cat in zip(self.n_cat_list, cat_list, strict=False):
            if n_cat and cat is None:
                raise ValueError("cat not provided while n_cat != 0 in init. params.")
        fusable = (
            x.dim() == 2
            and x.dtype == torch.float32
            and all(self._layer_is_fusable(layers) for layers in self.fc_layers)
            and all(not (n > 1 and c.size(1) == n) for n, c in zip(self.n_cat_list, cat_list, strict=False))
        )
        if not fusable:
            if _first_flags or _row_idx is not None:
                raise NotImplementedError("fused encoder pre-ops need a fusable FCLayers configuration")
            return self._forward_composed(x, cat_list, cont)

        # covariate columns of W: [x | cont | one-hot(cat_0) | one-hot(cat_1) ...]
        library = None
        for i, layers in enumerate(self.fc_layers):
            lin, bn, ln, act, drop = layers
            inject = self.inject_into_layer(i) and self.n_cov > 0
            n_dense = lin.in_features - (sum(self.n_cat_list) if inject else 0)
            if inject and cont is not None:
                x = torch.cat((x, cont.to(x.dtype)), dim=-1)
            cov_cols = None
            if inject and sum(self.n_cat_list) > 0:
                cols, off = [], n_dense
                for n_cat, cat in zip(self.n_cat_list, cat_list, strict=False):
                    if n_cat > 1:
                        cols.append(cat.reshape(-1, 1).to(torch.int32) + off)
                        off += n_cat
                cov_cols = torch.cat(cols, dim=1).contiguous()
            flags = _first_flags if i == 0 else 0
            if (getattr(self, "tc_linear", False) and (flags & _lib.LIN_LOG1P) and lin.out_features == 128
                    and cov_cols is None and x.dim() == 2):
                flags |= _lib.LIN_TC  # tcgen05 path of the encoder input layer
            training_bn = bn is not None and self.training
            y, part, lib = ops.FusedLinear.apply(
                x, lin.weight, lin.bias, cov_cols, n_dense, flags, _row_idx if i == 0 else None, training_bn
            )
            if flags & _lib.LIN_LIBRARY:
                library = lib
            use_drop = drop is not None and self.training and drop.p > 0
            if bn is None and ln is None and act is None and not use_drop:
                x = y
                continue
            mask = _next_mask(y.shape, drop.p, y.device) if use_drop else None
            scale = 1.0 / (1.0 - drop.p) if use_drop else 1.0
            if ln is not None:
                x = ops.LNActDrop.apply(y, mask, scale, act is not None, ln.eps)
                continue
            x = ops.BNActDrop.apply(
                y, part, bn.weight if bn is not None else None, bn.bias if bn is not None else None,
                bn.running_mean if bn is not None else None, bn.running_var if bn is not None else None,
                mask, scale, act is not None, self.training, bn is not None,
            )
            if training_bn and bn.num_batches_tracked is not None:
                bn.num_batches_tracked.add_(1)
        if _first_flags & _lib.LIN_LIBRARY:
            return x, library
        return x

    def _forward_composed(self, x, cat_list, cont):
        """Reference semantics with stock torch ops on the GPU (non-default options only)."""
        one_hots = []
        for n_cat, cat in zip(self.n_cat_list, cat_list, strict=False):
            if n_cat > 1:
                one_hots.append(cat if cat.size(1) == n_cat else nn.functional.one_hot(cat.squeeze(-1), n_cat))
        covs = ([cont] if cont is not None else []) + one_hots
        for i, layers in enumerate(self.fc_layers):
            for layer in layers:
                if layer is None:
                    continue
                if isinstance(layer, nn.BatchNorm1d) and x.dim() == 3:
                    x = torch.cat([layer(s).unsqueeze(0) for s in x], dim=0)
                    continue
                if isinstance(layer, nn.Dropout):  # same keep-mask source (and `rng_feed`) as the fused path
                    if self.training and layer.p > 0:
                        x = x * (_next_mask(x.shape, layer.p, x.device).to(x.dtype) / (1.0 - layer.p))
                    continue
                if self._is_linear_layer(layer) and self.inject_into_layer(i):
                    cl = [c.unsqueeze(0).expand((x.size(0), c.size(0), c.size(1))) for c in covs] if x.dim() == 3 else covs
                    x = torch.cat((x, *cl), dim=-1)
                x = layer(x)
        return x


class Encoder(nn.Module):
    """Encoder to a Normal latent (reference src/scvi/nn/_base_components.py:219-320)."""

    def __init__(
        self,
        n_input: int,
        n_output: int,
        n_cat_list: Iterable[int] = None,
        n_layers: int = 1,
        n_hidden: int = 128,
        dropout_rate: float = 0.1,
        distribution: str = "normal",
        var_eps: float = 1e-4,
        var_activation: Callable | None = None,
        return_dist: bool = False,
        **kwargs,
    ):
        super().__init__()
        self.distribution = distribution
        self.var_eps = var_eps
        self.encoder = FCLayers(
            n_in=n_input, n_out=n_hidden, n_cat_list=n_cat_list, n_layers=n_layers, n_hidden=n_hidden,
            dropout_rate=dropout_rate, **kwargs,
        )
        self.mean_encoder = nn.Linear(n_hidden, n_output)
        self.var_encoder = nn.Linear(n_hidden, n_output)
        self.return_dist = return_dist
        self.z_transformation = nn.Softmax(dim=-1) if distribution == "ln" else _identity
        self.var_activation = torch.exp if var_activation is None else var_activation

    def forward(self, x: torch.Tensor, *cat_list: int, _first_flags: int = 0, _row_idx=None):
        """Returns ``(dist, latent)`` if ``return_dist`` else ``(q_m, q_v, latent)``.

        With the private ``_first_flags`` (see `FCLayers.forward`) the library is returned as
        an extra last element.  The returned Normal carries ``_b200_kl`` = KL(q || N(0,1))
        summed over the latent dimensions, computed by the same kernel.
        """
        _require_cuda(x, "Encoder")
        q = self.encoder(x, *cat_list, _first_flags=_first_flags, _row_idx=_row_idx)
        library = None
        if _first_flags & _lib.LIN_LIBRARY:
            q, library = q
        kl = None
        if self.var_activation is torch.exp and q.dim() == 2 and self.mean_encoder.out_features <= 32:
            noise = _next_noise((q.shape[0], self.mean_encoder.out_features), q.device)
            q_m, q_v, untran, kl = ops.LatentHeads.apply(
                q, self.mean_encoder.weight, self.mean_encoder.bias, self.var_encoder.weight, self.var_encoder.bias,
                noise, self.var_eps,
            )
            dist = Normal(q_m, q_v.sqrt(), validate_args=False)
        else:  # custom var_activation / 3-D input: composed torch ops
            q_m = self.mean_encoder(q)
            q_v = self.var_activation(self.var_encoder(q)) + self.var_eps
            dist = Normal(q_m, q_v.sqrt(), validate_args=False)
            untran = dist.rsample()
        dist._b200_kl = kl
        latent = self.z_transformation(untran)
        out = (dist, latent) if self.return_dist else (q_m, q_v, latent)
        return out + (library,) if library is not None else out


class DecoderSCVI(nn.Module):
    """Decoder to the parameters of the (ZI)NB (reference src/scvi/nn/_base_components.py:324-442)."""

    def __init__(
        self,
        n_input: int,
        n_output: int,
        n_cat_list: Iterable[int] = None,
        n_layers: int = 1,
        n_hidden: int = 128,
        inject_covariates: bool = True,
        use_batch_norm: bool = False,
        use_layer_norm: bool = False,
        scale_activation: Literal["softmax", "softplus"] = "softmax",
        **kwargs,
    ):
        super().__init__()
        self.px_decoder = FCLayers(
            n_in=n_input, n_out=n_hidden, n_cat_list=n_cat_list, n_layers=n_layers, n_hidden=n_hidden,
            dropout_rate=0, inject_covariates=inject_covariates, use_batch_norm=use_batch_norm,
            use_layer_norm=use_layer_norm, **kwargs,
        )
        self.scale_activation = scale_activation
        act = nn.Softmax(dim=-1) if scale_activation == "softmax" else nn.Softplus()
        self.px_scale_decoder = nn.Sequential(nn.Linear(n_hidden, n_output), act)
        self.px_r_decoder = nn.Linear(n_hidden, n_output)  # only used for dispersion == "gene-cell"
        self.px_dropout_decoder = nn.Linear(n_hidden, n_output)

    def hidden(self, z: torch.Tensor, *cat_list: int) -> torch.Tensor:
        """The shared trunk ``px = px_decoder(z, *cat_list)`` (reference :436)."""
        return self.px_decoder(z, *cat_list)

    def supports_fused_nll(self, px: torch.Tensor) -> bool:
        return self.scale_activation == "softmax" and px.dim() == 2 and px.shape[1] in (64, 128, 256)

    def nll(self, px, library, x, px_r, gene_likelihood: str, row_idx=None, precision: int = _lib.PREC_FP32):
        """``-log_prob(x).sum(-1)`` of the (ZI)NB the heads parameterise, fused (no [B,G] tensor)."""
        lik = {"nb": _lib.LIK_NB, "zinb": _lib.LIK_ZINB}[gene_likelihood]
        sc = self.px_scale_decoder[0]
        dr = self.px_dropout_decoder
        return ops.DecoderNLL.apply(
            px, sc.weight, sc.bias, dr.weight if lik else None, dr.bias if lik else None, px_r, x, library, lik,
            row_idx, precision,
        )

    def forward(self, dispersion: str, z: torch.Tensor, library: torch.Tensor, *cat_list: int):
        """Materialised ``(px_scale, px_r, px_rate, px_dropout)`` as the reference (:400-442)."""
        _require_cuda(z, "DecoderSCVI")
        px = self.hidden(z, *cat_list)
        needs_grad = torch.is_grad_enabled() and (px.requires_grad or self.px_scale_decoder[0].weight.requires_grad)
        if self.scale_activation == "softmax" and px.dim() == 2 and not needs_grad:
            sc, dr = self.px_scale_decoder[0], self.px_dropout_decoder
            px_scale, px_rate, px_dropout = ops.decoder_params(px, sc.weight, sc.bias, dr.weight, dr.bias, library)
        else:  # autograd through materialised parameters / softplus / 3-D: composed torch ops
            px_scale = self.px_scale_decoder(px)
            px_dropout = self.px_dropout_decoder(px)
            px_rate = torch.exp(library) * px_scale
        px_r = self.px_r_decoder(px) if dispersion == "gene-cell" else None
        return px_scale, px_r, px_rate, px_dropout
