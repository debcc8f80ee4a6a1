"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the SCVI `VAE` minibatch training step.

A functional restatement (plain torch CPU ops on explicit parameter dicts, dtype-agnostic so
it also runs in float64) of the reference algorithm for the one hot path this repo
accelerates.  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it; nothing under ``scvi_tools_b200/`` does.

Parity status: PINNED.  ``oracle/make_golden.py`` runs the *unmodified* reference files
(``oracle/ref_shim.py``) in the build container and commits inputs + outputs under
``tests/golden/``; ``tests/test_oracle_pins.py`` checks this restatement against those
vectors (and, where ``/root/reference`` exists, against the live reference), and against the
reference's own known-answer table for ``_compute_kl_weight``
(``tests/train/test_trainingplans.py:11-51``).

Third-party arithmetic: the reference's math bottoms out in PyTorch (unpinned ``torch`` in
``pyproject.toml:44``; 2.11.0+cu128 here) -- ``torch.lgamma``, ``F.softplus``,
``nn.BatchNorm1d``, ``torch.optim.Adam``.  Each function below cites the reference
``file:line`` it follows (paths relative to ``/root/reference``).

Parameter naming follows the reference ``state_dict``.
"""

from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

ENC = "z_encoder.encoder.fc_layers.Layer 0."
DEC = "decoder.px_decoder.fc_layers.Layer 0."

BN_EPS = 1e-3  # src/scvi/nn/_base_components.py:111
BN_MOMENTUM = 0.01  # src/scvi/nn/_base_components.py:111
VAR_EPS = 1e-4  # src/scvi/nn/_base_components.py:263
LIK_EPS = 1e-8  # src/scvi/distributions/_negative_binomial.py:41,103


# --------------------------------------------------------------------------------------
# likelihoods
# --------------------------------------------------------------------------------------
def log_nb_positive(x, mu, theta, eps: float = LIK_EPS):
    """NB log-likelihood per element.  src/scvi/distributions/_negative_binomial.py:99-135."""
    l_tm = torch.log(theta + mu + eps)
    return (
        theta * (torch.log(theta + eps) - l_tm)
        + x * (torch.log(mu + eps) - l_tm)
        + torch.lgamma(x + theta)
        - torch.lgamma(theta)
        - torch.lgamma(x + 1)
    )


def log_zinb_positive(x, mu, theta, pi, eps: float = LIK_EPS):
    """ZINB log-likelihood per element.  src/scvi/distributions/_negative_binomial.py:36-96.

    ``pi`` is the zero-inflation logit.  The zero / non-zero branches are both evaluated and
    masked by ``x < eps`` / ``x > eps`` exactly as the reference does (:81-94).
    """
    if theta.dim() == 1:
        theta = theta.view(1, -1)
    sp_neg_pi = F.softplus(-pi)
    l_t = torch.log(theta + eps)
    l_tm = torch.log(theta + mu + eps)
    a = -pi + theta * (l_t - l_tm)
    zero_case = F.softplus(a) - sp_neg_pi
    nonzero_case = (
        -sp_neg_pi
        + a
        + x * (torch.log(mu + eps) - l_tm)
        + torch.lgamma(x + theta)
        - torch.lgamma(theta)
        - torch.lgamma(x + 1)
    )
    is_zero = (x < eps).to(zero_case.dtype)
    is_pos = (x > eps).to(zero_case.dtype)
    return is_zero * zero_case + is_pos * nonzero_case


# --------------------------------------------------------------------------------------
# FCLayers pieces
# --------------------------------------------------------------------------------------
def batchnorm(y, gamma, beta, run_mean, run_var, training: bool, eps: float = BN_EPS):
    """BatchNorm1d(momentum=0.01, eps=1e-3).  src/scvi/nn/_base_components.py:111.

    Training: batch mean / *biased* variance normalise; the running buffers receive
    ``(1-m)*old + m*new`` with the *unbiased* variance (torch semantics).  Returns the output
    and the new running buffers (functional: inputs are not modified).
    """
    if training:
        n = y.shape[0]
        mean = y.mean(0)
        var_b = ((y - mean) ** 2).mean(0)
        out = (y - mean) / torch.sqrt(var_b + eps) * gamma + beta
        var_u = var_b * (n / max(n - 1, 1))
        new_mean = (1 - BN_MOMENTUM) * run_mean + BN_MOMENTUM * mean.detach()
        new_var = (1 - BN_MOMENTUM) * run_var + BN_MOMENTUM * var_u.detach()
        return out, new_mean, new_var
    out = (y - run_mean) / torch.sqrt(run_var + eps) * gamma + beta
    return out, run_mean, run_var


def inject_one_hot(x, batch_index, n_batch: int):
    """Covariate injection by one-hot concat.  src/scvi/nn/_base_components.py:169-178,198-215.

    A category with ``n_cat <= 1`` contributes nothing (:80,174).
    """
    if n_batch > 1:
        oh = F.one_hot(batch_index.reshape(-1).long(), n_batch).to(x.dtype)
        return torch.cat((x, oh), dim=-1)
    return x


def dropout(a, mask, p: float, training: bool):
    """nn.Dropout(p) with an explicit keep-mask (1 = keep).  _base_components.py:114."""
    if training and p > 0:
        return a * mask.to(a.dtype) / (1.0 - p)
    return a


# --------------------------------------------------------------------------------------
# the VAE minibatch step
# --------------------------------------------------------------------------------------
def vae_forward(
    P: dict,
    x,
    batch_index,
    *,
    n_batch: int,
    gene_likelihood: str = "zinb",
    kl_weight: float = 1.0,
    drop_mask=None,
    noise=None,
    training: bool = True,
    dropout_rate: float = 0.1,
    log_variational: bool = True,
    relu_gate1=None,
    relu_gate2=None,
):
    """Forward of the default SCVI VAE (n_layers=1, one-hot batch, observed library, gene dispersion).

    Follows ``_generic_forward`` (src/scvi/module/base/_base_module.py:458-488):
    ``_regular_inference`` (src/scvi/module/_vae.py:360-417) -> ``Encoder.forward``
    (src/scvi/nn/_base_components.py:291-320) -> ``generative`` (_vae.py:445-549) ->
    ``DecoderSCVI.forward`` (_base_components.py:400-442) -> ``loss`` (_vae.py:551-600).

    ``drop_mask`` [B,H] keep-mask and ``noise`` [B,L] standard-normal draws replace the
    reference's RNG consumers (``nn.Dropout`` and ``Normal.rsample``).
    ``relu_gate1`` / ``relu_gate2`` (bool [B,H], default None = the reference's ``torch.relu``) fix the side of
    the ReLU kink per element for the encoder / decoder hidden layer: ``a = bn * gate``.  A pre-activation that
    sits within float32 rounding distance of zero lands on either side depending on the summation order of the
    GEMM that produced it, and the derivative jumps with it; full-size parity tests evaluate the oracle on the
    side the kernel under test took (and check that every such element is at rounding distance from zero).
    Returns a dict of every intermediate the parity tests compare, plus the new BN buffers.
    """
    dt = x.dtype
    # --- inference: _vae.py:370-374
    library = torch.log(x.sum(1)).unsqueeze(1)
    x_ = torch.log1p(x) if log_variational else x
    # --- encoder FCLayers layer 0 (encode_covariates=False => no injection; _vae.py:237)
    y1 = F.linear(x_, P[ENC + "0.weight"], P[ENC + "0.bias"])
    bn1, rm1, rv1 = batchnorm(
        y1, P[ENC + "1.weight"], P[ENC + "1.bias"], P[ENC + "1.running_mean"], P[ENC + "1.running_var"], training
    )
    a1 = dropout(torch.relu(bn1) if relu_gate1 is None else bn1 * relu_gate1.to(dt), drop_mask, dropout_rate, training)
    # --- Encoder heads: _base_components.py:311-317
    q_m = F.linear(a1, P["z_encoder.mean_encoder.weight"], P["z_encoder.mean_encoder.bias"])
    q_v = torch.exp(F.linear(a1, P["z_encoder.var_encoder.weight"], P["z_encoder.var_encoder.bias"])) + VAR_EPS
    q_s = torch.sqrt(q_v)
    if noise is None:
        noise = torch.zeros_like(q_m)
    z = q_m + q_s * noise.to(dt)
    # --- decoder FCLayers layer 0 with one-hot injection; no dropout (_base_components.py:377)
    d_in = inject_one_hot(z, batch_index, n_batch)
    y2 = F.linear(d_in, P[DEC + "0.weight"], P[DEC + "0.bias"])
    bn2, rm2, rv2 = batchnorm(
        y2, P[DEC + "1.weight"], P[DEC + "1.bias"], P[DEC + "1.running_mean"], P[DEC + "1.running_var"], training
    )
    h = torch.relu(bn2) if relu_gate2 is None else bn2 * relu_gate2.to(dt)
    # --- heads: _base_components.py:436-441
    scale_logits = F.linear(h, P["decoder.px_scale_decoder.0.weight"], P["decoder.px_scale_decoder.0.bias"])
    px_scale = torch.softmax(scale_logits, dim=-1)
    px_dropout = F.linear(h, P["decoder.px_dropout_decoder.weight"], P["decoder.px_dropout_decoder.bias"])
    px_rate = torch.exp(library) * px_scale
    theta = torch.exp(P["px_r"])  # _vae.py:516-518
    # --- loss: _vae.py:563-580; KL(N(q_m, q_s) || N(0,1)) as torch's kl_normal_normal
    var_ratio = q_s**2
    kl_z = (0.5 * (var_ratio + q_m**2 - 1 - torch.log(var_ratio))).sum(-1)
    if gene_likelihood == "zinb":
        ll = log_zinb_positive(x, px_rate, theta, px_dropout)
    elif gene_likelihood == "nb":
        ll = log_nb_positive(x, px_rate, theta.view(1, -1))
    else:
        raise ValueError(gene_likelihood)
    reconst = -ll.sum(-1)
    loss = torch.mean(reconst + kl_weight * kl_z)
    return {
        "library": library,
        "y1": y1,
        "bn1": bn1,
        "bn2": bn2,
        "a1": a1,
        "q_m": q_m,
        "q_v": q_v,
        "z": z,
        "y2": y2,
        "h": h,
        "px_scale": px_scale,
        "px_rate": px_rate,
        "px_dropout": px_dropout,
        "theta": theta,
        "kl_z": kl_z,
        "reconst": reconst,
        "loss": loss,
        "bn_buffers": {
            ENC + "1.running_mean": rm1,
            ENC + "1.running_var": rv1,
            DEC + "1.running_mean": rm2,
            DEC + "1.running_var": rv2,
        },
    }


TRAINED_KEYS_ZINB = [
    "px_r",
    ENC + "0.weight",
    ENC + "0.bias",
    ENC + "1.weight",
    ENC + "1.bias",
    "z_encoder.mean_encoder.weight",
    "z_encoder.mean_encoder.bias",
    "z_encoder.var_encoder.weight",
    "z_encoder.var_encoder.bias",
    DEC + "0.weight",
    DEC + "0.bias",
    DEC + "1.weight",
    DEC + "1.bias",
    "decoder.px_scale_decoder.0.weight",
    "decoder.px_scale_decoder.0.bias",
    "decoder.px_dropout_decoder.weight",
    "decoder.px_dropout_decoder.bias",
]
TRAINED_KEYS_NB = [k for k in TRAINED_KEYS_ZINB if "px_dropout" not in k]


def trained_keys(gene_likelihood: str):
    """Parameters that receive gradient (the rest -- l_encoder, px_r_decoder, and for `nb`
    px_dropout_decoder -- are unused; SURVEY.md section 3.2 row d)."""
    return TRAINED_KEYS_ZINB if gene_likelihood == "zinb" else TRAINED_KEYS_NB


def loss_and_grads(P: dict, x, batch_index, **kw):
    """Forward + autograd backward; returns (outputs, {name: grad}) for the trained keys."""
    keys = trained_keys(kw.get("gene_likelihood", "zinb"))
    Pg = {k: (v.detach().clone().requires_grad_(True) if k in keys else v) for k, v in P.items()}
    out = vae_forward(Pg, x, batch_index, **kw)
    out["loss"].backward()
    grads = {k: Pg[k].grad.detach() for k in keys}
    return out, grads


# --------------------------------------------------------------------------------------
# training-plan semantics
# --------------------------------------------------------------------------------------
def compute_kl_weight(epoch, step, n_epochs_kl_warmup=400, n_steps_kl_warmup=None, max_kl_weight=1.0, min_kl_weight=0.0):
    """KL warm-up weight.  src/scvi/train/_trainingplans.py:31-72."""
    if min_kl_weight > max_kl_weight:
        raise ValueError("min_kl_weight > max_kl_weight")
    span = max_kl_weight - min_kl_weight
    if n_epochs_kl_warmup:
        if epoch < n_epochs_kl_warmup:
            return span * (epoch / n_epochs_kl_warmup) + min_kl_weight
    elif n_steps_kl_warmup:
        if step < n_steps_kl_warmup:
            return span * (step / n_steps_kl_warmup) + min_kl_weight
    return max_kl_weight


def adam_step(p, g, m, v, step: int, lr=1e-3, beta1=0.9, beta2=0.999, eps=0.01, weight_decay=1e-6):
    """One `torch.optim.Adam` update (coupled L2), op order as torch's single-tensor path.

    Reference hyper-parameters: src/scvi/train/_trainingplans.py:150-153,475-482.
    Returns (p_new, m_new, v_new); ``step`` is the 1-based step count.
    """
    g = g + weight_decay * p
    m = m + (g - m) * (1 - beta1)  # exp_avg.lerp_(grad, 1-beta1)
    v = v * beta2 + (1 - beta2) * g * g  # exp_avg_sq.mul_(b2).addcmul_(g, g, value=1-b2)
    bc1 = 1 - beta1**step
    bc2 = 1 - beta2**step
    step_size = lr / bc1
    denom = v.sqrt() / math.sqrt(bc2) + eps
    p = p - step_size * (m / denom)
    return p, m, v


# --------------------------------------------------------------------------------------
# data path
# --------------------------------------------------------------------------------------
def gather_rows(X: np.ndarray, indexes) -> np.ndarray:
    """`AnnTorchDataset.__getitem__` for a dense matrix.  src/scvi/data/_anntorchdataset.py:143-147
    (`data[indexes].astype(float32)`; dtype default from src/scvi/data/_utils.py:37-47)."""
    return X[np.asarray(indexes)].astype(np.float32, copy=False)


def distributed_batches(n_obs: int, batch_size: int, rank: int, world: int, epoch: int, seed: int = 0,
                        shuffle: bool = True, drop_last: bool = False, drop_dataset_tail: bool = False):
    """Index stream of `BatchDistributedSampler`.  src/scvi/dataloaders/_samplers.py:44-89 on top of
    torch's `DistributedSampler.__iter__` (global permutation seeded `seed+epoch`, padded or
    tail-dropped to a multiple of `world`, rank-strided).  Returns a list of int64 arrays."""
    if shuffle:
        g = torch.Generator()
        g.manual_seed(seed + epoch)
        idx = torch.randperm(n_obs, generator=g).tolist()
    else:
        idx = list(range(n_obs))
    if drop_dataset_tail and n_obs % world != 0:
        num = math.ceil((n_obs - world) / world)
    else:
        num = math.ceil(n_obs / world)
    total = num * world
    if not drop_dataset_tail:
        pad = total - len(idx)
        if pad <= len(idx):
            idx += idx[:pad]
        else:
            idx += (idx * math.ceil(pad / len(idx)))[:pad]
    else:
        idx = idx[:total]
    mine = idx[rank:total:world]
    out = []
    for s in range(0, len(mine), batch_size):
        chunk = mine[s : s + batch_size]
        if len(chunk) < batch_size and drop_last:
            break
        out.append(np.asarray(chunk, dtype=np.int64))
    return out


def synthetic_iid(n_obs: int, n_genes: int, n_batches: int, seed: int = 0, dtype=np.float32):
    """Counts as `scvi.data.synthetic_iid`.  src/scvi/data/_built_in_data/_synthetic.py:51-59,74-76:
    NB(5, 0.3) * Bernoulli(0.7), contiguous batch blocks.  (numpy Generator instead of the
    reference's legacy global RandomState; the distribution is the same.)"""
    rng = np.random.default_rng(seed)
    per = n_obs // n_batches
    n = per * n_batches
    X = rng.negative_binomial(5, 0.3, size=(n, n_genes))
    X = X * rng.binomial(1, 0.7, size=(n, n_genes))
    batch = np.repeat(np.arange(n_batches), per).astype(np.int64)
    return X.astype(dtype), batch
