"""TEST INFRASTRUCTURE ONLY -- tests/golden/opt_*.npz: non-default `VAE` options
recorded from the UNMODIFIED reference module.

    python oracle/make_golden_options.py      (build container only: needs /root/reference)

Each case builds the reference `scvi.module.VAE` (src/scvi/module/_vae.py:146-285) with one group of
non-default keyword arguments -- dispersion variants (:205-212, :509-518), encoded library size
(:255-268, :392-400, :535-542, :566-569), encoded / deeply injected categorical and continuous covariates,
n_layers > 1, LayerNorm -- runs one training-mode forward + backward on seeded inputs and stores the inputs,
the complete state_dict, every dropout mask and reparameterisation draw in call order (the reference's
`nn.Dropout` / `Normal.rsample` are routed through a recording feed), the ELBO pieces and every gradient.
"""

from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import ref_shim  # noqa: E402
from oracle import scvi_oracle as O  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


class RecordingFeed:
    """Draw dropout keep-masks / N(0,1) noise from a seeded generator in call order and keep them."""

    def __init__(self, seed, p_keep_from_call=True):
        self.g = torch.Generator().manual_seed(seed)
        self.masks, self.noises = [], []

    def __enter__(self):
        import torch.nn.functional as F
        from torch.distributions import Normal

        self._F, self._Normal = F, Normal
        self._old_dropout, self._old_rsample = F.dropout, Normal.rsample
        feed = self

        def fed_dropout(inp, p=0.5, training=True, inplace=False):
            if not training or p == 0:
                return inp
            mask = (torch.rand(inp.shape, generator=feed.g) >= p).to(torch.uint8)
            feed.masks.append(mask)
            return inp * (mask.to(inp.dtype) / (1.0 - p))

        def fed_rsample(dist, sample_shape=torch.Size()):
            noise = torch.randn(dist.loc.shape, generator=feed.g)
            feed.noises.append(noise)
            return dist.loc + noise * dist.scale

        F.dropout = fed_dropout
        Normal.rsample = fed_rsample
        return self

    def __exit__(self, *exc):
        self._F.dropout = self._old_dropout
        self._Normal.rsample = self._old_rsample


def perturb_(module, seed):
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for _, p in module.named_parameters():
            if p.dim() == 1 and p.requires_grad:
                p.add_(0.1 * torch.randn(p.shape, generator=g))
        for k, b in module.named_buffers():
            if k.endswith("running_mean"):
                b.add_(0.05 * torch.randn(b.shape, generator=g))
            if k.endswith("running_var"):
                b.mul_(1 + 0.1 * torch.rand(b.shape, generator=g))


def make_inputs(n_obs, G, n_batch, B, seed, n_labels=0, n_cont=0, n_cats=()):
    X, batch = O.synthetic_iid(n_obs, G, max(n_batch, 1), seed=seed)
    rng = np.random.default_rng(seed + 1)
    idx = rng.permutation(X.shape[0])[:B]
    t = {"X": torch.from_numpy(O.gather_rows(X, idx)), "batch": torch.from_numpy(batch[idx]).view(-1, 1)}
    t["labels"] = torch.from_numpy(rng.integers(0, max(n_labels, 1), size=(B, 1)))
    if n_cont:
        t["extra_continuous_covs"] = torch.from_numpy(rng.normal(size=(B, n_cont)).astype(np.float32))
    if n_cats:
        t["extra_categorical_covs"] = torch.from_numpy(np.stack([rng.integers(0, c, size=B) for c in n_cats], axis=1))
    return t


def save_case(name, kwargs, tensors, module, feed, out, extra_in=None):
    save = {"in::kwargs": np.array(json.dumps(kwargs))}
    for k, v in tensors.items():
        save["in::t::" + k] = v.numpy()
    for i, m in enumerate(feed.masks):
        save[f"in::mask::{i}"] = m.numpy()
    for i, n in enumerate(feed.noises):
        save[f"in::noise::{i}"] = n.numpy()
    for k, v in (extra_in or {}).items():
        save["in::" + k] = v
    for k, v in out.items():
        save["out::" + k] = v
    path = os.path.join(GOLD, name + ".npz")
    np.savez_compressed(path, **save)
    print(f"{name}: loss={float(out['loss']):.6f} masks={len(feed.masks)} noises={len(feed.noises)} "
          f"grads={sum(k.startswith('grad::') for k in out)} -> {os.path.getsize(path)/1e3:.0f} kB")


def vae_option_case(name, G, n_batch, B, seed, kl_weight, vae_kwargs, n_labels=0, n_cont=0, n_cats=(), lib_prior=False):
    R = ref_shim.load()
    kwargs = dict(n_input=G, n_batch=n_batch, n_labels=n_labels, n_continuous_cov=n_cont,
                  n_cats_per_cov=list(n_cats) or None, **vae_kwargs)
    extra_in = {}
    ctor = dict(kwargs)
    if lib_prior:
        rng = np.random.default_rng(seed + 7)
        extra_in["library_log_means"] = rng.normal(6.0, 0.3, size=(1, n_batch)).astype(np.float32)
        extra_in["library_log_vars"] = rng.uniform(0.5, 1.5, size=(1, n_batch)).astype(np.float32)
        ctor.update(library_log_means=extra_in["library_log_means"], library_log_vars=extra_in["library_log_vars"])
    torch.manual_seed(seed)
    m = R.VAE(**ctor)
    perturb_(m, seed + 3)
    m.train()
    tensors = make_inputs(400, G, n_batch, B, seed, n_labels, n_cont, n_cats)
    P0 = {k: v.detach().clone() for k, v in m.state_dict().items()}
    with RecordingFeed(seed + 2) as feed:
        inf, gen, lo = m(tensors, loss_kwargs={"kl_weight": kl_weight})
    lo.loss.backward()
    out = {"loss": lo.loss.detach().numpy(),
           "reconst": lo.reconstruction_loss["reconstruction_loss"].detach().numpy(),
           "kl_z": lo.kl_local["kl_divergence_z"].detach().numpy(),
           "kl_l": lo.kl_local["kl_divergence_l"].detach().numpy(),
           "z": inf["z"].detach().numpy(), "library": inf["library"].detach().numpy()}
    for k, p in m.named_parameters():
        if p.grad is not None:
            out["grad::" + k] = p.grad.detach().clone().numpy()
    for k, b in m.named_buffers():
        if "running" in k:
            out["buf_after::" + k] = b.detach().clone().numpy()
    extra_in["kl_weight"] = np.float64(kl_weight)
    for k, v in P0.items():
        extra_in["param::" + k] = v.numpy()
    save_case(name, kwargs, tensors, m, feed, out, extra_in)


def main():
    os.makedirs(GOLD, exist_ok=True)
    torch.set_num_threads(1)
    only = sys.argv[1] if len(sys.argv) > 1 else None  # e.g. "embedding": (re)generate that group only
    if only is None:
        main_round1()
    if only in (None, "embedding"):
        main_embedding()


def main_embedding():
    only = "embedding"
    # batch_representation="embedding" (src/scvi/nn/_embedding.py, src/scvi/module/_vae.py:218-236,385-396,489-498)
    vae_option_case("opt_embedding_zinb", 88, 4, 60, 18, 0.6,
                    {"batch_representation": "embedding", "gene_likelihood": "zinb"})
    vae_option_case("opt_embedding_enc_nb", 76, 3, 52, 19, 0.8,
                    {"batch_representation": "embedding", "encode_covariates": True, "gene_likelihood": "nb",
                     "batch_embedding_kwargs": {"embedding_dim": 7}}, n_cont=1)


def main_round1():
    vae_option_case("opt_genebatch_zinb", 96, 3, 64, 10, 0.5, {"dispersion": "gene-batch", "gene_likelihood": "zinb"})
    vae_option_case("opt_genelabel_nb", 80, 2, 48, 11, 1.0, {"dispersion": "gene-label", "gene_likelihood": "nb"}, n_labels=4)
    vae_option_case("opt_genecell_zinb", 72, 2, 40, 12, 0.3, {"dispersion": "gene-cell", "gene_likelihood": "zinb"})
    vae_option_case("opt_libenc_nb", 100, 3, 64, 13, 0.7,
                    {"use_observed_lib_size": False, "gene_likelihood": "nb"}, lib_prior=True)
    vae_option_case("opt_covs_zinb", 90, 3, 56, 14, 0.4,
                    {"encode_covariates": True, "n_layers": 2, "gene_likelihood": "zinb", "n_hidden": 128},
                    n_cont=2, n_cats=(3, 2))
    vae_option_case("opt_layernorm_nb", 64, 2, 33, 15, 1.0,
                    {"use_layer_norm": "both", "use_batch_norm": "none", "gene_likelihood": "nb"})
    vae_option_case("opt_deep_noinject_zinb", 120, 4, 50, 16, 0.2,
                    {"n_layers": 3, "deeply_inject_covariates": False, "n_hidden": 64, "n_latent": 7,
                     "dropout_rate": 0.2, "gene_likelihood": "zinb"})
    vae_option_case("opt_nologvar_nb", 64, 2, 32, 17, 1.0, {"log_variational": False, "gene_likelihood": "nb"})


if __name__ == "__main__":
    main()
