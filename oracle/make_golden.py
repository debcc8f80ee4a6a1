"""TEST INFRASTRUCTURE ONLY -- generate tests/golden/*.npz from the UNMODIFIED reference.

Run in the build container (where /root/reference exists):

    python oracle/make_golden.py

It imports the reference's own hot-path files through ``oracle/ref_shim.py`` and records, for
seeded inputs, what the reference computes: per-element NB/ZINB log-likelihoods, the VAE
minibatch ELBO pieces, every trained parameter's gradient, the BatchNorm buffers after the
step and the parameters after real ``torch.optim.Adam`` steps.  The two RNG consumers on the
path (``nn.Dropout`` and ``Normal.rsample``) are fed explicit tensors so the CUDA path can
be given the same masks / noise (BASELINE.json: "identical weights, inputs and dropout
masks").  The vectors travel to the GPU box; the reference itself cannot.
"""

from __future__ import annotations

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


class _FeedRNG:
    """Route the reference's dropout / rsample through supplied tensors."""

    def __init__(self, drop_mask, noise):
        self.drop_mask, self.noise = drop_mask, noise

    def __enter__(self):
        import torch.nn.functional as F
        from torch.distributions import Normal

        self._F, self._Normal = F, Normal
        self._old_dropout, self._old_rsample = F.dropout, Normal.rsample
        mask, noise = self.drop_mask, self.noise

        def fed_dropout(inp, p=0.5, training=True, inplace=False):
            if not training or p == 0:
                return inp
            return inp * (mask.to(inp.dtype) / (1.0 - p))  # as aten: noise.div_(1-p); input*noise

        def fed_rsample(dist, sample_shape=torch.Size()):
            return dist.loc + noise.to(dist.loc.dtype) * dist.scale

        F.dropout = fed_dropout
        Normal.rsample = fed_rsample
        return self

    def __exit__(self, *exc):
        self._F.dropout = self._old_dropout
        self._Normal.rsample = self._old_rsample


def _needed_state(module, gene_likelihood):
    sd = module.state_dict()
    keep = {}
    for k, v in sd.items():
        if k.startswith("l_encoder") or k.startswith("decoder.px_r_decoder") or k.endswith("num_batches_tracked"):
            continue
        keep[k] = v.detach().clone()
    return keep


def vae_case(name, n_obs, G, n_batch, B, likelihood, seed, kl_weight, n_adam_steps=2, perturb=True):
    R = ref_shim.load()
    X, batch = O.synthetic_iid(n_obs, G, n_batch, seed=seed)
    torch.manual_seed(seed)
    m = R.VAE(n_input=G, n_batch=n_batch, gene_likelihood=likelihood)
    if perturb:
        # move away from init so BN affine / running stats / biases are not at trivial values
        with torch.no_grad():
            for k, p in m.named_parameters():
                if p.dim() == 1:
                    p.add_(0.1 * torch.randn_like(p))
            for k, b in m.named_buffers():
                if k.endswith("running_mean"):
                    b.add_(0.05 * torch.randn_like(b))
                if k.endswith("running_var"):
                    b.mul_(1 + 0.1 * torch.rand_like(b))
    m.train()
    rng = np.random.default_rng(seed + 1)
    idx = rng.permutation(X.shape[0])[:B]
    x = torch.from_numpy(O.gather_rows(X, idx))
    bi = torch.from_numpy(batch[idx]).view(-1, 1)
    H, L = m.n_hidden, m.n_latent
    g = torch.Generator().manual_seed(seed + 2)
    drop_mask = (torch.rand(B, H, generator=g) >= 0.1).to(torch.uint8)
    noise = torch.randn(B, L, generator=g)
    tensors = {"X": x, "batch": bi, "labels": torch.zeros(B, 1, dtype=torch.long)}

    P0 = _needed_state(m, likelihood)
    opt = torch.optim.Adam(filter(lambda p: p.requires_grad, m.parameters()), lr=1e-3, eps=0.01, weight_decay=1e-6)
    out = {}
    for step in range(n_adam_steps):
        opt.zero_grad()
        with _FeedRNG(drop_mask, noise):
            inf, gen, lo = m(tensors, loss_kwargs={"kl_weight": kl_weight})
        lo.loss.backward()
        if step == 0:
            out["loss"] = lo.loss.detach().numpy()
            out["reconst"] = lo.reconstruction_loss["reconstruction_loss"].detach().numpy()
            out["kl_z"] = lo.kl_local["kl_divergence_z"].detach().numpy()
            out["z"] = inf["z"].detach().numpy()
            out["q_m"] = inf["qz"].loc.detach().numpy()
            out["q_s"] = inf["qz"].scale.detach().numpy()
            out["library"] = inf["library"].detach().numpy()
            px = gen["px"]
            out["px_scale"] = px.scale.detach().numpy()
            out["px_rate"] = px.mu.detach().numpy()
            if likelihood == "zinb":
                out["px_dropout"] = px.zi_logits.detach().numpy()
            for k, p in m.named_parameters():
                if p.grad is not None:
                    out["grad::" + k] = p.grad.detach().clone().numpy()
            for k, b in m.named_buffers():
                if (k.startswith("z_encoder.encoder") or k.startswith("decoder.px_decoder")) and "running" in k:
                    out["buf_after::" + k] = b.detach().clone().numpy()
        opt.step()
    for k in O.trained_keys(likelihood):
        out[f"param_after{n_adam_steps}::" + k] = m.state_dict()[k].detach().clone().numpy()

    # eval-mode forward (running stats, no dropout, same noise) from the INITIAL state
    m2 = R.VAE(n_input=G, n_batch=n_batch, gene_likelihood=likelihood)
    sd = m2.state_dict()
    sd.update(P0)
    m2.load_state_dict(sd)
    m2.eval()
    with torch.no_grad(), _FeedRNG(drop_mask, noise):
        inf, gen, lo = m2(tensors, loss_kwargs={"kl_weight": kl_weight})
    out["eval_loss"] = lo.loss.numpy()
    out["eval_reconst"] = lo.reconstruction_loss["reconstruction_loss"].numpy()
    out["eval_kl_z"] = lo.kl_local["kl_divergence_z"].numpy()

    save = {"in::x": x.numpy(), "in::batch": bi.numpy(), "in::drop_mask": drop_mask.numpy(),
            "in::noise": noise.numpy(), "in::kl_weight": np.float64(kl_weight),
            "in::n_batch": np.int64(n_batch), "in::likelihood": np.array(likelihood),
            "in::n_adam_steps": np.int64(n_adam_steps)}
    for k, v in P0.items():
        save["param::" + k] = v.numpy()
    for k, v in out.items():
        save["out::" + k] = v
    path = os.path.join(GOLD, name + ".npz")
    np.savez_compressed(path, **save)
    print(f"{name}: loss={float(out['loss']):.6f}  -> {path} ({os.path.getsize(path)/1e3:.0f} kB)")


def likelihood_case():
    R = ref_shim.load()
    rng = np.random.default_rng(7)
    B, G = 48, 37
    x = rng.negative_binomial(5, 0.3, size=(B, G)) * rng.binomial(1, 0.7, size=(B, G))
    x = x.astype(np.float32)
    x[0, :5] = [0, 1, 2, 1000, 65535]  # small / large counts
    mu = np.exp(rng.normal(1.0, 2.0, size=(B, G))).astype(np.float32)
    mu[1, :4] = [1e-12, 1e-8, 1e-4, 1e6]  # around and below eps
    theta = np.exp(rng.normal(0.0, 1.5, size=(G,))).astype(np.float32)
    theta[:3] = [1e-3, 1.0, 500.0]
    pi = rng.normal(0.0, 3.0, size=(B, G)).astype(np.float32)
    pi[2, :4] = [-40.0, 40.0, -90.0, 90.0]
    tx, tmu, tth, tpi = map(torch.from_numpy, (x, mu, theta, pi))
    nb = R.log_nb_positive(tx, tmu, tth.view(1, -1))
    zinb = R.log_zinb_positive(tx, tmu, tth, tpi)
    nb_dist = R.NegativeBinomial(mu=tmu, theta=tth).log_prob(tx)
    zinb_dist = R.ZeroInflatedNegativeBinomial(mu=tmu, theta=tth, zi_logits=tpi).log_prob(tx)
    assert torch.equal(nb, nb_dist) and torch.equal(zinb, zinb_dist)
    # gradients of the summed log-likelihood
    tmu_g, tth_g, tpi_g = (t.clone().requires_grad_(True) for t in (tmu, tth, tpi))
    w = torch.from_numpy(rng.normal(size=(B, G)).astype(np.float32))
    (R.log_nb_positive(tx, tmu_g, tth_g.view(1, -1)) * w).sum().backward()
    nb_dmu, nb_dth = tmu_g.grad.clone(), tth_g.grad.clone()
    tmu_g.grad = None
    tth_g.grad = None
    (R.log_zinb_positive(tx, tmu_g, tth_g, tpi_g) * w).sum().backward()
    path = os.path.join(GOLD, "likelihood.npz")
    np.savez_compressed(
        path, x=x, mu=mu, theta=theta, pi=pi, w=w.numpy(), nb=nb.numpy(), zinb=zinb.numpy(),
        nb_dmu=nb_dmu.numpy(), nb_dtheta=nb_dth.numpy(),
        zinb_dmu=tmu_g.grad.numpy(), zinb_dtheta=tth_g.grad.numpy(), zinb_dpi=tpi_g.grad.numpy(),
    )
    print("likelihood ->", path)


def main():
    os.makedirs(GOLD, exist_ok=True)
    torch.set_num_threads(1)  # deterministic reduction order for the stored vectors
    likelihood_case()
    # C1: BASELINE.json configs[0] (400 x 100, 2 batches, zinb, n_latent=10), reference B=128
    vae_case("c1_zinb", 400, 100, 2, 128, "zinb", seed=0, kl_weight=0.25)
    vae_case("c1_nb", 400, 100, 2, 128, "nb", seed=1, kl_weight=1.0)
    # ragged shapes: G not a multiple of any tile, B not a multiple of 32, 8 batches
    vae_case("ragged_zinb", 800, 333, 8, 77, "zinb", seed=2, kl_weight=0.5)
    # single batch category (n_cat==1 is dropped from the one-hot, _base_components.py:80,174)
    vae_case("onebatch_nb", 200, 64, 1, 50, "nb", seed=3, kl_weight=0.0)


if __name__ == "__main__":
    main()
