"""TEST INFRASTRUCTURE ONLY -- tests/golden/totalvi_*.npz from the UNMODIFIED reference `TOTALVAE`.

    python oracle/make_golden_totalvi.py      (build container only: needs /root/reference)

One `AdversarialTrainingPlan.training_step` worth of module work (src/scvi/train/_trainingplans.py:694-744, without
the adversarial classifier, which is off by default for one batch and scaled by `scale_adversarial_loss`):
`module(batch, loss_kwargs={"kl_weight": w})` in training mode on seeded gene + protein counts, then `backward()`.
Stored: inputs, the complete state_dict, every dropout mask / normal draw in call order, loss pieces, decoder
outputs and every gradient.
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
from oracle.make_golden_options import RecordingFeed, make_inputs, perturb_, save_case  # noqa: E402


def totalvi_case(name, G, n_prot, n_batch, B, seed, kl_weight, kwargs_extra, n_labels=0, mask=False, lib_prior=False):
    R = ref_shim.load()
    # n_panel is always passed by `scvi.model.TOTALVI` (the module's `n_panel=None` default does not construct)
    kwargs = dict(n_input_genes=G, n_input_proteins=n_prot, n_batch=n_batch, n_labels=n_labels, n_panel=n_batch,
                  **kwargs_extra)
    ctor = dict(kwargs)
    extra_in = {}
    rng = np.random.default_rng(seed + 9)
    if mask:
        pm = {str(b): (rng.random(n_prot) < 0.7) for b in range(n_batch)}
        ctor["protein_batch_mask"] = pm
        for b, v in pm.items():
            extra_in["protein_batch_mask::" + b] = v
    if lib_prior:
        extra_in["library_log_means"] = rng.normal(6.0, 0.3, size=(1, n_batch)).astype(np.float32)
        extra_in["library_log_vars"] = rng.uniform(0.5, 1.5, size=(1, n_batch)).astype(np.float32)
        ctor.update(library_log_means=extra_in["library_log_means"], library_log_vars=extra_in["library_log_vars"])
    torch.manual_seed(seed)
    m = R.TOTALVAE(**ctor)
    perturb_(m, seed + 3)
    with torch.no_grad():  # non-trivial batch efficiencies / background prior
        m.log_per_batch_efficiency.add_(0.2 * torch.randn_like(m.log_per_batch_efficiency))
        m.background_pro_alpha.add_(0.3 * torch.randn_like(m.background_pro_alpha))
        m.background_pro_log_beta.add_(0.2 * torch.randn_like(m.background_pro_log_beta))
    m.train()
    tensors = make_inputs(400, G, n_batch, B, seed, n_labels=n_labels)
    y = rng.negative_binomial(2, 0.02, size=(B, n_prot)) * rng.binomial(1, 0.9, size=(B, n_prot))
    tensors["proteins"] = torch.from_numpy(y.astype(np.float32))
    P0 = {k: v.detach().clone() for k, v in m.state_dict().items()}
    with RecordingFeed(seed + 2) as feed:
        inf, gen, lo = m(tensors, loss_kwargs={"kl_weight": kl_weight})
    lo.loss.backward()
    out = {"loss": lo.loss.detach().numpy(), "z": inf["z"].detach().numpy(),
           "library_gene": inf["library_gene"].detach().numpy(),
           "px_rate": gen["px_"]["rate"].detach().numpy(), "py_rate_fore": gen["py_"]["rate_fore"].detach().numpy(),
           "py_mixing": gen["py_"]["mixing"].detach().numpy(), "py_scale": gen["py_"]["scale"].detach().numpy()}
    for k, v in lo.reconstruction_loss.items():
        out["rec::" + k] = v.detach().numpy()
    for k, v in lo.kl_local.items():
        out["kl::" + k] = v.detach().numpy()
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


def totalvi_trajectory(name, G, n_prot, n_batch, B, seed, n_steps, kwargs_extra, first_epoch=100):
    """`AdversarialTrainingPlan.training_step` x n_steps on the unmodified reference `TOTALVAE` + reference
    `Classifier(n_latent, 32, n_batch, n_layers=2, logits=True)` (src/scvi/train/_trainingplans.py:631-744,751-777):
    ELBO + kappa * fool loss -> backward -> Adam(lr 4e-3, eps 0.01, wd 1e-6) on the module; then kappa * true-batch
    cross-entropy on z.detach() -> Adam(lr 1e-3, eps 0.01, wd 1e-6) on the classifier.  kl_weight = epoch / 400
    with epochs first_epoch, first_epoch + 1, ...; kappa = 1 - kl_weight."""
    import json

    import torch.nn.functional as F

    R = ref_shim.load()
    kwargs = dict(n_input_genes=G, n_input_proteins=n_prot, n_batch=n_batch, n_labels=0, n_panel=n_batch, **kwargs_extra)
    rng = np.random.default_rng(seed + 9)
    torch.manual_seed(seed)
    m = R.TOTALVAE(**kwargs)
    perturb_(m, seed + 3)
    clf = R.Classifier(n_input=m.n_latent, n_hidden=32, n_labels=n_batch, n_layers=2, logits=True)
    perturb_(clf, seed + 4)
    m.train()
    clf.train()
    P0 = {k: v.detach().clone() for k, v in m.state_dict().items()}
    C0 = {k: v.detach().clone() for k, v in clf.state_dict().items()}
    opt1 = torch.optim.Adam(filter(lambda p: p.requires_grad, m.parameters()), lr=4e-3, eps=0.01, weight_decay=1e-6)
    opt2 = torch.optim.Adam(filter(lambda p: p.requires_grad, clf.parameters()), lr=1e-3, eps=0.01, weight_decay=1e-6)

    def adv_loss(z, b, true_class):
        logp = torch.nn.LogSoftmax(dim=1)(clf(z))
        oh = F.one_hot(b.squeeze(-1), n_batch)
        tgt = oh if true_class else (~oh.bool()).float() / (n_batch - 1)
        return -(logp * tgt).sum(dim=1).mean()

    tensors, out = {}, {}
    with RecordingFeed(seed + 2) as feed:
        for t in range(n_steps):
            tt = make_inputs(400, G, n_batch, B, seed + 10 * t)
            y = rng.negative_binomial(2, 0.02, size=(B, n_prot)) * rng.binomial(1, 0.9, size=(B, n_prot))
            tt["proteins"] = torch.from_numpy(y.astype(np.float32))
            for k, v in tt.items():
                tensors[f"s{t}::{k}"] = v
            klw = (first_epoch + t) / 400.0
            kappa = 1 - klw
            inf, _, lo = m(tt, loss_kwargs={"kl_weight": klw})
            z = inf["z"]
            elbo = lo.loss.detach().clone()
            fool = adv_loss(z, tt["batch"], False)
            loss = lo.loss + fool * kappa
            opt1.zero_grad()
            loss.backward()
            opt1.step()
            adv = adv_loss(z.detach(), tt["batch"], True) * kappa
            opt2.zero_grad()
            adv.backward()
            opt2.step()
            out[f"elbo{t}"] = elbo.numpy()
            out[f"fool{t}"] = fool.detach().numpy()
            out[f"train_loss{t}"] = loss.detach().numpy()
            out[f"adv{t}"] = adv.detach().numpy()
    out["loss"] = out[f"train_loss{n_steps - 1}"]
    for k, v in m.state_dict().items():
        out["param_after::" + k] = v.detach().clone().numpy()
    for k, v in clf.state_dict().items():
        out["clf_after::" + k] = v.detach().clone().numpy()
    trained = [k for k, p in m.named_parameters() if p.grad is not None]
    extra_in = {"n_steps": np.int64(n_steps), "first_epoch": np.int64(first_epoch), "trained": np.array(json.dumps(trained))}
    for k, v in P0.items():
        extra_in["param::" + k] = v.numpy()
    for k, v in C0.items():
        extra_in["clf::" + k] = v.numpy()
    save_case(name, kwargs, tensors, m, feed, out, extra_in)


def main():
    torch.set_num_threads(1)
    if len(sys.argv) > 1 and sys.argv[1] == "trajectory":
        totalvi_trajectory("totalvi_adversarial_trajectory", 100, 12, 3, 48, 60, 3, {"n_hidden": 64, "n_latent": 8})
        return
    totalvi_case("totalvi_nb", 120, 14, 2, 48, 40, 0.6, {"n_hidden": 64, "n_latent": 8})
    totalvi_case("totalvi_zinb_mask_ln", 90, 10, 3, 40, 41, 1.0,
                 {"n_hidden": 128, "gene_likelihood": "zinb", "latent_distribution": "ln",
                  "protein_dispersion": "protein-batch", "gene_dispersion": "gene-batch"}, mask=True)
    totalvi_case("totalvi_libenc_default_width", 64, 6, 2, 24, 42, 0.3,
                 {"use_observed_lib_size": False, "n_layers_decoder": 2}, lib_prior=True)


if __name__ == "__main__":
    main()
