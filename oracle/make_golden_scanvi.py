"""TEST INFRASTRUCTURE ONLY -- tests/golden/scanvi_*.npz from the UNMODIFIED reference `SCANVAE`.

    python oracle/make_golden_scanvi.py      (build container only: needs /root/reference)

One `SemiSupervisedTrainingPlan.training_step` worth of module work (src/scvi/train/_trainingplans.py:905-942):
`module(full_dataset, loss_kwargs={kl_weight, labelled_tensors, classification_ratio})` on seeded inputs in
training mode, then `backward()`.  Stored: inputs (full + labelled minibatch), the complete state_dict, every
dropout mask / reparameterisation draw in call order, the loss pieces, logits and every gradient.  Also an
eval-mode `classify` (the `SCANVI.predict` path, posterior mean) on the same cells.
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
from oracle.make_golden_options import RecordingFeed, make_inputs, perturb_, save_case  # noqa: E402


def scanvi_case(name, G, n_batch, n_labels, B, B_lab, seed, kl_weight, ratio, kwargs_extra):
    R = ref_shim.load()
    kwargs = dict(n_input=G, n_batch=n_batch, n_labels=n_labels, **kwargs_extra)
    torch.manual_seed(seed)
    m = R.SCANVAE(**kwargs)
    perturb_(m, seed + 3)
    m.train()
    full = make_inputs(400, G, n_batch, B, seed, n_labels=n_labels)
    tensors = dict(full)
    lab = None
    if B_lab:
        lab = make_inputs(400, G, n_batch, B_lab, seed + 50, n_labels=n_labels)
        for k, v in lab.items():
            tensors["lab::" + k] = v
    P0 = {k: v.detach().clone() for k, v in m.state_dict().items()}
    with RecordingFeed(seed + 2) as feed:
        inf, gen, lo = m(full, loss_kwargs={"kl_weight": kl_weight, "labelled_tensors": lab,
                                            "classification_ratio": ratio})
    lo.loss.backward()
    out = {"loss": lo.loss.detach().numpy(),
           "reconst": lo.reconstruction_loss["reconstruction_loss"].detach().numpy(),
           "kl_local": lo.kl_local["kl_local"].detach().numpy(),
           "z": inf["z"].detach().numpy()}
    if lab is not None:
        out["classification_loss"] = lo.classification_loss.detach().numpy()
        out["logits"] = lo.logits.detach().numpy()
        out["true_labels"] = lo.true_labels.numpy()
    for k, p in m.named_parameters():
        if p.grad is not None:
            out["grad::" + k] = p.grad.detach().clone().numpy()
    for k, b in m.named_buffers():
        if "running" in k:
            out["buf_after::" + k] = b.detach().clone().numpy()
    # eval-mode classify from the initial state (predict path)
    m2 = R.SCANVAE(**kwargs)
    m2.load_state_dict(P0)
    m2.eval()
    with torch.no_grad():
        out["eval_classify_logits"] = m2.classify(full["X"], batch_index=full["batch"]).numpy()
    extra_in = {"kl_weight": np.float64(kl_weight), "classification_ratio": np.float64(ratio)}
    for k, v in P0.items():
        extra_in["param::" + k] = v.numpy()
    save_case(name, kwargs, tensors, m, feed, out, extra_in)


def scanvi_trajectory(name, G, n_batch, n_labels, B, B_lab, seed, n_steps, kwargs_extra):
    """`SemiSupervisedTrainingPlan.training_step` x n_steps on the unmodified reference module
    (src/scvi/train/_trainingplans.py:936-969): loss(full minibatch, labelled minibatch, classification_ratio 50,
    kl_weight epoch / 400) -> backward -> torch.optim.Adam(lr 1e-3, eps 0.01, weight_decay 1e-6) over the
    parameters that require grad (:475-482)."""
    R = ref_shim.load()
    kwargs = dict(n_input=G, n_batch=n_batch, n_labels=n_labels, **kwargs_extra)
    torch.manual_seed(seed)
    m = R.SCANVAE(**kwargs)
    perturb_(m, seed + 3)
    m.train()
    P0 = {k: v.detach().clone() for k, v in m.state_dict().items()}
    opt = torch.optim.Adam(filter(lambda p: p.requires_grad, m.parameters()), lr=1e-3, eps=0.01, weight_decay=1e-6)
    tensors, out = {}, {}
    with RecordingFeed(seed + 2) as feed:
        for t in range(n_steps):
            full = make_inputs(400, G, n_batch, B, seed + 10 * t, n_labels=n_labels)
            lab = make_inputs(400, G, n_batch, B_lab, seed + 10 * t + 5, n_labels=n_labels)
            for k, v in full.items():
                tensors[f"s{t}::full::{k}"] = v
            for k, v in lab.items():
                tensors[f"s{t}::lab::{k}"] = v
            klw = t / 400.0  # _compute_kl_weight(epoch=t, step, n_epochs_kl_warmup=400)
            opt.zero_grad()
            _, _, lo = m(full, loss_kwargs={"kl_weight": klw, "labelled_tensors": lab, "classification_ratio": 50})
            lo.loss.backward()
            opt.step()
            out[f"loss{t}"] = lo.loss.detach().numpy()
            out[f"classification_loss{t}"] = lo.classification_loss.detach().numpy()
    out["loss"] = out[f"loss{n_steps - 1}"]
    for k, v in m.state_dict().items():
        out["param_after::" + k] = v.detach().clone().numpy()
    trained = [k for k, p in m.named_parameters() if p.grad is not None]
    extra_in = {"n_steps": np.int64(n_steps), "trained": np.array(json.dumps(trained))}
    for k, v in P0.items():
        extra_in["param::" + k] = v.numpy()
    save_case(name, kwargs, tensors, m, feed, out, extra_in)


def main():
    torch.set_num_threads(1)
    if len(sys.argv) > 1 and sys.argv[1] == "trajectory":
        scanvi_trajectory("scanvi_trajectory", 90, 2, 5, 64, 32, 30, 3, {"gene_likelihood": "zinb"})
        return
    scanvi_case("scanvi_zinb_labelled", 100, 2, 4, 64, 24, 20, 0.5, 50.0, {"gene_likelihood": "zinb"})
    scanvi_case("scanvi_nb_unlabelled", 80, 3, 20, 40, 0, 21, 1.0, 50.0, {"gene_likelihood": "nb"})
    scanvi_case("scanvi_linear_classifier", 64, 1, 3, 48, 16, 22, 0.25, 50.0,
                {"gene_likelihood": "zinb", "linear_classifier": True, "n_layers": 2})


if __name__ == "__main__":
    main()
