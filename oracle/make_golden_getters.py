"""TEST INFRASTRUCTURE ONLY -- tests/golden/getters_*.npz from the UNMODIFIED reference module.

    python oracle/make_golden_getters.py      (build container only: needs /root/reference)

Eval-mode quantities behind the post-training getters of `scvi.model.SCVI` on the inputs and initial weights of
an existing VAE fixture: q(z|x) mean / variance (get_latent_representation, _vaemixin.py:270-383), per-cell
reconstruction loss and KL (get_elbo / get_reconstruction_error, _log_likelihood.py:20-136) and decoded
expression `px.scale` / `px.mu` for each cell's own batch and for `transform_batch` (get_normalized_expression,
_rnamixin.py:155-330).  The reparameterisation noise is fed explicitly (same mechanism as make_golden.py).
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
from oracle.make_golden import GOLD, _FeedRNG  # noqa: E402


def getters_case(src: str):
    R = ref_shim.load()
    z = np.load(os.path.join(GOLD, src + ".npz"))
    P0 = {k[len("param::"):]: torch.from_numpy(z[k]) for k in z.files if k.startswith("param::")}
    x, bi = torch.from_numpy(z["in::x"]), torch.from_numpy(z["in::batch"])
    noise = torch.from_numpy(z["in::noise"])
    n_batch, lik = int(z["in::n_batch"]), str(z["in::likelihood"])
    m = R.VAE(n_input=x.shape[1], n_batch=n_batch, gene_likelihood=lik)
    sd = m.state_dict()
    sd.update(P0)
    m.load_state_dict(sd)
    m.eval()
    tensors = {"X": x, "batch": bi, "labels": torch.zeros_like(bi)}
    out = {}
    with torch.no_grad(), _FeedRNG(torch.ones(x.shape[0], 128), noise):
        inf, gen, lo = m(tensors, loss_kwargs={"kl_weight": 1})
        out["qz_mean"] = inf["qz"].loc.numpy()
        out["qz_var"] = inf["qz"].scale.square().numpy()
        out["z"] = inf["z"].numpy()
        out["reconst"] = lo.reconstruction_loss["reconstruction_loss"].numpy()
        out["kl_z"] = lo.kl_local["kl_divergence_z"].numpy()
        out["px_scale"] = gen["px"].get_normalized("scale").numpy()
        out["px_mu"] = gen["px"].get_normalized("mu").numpy()
        for b in range(n_batch):
            _, gen_b = m(tensors, generative_kwargs={"transform_batch": b}, compute_loss=False)
            out[f"px_scale_tb{b}"] = gen_b["px"].get_normalized("scale").numpy()
    path = os.path.join(GOLD, f"getters_{src}.npz")
    np.savez_compressed(path, **out)
    print(f"getters_{src}: elbo={-float((out['reconst'] + out['kl_z']).mean()):.6f} -> {path} ({os.path.getsize(path)/1e3:.0f} kB)")


if __name__ == "__main__":
    torch.set_num_threads(1)
    getters_case("c1_zinb")
    getters_case("onebatch_nb")
