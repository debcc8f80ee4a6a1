"""TEST INFRASTRUCTURE ONLY -- tests/golden/ref_checkpoint/model.pt: a `model.pt` holding the weights of the
UNMODIFIED reference `scvi.module.VAE`, in the dictionary layout `BaseModelClass.save` writes
(src/scvi/model/base/_base_model.py:792-905; keys from src/scvi/model/base/_constants.py SAVE_KEYS), together with
outputs of that reference module on recorded cells (tests/golden/ref_checkpoint_expect.npz).

    python oracle/make_golden_checkpoint.py      (build container only: needs /root/reference)

`scvi.model.SCVI.save` itself cannot run here (it needs AnnData, absent from this image): the file is assembled
from the reference module's own `state_dict()`, the reference's key constants and the `init_params_` record
`BaseModelClass._get_init_params` would store for `SCVI(adata, n_hidden=128, n_latent=10, gene_likelihood="nb")`.
"""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import ref_shim  # noqa: E402
from oracle import scvi_oracle as O  # noqa: E402


def main():
    R = ref_shim.load()
    spec = importlib.util.spec_from_file_location(
        "_ref_model_constants", "/root/reference/src/scvi/model/base/_constants.py")
    consts = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(consts)
    K = consts.SAVE_KEYS
    G, nb = 60, 2
    torch.manual_seed(21)
    m = R.VAE(n_input=G, n_batch=nb, n_hidden=128, n_latent=10, gene_likelihood="nb")
    with torch.no_grad():  # a "trained" look: non-trivial BatchNorm buffers and dispersions
        g = torch.Generator().manual_seed(22)
        for k, b in m.named_buffers():
            if k.endswith("running_mean"):
                b.copy_(0.2 * torch.randn(b.shape, generator=g))
            if k.endswith("running_var"):
                b.copy_(0.5 + torch.rand(b.shape, generator=g))
        m.px_r.copy_(torch.randn(G, generator=g))
    m.eval()
    X, batch = O.synthetic_iid(200, G, nb, seed=23)
    sd = {k: v.detach().clone() for k, v in m.state_dict().items()}
    sd["pyro_param_store"] = {"params": {}, "constraints": {}}  # BaseModelClass.save adds it for every model
    attr = {
        "init_params_": {"non_kwargs": {"n_hidden": 128, "n_latent": 10, "n_layers": 1, "dropout_rate": 0.1,
                                        "dispersion": "gene", "gene_likelihood": "nb", "latent_distribution": "normal",
                                        "use_observed_lib_size": True},
                         "kwargs": {"kwargs": {}}},
        "is_trained_": True,
        "registry_": {"scvi_version": "1.5.0.post1", "model_name": "SCVI", "setup_args": {"batch_key": "batch"}},
        "summary_stats_": {"n_cells": 200, "n_vars": G, "n_batch": nb},
    }
    var_names = np.array([f"gene_{i}" for i in range(G)])
    blob = {K.MODEL_STATE_DICT_KEY: sd, K.VAR_NAMES_KEY: var_names, K.ATTR_DICT_KEY: attr}
    d = os.path.join(ROOT, "tests", "golden", "ref_checkpoint")
    os.makedirs(d, exist_ok=True)
    torch.save(blob, os.path.join(d, "model.pt"))
    idx = np.arange(0, 200, 7)
    t = {"X": torch.from_numpy(X[idx]), "batch": torch.from_numpy(batch[idx]).view(-1, 1),
         "labels": torch.zeros(len(idx), 1, dtype=torch.long)}
    with torch.no_grad():
        inf = m.inference(**m._get_inference_input(t))
        gi = m._get_generative_input(t, inf)
        gi["z"] = inf["qz"].loc  # decode the posterior MEAN: no random draw in the expected values
        gen = m.generative(**gi)
        qzm = inf["qz"].loc.numpy()
        scale = gen["px"].scale.numpy()
        rec = -gen["px"].log_prob(t["X"]).sum(-1).numpy()
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ref_checkpoint_expect.npz"), X=X, batch=batch, idx=idx,
                        qzm=qzm, px_scale=scale, reconst=rec, var_names=var_names)
    print("wrote", d, os.path.getsize(os.path.join(d, "model.pt")), "bytes; keys", list(blob))


if __name__ == "__main__":
    main()
