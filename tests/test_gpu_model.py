"""GPU: the `SCVI`-shaped driver (train + post-training getters) against quantities recorded from the
unmodified reference module in eval mode (tests/golden/getters_*.npz, oracle/make_golden_getters.py)."""
import os

import numpy as np
import pytest
import torch

from oracle import scvi_oracle as O
from tests._golden import GOLD, load_vae_case, max_rel, rel_err

pytestmark = pytest.mark.gpu


def _model_from_case(name):
    import scvi_tools_b200 as S
    from scvi_tools_b200.model import SCVI

    P, inp, _ = load_vae_case(name)
    model = SCVI(inp["x"], batch=inp["batch"].reshape(-1), gene_likelihood=inp["likelihood"])
    assert model.summary["n_batch"] == inp["n_batch"]
    sd = model.state_dict()
    sd.update({k: v.clone() for k, v in P.items()})
    model.load_state_dict(sd)
    gold = np.load(os.path.join(GOLD, f"getters_{name}.npz"))
    return S, model, inp, gold


@pytest.mark.parametrize("name", ["c1_zinb", "onebatch_nb"])
def test_getters_match_reference_eval_mode(name):
    S, model, inp, gold = _model_from_case(name)
    noise = torch.from_numpy(inp["noise"])
    qm, qv = model.get_latent_representation(return_dist=True)
    assert rel_err(qm, gold["qz_mean"]) < 1e-5 and rel_err(qv, gold["qz_var"]) < 1e-5
    assert rel_err(model.get_latent_representation(), gold["qz_mean"]) < 1e-5
    # indices + small batches walk the same cells
    sub = np.array([5, 3, 47, 0, 41])
    assert rel_err(model.get_latent_representation(indices=sub, batch_size=2), gold["qz_mean"][sub]) < 1e-5
    with S.rng_feed(noise=[noise]):
        elbo = model.get_elbo(return_mean=False)
    assert max_rel(elbo, -(gold["reconst"] + gold["kl_z"])) < 1e-5
    with S.rng_feed(noise=[noise]):
        assert abs(model.get_elbo() + float((gold["reconst"] + gold["kl_z"]).mean())) < 1e-5 * abs(float(gold["reconst"].mean()))
    with S.rng_feed(noise=[noise]):
        rec = model.get_reconstruction_error(return_mean=False)["reconstruction_loss"]
    assert max_rel(rec, gold["reconst"]) < 1e-5
    with S.rng_feed(noise=[noise]):
        scale = model.get_normalized_expression()
    assert rel_err(scale, gold["px_scale"]) < 1e-5
    np.testing.assert_allclose(scale.sum(1), 1.0, rtol=1e-5)
    with S.rng_feed(noise=[noise]):
        assert rel_err(model.get_normalized_expression(library_size=1e4), 1e4 * gold["px_scale"]) < 1e-5
    with S.rng_feed(noise=[noise]):
        assert rel_err(model.get_normalized_expression(library_size="latent"), gold["px_mu"]) < 1e-5
    nb = inp["n_batch"]
    with S.rng_feed(noise=[noise] * nb):
        tb = model.get_normalized_expression(transform_batch=list(range(nb)), gene_indices=[0, 7, 9])
    want = np.mean([gold[f"px_scale_tb{b}"] for b in range(nb)], axis=0)[:, [0, 7, 9]]
    assert rel_err(tb, want) < 1e-5
    with S.rng_feed(noise=[noise]):
        assert rel_err(model.get_normalized_expression(transform_batch=nb - 1), gold[f"px_scale_tb{nb - 1}"]) < 1e-5


@pytest.mark.parametrize("fx", ["marginal_c1_zinb_6x3", "marginal_c1_zinb_4x1", "marginal_onebatch_nb_5x5"])
def test_marginal_ll_matches_reference(fx):
    """`get_marginal_ll` / `VAE.marginal_ll`: the importance samples flattened into the fused kernel's row
    dimension give the reference's per-cell log p(x) for the same normal draws."""
    gold = np.load(os.path.join(GOLD, fx + ".npz"))
    S, model, inp, _ = _model_from_case(str(gold["src"]))
    noises = [torch.from_numpy(gold[f"noise::{i}"]) for i in range(sum(k.startswith("noise::") for k in gold.files))]
    n_mc, n_per = int(gold["n_mc"]), int(gold["n_per"])
    with S.rng_feed(noise=noises):
        ll = model.get_marginal_ll(n_mc_samples=n_mc, n_mc_samples_per_pass=n_per, return_mean=False)
    assert ll.shape == gold["marginal_ll"].shape
    assert max_rel(ll, gold["marginal_ll"]) < 1e-5
    with S.rng_feed(noise=noises):
        mean = model.get_marginal_ll(n_mc_samples=n_mc, n_mc_samples_per_pass=n_per)
    assert abs(mean - float(gold["marginal_ll"].mean())) < 1e-5 * abs(float(gold["marginal_ll"].mean()))
    # training mode takes the generic per-sample route (BatchNorm per importance-sample slice) and stays finite
    model.module.train()
    assert np.isfinite(model.module.marginal_ll(model.data.minibatch(torch.arange(16, device="cuda:0")), 4, True, 2))
    model.module.eval()


def test_save_load_round_trip_and_reference_shaped_checkpoint(tmp_path):
    """`model.pt` in the reference's layout: our own save -> load reproduces the getters bit for bit; a checkpoint
    shaped like the reference's own (`model_state_dict` incl. `pyro_param_store`, `var_names`, `attr_dict` with
    `init_params_` and a `registry_` this driver ignores) loads too."""
    from scvi_tools_b200.model import SCVI

    S, model, inp, gold = _model_from_case("c1_zinb")
    names = np.array([f"gene_{i}" for i in range(inp["x"].shape[1])])
    model.var_names = names
    model.is_trained_ = True
    d = str(tmp_path / "m")
    model.save(d)
    with pytest.raises(ValueError):
        model.save(d)
    model.save(d, overwrite=True)
    blob = torch.load(os.path.join(d, "model.pt"), weights_only=False)
    assert set(blob) == {"model_state_dict", "var_names", "attr_dict"}
    assert "pyro_param_store" in blob["model_state_dict"] and "init_params_" in blob["attr_dict"]
    m2 = SCVI.load(d, inp["x"], batch=inp["batch"].reshape(-1), var_names=names)
    assert m2.is_trained_ and np.array_equal(m2.var_names, names)
    assert np.array_equal(m2.get_latent_representation(), model.get_latent_representation())
    with pytest.raises(ValueError):
        SCVI.load(d, inp["x"], batch=inp["batch"].reshape(-1), var_names=names[::-1])
    with pytest.raises(ValueError):
        SCVI.load(str(tmp_path / "nowhere"), inp["x"])
    # a reference-shaped file: the golden state dict (reference keys), SCVI.__init__'s recorded arguments
    P, _, _ = load_vae_case("c1_zinb")
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    sd.update({k: v.clone() for k, v in P.items()})
    sd["pyro_param_store"] = {"params": {}, "constraints": {}}
    ref_blob = {
        "model_state_dict": sd, "var_names": names,
        "attr_dict": {"init_params_": {"non_kwargs": {"n_hidden": 128, "n_latent": 10, "n_layers": 1,
                                                        "dropout_rate": 0.1, "dispersion": "gene",
                                                        "gene_likelihood": "zinb", "use_observed_lib_size": True,
                                                        "latent_distribution": "normal"},
                                       "kwargs": {"kwargs": {}}},
                      "is_trained_": True, "registry_": {"scvi_version": "1.5.0", "model_name": "SCVI"},
                      "history_": None, "train_indices_": np.arange(3)},
    }
    d2 = tmp_path / "ref"
    d2.mkdir()
    torch.save(ref_blob, str(d2 / "model.pt"))
    m3 = SCVI.load(str(d2), inp["x"], batch=inp["batch"].reshape(-1))
    assert rel_err(m3.get_latent_representation(), gold["qz_mean"]) < 1e-5


def test_checkpoint_with_reference_module_weights_loads_and_reproduces_reference_outputs():
    """tests/golden/ref_checkpoint/model.pt (oracle/make_golden_checkpoint.py): weights, key constants and expected
    outputs all come from the unmodified reference module.  `SCVI.load` on it reproduces the reference's latent
    means, px_scale and per-cell reconstruction loss for the recorded cells."""
    from scvi_tools_b200.model import SCVI
    from tests._golden import GOLD

    e = np.load(os.path.join(GOLD, "ref_checkpoint_expect.npz"))
    model = SCVI.load(os.path.join(GOLD, "ref_checkpoint"), e["X"], batch=e["batch"], var_names=e["var_names"])
    assert model.is_trained_ and model.module.gene_likelihood == "nb"
    idx = e["idx"]
    assert rel_err(model.get_latent_representation(indices=idx), e["qzm"]) < 1e-5
    # decode the posterior mean, as the recording did (no random draw)
    mod = model.module
    t = model.data.minibatch(torch.as_tensor(idx, device=model.device))
    with torch.inference_mode():
        inf = mod.inference(**mod._get_inference_input(t))
        gi = mod._get_generative_input(t, inf)
        gi["z"] = inf["qz"].loc
        px = mod.generative(**gi)["px"]
        rec = px.nll(t["X"]) if hasattr(px, "nll") else -px.log_prob(t["X"]).sum(-1)
        scale = px.scale
    assert rel_err(scale.cpu(), e["px_scale"]) < 1e-5
    assert float(((rec.cpu().double() - torch.from_numpy(e["reconst"]).double()).abs()
                  / torch.from_numpy(e["reconst"]).double().abs()).max()) < 1e-5


def test_train_then_getters_end_to_end():
    """`SCVI(...).train()` lowers the ELBO; getters run on the trained, arena-backed parameters; a state-dict
    round trip into a fresh model reproduces them."""
    from scvi_tools_b200.model import SCVI, get_max_epochs_heuristic

    assert get_max_epochs_heuristic(400) == 400 and get_max_epochs_heuristic(1_000_000) == 8
    X, batch = O.synthetic_iid(1200, 200, 3, seed=2)
    torch.manual_seed(0)
    model = SCVI(X, batch=batch, gene_likelihood="nb")
    before = model.get_reconstruction_error()["reconstruction_loss"]
    hist = model.train(max_epochs=8, batch_size=256)
    assert model.is_trained_ and len(hist) == 8 and hist[-1]["train_loss"] < hist[0]["train_loss"]
    after = model.get_reconstruction_error()["reconstruction_loss"]
    assert np.isfinite(after) and after < before
    z = model.get_latent_representation()
    assert z.shape == (1200, 10) and np.isfinite(z).all()
    assert np.isfinite(model.get_elbo())
    fresh = SCVI(X, batch=batch, gene_likelihood="nb")
    fresh.load_state_dict({k: v.clone() for k, v in model.state_dict().items()})
    np.testing.assert_allclose(fresh.get_latent_representation(), z, rtol=1e-6, atol=1e-6)
    hist2 = model.train(max_epochs=2, batch_size=256)  # continues from the same plan / optimizer state
    assert hist2[0]["epoch"] == 8
