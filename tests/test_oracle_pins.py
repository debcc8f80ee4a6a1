"""Pin oracle/scvi_oracle.py (the CPU restatement) against vectors recorded from the
UNMODIFIED reference (tests/golden, oracle/make_golden.py) and the reference's own
known-answer tests."""
import numpy as np
import pytest
import torch

from oracle import scvi_oracle as O
from tests._golden import GOLD, VAE_CASES, grad_close, grad_scale, load_vae_case, max_rel, rel_err


def test_likelihood_matches_reference_vectors():
    z = np.load(f"{GOLD}/likelihood.npz")
    x, mu, th, pi, w = (torch.from_numpy(z[k]) for k in ("x", "mu", "theta", "pi", "w"))
    nb = O.log_nb_positive(x, mu, th.view(1, -1))
    zinb = O.log_zinb_positive(x, mu, th, pi)
    assert torch.equal(nb, torch.from_numpy(z["nb"]))  # same torch ops, same order: bit-exact
    assert torch.equal(zinb, torch.from_numpy(z["zinb"]))
    mu_g, th_g, pi_g = (t.clone().requires_grad_(True) for t in (mu, th, pi))
    (O.log_zinb_positive(x, mu_g, th_g, pi_g) * w).sum().backward()
    assert rel_err(mu_g.grad, z["zinb_dmu"]) < 1e-6
    assert rel_err(th_g.grad, z["zinb_dtheta"]) < 1e-6
    assert rel_err(pi_g.grad, z["zinb_dpi"]) < 1e-6


@pytest.mark.parametrize("name", VAE_CASES)
def test_vae_step_matches_reference_vectors(name):
    P, inp, out = load_vae_case(name)
    torch.set_num_threads(1)
    x = torch.from_numpy(inp["x"])
    bi = torch.from_numpy(inp["batch"])
    kw = dict(n_batch=inp["n_batch"], gene_likelihood=inp["likelihood"], kl_weight=inp["kl_weight"],
              drop_mask=torch.from_numpy(inp["drop_mask"]), noise=torch.from_numpy(inp["noise"]))
    o, grads = O.loss_and_grads(P, x, bi, **kw)
    assert abs(float(o["loss"].detach()) - float(out["loss"])) <= 1e-6 * abs(float(out["loss"]))
    assert max_rel(o["reconst"].detach(), out["reconst"]) < 2e-6
    assert max_rel(o["kl_z"].detach(), out["kl_z"], floor=1e-3) < 2e-5
    assert rel_err(o["z"].detach(), out["z"]) < 1e-6
    assert rel_err(o["px_rate"].detach(), out["px_rate"]) < 1e-6
    for k in O.trained_keys(inp["likelihood"]):
        assert grad_close(grads[k], out["grad::" + k], 2e-5, grad_scale(out)), k
    for k, v in o["bn_buffers"].items():
        assert rel_err(v, out["buf_after::" + k]) < 1e-6, k
    # eval mode (running stats, no dropout)
    with torch.no_grad():
        e = O.vae_forward(P, x, bi, training=False, **kw)
    assert abs(float(e["loss"]) - float(out["eval_loss"])) <= 2e-6 * abs(float(out["eval_loss"]))
    assert max_rel(e["reconst"], out["eval_reconst"]) < 3e-6


@pytest.mark.parametrize("name", ["c1_zinb", "onebatch_nb"])
def test_adam_steps_match_reference_vectors(name):
    """Two full oracle train steps (fwd, bwd, restated Adam) reproduce the parameters the
    reference module + real torch.optim.Adam(lr=1e-3, eps=0.01, weight_decay=1e-6) reach."""
    P, inp, out = load_vae_case(name)
    torch.set_num_threads(1)
    x = torch.from_numpy(inp["x"])
    bi = torch.from_numpy(inp["batch"])
    kw = dict(n_batch=inp["n_batch"], gene_likelihood=inp["likelihood"], kl_weight=inp["kl_weight"],
              drop_mask=torch.from_numpy(inp["drop_mask"]), noise=torch.from_numpy(inp["noise"]))
    keys = O.trained_keys(inp["likelihood"])
    P = {k: v.clone() for k, v in P.items()}
    M = {k: torch.zeros_like(P[k]) for k in keys}
    V = {k: torch.zeros_like(P[k]) for k in keys}
    n = inp["n_adam_steps"]
    for step in range(1, n + 1):
        o, g = O.loss_and_grads(P, x, bi, **kw)
        for k in keys:
            P[k], M[k], V[k] = O.adam_step(P[k], g[k], M[k], V[k], step)
        P.update(o["bn_buffers"])
    for k in keys:
        assert rel_err(P[k], out[f"param_after{n}::" + k]) < 1e-6, k


@pytest.mark.parametrize(
    ("current", "n_warm_up", "min_kl_weight", "max_kl_weight", "expected"),
    [(0, 400, 0.0, 1.0, 0.0), (200, 400, 0.0, 1.0, 0.5), (400, 400, 0.0, 1.0, 1.0),
     (0, 400, 0.5, 1.0, 0.5), (200, 400, 0.5, 1.0, 0.75), (400, 400, 0.0, 2.0, 2.0)],
)
def test_kl_weight_known_answers(current, n_warm_up, min_kl_weight, max_kl_weight, expected):
    # the reference's own table: tests/train/test_trainingplans.py:11-28
    assert O.compute_kl_weight(current, 1, n_warm_up, None, max_kl_weight, min_kl_weight) == pytest.approx(expected)
    assert O.compute_kl_weight(1, current, None, n_warm_up, max_kl_weight, min_kl_weight) == pytest.approx(expected)


def test_kl_weight_precedence_and_errors():
    # tests/train/test_trainingplans.py:31-51
    assert O.compute_kl_weight(1, 1, None, None, 2.0, 0.0) == 2.0
    with pytest.raises(ValueError):
        O.compute_kl_weight(1, 1, 400, None, 0.5, 1.0)
    assert O.compute_kl_weight(0, 100, 100, 100) == 0.0
    assert O.compute_kl_weight(50, 200, 100, 1000) == 0.5
    assert O.compute_kl_weight(100, 200, 100, 1000) == 1.0


def test_distributed_batches_match_torch_sampler():
    """oracle index stream == torch DistributedSampler + the reference's batching
    (src/scvi/dataloaders/_samplers.py:44-89)."""
    from torch.utils.data import DistributedSampler

    class DS:
        def __len__(self):
            return 103

    for world in (1, 2, 4):
        for tail in (False, True):
            seen = []
            for rank in range(world):
                s = DistributedSampler(DS(), num_replicas=world, rank=rank, shuffle=True, seed=3, drop_last=tail)
                s.set_epoch(5)
                ref = list(iter(s))
                got = O.distributed_batches(103, 16, rank, world, epoch=5, seed=3, drop_dataset_tail=tail)
                assert np.concatenate(got).tolist() == ref
                assert all(len(b) == 16 for b in got[:-1])
                seen += ref
            if tail:
                assert len(set(seen)) == len(seen)  # disjoint across replicas


def test_live_reference_if_present():
    """Where /root/reference exists (build container), cross-check against the live reference."""
    from oracle import ref_shim

    if not ref_shim.available():
        pytest.skip("reference sources not on this box")
    R = ref_shim.load()
    g = torch.Generator().manual_seed(11)
    x = torch.poisson(torch.rand(33, 21, generator=g) * 6, generator=g)
    mu = torch.rand(33, 21, generator=g) * 8 + 1e-3
    th = torch.rand(21, generator=g) * 3 + 0.1
    pi = torch.randn(33, 21, generator=g)
    assert torch.equal(R.log_zinb_positive(x, mu, th, pi), O.log_zinb_positive(x, mu, th, pi))
    assert torch.equal(R.log_nb_positive(x, mu, th), O.log_nb_positive(x, mu, th.view(1, -1)))
