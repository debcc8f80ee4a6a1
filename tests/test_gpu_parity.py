"""GPU parity: the CUDA path (through the C ABI) against the golden vectors recorded from the
unmodified reference and against the oracle on seeded inputs.

Tolerances (BASELINE.json north_star, fp32 path): per-minibatch ELBO 1e-5 relative, gradients
1e-4 relative, gather bit-exact.
"""
import numpy as np
import pytest
import torch

from oracle import scvi_oracle as O
from tests._golden import GOLD, VAE_CASES, grad_close, grad_scale, load_vae_case, max_rel, rel_err

pytestmark = pytest.mark.gpu

ELBO_RTOL = 1e-5
GRAD_RTOL = 1e-4


def dev():
    return torch.device("cuda:0")


def _module_from_case(P, inp):
    import scvi_tools_b200 as S

    G = P["px_r"].shape[0]
    m = S.VAE(n_input=G, n_batch=inp["n_batch"], gene_likelihood=inp["likelihood"])
    sd = m.state_dict()
    for k, v in P.items():
        sd[k] = v.clone()
    m.load_state_dict(sd)
    return m.to(dev())


def _tensors(inp):
    B = inp["x"].shape[0]
    return {
        "X": torch.from_numpy(inp["x"]).to(dev()),
        "batch": torch.from_numpy(inp["batch"]).to(dev()),
        "labels": torch.zeros(B, 1, dtype=torch.long, device=dev()),
    }


@pytest.mark.parametrize("name", VAE_CASES)
def test_vae_step_vs_reference_golden(name):
    import scvi_tools_b200 as S

    P, inp, out = load_vae_case(name)
    m = _module_from_case(P, inp).train()
    mask = torch.from_numpy(inp["drop_mask"])
    noise = torch.from_numpy(inp["noise"])
    with S.rng_feed(drop_masks=[mask], noise=[noise]):
        inf, gen, lo = m(_tensors(inp), loss_kwargs={"kl_weight": inp["kl_weight"]})
    lo.loss.backward()
    assert abs(float(lo.loss) - float(out["loss"])) <= ELBO_RTOL * abs(float(out["loss"]))
    rec = lo.reconstruction_loss["reconstruction_loss"].detach().cpu()
    assert max_rel(rec, out["reconst"]) < ELBO_RTOL
    assert max_rel(lo.kl_local["kl_divergence_z"].detach().cpu(), out["kl_z"], floor=1e-2) < 1e-4
    assert rel_err(inf["z"].detach().cpu(), out["z"]) < 1e-5
    assert rel_err(inf["library"].detach().cpu(), out["library"]) < 1e-6
    gs = grad_scale(out)
    named = dict(m.named_parameters())
    for k in O.trained_keys(inp["likelihood"]):
        assert named[k].grad is not None, k
        assert grad_close(named[k].grad.cpu(), out["grad::" + k], GRAD_RTOL, gs), k
    # parameters the reference leaves without gradient stay without gradient
    for k, p in named.items():
        if k not in O.trained_keys(inp["likelihood"]):
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, k
    sd = m.state_dict()
    for k in out:
        if k.startswith("buf_after::"):
            assert rel_err(sd[k.split("::", 1)[1]].cpu(), out[k]) < 1e-5, k
    assert int(lo.n_obs_minibatch) == inp["x"].shape[0]


@pytest.mark.parametrize("name", ["c1_zinb", "ragged_zinb"])
def test_vae_eval_forward_vs_reference_golden(name):
    import scvi_tools_b200 as S

    P, inp, out = load_vae_case(name)
    m = _module_from_case(P, inp).eval()
    with torch.no_grad(), S.rng_feed(noise=[torch.from_numpy(inp["noise"])]):
        # eval mode: auto_move_data moves CPU inputs to the module's device (reference _decorators.py:30-41)
        cpu_t = {k: v.cpu() for k, v in _tensors(inp).items()}
        _, gen, lo = m(cpu_t, loss_kwargs={"kl_weight": inp["kl_weight"]})
    assert abs(float(lo.loss) - float(out["eval_loss"])) <= ELBO_RTOL * abs(float(out["eval_loss"]))
    assert max_rel(lo.reconstruction_loss["reconstruction_loss"].cpu(), out["eval_reconst"]) < ELBO_RTOL


def test_materialised_px_matches_reference_golden():
    """generative()'s lazy px materialises px.scale / px.mu / px.zi_logits for the getters."""
    import scvi_tools_b200 as S

    P, inp, out = load_vae_case("c1_zinb")
    m = _module_from_case(P, inp).train()
    with torch.no_grad(), S.rng_feed(drop_masks=[torch.from_numpy(inp["drop_mask"])], noise=[torch.from_numpy(inp["noise"])]):
        _, gen = m(_tensors(inp), compute_loss=False)
    px = gen["px"]
    assert rel_err(px.scale.cpu(), out["px_scale"]) < 1e-5
    assert rel_err(px.mu.cpu(), out["px_rate"]) < 1e-5
    assert rel_err(px.zi_logits.cpu(), out["px_dropout"]) < 1e-5
    assert px.get_normalized("scale").shape == out["px_scale"].shape
    lp = px.log_prob(_tensors(inp)["X"]).sum(-1)
    # out["reconst"] was recorded from the same training-mode forward
    assert max_rel(-lp.cpu(), out["reconst"]) < ELBO_RTOL


def _fp64_truth(z):
    x, mu, th, pi = (torch.from_numpy(z[k]).double() for k in ("x", "mu", "theta", "pi"))
    nb = O.log_nb_positive(x, mu, th.view(1, -1)).numpy()
    zinb = O.log_zinb_positive(x, mu, th, pi).numpy()
    t = th.view(1, -1)
    l_tm = torch.log(t + mu + 1e-8)
    M = (t * (torch.log(t + 1e-8) - l_tm)).abs() + (x * (torch.log(mu + 1e-8) - l_tm)).abs() + torch.lgamma(x + t).abs()
    M = M + torch.lgamma(t).abs() + torch.lgamma(x + 1) + pi.abs() + 1.0
    return nb, zinb, M.numpy()


@pytest.mark.parametrize("ncols", [37, 36])
def test_likelihood_kernels_vs_reference_golden(ncols):
    """ncols = 37: scalar kernels (libdevice math); ncols = 36: the vectorised fast-math kernels (G % 4 == 0)."""
    import scvi_tools_b200 as S

    z = {k: (v[..., :ncols].copy() if v.shape[-1] == 37 else v) for k, v in np.load(f"{GOLD}/likelihood.npz").items()}
    x, mu, th, pi, w = (torch.from_numpy(z[k]).to(dev()) for k in ("x", "mu", "theta", "pi", "w"))
    mu_g, th_g, pi_g = (t.clone().requires_grad_(True) for t in (mu, th, pi))
    nb = S.log_nb_positive(x, mu_g, th_g)
    (nb * w).sum().backward()
    # per element: error vs the fp64 oracle, relative to the magnitude M of the terms being summed
    # (lgamma(65535 + theta) ~ 6e5 cancels against lgamma(x + 1): fp32 cannot do better than eps * M)
    truth_nb, truth_zinb, M = _fp64_truth(z)
    assert float(np.max(np.abs(nb.detach().cpu().numpy() - truth_nb) / M)) < 1e-6
    assert float(np.max(np.abs(z["nb"] - truth_nb) / M)) < 1e-6  # the fp32 reference meets the same bound
    assert rel_err(mu_g.grad.cpu(), z["nb_dmu"]) < GRAD_RTOL
    assert rel_err(th_g.grad.cpu(), z["nb_dtheta"]) < GRAD_RTOL
    mu_g.grad = th_g.grad = None
    zinb = S.log_zinb_positive(x, mu_g, th_g, pi_g)
    (zinb * w).sum().backward()
    assert float(np.max(np.abs(zinb.detach().cpu().numpy() - truth_zinb) / M)) < 1e-6
    assert rel_err(mu_g.grad.cpu(), z["zinb_dmu"]) < GRAD_RTOL
    assert rel_err(th_g.grad.cpu(), z["zinb_dtheta"]) < GRAD_RTOL
    assert rel_err(pi_g.grad.cpu(), z["zinb_dpi"]) < GRAD_RTOL


def test_distribution_log_prob_equals_function_and_warns():
    """Reference tests/distributions/test_negative_binomial.py:8-67: dist.log_prob == function;
    out-of-support values raise a UserWarning."""
    import scvi_tools_b200 as S

    g = torch.Generator().manual_seed(0)
    x = torch.poisson(torch.rand(10, 7, generator=g) * 5).to(dev())
    mu = (torch.rand(10, 7, generator=g) * 4 + 0.1).to(dev())
    theta = (torch.rand(7, generator=g) + 0.2).to(dev())
    pi = torch.randn(10, 7, generator=g).to(dev())
    d1 = S.NegativeBinomial(mu=mu, theta=theta).log_prob(x)
    assert torch.equal(d1, S.log_nb_positive(x, mu, theta))
    d2 = S.ZeroInflatedNegativeBinomial(mu=mu, theta=theta, zi_logits=pi).log_prob(x)
    assert torch.equal(d2, S.log_zinb_positive(x, mu, theta, pi))
    with pytest.warns(UserWarning):
        S.ZeroInflatedNegativeBinomial(mu=mu, theta=theta, zi_logits=pi).log_prob(-x - 1)
    with pytest.warns(UserWarning):
        S.NegativeBinomial(mu=mu, theta=theta, validate_args=True).log_prob(x + 0.5)
    s = S.ZeroInflatedNegativeBinomial(mu=mu, theta=theta, zi_logits=pi).sample((3,))
    assert s.shape == (3, 10, 7)


def test_gather_bit_exact():
    from scvi_tools_b200 import ops

    X, batch = O.synthetic_iid(1000, 333, 4, seed=5)
    rng = np.random.default_rng(0)
    for B in (1, 77, 128):
        idx = rng.integers(0, X.shape[0], size=B)
        ref = O.gather_rows(X, idx)
        got = ops.gather_rows(torch.from_numpy(X).to(dev()), torch.from_numpy(idx).to(dev()))
        assert np.array_equal(got.cpu().numpy(), ref)
        gb = ops.gather_i64(torch.from_numpy(batch).to(dev()), torch.from_numpy(idx).to(dev()))
        assert np.array_equal(gb.cpu().numpy(), batch[idx])
    X4, _ = O.synthetic_iid(512, 2000, 2, seed=6)  # vectorised path
    idx = rng.permutation(512)[:200]
    got = ops.gather_rows(torch.from_numpy(X4).to(dev()), torch.from_numpy(idx).to(dev()))
    assert np.array_equal(got.cpu().numpy(), O.gather_rows(X4, idx))
    Xu = torch.from_numpy(X4.astype(np.uint16)).to(dev())
    got = ops.gather_rows(Xu, torch.from_numpy(idx).to(dev()))
    assert np.array_equal(got.cpu().numpy(), O.gather_rows(X4, idx))
    empty = ops.gather_rows(torch.from_numpy(X4).to(dev()), torch.empty(0, dtype=torch.int64, device=dev()))
    assert empty.shape == (0, 2000)


@pytest.mark.parametrize("G", [2000, 333, 16, 5])
@pytest.mark.parametrize("dt", ["u8", "u16"])
def test_compact_count_gather_bit_exact(G, dt):
    """uint8 / uint16 count matrices (resident or staged from pinned host memory) expand to exactly the
    fp32 rows `data[indexes].astype(float32)` gives, with an index (gather) and without (in-order expand);
    vectorised and scalar tails, unaligned row pitches, extreme values."""
    from scvi_tools_b200 import ops
    from scvi_tools_b200.data import narrowest_count_dtype

    rng = np.random.default_rng(G)
    hi = 256 if dt == "u8" else 65536
    npdt = np.uint8 if dt == "u8" else np.uint16
    X = rng.integers(0, hi, size=(257, G)).astype(npdt)
    X[0, :] = hi - 1
    X[1, :] = 0
    Xd = torch.from_numpy(X).to(dev())
    assert narrowest_count_dtype(torch.from_numpy(X.astype(np.float32))) == (torch.uint8 if dt == "u8" else torch.uint16)
    for B in (1, 77, 257):
        idx = rng.integers(0, X.shape[0], size=B)
        got = ops.gather_rows(Xd, torch.from_numpy(idx).to(dev()))
        assert got.dtype == torch.float32 and np.array_equal(got.cpu().numpy(), X[idx].astype(np.float32))
    assert np.array_equal(ops.gather_rows(Xd, None).cpu().numpy(), X.astype(np.float32))
    # a row-sliced view (rows 3..) and a column-sliced view (pitch > G, odd offsets -> scalar path)
    assert np.array_equal(ops.gather_rows(Xd[3:40], None).cpu().numpy(), X[3:40].astype(np.float32))
    if G > 8:
        assert np.array_equal(ops.gather_rows(Xd[:, 1 : G - 2], None).cpu().numpy(), X[:, 1 : G - 2].astype(np.float32))
    assert narrowest_count_dtype(torch.tensor([[0.5, 1.0]])) == torch.float32
    assert narrowest_count_dtype(torch.tensor([[70000.0]])) == torch.float32


@pytest.mark.parametrize(("G", "density", "vmax"), [(2000, 0.1, 40), (333, 0.5, 3000), (70000, 0.001, 5), (17, 0.0, 1)])
def test_csr_resident_gather_bit_exact(G, density, vmax):
    """CSR-resident counts densify on the device to exactly `X[idx].toarray().astype(float32)` (uint16 / int32
    column ids, uint8 / uint16 / float32 values, empty rows, empty matrix, unaligned row pitches)."""
    import scipy.sparse as sp

    from scvi_tools_b200 import ops
    from scvi_tools_b200.data import DeviceResidentCSR

    rng = np.random.default_rng(G)
    n = 300
    X = sp.random(n, G, density=density, format="csr", random_state=rng,
                  data_rvs=lambda k: rng.integers(1, vmax + 1, size=k).astype(np.float32))
    X = X.astype(np.float32)
    batch = rng.integers(0, 3, size=n)
    csr = DeviceResidentCSR(X, batch=batch, device=dev())
    assert csr.indices.dtype == (torch.uint16 if G <= 65536 else torch.int32)
    if X.nnz:  # an empty matrix keeps float32 values (nothing to narrow)
        assert csr.values.dtype == (torch.uint8 if vmax < 256 else torch.uint16)
    dense = np.asarray(X.toarray(), dtype=np.float32)
    for B in (1, 77, 300):
        idx = rng.integers(0, n, size=B)
        mb = csr.minibatch(torch.from_numpy(idx))
        assert mb["X"].dtype == torch.float32 and np.array_equal(mb["X"].cpu().numpy(), dense[idx])
        assert np.array_equal(mb["batch"].cpu().numpy().ravel(), batch[idx])
    assert np.array_equal(csr.densify(None).cpu().numpy(), dense)
    # into a padded staging buffer (row pitch > G, odd pitch -> scalar zero fill); stale contents are overwritten
    stage = torch.full((64, G + 3), 7.0, device=dev())
    idx = torch.from_numpy(rng.integers(0, n, size=64)).to(dev())
    csr.densify(idx, out=stage[:, :G])
    assert np.array_equal(stage[:, :G].cpu().numpy(), dense[idx.cpu().numpy()]) and bool((stage[:, G:] == 7.0).all())
    # non-integer values stay float32
    Xf = X.copy()
    Xf.data = Xf.data + 0.5
    if Xf.nnz:
        cf = DeviceResidentCSR(Xf, device=dev())
        assert cf.values.dtype == torch.float32
        assert np.array_equal(cf.densify(None).cpu().numpy(), np.asarray(Xf.toarray(), dtype=np.float32))
    with pytest.raises(ValueError):
        DeviceResidentCSR((np.array([0, 1]), np.array([G]), np.array([1.0]), (1, G)), device=dev())


@pytest.mark.parametrize(("B", "G", "nb", "lik"), [(300, 2000, 8, "zinb"), (129, 4000, 32, "nb"), (64, 512, 2, "zinb")])
def test_vae_step_vs_oracle_seeded(B, G, nb, lik):
    """BASELINE.json configs[1]/[2] shapes (G=2000 zinb 8 batches, G=4000 nb 32 batches) at an
    oracle-sized B, against the oracle run on the same seeded inputs."""
    import scvi_tools_b200 as S

    X, batch = O.synthetic_iid(max(B, nb * 40), G, nb, seed=B)
    torch.manual_seed(B)
    m = S.VAE(n_input=G, n_batch=nb, gene_likelihood=lik)
    P = {k: v.detach().clone() for k, v in m.state_dict().items()}
    m = m.to(dev()).train()
    rng = np.random.default_rng(1)
    idx = rng.permutation(X.shape[0])[:B]
    x = torch.from_numpy(X[idx])
    bi = torch.from_numpy(batch[idx]).view(-1, 1)
    g = torch.Generator().manual_seed(2)
    mask = (torch.rand(B, 128, generator=g) >= 0.1).to(torch.uint8)
    noise = torch.randn(B, 10, generator=g)
    t = {"X": x.to(dev()), "batch": bi.to(dev()), "labels": torch.zeros(B, 1, dtype=torch.long, device=dev())}
    with S.rng_feed(drop_masks=[mask], noise=[noise]):
        _, _, lo = m(t, loss_kwargs={"kl_weight": 0.7})
    lo.loss.backward()
    torch.set_num_threads(8)
    o, grads = O.loss_and_grads(P, x, bi, n_batch=nb, gene_likelihood=lik, kl_weight=0.7, drop_mask=mask, noise=noise)
    assert abs(float(lo.loss) - float(o["loss"])) <= ELBO_RTOL * abs(float(o["loss"]))
    assert max_rel(lo.reconstruction_loss["reconstruction_loss"].detach().cpu(), o["reconst"].detach()) < ELBO_RTOL
    gs = max(float(v.abs().max()) for v in grads.values())
    named = dict(m.named_parameters())
    for k in O.trained_keys(lik):
        assert grad_close(named[k].grad.cpu(), grads[k], GRAD_RTOL, gs), k
