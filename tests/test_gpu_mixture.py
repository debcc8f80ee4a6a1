"""GPU parity of the NB-mixture likelihood (TOTALVI proteins) against vectors from the
reference's `log_mixture_nb` / `NegativeBinomialMixture.log_prob` (oracle/make_golden_mixture.py).  Tolerance:
1e-5 relative to the magnitude of the terms that are summed (fp32), gradients 1e-4 norm-wise."""
import os
import warnings

import numpy as np
import pytest
import torch

from tests._golden import GOLD, rel_err

pytestmark = pytest.mark.gpu


def _load():
    z = np.load(os.path.join(GOLD, "mixture_nb.npz"))
    return {k: z[k] for k in z.files}


def _magnitude(g, theta):
    x, m1, m2, pi = (torch.from_numpy(g[k]).double() for k in ("x", "mu1", "mu2", "pi"))
    t = torch.from_numpy(theta).double()
    M = 0
    for m in (m1, m2):
        l_tm = torch.log(t + m + 1e-8)
        M = M + (t * (torch.log(t + 1e-8) - l_tm)).abs() + (x * (torch.log(m + 1e-8) - l_tm)).abs()
    return (M + torch.lgamma(x + t).abs() + torch.lgamma(t).abs() + torch.lgamma(x + 1) + pi.abs() + 1.0).numpy()


@pytest.mark.parametrize("tag", ["p", "bp"])
def test_mixture_nb_fwd_bwd_vs_reference_golden(tag):
    import scvi_tools_b200 as S

    g = _load()
    dev = torch.device("cuda:0")
    theta = g["theta"] if tag == "p" else g["theta_bp"]
    x, w = torch.from_numpy(g["x"]).to(dev), torch.from_numpy(g["w"]).to(dev)
    m1, m2, th, pi = (torch.from_numpy(a).to(dev).requires_grad_(True) for a in (g["mu1"], g["mu2"], theta, g["pi"]))
    out = S.log_mixture_nb(x, m1, m2, th, None, pi)
    err = np.abs(out.detach().cpu().numpy().astype(np.float64) - g[f"out_{tag}"].astype(np.float64))
    assert float((err / _magnitude(g, theta)).max()) < 1e-5
    (out * w).sum().backward()
    for name, t in (("dmu1", m1), ("dmu2", m2), ("dtheta", th), ("dpi", pi)):
        assert rel_err(t.grad.cpu(), g[f"{name}_{tag}"]) < 1e-4, name
    # the distribution object calls the same kernel (and agrees with the function bit for bit)
    with warnings.catch_warnings():
        warnings.simplefilter("error")
        d = S.NegativeBinomialMixture(mu1=m1.detach(), mu2=m2.detach(), theta1=th.detach(), mixture_logits=pi.detach())
        assert torch.equal(d.log_prob(x), out.detach())
        assert d.mean.shape == x.shape and d.sample().shape == x.shape
    with pytest.warns(UserWarning):
        d.log_prob(x - 0.5)


def test_mixture_nb_two_dispersions_and_broadcast():
    import scvi_tools_b200 as S

    g = _load()
    dev = torch.device("cuda:0")
    t = lambda k: torch.from_numpy(g[k]).to(dev)  # noqa: E731
    out = S.log_mixture_nb(t("x"), t("mu1"), t("mu2"), t("theta"), t("theta2"), t("pi"))
    err = np.abs(out.cpu().numpy().astype(np.float64) - g["out_two_theta"].astype(np.float64))
    assert float((err / _magnitude(g, g["theta"])).max()) < 1e-5
    # 3-D (n_samples, B, P) parameters against 2-D counts broadcast like the reference
    m1 = t("mu1").unsqueeze(0).expand(3, -1, -1)
    out3 = S.log_mixture_nb(t("x"), m1, t("mu2"), t("theta"), None, t("pi"))
    assert out3.shape == (3,) + tuple(g["x"].shape)
    assert torch.equal(out3[1], S.log_mixture_nb(t("x"), t("mu1"), t("mu2"), t("theta"), None, t("pi")))
    empty = S.log_mixture_nb(t("x")[:0], t("mu1")[:0], t("mu2")[:0], t("theta"), None, t("pi")[:0])
    assert empty.shape == (0, g["x"].shape[1])
