"""TEST INFRASTRUCTURE ONLY -- tests/golden/mixture_nb.npz: the reference's `log_mixture_nb` /
`NegativeBinomialMixture.log_prob` (src/scvi/distributions/_negative_binomial.py:138-216, :589-717; the TOTALVI
protein likelihood) and its gradients on seeded inputs, including extreme means / logits and both layouts of
the shared inverse dispersion ([P] and [B, P]); plus the two-dispersion branch.

    python oracle/make_golden_mixture.py      (build container only: needs /root/reference)
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

GOLD = os.path.join(ROOT, "tests", "golden")


def main():
    R = ref_shim.load()
    torch.set_num_threads(1)
    rng = np.random.default_rng(11)
    B, P = 70, 37
    x = (rng.negative_binomial(2, 0.05, size=(B, P)) * rng.binomial(1, 0.8, size=(B, P))).astype(np.float32)
    x[0, :4] = [0, 1, 3000, 65535]
    mu1 = np.exp(rng.normal(1.0, 1.5, size=(B, P))).astype(np.float32)           # background
    mu2 = (mu1 * (1 + np.exp(rng.normal(1.0, 1.0, size=(B, P))))).astype(np.float32)  # foreground = (1 + alpha) background
    mu1[1, :3] = [1e-10, 1e-6, 1e5]
    theta = np.exp(rng.normal(0.5, 1.2, size=(P,))).astype(np.float32)
    theta[:3] = [1e-2, 1.0, 300.0]
    theta_bp = np.exp(rng.normal(0.5, 1.0, size=(B, P))).astype(np.float32)
    theta2 = np.exp(rng.normal(0.0, 1.0, size=(P,))).astype(np.float32)
    pi = rng.normal(0.0, 3.0, size=(B, P)).astype(np.float32)
    pi[2, :4] = [-50.0, 50.0, -90.0, 90.0]
    w = rng.normal(size=(B, P)).astype(np.float32)
    save = dict(x=x, mu1=mu1, mu2=mu2, theta=theta, theta_bp=theta_bp, theta2=theta2, pi=pi, w=w)
    tx, tw = torch.from_numpy(x), torch.from_numpy(w)
    for tag, th in (("p", theta), ("bp", theta_bp)):
        t1, t2, tt, tp = (torch.from_numpy(a).clone().requires_grad_(True) for a in (mu1, mu2, th, pi))
        out = R.log_mixture_nb(tx, t1, t2, tt, None, tp)
        dist = R.NegativeBinomialMixture(mu1=t1, mu2=t2, theta1=tt, mixture_logits=tp).log_prob(tx)
        assert torch.equal(out, dist)
        (out * tw).sum().backward()
        save.update({f"out_{tag}": out.detach().numpy(), f"dmu1_{tag}": t1.grad.numpy(), f"dmu2_{tag}": t2.grad.numpy(),
                     f"dtheta_{tag}": tt.grad.numpy(), f"dpi_{tag}": tp.grad.numpy()})
    out2 = R.log_mixture_nb(tx, torch.from_numpy(mu1), torch.from_numpy(mu2), torch.from_numpy(theta),
                            torch.from_numpy(theta2), torch.from_numpy(pi))
    save["out_two_theta"] = out2.numpy()
    path = os.path.join(GOLD, "mixture_nb.npz")
    np.savez_compressed(path, **save)
    print("mixture_nb ->", path, f"({os.path.getsize(path)/1e3:.0f} kB)", float(save["out_p"].mean()))


if __name__ == "__main__":
    main()
