"""Helpers to load tests/golden/*.npz (made by oracle/make_golden.py from the unmodified reference)."""
import os

import numpy as np
import torch

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
VAE_CASES = ["c1_zinb", "c1_nb", "ragged_zinb", "onebatch_nb"]


def load_vae_case(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    P, inp, out = {}, {}, {}
    for k in z.files:
        kind, key = k.split("::", 1)
        if kind == "param":
            P[key] = torch.from_numpy(z[k])
        elif kind == "in":
            inp[key] = z[k]
        elif kind == "out":
            out[key] = z[k]
    inp["likelihood"] = str(inp["likelihood"])
    inp["n_batch"] = int(inp["n_batch"])
    inp["kl_weight"] = float(inp["kl_weight"])
    inp["n_adam_steps"] = int(inp["n_adam_steps"])
    return P, inp, out


def rel_err(a, b):
    a = torch.as_tensor(a, dtype=torch.float64).flatten()
    b = torch.as_tensor(b, dtype=torch.float64).flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def max_rel(a, b, floor=1e-6):
    a = torch.as_tensor(a, dtype=torch.float64)
    b = torch.as_tensor(b, dtype=torch.float64)
    return float(((a - b).abs() / b.abs().clamp_min(floor)).max())


def grad_close(a, b, rtol, scale=1.0, atol=1e-6):
    """||a-b|| <= rtol*||b|| + atol*scale*sqrt(n).

    `scale` is the largest gradient magnitude of the whole step; the absolute floor covers
    gradients that are mathematically zero (Linear biases feeding a BatchNorm) and therefore
    hold only rounding noise of that scale in the reference as well."""
    a = torch.as_tensor(a, dtype=torch.float64).flatten()
    b = torch.as_tensor(b, dtype=torch.float64).flatten()
    return float((a - b).norm()) <= rtol * float(b.norm()) + atol * scale * (a.numel() ** 0.5)


def grad_scale(out):
    return max(float(abs(v).max()) for k, v in out.items() if k.startswith("grad::"))
