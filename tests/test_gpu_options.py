"""GPU parity of non-default `VAE` options against vectors recorded from the
unmodified reference (oracle/make_golden_options.py): dispersion variants, encoded library size, encoded /
injected covariates, deeper stacks, LayerNorm, raw-count encoder input.  Same tolerances as the default path:
ELBO 1e-5 relative, gradients 1e-4 relative."""
import json
import os

import numpy as np
import pytest
import torch

from tests._golden import GOLD, grad_close, max_rel, rel_err

pytestmark = pytest.mark.gpu

CASES = ["opt_genebatch_zinb", "opt_genelabel_nb", "opt_genecell_zinb", "opt_libenc_nb", "opt_covs_zinb",
         "opt_layernorm_nb", "opt_deep_noinject_zinb", "opt_nologvar_nb", "opt_embedding_zinb", "opt_embedding_enc_nb"]


def load_option_case(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    case = {"kwargs": json.loads(str(z["in::kwargs"])), "tensors": {}, "masks": {}, "noises": {}, "params": {},
            "out": {}, "extra": {}}
    for k in z.files:
        if k.startswith("in::t::"):
            case["tensors"][k[7:]] = torch.from_numpy(z[k])
        elif k.startswith("in::mask::"):
            case["masks"][int(k[10:])] = torch.from_numpy(z[k])
        elif k.startswith("in::noise::"):
            case["noises"][int(k[11:])] = torch.from_numpy(z[k])
        elif k.startswith("in::param::"):
            case["params"][k[11:]] = torch.from_numpy(z[k])
        elif k.startswith("out::"):
            case["out"][k[5:]] = z[k]
        elif k.startswith("in::") and k != "in::kwargs":
            case["extra"][k[4:]] = z[k]
    case["masks"] = [case["masks"][i] for i in sorted(case["masks"])]
    case["noises"] = [case["noises"][i] for i in sorted(case["noises"])]
    return case


@pytest.mark.parametrize("name", CASES)
def test_vae_option_step_vs_reference_golden(name):
    import scvi_tools_b200 as S

    c = load_option_case(name)
    dev = torch.device("cuda:0")
    kw = dict(c["kwargs"])
    for k in ("library_log_means", "library_log_vars"):
        if k in c["extra"]:
            kw[k] = c["extra"][k]
    m = S.VAE(**kw)
    missing = m.load_state_dict(c["params"], strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    m = m.to(dev).train()
    tensors = {k: v.to(dev) for k, v in c["tensors"].items()}
    with S.rng_feed(drop_masks=c["masks"], noise=c["noises"]):
        inf, gen, lo = m(tensors, loss_kwargs={"kl_weight": float(c["extra"]["kl_weight"])})
    lo.loss.backward()
    out = c["out"]
    assert abs(float(lo.loss) - float(out["loss"])) <= 1e-5 * abs(float(out["loss"]))
    assert max_rel(lo.reconstruction_loss["reconstruction_loss"].detach().cpu(), out["reconst"]) < 1e-5
    assert max_rel(lo.kl_local["kl_divergence_z"].detach().cpu(), out["kl_z"], floor=1e-2) < 1e-4
    assert max_rel(lo.kl_local["kl_divergence_l"].detach().cpu(), out["kl_l"], floor=1e-2) < 1e-4
    assert rel_err(inf["z"].detach().cpu(), out["z"]) < 1e-5
    assert rel_err(inf["library"].detach().cpu(), out["library"]) < 1e-5
    gs = max(float(abs(v).max()) for k, v in out.items() if k.startswith("grad::"))
    named = dict(m.named_parameters())
    n_checked = 0
    for k, p in named.items():
        if "grad::" + k in out:
            assert p.grad is not None, k
            assert grad_close(p.grad.cpu(), out["grad::" + k], 1e-4, gs), (k, rel_err(p.grad.cpu(), out["grad::" + k]))
            n_checked += 1
        else:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, k
    assert n_checked == sum(k.startswith("grad::") for k in out)
    sd = m.state_dict()
    for k in out:
        if k.startswith("buf_after::"):
            assert rel_err(sd[k[11:]].cpu(), out[k]) < 1e-5, k
