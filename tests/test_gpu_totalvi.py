"""GPU parity of `TOTALVAE` against vectors recorded from the unmodified reference
module (oracle/make_golden_totalvi.py): one training-mode step -- gene NB/ZINB + protein NB-mixture likelihoods,
latent / library / background-prior KL terms, every gradient -- with identical weights, inputs, dropout masks and
normal draws.  Tolerances as the VAE path: loss 1e-5 relative, gradients 1e-4 relative."""
import numpy as np
import pytest
import torch

from tests._golden import grad_close, max_rel, rel_err
from tests.test_gpu_options import load_option_case

pytestmark = pytest.mark.gpu

CASES = ["totalvi_nb", "totalvi_zinb_mask_ln", "totalvi_libenc_default_width"]


def _module(c, dev):
    from scvi_tools_b200.totalvi import TOTALVAE

    kw = dict(c["kwargs"])
    for k in ("library_log_means", "library_log_vars"):
        if k in c["extra"]:
            kw[k] = c["extra"][k]
    mask = {k.split("::", 1)[1]: v for k, v in c["extra"].items() if k.startswith("protein_batch_mask::")}
    if mask:
        kw["protein_batch_mask"] = mask
    m = TOTALVAE(**kw)
    r = m.load_state_dict(c["params"], strict=True)
    assert not r.missing_keys and not r.unexpected_keys
    return m.to(dev)


@pytest.mark.parametrize("name", CASES)
def test_totalvi_step_vs_reference_golden(name):
    import scvi_tools_b200 as S

    c = load_option_case(name)
    dev = torch.device("cuda:0")
    m = _module(c, dev).train()
    tensors = {k: v.to(dev) for k, v in c["tensors"].items()}
    with S.rng_feed(drop_masks=c["masks"], noise=c["noises"]):
        inf, gen, lo = m(tensors, loss_kwargs={"kl_weight": float(c["extra"]["kl_weight"])})
    lo.loss.backward()
    out = c["out"]
    assert abs(float(lo.loss.detach()) - float(out["loss"])) <= 1e-5 * abs(float(out["loss"]))
    for k, v in lo.reconstruction_loss.items():
        assert max_rel(v.detach().cpu(), out["rec::" + k], floor=1.0) < 1e-5, k
    for k, v in lo.kl_local.items():
        assert max_rel(v.detach().cpu(), out["kl::" + k], floor=1e-2) < 1e-4, k
    assert rel_err(inf["z"].detach().cpu(), out["z"]) < 1e-5
    assert rel_err(inf["library_gene"].detach().cpu(), out["library_gene"]) < 1e-5
    assert rel_err(gen["px_"]["rate"].detach().cpu(), out["px_rate"]) < 1e-5
    assert rel_err(gen["py_"]["rate_fore"].detach().cpu(), out["py_rate_fore"]) < 1e-5
    assert rel_err(gen["py_"]["mixing"].detach().cpu(), out["py_mixing"]) < 1e-5
    assert rel_err(gen["py_"]["scale"].detach().cpu(), out["py_scale"]) < 1e-5
    gs = max(float(abs(v).max()) for k, v in out.items() if k.startswith("grad::"))
    n_checked = 0
    for k, p in m.named_parameters():
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


def test_totalvi_eval_forward_runs_and_samples():
    c = load_option_case("totalvi_nb")
    dev = torch.device("cuda:0")
    m = _module(c, dev).eval()
    with torch.no_grad():  # CPU tensors are moved by auto_move_data in eval mode
        inf, gen, lo = m(c["tensors"], loss_kwargs={"kl_weight": 1.0})
        inf3, gen3 = m(c["tensors"], inference_kwargs={"n_samples": 3}, compute_loss=False)
    assert np.isfinite(float(lo.loss))
    B = c["tensors"]["X"].shape[0]
    assert inf3["z"].shape == (3, B, c["kwargs"]["n_latent"])
    assert gen3["py_"]["rate_fore"].shape == (3, B, c["kwargs"]["n_input_proteins"])
