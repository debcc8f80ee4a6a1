"""GPU parity of `SCANVAE` against vectors recorded from the unmodified reference
module (oracle/make_golden_scanvi.py): one semi-supervised training step (ELBO of the M1+M2 model + weighted
cross-entropy on the labelled minibatch) with identical weights, inputs, dropout masks and noise; and the
eval-mode `classify` path.  Tolerances as the VAE path: loss 1e-5 relative, gradients 1e-4 relative."""
import numpy as np
import pytest
import torch

from tests._golden import grad_close, max_rel, rel_err
from tests.test_gpu_options import load_option_case

pytestmark = pytest.mark.gpu

CASES = ["scanvi_zinb_labelled", "scanvi_nb_unlabelled", "scanvi_linear_classifier"]


def _module(c, dev):
    from scvi_tools_b200.scanvi import SCANVAE

    m = SCANVAE(**c["kwargs"])
    r = m.load_state_dict(c["params"], strict=True)
    assert not r.missing_keys and not r.unexpected_keys
    return m.to(dev)


@pytest.mark.parametrize("name", CASES)
def test_scanvi_step_vs_reference_golden(name):
    import scvi_tools_b200 as S

    c = load_option_case(name)
    dev = torch.device("cuda:0")
    m = _module(c, dev).train()
    full = {k: v.to(dev) for k, v in c["tensors"].items() if not k.startswith("lab::")}
    lab = {k[5:]: v.to(dev) for k, v in c["tensors"].items() if k.startswith("lab::")} or None
    with S.rng_feed(drop_masks=c["masks"], noise=c["noises"]):
        inf, gen, lo = m(full, loss_kwargs={"kl_weight": float(c["extra"]["kl_weight"]), "labelled_tensors": lab,
                                           "classification_ratio": float(c["extra"]["classification_ratio"])})
    lo.loss.backward()
    out = c["out"]
    assert abs(float(lo.loss) - float(out["loss"])) <= 1e-5 * abs(float(out["loss"]))
    assert max_rel(lo.reconstruction_loss["reconstruction_loss"].detach().cpu(), out["reconst"]) < 1e-5
    assert max_rel(lo.kl_local["kl_local"].detach().cpu(), out["kl_local"], floor=1e-2) < 1e-4
    assert rel_err(inf["z"].detach().cpu(), out["z"]) < 1e-5
    if lab is not None:
        assert abs(float(lo.classification_loss) - float(out["classification_loss"])) <= 1e-5 * abs(float(out["classification_loss"]))
        assert rel_err(lo.logits.detach().cpu(), out["logits"]) < 1e-5
        assert np.array_equal(lo.true_labels.cpu().numpy(), out["true_labels"])
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


@pytest.mark.parametrize("name", CASES)
def test_scanvi_classify_eval_vs_reference_golden(name):
    c = load_option_case(name)
    dev = torch.device("cuda:0")
    m = _module(c, dev).eval()
    with torch.no_grad():  # CPU inputs: auto_move_data moves them in eval mode
        logits = m.classify(c["tensors"]["X"], batch_index=c["tensors"]["batch"])
    assert rel_err(logits.cpu(), c["out"]["eval_classify_logits"]) < 1e-5
    assert np.array_equal(logits.argmax(-1).cpu().numpy(), c["out"]["eval_classify_logits"].argmax(-1))
