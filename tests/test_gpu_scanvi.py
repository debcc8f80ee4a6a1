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


def test_semisupervised_plan_three_steps_match_reference_trajectory():
    """`SemiSupervisedTrainingPlan.training_step` x 3 (full + labelled minibatch, classification_ratio 50, KL warm-up
    epoch / 400, the library's multi-tensor Adam over the parameters that receive gradient) against the unmodified
    reference module stepped with torch.optim.Adam (oracle/make_golden_scanvi.py trajectory)."""
    import json

    import scvi_tools_b200 as S
    from scvi_tools_b200.semisupervised import SemiSupervisedTrainingPlan

    c = load_option_case("scanvi_trajectory")
    dev = torch.device("cuda:0")
    m = _module(c, dev).train()
    n_steps = int(c["extra"]["n_steps"])
    plan = SemiSupervisedTrainingPlan(m, n_classes=c["kwargs"]["n_labels"], classification_ratio=50,
                                      n_epochs_kl_warmup=400)
    out = c["out"]
    with S.rng_feed(drop_masks=c["masks"], noise=c["noises"]):
        for t in range(n_steps):
            plan.current_epoch = t
            full = {k.split("::", 2)[2]: v.to(dev) for k, v in c["tensors"].items() if k.startswith(f"s{t}::full::")}
            lab = {k.split("::", 2)[2]: v.to(dev) for k, v in c["tensors"].items() if k.startswith(f"s{t}::lab::")}
            lo = plan.training_step((full, lab))
            assert abs(float(lo.loss) - float(out[f"loss{t}"])) <= 2e-5 * abs(float(out[f"loss{t}"])), t
            ce = float(out[f"classification_loss{t}"])
            assert abs(float(lo.classification_loss) - ce) <= 2e-5 * abs(ce), t
    trained = set(json.loads(str(c["extra"]["trained"])))
    assert set(plan.optimizer.names) == trained
    sd = m.state_dict()
    for k, v0 in c["params"].items():
        after = out["param_after::" + k]
        if k in trained:
            upd_ref = after - v0.numpy()
            upd = sd[k].cpu().numpy() - v0.numpy()
            assert np.linalg.norm(upd - upd_ref) <= 2e-3 * np.linalg.norm(upd_ref) + 1e-6 * np.sqrt(upd.size), k
        assert rel_err(sd[k].cpu().float(), torch.from_numpy(np.asarray(after)).float()) < 2e-5 or np.abs(after).max() == 0, k


def test_semisupervised_fit_learns_labels_on_synthetic_data():
    """Splitter -> loader -> plan end to end on device-resident counts: the classifier learns labels that are
    recoverable from the counts, validation accuracy beats chance by a wide margin."""
    from scvi_tools_b200.data import DeviceResidentCounts
    from scvi_tools_b200.scanvi import SCANVAE
    from scvi_tools_b200.semisupervised import SemiSupervisedDataSplitter, SemiSupervisedTrainingPlan

    rng = np.random.default_rng(0)
    n, G, n_labels = 3000, 120, 4
    y = rng.integers(0, n_labels, size=n)
    prof = rng.gamma(2.0, 1.0, size=(n_labels, G))
    X = rng.poisson(prof[y] * 3.0).astype(np.float32)
    labels = y.copy()
    labels[rng.random(n) < 0.6] = n_labels  # 60 % of the cells carry the "unlabeled" code
    dev = torch.device("cuda:0")
    data = DeviceResidentCounts(X, batch=np.zeros(n, dtype=np.int64), labels=labels, device=dev)
    split = SemiSupervisedDataSplitter(labels, n_labels, train_size=0.8, batch_size=128, seed=1)
    assert len(np.intersect1d(split.train_idx, split.val_idx)) == 0
    torch.manual_seed(0)
    m = SCANVAE(n_input=G, n_batch=1, n_labels=n_labels, gene_likelihood="nb").to(dev)
    plan = SemiSupervisedTrainingPlan(m, n_classes=n_labels)
    tr, va = split.train_dataloader(data), split.val_dataloader(data)
    hist = plan.fit(tr, max_epochs=6, val_loader=va)
    assert np.isfinite(hist[-1]["train_loss"]) and hist[-1]["train_loss"] < hist[0]["train_loss"]
    assert hist[-1]["validation_accuracy"] > 0.8, hist[-1]
    # predictions on unlabelled training cells agree with the hidden truth
    un = np.where(labels == n_labels)[0][:512]
    m.eval()
    with torch.no_grad():
        t = data.minibatch(torch.as_tensor(un, device=dev))
        pred = m.classify(t["X"], batch_index=t["batch"]).argmax(-1).cpu().numpy()
    assert (pred == y[un]).mean() > 0.8
