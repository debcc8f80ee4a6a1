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


def test_adversarial_plan_three_steps_match_reference_trajectory():
    """`AdversarialTrainingPlan.training_step` x 3 (ELBO + kappa * fool loss -> module Adam lr 4e-3; kappa * true-batch
    cross-entropy on z.detach() -> classifier Adam lr 1e-3) against the unmodified reference TOTALVAE + Classifier
    stepped with two torch.optim.Adam (oracle/make_golden_totalvi.py trajectory)."""
    import json

    import scvi_tools_b200 as S
    from scvi_tools_b200.adversarial import AdversarialTrainingPlan

    c = load_option_case("totalvi_adversarial_trajectory")
    dev = torch.device("cuda:0")
    m = _module(c, dev).train()
    nb = c["kwargs"]["n_batch"]
    clf = S.Classifier(n_input=m.n_latent, n_hidden=32, n_labels=nb, n_layers=2, logits=True)
    r = clf.load_state_dict({k[5:]: torch.from_numpy(v) for k, v in c["extra"].items() if k.startswith("clf::")})
    assert not r.missing_keys and not r.unexpected_keys
    clf.to(dev)
    plan = AdversarialTrainingPlan(m, lr=4e-3, adversarial_classifier=clf, n_epochs_kl_warmup=400)
    out = c["out"]
    n_steps, e0 = int(c["extra"]["n_steps"]), int(c["extra"]["first_epoch"])
    with S.rng_feed(drop_masks=c["masks"], noise=c["noises"]):
        for t in range(n_steps):
            plan.current_epoch = e0 + t
            b = {k.split("::", 1)[1]: v.to(dev) for k, v in c["tensors"].items() if k.startswith(f"s{t}::")}
            lo = plan.training_step(b)
            for got, key in ((lo.loss, "elbo"), (plan.last["fool_loss"], "fool"), (plan.last["train_loss"], "train_loss"),
                             (plan.last["classifier_loss"], "adv")):
                want = float(out[f"{key}{t}"])
                assert abs(float(got) - want) <= 2e-5 * abs(want), (t, key, float(got), want)
            assert abs(plan.last["kappa"] - (1 - (e0 + t) / 400.0)) < 1e-12
    trained = set(json.loads(str(c["extra"]["trained"])))
    assert set(plan.opt1.names) == trained
    assert set(plan.opt2.names) == {k for k, _ in clf.named_parameters()}
    for sd, p0, prefix in ((m.state_dict(), c["params"], "param_after::"),
                           (clf.state_dict(), {k[5:]: torch.from_numpy(v) for k, v in c["extra"].items()
                                               if k.startswith("clf::")}, "clf_after::")):
        for k, v0 in p0.items():
            after = np.asarray(out[prefix + k])
            if after.dtype.kind != "f":
                assert int(sd[k]) == int(after), k
                continue
            upd_ref = after - v0.numpy()
            upd = sd[k].cpu().numpy() - v0.numpy()
            assert np.linalg.norm(upd - upd_ref) <= 3e-3 * np.linalg.norm(upd_ref) + 1e-6 * np.sqrt(upd.size), (prefix, k)
            assert rel_err(sd[k].cpu().float(), torch.from_numpy(after).float()) < 2e-5 or np.abs(after).max() == 0, k


def test_adversarial_plan_fit_mixes_batches_and_schedules_lr():
    """End to end on synthetic two-batch data with a batch-specific shift: training lowers the ELBO; the classifier is
    disabled (with the reference's warning) for one batch; the plateau scheduler lowers optimizer 1's rate when the
    monitored metric stops improving."""
    from scvi_tools_b200.adversarial import AdversarialTrainingPlan, _Plateau
    from scvi_tools_b200.optim import ArenaAdam
    from scvi_tools_b200.totalvi import TOTALVAE

    rng = np.random.default_rng(0)
    n, G, P, nb, B = 1024, 80, 8, 2, 256
    batch = (np.arange(n) * nb // n).astype(np.int64)
    X = rng.poisson(rng.gamma(2.0, 1.0, size=(1, G)) * (1.0 + 0.5 * batch[:, None])).astype(np.float32)
    Y = rng.poisson(20.0 * rng.gamma(2.0, 1.0, size=(1, P)) * np.ones((n, 1))).astype(np.float32)
    dev = torch.device("cuda:0")
    Xd, Yd, bd = torch.from_numpy(X).to(dev), torch.from_numpy(Y).to(dev), torch.from_numpy(batch).to(dev).view(-1, 1)
    torch.manual_seed(0)
    m = TOTALVAE(n_input_genes=G, n_input_proteins=P, n_batch=nb, n_panel=nb, n_hidden=64).to(dev)
    plan = AdversarialTrainingPlan(m, lr=4e-3, adversarial_classifier=True, reduce_lr_on_plateau=True,
                                   lr_scheduler_metric="elbo_train", n_epochs_kl_warmup=20)
    assert plan.opt2 is not None and plan.adversarial_classifier is not False
    g = torch.Generator().manual_seed(1)

    def epoch():
        perm = torch.randperm(n, generator=g).to(dev)
        for s in range(0, n, B):
            ix = perm[s : s + B]
            yield {"X": Xd[ix], "proteins": Yd[ix], "batch": bd[ix], "labels": torch.zeros(len(ix), 1, dtype=torch.long, device=dev)}

    hist = plan.fit(epoch, max_epochs=8)
    assert all(np.isfinite(h["train_loss"]) for h in hist)
    assert hist[-1]["reconstruction_loss_train"] < hist[0]["reconstruction_loss_train"]
    assert plan.global_step == 8 * (n // B) and plan.opt1.arena["step"] == plan.opt2.arena["step"] == plan.global_step
    with pytest.warns(UserWarning, match="Disabling adversarial classifier"):
        p1 = AdversarialTrainingPlan(TOTALVAE(n_input_genes=G, n_input_proteins=P, n_batch=1, n_panel=1).to(dev),
                                     adversarial_classifier=True)
    assert p1.adversarial_classifier is False and p1.opt2 is None
    opt = ArenaAdam(m.named_parameters, lr=1.0)
    sch = _Plateau(opt, factor=0.6, patience=2, threshold=0.0, min_lr=0.3)
    for metric in (5.0, 4.0, 4.0, 4.0):
        sch.step(metric)
    assert opt.lr == 1.0
    sch.step(4.0)  # third epoch in a row without improvement > patience 2
    assert abs(opt.lr - 0.6) < 1e-12
    for _ in range(6):
        sch.step(4.0)
    assert abs(opt.lr - 0.36) < 1e-12 or abs(opt.lr - 0.3) < 1e-12
