"""GPU: whole-step CUDA graphs of the autograd-driven plans (scvi_tools_b200/graphed.py).

A captured step must be the eager step: same kernels, same order, device-resident KL weight and Adam counters.
torch's Philox offsets differ between eager launches and graph replays, so for this comparison the dropout masks
and the reparameterisation noise are replaced (monkeypatch) by fixed, capture-safe functions of the shape; the two
plans then have to end with identical parameters after a run that changes the KL weight every step and includes a
ragged minibatch (which runs eagerly between replays).
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture
def fixed_rng(monkeypatch):
    from scvi_tools_b200 import nn as N

    def mask(shape, p, device):
        n = int(np.prod(shape))
        return ((torch.arange(n, device=device) * 7919) % 10 != 0).to(torch.uint8).view(shape)

    def noise(shape, device):
        n = int(np.prod(shape))
        return torch.sin(torch.arange(n, device=device, dtype=torch.float32) * 0.37).view(shape)

    import scvi_tools_b200.module as M_
    import scvi_tools_b200.totalvi as T_

    monkeypatch.setattr(N, "_next_mask", mask)
    monkeypatch.setattr(N, "_next_noise", noise)
    monkeypatch.setattr(M_, "_next_noise", noise)  # bound by name at import
    monkeypatch.setattr(T_, "_next_noise", noise)


def _close(a, b, what):
    assert torch.allclose(a, b, rtol=1e-6, atol=1e-7), (what, float((a - b).abs().max()))


def test_semisupervised_graph_equals_eager(fixed_rng):
    from scvi_tools_b200.scanvi import SCANVAE
    from scvi_tools_b200.semisupervised import SemiSupervisedTrainingPlan

    dev = torch.device("cuda:0")
    rng = np.random.default_rng(0)
    n, G, L, B = 2048, 96, 5, 128
    X = torch.from_numpy(rng.poisson(3.0, size=(n, G)).astype(np.float32)).to(dev)
    lab = torch.from_numpy(rng.integers(0, L, size=(n, 1))).to(dev)
    bat = torch.from_numpy(rng.integers(0, 2, size=(n, 1))).to(dev)

    def mb(lo, hi):
        return {"X": X[lo:hi], "batch": bat[lo:hi], "labels": lab[lo:hi]}

    # 10 full-shape steps with one ragged step in the middle
    sched = [(i * B, i * B + B, (i + 3) * B, (i + 3) * B + 64) for i in range(5)] + [(0, 77, 300, 340)] + \
            [(i * B, i * B + B, (i + 2) * B, (i + 2) * B + 64) for i in range(5, 10)]
    end = []
    for graph in (False, True):
        torch.manual_seed(0)
        m = SCANVAE(n_input=G, n_batch=2, n_labels=L, gene_likelihood="zinb").to(dev)
        plan = SemiSupervisedTrainingPlan(m, n_classes=L, n_epochs_kl_warmup=20, use_cuda_graph=graph)
        losses = []
        for t, (a, b, c, d) in enumerate(sched):
            plan.current_epoch = t
            lo = plan.training_step((mb(a, b), mb(c, d)))
            losses.append(lo.loss.detach().clone())
        torch.cuda.synchronize()
        if graph:
            assert plan._graph.n_replays == 7 and plan._graph.n_eager == 4, (plan._graph.n_replays, plan._graph.n_eager)
        assert plan.optimizer.n_steps == len(sched)
        end.append((torch.stack(losses), plan.optimizer.arena["p"].clone(), plan.optimizer.arena["m"].clone(),
                    {k: v.clone() for k, v in m.state_dict().items() if "running" in k}))
    _close(end[0][0], end[1][0], "losses")
    _close(end[0][1], end[1][1], "parameters")
    _close(end[0][2], end[1][2], "first moments")
    for k in end[0][3]:
        _close(end[0][3][k], end[1][3][k], k)
    assert torch.isfinite(end[1][0]).all()


def test_adversarial_graph_equals_eager(fixed_rng):
    from scvi_tools_b200.adversarial import AdversarialTrainingPlan
    from scvi_tools_b200.totalvi import TOTALVAE

    dev = torch.device("cuda:0")
    rng = np.random.default_rng(1)
    n, G, P, nb, B = 1024, 80, 10, 3, 64
    X = torch.from_numpy(rng.poisson(3.0, size=(n, G)).astype(np.float32)).to(dev)
    Y = torch.from_numpy(rng.poisson(30.0, size=(n, P)).astype(np.float32)).to(dev)
    bat = torch.from_numpy(rng.integers(0, nb, size=(n, 1))).to(dev)
    lab = torch.zeros(n, 1, dtype=torch.long, device=dev)
    end = []
    for graph in (False, True):
        torch.manual_seed(0)
        m = TOTALVAE(n_input_genes=G, n_input_proteins=P, n_batch=nb, n_panel=nb, n_hidden=64).to(dev)
        torch.manual_seed(1)
        plan = AdversarialTrainingPlan(m, lr=4e-3, adversarial_classifier=True, n_epochs_kl_warmup=16,
                                       use_cuda_graph=graph)
        tr = []
        for t in range(9):
            plan.current_epoch = t
            s = slice(t * B, t * B + B)
            plan.training_step({"X": X[s], "proteins": Y[s], "batch": bat[s], "labels": lab[s]})
            tr.append(torch.stack([plan.last["train_loss"], plan.last["fool_loss"], plan.last["classifier_loss"]]).clone())
        torch.cuda.synchronize()
        if graph:
            assert plan._graph.n_replays == 6
        assert plan.opt1.n_steps == plan.opt2.n_steps == 9
        end.append((torch.stack(tr), plan.opt1.arena["p"].clone(), plan.opt2.arena["p"].clone()))
    _close(end[0][0], end[1][0], "loss pieces")
    _close(end[0][1], end[1][1], "module parameters")
    _close(end[0][2], end[1][2], "classifier parameters")
