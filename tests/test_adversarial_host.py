"""Host logic of the adversarial plan (no GPU): the plateau scheduler against torch's ReduceLROnPlateau, the fool /
true-class targets of `loss_adversarial_classifier` against the reference's formula (_trainingplans.py:652-668)."""
import random

import torch
import torch.nn.functional as F

from scvi_tools_b200.adversarial import AdversarialTrainingPlan, _Plateau


def test_plateau_matches_torch_scheduler():
    class Opt:
        lr = 1.0

    o = Opt()
    s = _Plateau(o, factor=0.6, patience=2, threshold=0.0, min_lr=0.3)
    p = torch.nn.Parameter(torch.zeros(1))
    to = torch.optim.SGD([p], lr=1.0)
    ts = torch.optim.lr_scheduler.ReduceLROnPlateau(to, patience=2, factor=0.6, threshold=0.0, min_lr=0.3,
                                                    threshold_mode="abs")
    random.seed(0)
    for i in range(40):
        metric = 5.0 - 0.1 * i if i < 8 else 4.2 + random.random() * 0.01
        s.step(metric)
        ts.step(metric)
        assert abs(o.lr - to.param_groups[0]["lr"]) < 1e-12, i


def test_adversarial_targets():
    class Plan(AdversarialTrainingPlan):
        def __init__(self, clf, n):  # no module, no CUDA: only the loss formula
            self.adversarial_classifier, self.n_output_classifier = clf, n

    torch.manual_seed(0)
    lin = torch.nn.Linear(5, 4)
    plan = Plan(lin, 4)
    z = torch.randn(7, 5)
    b = torch.randint(0, 4, (7, 1))
    logp = F.log_softmax(lin(z), dim=1)
    true = -logp.gather(1, b).mean()
    assert torch.allclose(plan.loss_adversarial_classifier(z, b, True), true, atol=1e-6)
    oh = F.one_hot(b.squeeze(-1), 4).bool()
    fool = -((logp * (~oh).float() / 3).sum(1)).mean()
    assert torch.allclose(plan.loss_adversarial_classifier(z, b, False), fool, atol=1e-6)
