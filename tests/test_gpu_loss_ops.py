"""The fused per-row loss reductions (scvi_normal_logprob_sum_*, scvi_categorical_kl) and the gradient gather of
`ArenaAdam` against torch.distributions / torch.optim.Adam on the same inputs (fp32; tolerance 1e-6 relative on the
values, 1e-5 on the gradients -- the chain rule differs only in rounding order)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _close(a, b, rtol, atol=1e-6):
    return torch.allclose(a, b, rtol=rtol, atol=atol)


@pytest.mark.parametrize("shape", [(1024, 10), (3, 257, 7), (1, 1)])
@pytest.mark.parametrize("is_var", [False, True])
def test_normal_log_prob_sum(shape, is_var):
    from torch.distributions import Normal

    from scvi_tools_b200 import ops

    g = torch.Generator(device="cuda").manual_seed(0)
    mk = lambda: torch.randn(shape, device="cuda", generator=g)  # noqa: E731
    v, m = mk().requires_grad_(), mk().requires_grad_()
    s = (mk().abs() + 0.05).requires_grad_()
    w = torch.randn(shape[:-1], device="cuda", generator=g)
    ref = Normal(m, torch.sqrt(s) if is_var else s).log_prob(v).sum(-1)
    gr = torch.autograd.grad((ref * w).sum(), [v, m, s])
    got = ops.normal_log_prob_sum(v, m, var=s) if is_var else ops.normal_log_prob_sum(v, m, scale=s)
    gg = torch.autograd.grad((got * w).sum(), [v, m, s])
    assert got.shape == ref.shape and _close(got, ref, 1e-6, 1e-5)
    for a, b in zip(gg, gr, strict=True):
        assert _close(a, b, 1e-5, 1e-5)


def test_normal_log_prob_sum_broadcasts():
    from torch.distributions import Normal

    from scvi_tools_b200 import ops

    g = torch.Generator(device="cuda").manual_seed(1)
    v = torch.randn(4, 64, 10, device="cuda", generator=g).requires_grad_()
    m = torch.randn(64, 10, device="cuda", generator=g).requires_grad_()
    s = (torch.rand(64, 10, device="cuda", generator=g) + 0.1).requires_grad_()
    ref = Normal(m, s).log_prob(v).sum(-1)
    got = ops.normal_log_prob_sum(v, m, scale=s)
    gr = torch.autograd.grad(ref.sum(), [v, m, s])
    gg = torch.autograd.grad(got.sum(), [v, m, s])
    assert _close(got, ref, 1e-6, 1e-5)
    for a, b in zip(gg, gr, strict=True):
        assert a.shape == b.shape and _close(a, b, 1e-5, 1e-5)


@pytest.mark.parametrize("shape", [(1024, 20), (2, 33, 5)])
@pytest.mark.parametrize("zeros", [False, True])
def test_categorical_kl(shape, zeros):
    from torch.distributions import Categorical
    from torch.distributions import kl_divergence as kl

    from scvi_tools_b200 import ops

    g = torch.Generator(device="cuda").manual_seed(2)
    logits = (3 * torch.randn(shape, device="cuda", generator=g)).requires_grad_()
    C = shape[-1]
    prior = torch.rand(1, C, device="cuda", generator=g) + 0.1
    w = torch.randn(shape[:-1], device="cuda", generator=g)

    def probs_of(lg):
        p = torch.softmax(lg, dim=-1)
        if zeros:  # exact zeros in p (t = 0 there) and saturated entries (clamp branches)
            mask = torch.ones_like(p)
            mask[..., 0] = 0
            p = p * mask
        return p

    p = probs_of(logits)
    pr = prior.repeat(*shape[:-1], 1) if len(shape) == 3 else prior.repeat(shape[0], 1)
    ref = kl(Categorical(probs=p), Categorical(probs=pr))
    gr = torch.autograd.grad((ref * w).sum(), [logits])[0]
    got = ops.categorical_kl(probs_of(logits), prior)
    gg = torch.autograd.grad((got * w).sum(), [logits])[0]
    assert got.shape == ref.shape and _close(got, ref, 1e-5, 1e-6)
    assert _close(gg, gr, 1e-4, 1e-6)
    # a full-shape prior takes the per-row path of the kernel
    got2 = ops.categorical_kl(probs_of(logits), pr)
    assert torch.equal(got2, got)


def test_categorical_kl_zero_prior_is_inf():
    from scvi_tools_b200 import ops

    p = torch.tensor([[0.5, 0.5, 0.0], [0.2, 0.3, 0.5]], device="cuda")
    q = torch.tensor([[0.5, 0.0, 0.5]], device="cuda")
    out = ops.categorical_kl(p, q)
    assert torch.isinf(out).all()


def test_arena_adam_matches_torch_adam():
    """Several steps of ArenaAdam (grads handed over by autograd, gathered with one launch) against torch.optim.Adam
    on a copy: parameters agree to fp32 rounding; a parameter that receives no gradient in a later step is treated
    as a zero gradient (its moments still decay), as torch does for a zero `.grad`."""
    from scvi_tools_b200.optim import ArenaAdam

    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(7, 5), torch.nn.Tanh(), torch.nn.Linear(5, 3)).cuda()
    ref = torch.nn.Sequential(torch.nn.Linear(7, 5), torch.nn.Tanh(), torch.nn.Linear(5, 3)).cuda()
    ref.load_state_dict(net.state_dict())
    opt = ArenaAdam(net.named_parameters, lr=1e-2, eps=0.01, weight_decay=1e-6)
    ropt = torch.optim.Adam(ref.parameters(), lr=1e-2, eps=0.01, weight_decay=1e-6)
    x = torch.randn(16, 7, device="cuda")
    for _ in range(5):
        opt.zero_grad()
        net(x).square().sum().backward()
        assert all(p.grad is not None for p in net.parameters())
        opt.step()
        ropt.zero_grad()
        ref(x).square().sum().backward()
        ropt.step()
    for a, b in zip(net.parameters(), ref.parameters(), strict=True):
        assert torch.allclose(a, b, rtol=1e-5, atol=1e-6)
    # gradients were moved into the arena in parameter order
    offs = opt.arena["offs"]
    for (_, p), o in zip(net.named_parameters(), offs, strict=True):
        assert torch.equal(opt.arena["g"][o : o + p.numel()], p.grad.reshape(-1))
    # a step in which one parameter gets no gradient
    opt.zero_grad()
    net[0](x).sum().backward()
    opt.step()
    o = offs[2]
    assert float(opt.arena["g"][o : o + net[2].weight.numel()].abs().max()) == 0.0
