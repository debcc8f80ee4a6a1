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


@pytest.mark.parametrize("B,K,N,ncov", [(256, 4100, 256, 0), (256, 276, 4000, 3), (100, 2049, 130, 2), (64, 7, 5, 1)])
def test_fused_linear_few_cells_long_reduction(B, K, N, ncov):
    """FusedLinear against torch on shapes whose 64 x 64 tiles alone would not fill the GPU: the forward splits the
    K reduction (partial tiles + finish kernel), dx splits the N reduction (slabs + ordered sum).  BatchNorm tile
    statistics, covariate columns and all three gradients included."""
    from scvi_tools_b200 import _lib, ops

    g = torch.Generator(device="cuda").manual_seed(B + K)
    x = torch.randn(B, K, device="cuda", generator=g).requires_grad_()
    W = (torch.randn(N, K + ncov, device="cuda", generator=g) / K**0.5).requires_grad_()
    b = torch.randn(N, device="cuda", generator=g).requires_grad_()
    cov = torch.randint(0, max(ncov, 1), (B, 1), device="cuda", generator=g)
    cols = (K + cov).to(torch.int32) if ncov else None
    w = torch.randn(B, N, device="cuda", generator=g)
    y, part, _ = ops.FusedLinear.apply(x, W, b, cols, K, 0, None, True)
    gx, gW, gb = torch.autograd.grad((y * w).sum(), [x, W, b])
    x64, W64 = x.detach().double(), W.detach().double()
    ref = x64 @ W64[:, :K].t() + b.detach().double()
    if ncov:
        ref = ref + W64[:, K:].t()[cov.view(-1)]
    scale = float(ref.abs().max())
    assert float((y.double() - ref).abs().max()) <= 2e-6 * scale * K**0.5
    w64 = w.double()
    rx = w64 @ W64[:, :K]
    assert float((gx.double() - rx).abs().max()) <= 2e-6 * float(rx.abs().max()) * N**0.5
    rW = torch.zeros_like(W64)
    rW[:, :K] = w64.t() @ x64
    if ncov:
        oh = torch.nn.functional.one_hot(cov.view(-1), ncov).double()
        rW[:, K:] = w64.t() @ oh
    assert float((gW.double() - rW).abs().max()) <= 2e-6 * float(rW.abs().max()) * B**0.5
    assert float((gb.double() - w64.sum(0)).abs().max()) <= 2e-6 * float(w64.sum(0).abs().max()) * B**0.5
    # tile statistics of the (summed) output: mean and M2 per 64-row tile
    T = (B + _lib.BN_TILE - 1) // _lib.BN_TILE
    assert part.shape == (T, 2, N)
    for t in range(T):
        rows = y[t * _lib.BN_TILE : (t + 1) * _lib.BN_TILE].double()
        m = rows.mean(0)
        assert torch.allclose(part[t, 0].double(), m, rtol=1e-5, atol=1e-5 * scale)
        assert torch.allclose(part[t, 1].double(), ((rows - m) ** 2).sum(0), rtol=1e-4, atol=1e-5 * scale**2)


@pytest.mark.parametrize("B,P", [(256, 100), (70, 130), (1, 3)])
def test_small_batch_likelihood_tiles(B, P):
    """The few-cell launch geometry of the stand-alone likelihood kernels (short row tiles in the forward, row groups
    in the mixture backward) against torch.distributions-style float64 formulas via the existing wide-batch path:
    the same op evaluated on the rows repeated to a large batch must agree bit for bit in the forward, and the
    per-column dispersion gradient to rounding."""
    from scvi_tools_b200 import ops

    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.poisson(3 * torch.rand(B, P, device="cuda", generator=g))
    m1 = (torch.rand(B, P, device="cuda", generator=g) * 5 + 0.1).requires_grad_()
    m2 = (torch.rand(B, P, device="cuda", generator=g) * 50 + 1).requires_grad_()
    th = (torch.rand(P, device="cuda", generator=g) * 3 + 0.2).requires_grad_()
    pi = torch.randn(B, P, device="cuda", generator=g).requires_grad_()
    rep = 20000 // B + 1  # tall enough for the 64-row tiles

    def run(fn, args, reps):
        big = [a.repeat(reps, 1) if a.dim() == 2 else a for a in args]
        return fn(*big)

    for fn, args in ((ops.mixture_nb_log_prob, (x, m1, m2, th, pi)), (ops.nb_log_prob, (x, m1, th)),
                     (ops.zinb_log_prob, (x, m1, th, pi))):
        if fn is not ops.mixture_nb_log_prob and P % 4:
            continue  # the vectorised kernels need 16-byte rows; the scalar path has one geometry
        small = fn(*args)
        tall = run(fn, args, rep)
        assert torch.equal(tall[:B], small), fn.__name__
        gs = torch.autograd.grad(small.sum(), [a for a in args if a.requires_grad])
        gt = torch.autograd.grad(tall[:B].sum(), [a for a in args if a.requires_grad])
        for a, b_ in zip(gs, gt, strict=True):
            assert torch.allclose(a, b_, rtol=1e-5, atol=1e-6), fn.__name__
