"""GPU: size-independent properties at the bench workload's full shapes (BASELINE.json configs[1]: 100k cells x
2000 genes, 8 batches, ZINB, n_hidden 128, minibatch 8192).  The direct oracle comparison at these shapes is
tests/test_gpu_benchconfig.py; the properties below need no oracle.

* the minibatch gather (fp32 / uint8 / uint16 / CSR residency) equals an independent indexing of the same matrix
  bit for bit, and round-trips through the compact formats;
* in eval mode a cell's reconstruction loss and KL do not depend on which other cells share its minibatch
  (permutation equivariance, splitting the minibatch), for the FFMA and the tcgen05 paths;
* `loss == mean(reconst + kl_weight * kl)` (linearity of the ELBO reduction) and the two precisions agree.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

N, G, NB, B = 100_000, 2000, 8, 8192


def _counts(dev, n=N):
    g = torch.Generator(device=dev).manual_seed(0)
    lam = torch.rand(n, 1, device=dev, generator=g) * 20 + 0.5
    x = torch.poisson(lam.expand(n, G), generator=g)
    x = x * (torch.rand(n, G, device=dev, generator=g) < 0.7)
    return x.contiguous()


def test_fullsize_gather_all_residencies_bit_exact():
    import scipy.sparse as sp

    from scvi_tools_b200 import ops
    from scvi_tools_b200.data import DeviceResidentCounts, DeviceResidentCSR

    dev = torch.device("cuda:0")
    X = _counts(dev)
    idx = torch.randperm(N, device=dev)[:B]
    want = X[idx]
    got = ops.gather_rows(X, idx)
    assert torch.equal(got, want)
    assert float(got.double().sum()) == float(want.double().sum())  # checksum of the whole minibatch
    assert float(X.max()) < 256
    for dt in (torch.uint8, torch.uint16):
        assert torch.equal(ops.gather_rows(X.to(dt), idx), want), dt
    auto = DeviceResidentCounts(X, device=dev, compress="auto")
    assert auto.X.dtype == torch.uint8 and torch.equal(auto.minibatch(idx)["X"], want)
    sub = X[:20_000]
    csr = DeviceResidentCSR(sp.csr_matrix(sub.cpu().numpy()), device=dev)
    j = torch.randperm(20_000, device=dev)[:B]
    assert torch.equal(csr.minibatch(j)["X"], sub[j])
    assert csr.nbytes < sub.numel() * 4 * 0.6


@pytest.mark.parametrize("precision", [0, 1])
def test_fullsize_eval_losses_do_not_depend_on_the_minibatch(precision):
    import scvi_tools_b200 as S

    dev = torch.device("cuda:0")
    X = _counts(dev, 2 * B)
    torch.manual_seed(1)
    m = S.VAE(n_input=G, n_batch=NB, gene_likelihood="zinb", precision=precision).to(dev).eval()
    with torch.no_grad():  # non-trivial BatchNorm running statistics
        for k, b in m.named_buffers():
            if k.endswith("running_mean"):
                b.normal_(0, 0.3)
            if k.endswith("running_var"):
                b.uniform_(0.5, 1.5)
    batch = torch.randint(0, NB, (2 * B, 1), device=dev)
    noise = torch.randn(2 * B, m.n_latent, device=dev)

    def run(rows):
        t = {"X": X[rows].contiguous(), "batch": batch[rows].contiguous(), "labels": torch.zeros_like(batch[rows])}
        with torch.no_grad(), S.rng_feed(noise=[noise[rows]]):
            _, _, lo = m(t, loss_kwargs={"kl_weight": 0.3})
        return lo

    rows = torch.arange(B, device=dev)
    full = run(rows)
    rec, klz = full.reconstruction_loss["reconstruction_loss"], full.kl_local["kl_divergence_z"]
    assert torch.isfinite(rec).all() and torch.isfinite(klz).all()
    # linearity of the reduction
    want = float((rec.double() + 0.3 * klz.double()).mean())
    assert abs(float(full.loss) - want) <= 1e-6 * abs(want)
    # permutation equivariance
    perm = torch.randperm(B, device=dev)
    p = run(rows[perm])
    tol = 2e-6
    assert float(((p.reconstruction_loss["reconstruction_loss"] - rec[perm]).abs() / rec[perm].abs()).max()) < tol
    assert float((p.kl_local["kl_divergence_z"] - klz[perm]).abs().max()) < 1e-4
    # a cell evaluated inside a different, smaller, ragged minibatch
    for lo_, hi_ in ((0, 3000), (3000, 8192 - 77), (8192 - 77, 8192)):
        q = run(rows[lo_:hi_])
        assert float(((q.reconstruction_loss["reconstruction_loss"] - rec[lo_:hi_]).abs() / rec[lo_:hi_].abs()).max()) < tol
    # ... and mixed with cells it was not batched with before
    mix = torch.cat([rows[:100], torch.arange(B, 2 * B - 100, device=dev)])
    q = run(mix)
    assert float(((q.reconstruction_loss["reconstruction_loss"][:100] - rec[:100]).abs() / rec[:100].abs()).max()) < tol


def test_fullsize_precisions_agree_in_training_mode():
    import scvi_tools_b200 as S

    dev = torch.device("cuda:0")
    X = _counts(dev, B)
    batch = torch.randint(0, NB, (B, 1), device=dev)
    mask = (torch.rand(B, 128, device=dev) >= 0.1).to(torch.uint8)
    noise = torch.randn(B, 10, device=dev)
    out = []
    for precision in (0, 1):
        torch.manual_seed(2)
        m = S.VAE(n_input=G, n_batch=NB, gene_likelihood="zinb", precision=precision).to(dev).train()
        t = {"X": X, "batch": batch, "labels": torch.zeros_like(batch)}
        with S.rng_feed(drop_masks=[mask], noise=[noise]):
            _, _, lo = m(t, loss_kwargs={"kl_weight": 1.0})
        lo.loss.backward()
        out.append((float(lo.loss), {k: p.grad.clone() for k, p in m.named_parameters() if p.grad is not None}))
    assert abs(out[0][0] - out[1][0]) <= 1e-5 * abs(out[0][0])
    gs = max(float(g.abs().max()) for g in out[0][1].values())
    for k, g in out[0][1].items():
        d = float((out[1][1][k] - g).norm())
        assert d <= 1e-4 * float(g.norm()) + 1e-6 * gs * g.numel() ** 0.5, k
