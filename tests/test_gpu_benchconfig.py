"""GPU: the configuration `bench.py` times, pinned against the CPU oracle at full size.

`FusedTrainingPlan` exactly as the bench builds it -- tcgen05 kernels (`precision=1`), minibatch of 8192 cells
gathered in place from a >= 100k-row HBM-resident matrix, 2000 genes / ZINB / 8 batches (BASELINE.json
configs[1]) and 4000 genes / NB / 32 batches (configs[2]) -- with a fed dropout mask and reparameterisation
noise, against `oracle.loss_and_grads` + `oracle.adam_step` on the host (a fraction of a second per step at
this size).  Tolerances are the north-star's: ELBO and per-cell reconstruction loss 1e-5 relative, every
gradient of the arena 1e-4 relative (norm-wise, absolute floor for the mathematically-zero gradients of the
Linear biases feeding a BatchNorm), parameters after two Adam steps.

ReLU kinks.  At these sizes a handful of the 8192 x 128 hidden pre-activations sit within float32 rounding
distance of zero (the encoder's are sums of thousands of log1p(x) w terms: rounding noise ~1e-7 x sum |terms|), and
which side of the kink such an element lands on depends on the GEMM's summation order -- in this library, in
cuBLAS, on the CPU.  One flipped gate moves the BatchNorm bias gradient by ~4e-4 relative and the encoder weight
gradient by ~1e-3 (measured; tools/debug_c3.py), so two correct float32 implementations cannot agree to 1e-4 on
those tensors unless they take the same side.  The test therefore (1) reads the side the kernels took from their
activations, (2) asserts that wherever it differs from the oracle's the oracle's pre-activation is at rounding
distance from zero, and (3) evaluates the oracle on the kernels' side (`relu_gate1/2`); everything else is the
plain oracle.

Also here: CUDA-graph replay == the eager kernel sequence, bit for bit, with the library's own Philox streams
(reference contract: `tests/model/test_scvi.py:1245-1259`, same seed => same weights).
"""
import numpy as np
import pytest
import torch

from oracle import scvi_oracle as O
from tests._golden import grad_close, rel_err

pytestmark = pytest.mark.gpu

N_ROWS, B = 100_352, 8192


def _resident_counts(dev, G, seed=0, zero_frac=0.3):
    """NB(5, .3) x Bernoulli(1 - zero_frac), the synthetic_iid recipe, generated on the device."""
    gen = torch.Generator(device=dev).manual_seed(seed)
    conc = torch.full((1,), 5.0, device=dev)
    lam = torch._standard_gamma(conc.expand(N_ROWS, G), generator=gen) * (0.7 / 0.3)
    x = torch.poisson(lam, generator=gen) * (torch.rand(N_ROWS, G, device=dev, generator=gen) >= zero_frac)
    return x.contiguous()


def _build(dev, G, nb, lik, kl_weight, graph=False, seed=0, counts=None, **kw):
    import scvi_tools_b200 as S
    from scvi_tools_b200.data import DeviceResidentCounts
    from scvi_tools_b200.train import FusedTrainingPlan

    X = _resident_counts(dev, G) if counts is None else counts
    batch = torch.arange(N_ROWS, device=dev) * nb // N_ROWS
    torch.manual_seed(seed)
    m = S.VAE(n_input=G, n_batch=nb, gene_likelihood=lik, n_hidden=128, n_latent=10, precision=1).train()
    P0 = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    plan = FusedTrainingPlan(m, DeviceResidentCounts(X, batch=batch, device=dev), batch_size=B,
                             n_epochs_kl_warmup=None, max_kl_weight=kl_weight, use_cuda_graph=graph, **kw)
    plan.set_epoch(0)
    return plan, m, P0, X, batch


def _oracle_on_kernel_side(plan, P, x, bi, nb, lik, klw, mask, noise, max_flips=64):
    """`oracle.loss_and_grads` with the ReLU kink side of every hidden unit taken from the kernels' activations."""
    kw = dict(n_batch=nb, gene_likelihood=lik, kl_weight=klw, drop_mask=mask, noise=noise)
    out, grads = O.loss_and_grads(P, x, bi, **kw)
    keep = mask.bool()
    gate1 = torch.where(keep, plan.a1.cpu() > 0, out["bn1"] > 0)  # a dropped unit carries no gradient either way
    gate2 = plan.h.cpu() > 0
    flip1, flip2 = gate1 != (out["bn1"] > 0), gate2 != (out["bn2"] > 0)
    n_flips = int(flip1.sum()) + int(flip2.sum())
    if n_flips == 0:
        return out, grads
    assert n_flips <= max_flips, f"{n_flips} ReLU gates differ from the oracle's"
    # BatchNorm output is O(1): a float32 GEMM of K terms leaves it uncertain by ~1e-7 * sum|terms| / std ~ 1e-5
    assert float(out["bn1"][flip1].abs().max() if flip1.any() else 0.0) < 2e-4
    assert float(out["bn2"][flip2].abs().max() if flip2.any() else 0.0) < 2e-5
    return O.loss_and_grads(P, x, bi, relu_gate1=gate1, relu_gate2=gate2, **kw)


@pytest.mark.parametrize("G,nb,lik", [(2000, 8, "zinb"), (4000, 32, "nb")], ids=["c2-2000-zinb-8b", "c3-4000-nb-32b"])
def test_bench_configuration_step_matches_oracle(G, nb, lik):
    dev = torch.device("cuda:0")
    klw = 0.37
    plan, m, P0, X, batch = _build(dev, G, nb, lik, klw)
    g = torch.Generator().manual_seed(11)
    n_steps = 2
    idxs = [torch.randperm(N_ROWS, generator=g)[:B].sort().values for _ in range(n_steps)]
    masks = [(torch.rand(B, 128, generator=g) >= 0.1).to(torch.uint8) for _ in range(n_steps)]
    noises = [torch.randn(B, 10, generator=g) for _ in range(n_steps)]
    keys = O.trained_keys(lik)
    torch.set_num_threads(max(1, torch.get_num_threads()))
    P = {k: v.clone() for k, v in P0.items()}
    M = {k: torch.zeros_like(P[k]) for k in keys}
    V = {k: torch.zeros_like(P[k]) for k in keys}
    for t in range(n_steps):
        o3 = plan.training_step(idxs[t].to(dev), drop_mask=masks[t], noise=noises[t])
        torch.cuda.synchronize()
        x = X[idxs[t].to(dev)].cpu()
        bi = batch[idxs[t].to(dev)].cpu().view(-1, 1)
        out, grads = _oracle_on_kernel_side(plan, P, x, bi, nb, lik, klw, masks[t], noises[t])
        # ---- ELBO, per-cell pieces
        loss_ref = float(out["loss"])
        assert abs(float(o3[0]) - loss_ref) <= 1e-5 * abs(loss_ref), (t, float(o3[0]), loss_ref)
        rec, rec_ref = plan.reconst.cpu().double(), out["reconst"].detach().double()
        assert float(((rec - rec_ref).abs() / rec_ref.abs()).max()) <= 1e-5, t
        assert float((plan.kl.cpu().double() - out["kl_z"].detach().double()).abs().max()) <= 1e-4
        assert abs(float(o3[1]) - float(rec_ref.sum())) <= 1e-5 * abs(float(rec_ref.sum()))
        # ---- every gradient of the arena
        scale = max(float(v.abs().max()) for v in grads.values())
        for k in keys:
            assert grad_close(plan.Gd[k].cpu(), grads[k], 1e-4, scale), (
                t, k, rel_err(plan.Gd[k].cpu(), grads[k]))
        # ---- the reference optimizer on the oracle's gradients
        for k in keys:
            P[k], M[k], V[k] = O.adam_step(P[k], grads[k], M[k], V[k], t + 1)
        P.update(out["bn_buffers"])
    sd = m.state_dict()
    for k in keys:
        upd_ref = (P[k] - P0[k]).numpy()
        upd = sd[k].cpu().numpy() - P0[k].numpy()
        assert np.linalg.norm(upd - upd_ref) <= 2e-3 * np.linalg.norm(upd_ref) + 1e-6 * np.sqrt(upd.size), k
        assert rel_err(sd[k].cpu(), P[k]) < 1e-5, k
    for k in out["bn_buffers"]:
        # running statistics of pre-activations that are float32 sums of thousands of terms (batch mean ~ 0 by
        # cancellation): 1e-4 on the buffer, i.e. 1e-6 of the pre-activation scale at momentum 0.01
        assert rel_err(sd[k].cpu(), P[k]) < 1e-4, k


def test_bench_configuration_on_sparse_counts_matches_oracle():
    """The ~90 %-zeros count distribution of SURVEY 8(d) (real scRNA sparsity): the zero-count fast paths of the
    fused likelihood kernels against the oracle, full size."""
    dev = torch.device("cuda:0")
    G, nb, lik, klw = 2000, 8, "zinb", 1.0
    X = _resident_counts(dev, G, seed=3, zero_frac=0.9)
    plan, m, P0, X, batch = _build(dev, G, nb, lik, klw, counts=X)
    g = torch.Generator().manual_seed(5)
    idx = torch.randperm(N_ROWS, generator=g)[:B].sort().values
    mask = (torch.rand(B, 128, generator=g) >= 0.1).to(torch.uint8)
    noise = torch.randn(B, 10, generator=g)
    o3 = plan.training_step(idx.to(dev), drop_mask=mask, noise=noise)
    torch.cuda.synchronize()
    out, grads = _oracle_on_kernel_side(plan, P0, X[idx.to(dev)].cpu(), batch[idx.to(dev)].cpu().view(-1, 1), nb, lik,
                                        klw, mask, noise)
    assert abs(float(o3[0]) - float(out["loss"])) <= 1e-5 * abs(float(out["loss"]))
    rec, rec_ref = plan.reconst.cpu().double(), out["reconst"].detach().double()
    assert float(((rec - rec_ref).abs() / rec_ref.abs()).max()) <= 1e-5
    scale = max(float(v.abs().max()) for v in grads.values())
    for k in O.trained_keys(lik):
        assert grad_close(plan.Gd[k].cpu(), grads[k], 1e-4, scale), (k, rel_err(plan.Gd[k].cpu(), grads[k]))


@pytest.mark.parametrize("G,nb,lik", [(2000, 8, "zinb"), (4000, 32, "nb")], ids=["c2", "c3"])
def test_graph_replay_equals_eager_sequence_bitwise(G, nb, lik):
    """Same seed, same minibatches: the captured graph (side-stream prepack included) and the eager launch
    sequence end with bit-identical parameter / Adam arenas and BatchNorm buffers."""
    dev = torch.device("cuda:0")
    X = _resident_counts(dev, G, seed=1)
    g = torch.Generator().manual_seed(2)
    idxs = [torch.randperm(N_ROWS, generator=g)[:B].sort().values.to(dev) for _ in range(4)]
    res = []
    for graph in (False, True):
        plan, m, _, _, _ = _build(dev, G, nb, lik, 0.5, graph=graph, seed=7, counts=X)
        losses = []
        for ix in idxs:
            losses.append(plan.training_step(ix).clone())
        torch.cuda.synchronize()
        assert (plan._graphs is not None) == graph
        res.append((plan.p.clone(), plan.m.clone(), plan.v.clone(), plan.bn1_rm.clone(), plan.bn1_rv.clone(),
                    plan.bn2_rm.clone(), plan.bn2_rv.clone(), plan.step_counter.clone(), torch.stack(losses)))
    for a, b in zip(res[0], res[1], strict=True):
        assert torch.equal(a, b)
    assert int(res[0][7][0]) == 4 and torch.isfinite(res[0][8]).all()


@pytest.mark.parametrize("B_,graph", [(8192, True), (8192, False), (1000, False), (100, False), (20000, False)])
def test_cooperative_middle_equals_multi_launch_middle_bitwise(B_, graph, monkeypatch):
    """`scvi_mid_fwd` / `scvi_mid_bwd` as one cooperative launch each (grid barriers between the phases) against the
    multi-launch path they replace (SCVI_MID_COOP=0): same tile bodies, same merge arithmetic, so parameters, Adam
    moments, BatchNorm buffers and losses are identical bit for bit -- at the benched size (with and without graph
    replay), ragged sizes, and a minibatch with more row tiles than co-resident CTAs (20000 rows = 625 tiles)."""
    from scvi_tools_b200.data import DeviceResidentCounts
    from scvi_tools_b200.train import FusedTrainingPlan
    import scvi_tools_b200 as S

    dev = torch.device("cuda:0")
    G, nb = 512, 5
    n_rows = max(4 * B_, 4096)
    gen = torch.Generator(device=dev).manual_seed(0)
    X = torch.poisson(torch.rand(n_rows, G, device=dev, generator=gen) * 6.0, generator=gen)
    batch = torch.arange(n_rows, device=dev) * nb // n_rows
    g = torch.Generator().manual_seed(2)
    idxs = [torch.randperm(n_rows, generator=g)[:B_].sort().values.to(dev) for _ in range(4)]
    res = []
    for coop in ("0", "1"):
        monkeypatch.setenv("SCVI_MID_COOP", coop)
        torch.manual_seed(7)
        m = S.VAE(n_input=G, n_batch=nb, gene_likelihood="zinb", n_hidden=128, n_latent=10, precision=1).train()
        plan = FusedTrainingPlan(m, DeviceResidentCounts(X, batch=batch, device=dev), batch_size=B_,
                                 n_epochs_kl_warmup=None, max_kl_weight=0.5, use_cuda_graph=graph, seed=3)
        plan.set_epoch(0)
        losses = [plan.training_step(ix).clone() for ix in idxs]
        torch.cuda.synchronize()
        res.append((plan.p.clone(), plan.m.clone(), plan.v.clone(), plan.bn1_rm.clone(), plan.bn1_rv.clone(),
                    plan.bn2_rm.clone(), plan.bn2_rv.clone(), torch.stack(losses), plan.g.clone()))
    for a, b in zip(res[0], res[1], strict=True):
        assert torch.equal(a, b), float((a - b).abs().max())
    assert torch.isfinite(res[1][7]).all() and float(res[1][0].abs().sum()) > 0


@pytest.mark.parametrize("G,B_,lik", [(2000, 8192, "zinb"), (500, 1000, "nb"), (4000, 300, "zinb")])
def test_uint8_resident_counts_equal_fp32_resident_counts_bitwise(G, B_, lik):
    """uint8-resident counts (SCVI_X_U8 / SCVI_LIN_X_U8: 1 byte per count through E1 / K2 / K3 / E2, log1p from a
    256-entry table built with the fp32 path's own arithmetic) against the same counts resident as fp32: identical
    parameters, Adam moments, BatchNorm buffers and losses after 3 steps.  G = 500 exercises padded rows (500 % 16
    != 0) and a partial last 128-gene chunk; B = 1000 / 300 ragged row tiles."""
    import scvi_tools_b200 as S
    from scvi_tools_b200.data import DeviceResidentCounts
    from scvi_tools_b200.train import FusedTrainingPlan

    dev = torch.device("cuda:0")
    nb = 4
    n_rows = max(4 * B_, 4096)
    gen = torch.Generator(device=dev).manual_seed(0)
    X = torch.poisson(torch.rand(n_rows, G, device=dev, generator=gen) * 9.0, generator=gen)
    X[0, 0], X[1, G - 1] = 255.0, 255.0  # the table's last entry, first and last gene
    batch = torch.arange(n_rows, device=dev) * nb // n_rows
    g = torch.Generator().manual_seed(2)
    idxs = [torch.randperm(n_rows, generator=g)[:B_].sort().values.to(dev) for _ in range(3)]
    idxs[0][0] = 0
    idxs[0][1] = 1
    idxs[0] = idxs[0].sort().values
    res = []
    for compress in (None, "u8"):
        data = DeviceResidentCounts(X, batch=batch, device=dev, compress=compress)
        assert data.X.dtype == (torch.uint8 if compress else torch.float32)
        if compress:
            assert data.X.stride(0) % 16 == 0
        torch.manual_seed(7)
        m = S.VAE(n_input=G, n_batch=nb, gene_likelihood=lik, n_hidden=128, n_latent=10, precision=1).train()
        plan = FusedTrainingPlan(m, data, batch_size=B_, n_epochs_kl_warmup=None, max_kl_weight=0.5,
                                 use_cuda_graph=False, seed=3)
        plan.set_epoch(0)
        losses = [plan.training_step(ix).clone() for ix in idxs]
        torch.cuda.synchronize()
        res.append((plan.p.clone(), plan.m.clone(), plan.v.clone(), plan.bn1_rm.clone(), plan.bn2_rv.clone(),
                    torch.stack(losses), plan.g.clone(), plan.library.clone()))
    # fp32 resident counts take the integer-count element path (SCVI_X_INTEGER): the same arithmetic as uint8
    for a, b in zip(res[0], res[1], strict=True):
        assert torch.equal(a, b), float((a - b).abs().max())
    assert torch.isfinite(res[1][5]).all()


@pytest.mark.parametrize("G,nb,lik", [(2000, 8, "zinb"), (4000, 32, "nb")], ids=["c2", "c3"])
def test_kept_operand_gemm_backward_matches_recompute_backward(G, nb, lik, monkeypatch):
    """SCVI_PREC_AOPS (opt-in, SCVI_AOPS=1): the fused likelihood kernel keeps its per-element operands and the
    head-weight gradients are a pure tcgen05 GEMM over them (+ a count-only digamma sum) instead of the recompute
    kernel.  Same inputs, same masks / noise: every gradient of the arena agrees with the default path to 2e-5
    (norm-wise; both paths are within 1e-4 of the oracle), the loss is identical."""
    dev = torch.device("cuda:0")
    X = _resident_counts(dev, G, seed=4)
    g = torch.Generator().manual_seed(3)
    idx = torch.randperm(N_ROWS, generator=g)[:B].sort().values.to(dev)
    mask = (torch.rand(B, 128, generator=g) >= 0.1).to(torch.uint8)
    noise = torch.randn(B, 10, generator=g)
    res = []
    for flag in ("0", "1"):
        monkeypatch.setenv("SCVI_AOPS", flag)
        plan, m, _, _, _ = _build(dev, G, nb, lik, 0.7, seed=5, counts=X)
        assert bool(plan._aops) == (flag == "1")
        o3 = plan.training_step(idx, drop_mask=mask, noise=noise).clone()
        torch.cuda.synchronize()
        res.append((o3, {k: v.clone() for k, v in plan.Gd.items()}))
    assert torch.equal(res[0][0], res[1][0])
    scale = max(float(v.abs().max()) for v in res[0][1].values())
    for k, v in res[0][1].items():
        assert grad_close(res[1][1][k].cpu(), v.cpu(), 2e-5, scale), (k, rel_err(res[1][1][k].cpu(), v.cpu()))
