"""GPU: the fused training plan (explicit kernel sequence + flat-arena Adam + CUDA graph) against
the reference's recorded two-step trajectory and against the autograd path."""
import numpy as np
import pytest
import torch

from oracle import scvi_oracle as O
from tests._golden import load_vae_case, rel_err

pytestmark = pytest.mark.gpu


def _plan_from_case(name, precision=0, use_mid=True, **kw):
    import scvi_tools_b200 as S
    from scvi_tools_b200.data import DeviceResidentCounts
    from scvi_tools_b200.train import FusedTrainingPlan

    P, inp, out = load_vae_case(name)
    G = P["px_r"].shape[0]
    m = S.VAE(n_input=G, n_batch=inp["n_batch"], gene_likelihood=inp["likelihood"], precision=precision)
    sd = m.state_dict()
    sd.update({k: v.clone() for k, v in P.items()})
    m.load_state_dict(sd)
    m.train()
    data = DeviceResidentCounts(inp["x"], batch=inp["batch"].reshape(-1), device="cuda:0")
    plan = FusedTrainingPlan(m, data, batch_size=inp["x"].shape[0], n_epochs_kl_warmup=None, max_kl_weight=inp["kl_weight"], **kw)
    plan.use_mid = plan.use_mid and use_mid
    return plan, m, P, inp, out


@pytest.mark.parametrize("variant", ["fp32", "tc", "fp32-unfused-middle"])
@pytest.mark.parametrize("name", ["c1_zinb", "c1_nb", "ragged_zinb", "onebatch_nb"])
def test_two_adam_steps_match_reference_trajectory(name, variant):
    """Parameters after 2 x (fwd, bwd, Adam) == what the reference module + torch.optim.Adam reach,
    for the FFMA kernels, the tcgen05 kernels (decoder + encoder) and the unfused middle sequence."""
    plan, m, P, inp, out = _plan_from_case(name, precision=1 if variant == "tc" else 0,
                                           use_mid=variant != "fp32-unfused-middle", use_cuda_graph=False)
    plan.set_epoch(0)
    B = inp["x"].shape[0]
    idx = torch.arange(B, device="cuda:0")
    mask, noise = torch.from_numpy(inp["drop_mask"]), torch.from_numpy(inp["noise"])
    for step in range(inp["n_adam_steps"]):
        o3 = plan.training_step(idx, drop_mask=mask, noise=noise)
        if step == 0:
            assert abs(float(o3[0]) - float(out["loss"])) <= 1e-5 * abs(float(out["loss"]))
            assert abs(float(o3[1]) - float(out["reconst"].sum())) <= 1e-5 * abs(float(out["reconst"].sum()))
    sd = m.state_dict()
    n = inp["n_adam_steps"]
    for k in O.trained_keys(inp["likelihood"]):
        # Adam with eps=0.01 moves each weight by <= lr per step: compare the *update*, not only the value
        upd_ref = out[f"param_after{n}::" + k] - P[k].numpy()
        upd = sd[k].cpu().numpy() - P[k].numpy()
        # floor: 1e-6 per element = 0.1% of lr (biases feeding a BatchNorm have pure-noise gradients)
        assert np.linalg.norm(upd - upd_ref) <= 2e-3 * np.linalg.norm(upd_ref) + 1e-6 * np.sqrt(upd.size), k
        assert rel_err(sd[k].cpu(), out[f"param_after{n}::" + k]) < 1e-5, k
    assert int(sd["z_encoder.encoder.fc_layers.Layer 0.1.num_batches_tracked"]) == n


def test_plan_gradients_equal_autograd_path():
    """The explicit backward sequence and torch autograd over the same kernels agree."""
    import scvi_tools_b200 as S

    plan, m, P, inp, out = _plan_from_case("ragged_zinb", use_cuda_graph=False)
    plan.set_epoch(0)
    B = inp["x"].shape[0]
    mask, noise = torch.from_numpy(inp["drop_mask"]), torch.from_numpy(inp["noise"])
    t = {"X": torch.from_numpy(inp["x"]).cuda(), "batch": torch.from_numpy(inp["batch"]).cuda(),
         "labels": torch.zeros(B, 1, dtype=torch.long, device="cuda")}
    with S.rng_feed(drop_masks=[mask], noise=[noise]):
        _, _, lo = m(t, loss_kwargs={"kl_weight": inp["kl_weight"]})
    lo.loss.backward()
    auto = {k: p.grad.clone() for k, p in m.named_parameters() if p.grad is not None}
    st = torch.cuda.current_stream().cuda_stream
    plan.idx.copy_(torch.arange(B, device="cuda"))
    plan.mask.copy_(mask.cuda())
    plan.noise.copy_(noise.cuda())
    plan.cov_cols.copy_((torch.from_numpy(inp["batch"]).cuda() + plan.L).to(torch.int32))
    import ctypes

    plan._prologue(ctypes.c_void_p(st), feed_rng=False)  # ELBO backward seeds (1/B, kl_weight/B); keeps the fed mask / noise
    plan._forward_backward(ctypes.c_void_p(st))
    for k in plan.keys:
        assert rel_err(plan.Gd[k].cpu(), auto[k].cpu()) < 1e-5 or float(auto[k].abs().max()) < 1e-5, k


def test_graph_training_learns_and_is_finite():
    """CUDA-graph replayed steps (on-device RNG) reduce the ELBO on synthetic_iid-like counts."""
    import scvi_tools_b200 as S
    from scvi_tools_b200.data import DeviceResidentCounts
    from scvi_tools_b200.train import FusedTrainingPlan

    X, batch = O.synthetic_iid(2048, 256, 4, seed=0)
    torch.manual_seed(0)
    m = S.VAE(n_input=256, n_batch=4).train()
    data = DeviceResidentCounts(X, batch=batch, device="cuda:0")
    plan = FusedTrainingPlan(m, data, batch_size=256, use_cuda_graph=True)
    hist = plan.fit(max_epochs=6)
    losses = [h["train_loss"] for h in hist]
    assert all(np.isfinite(losses)), losses
    assert losses[-1] < losses[0] - 1.0, losses
    assert hist[0]["kl_weight"] == 0.0 and abs(hist[4]["kl_weight"] - 4 / 400) < 1e-9
    assert hist[0]["n_cells"] == 2048
    # the module is usable afterwards (eval-mode forward on the trained weights, arena-backed params)
    m.eval()
    with torch.no_grad():
        _, _, lo = m(data.minibatch(torch.arange(100, device="cuda:0")))
    assert np.isfinite(float(lo.loss))


@pytest.mark.parametrize("count_max", [40, 1000])
def test_streamed_host_counts_train_like_resident_counts(count_max):
    """`HostStreamedCounts` (pinned host matrix in uint8 / uint16, double-buffered H2D, on-device expansion)
    feeds the plan the same fp32 rows as the HBM-resident matrix: identical weights after two epochs,
    ragged last minibatch included."""
    import scvi_tools_b200 as S
    from scvi_tools_b200.data import DeviceResidentCounts, HostStreamedCounts
    from scvi_tools_b200.train import FusedTrainingPlan

    X, batch = O.synthetic_iid(700, 192, 3, seed=2)
    X = np.minimum(X * (count_max / 40.0), count_max).round().astype(np.float32)
    out = []
    for streamed in (False, True):
        torch.manual_seed(0)
        m = S.VAE(n_input=192, n_batch=3).train()
        if streamed:
            host = HostStreamedCounts(X, batch=batch, batch_size=256, device="cuda:0")
            assert host.dtype == (torch.uint8 if count_max < 256 else torch.uint16)
            plan = FusedTrainingPlan(m, host.stage, batch_size=256, seed=3)
            hist = plan.fit_streamed(host, max_epochs=2)
        else:
            plan = FusedTrainingPlan(m, DeviceResidentCounts(X, batch=batch, device="cuda:0"), batch_size=256, seed=3)
            hist = plan.fit(max_epochs=2)
        assert hist[-1]["n_cells"] == X.shape[0]  # 2 full minibatches + a ragged one
        out.append(({k: v.detach().clone() for k, v in m.state_dict().items()}, hist))
    for k in out[0][0]:
        assert torch.equal(out[0][0][k], out[1][0][k]), k
    assert out[0][1][-1]["train_loss"] == out[1][1][-1]["train_loss"]


def test_csr_resident_counts_train_like_dense_counts():
    """`SCVI` on a scipy CSR matrix (kept sparse in HBM, densified per minibatch on the device) ends with
    exactly the weights of training on the dense matrix; the getters walk the sparse matrix too."""
    import scipy.sparse as sp

    from scvi_tools_b200.model import SCVI

    X, batch = O.synthetic_iid(600, 160, 2, seed=4)
    X = X * (np.random.default_rng(0).random(X.shape) < 0.3)  # sparser, as real scRNA counts
    out = []
    for sparse in (False, True):
        torch.manual_seed(0)
        model = SCVI(sp.csr_matrix(X) if sparse else X, batch=batch, gene_likelihood="nb")
        assert model.sparse == sparse
        hist = model.train(max_epochs=2, batch_size=256, seed=5)
        assert hist[-1]["n_cells"] == X.shape[0]
        out.append((model.state_dict(), hist, model.get_latent_representation(indices=[3, 599, 0, 41])))
    for k in out[0][0]:
        assert torch.equal(out[0][0][k], out[1][0][k]), k
    assert out[0][1][-1]["train_loss"] == out[1][1][-1]["train_loss"]
    assert np.array_equal(out[0][2], out[1][2])
    assert model.data.nbytes < X.size * 4 / 3


def test_ragged_last_batch_and_epoch_coverage():
    import scvi_tools_b200 as S
    from scvi_tools_b200.data import DeviceResidentCounts
    from scvi_tools_b200.train import FusedTrainingPlan

    X, batch = O.synthetic_iid(300, 64, 2, seed=1)
    m = S.VAE(n_input=64, n_batch=2, gene_likelihood="nb").train()
    plan = FusedTrainingPlan(m, DeviceResidentCounts(X, batch=batch, device="cuda:0"), batch_size=128)
    hist = plan.fit(max_epochs=1)
    assert hist[0]["n_cells"] == 300  # 128 + 128 + 44 (last batch smaller, as drop_last=False)
    assert np.isfinite(hist[0]["train_loss"])


def test_step_prologue_random_streams_and_seeds():
    """Library Philox streams: keep-rate of the dropout mask, N(0,1) moments of the noise, fresh draws per
    step / rank, reproducible for the same (seed, step, rank); covariate columns and ELBO seeds exact."""
    import ctypes as C

    from scvi_tools_b200._lib import call

    dev = "cuda:0"
    B, H, L, p = 4096, 128, 10, 0.1
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    ptr = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    counter = torch.zeros(2, dtype=torch.int64, device=dev)
    idx = torch.randperm(10000, device=dev)[:B]
    batch_cols = torch.randint(10, 18, (10000,), dtype=torch.int32, device=dev)
    klw = torch.tensor([0.25], device=dev)
    nbt = torch.zeros(2, dtype=torch.int64, device=dev)

    def draw(seed, step, rank):
        counter[0] = step
        mask = torch.zeros(B, H, dtype=torch.uint8, device=dev)
        noise = torch.zeros(B, L, device=dev)
        cov = torch.zeros(B, dtype=torch.int32, device=dev)
        gr, gk = torch.zeros(B, device=dev), torch.zeros(B, device=dev)
        call("scvi_step_prologue", seed, ptr(counter), rank, B, H, L, p, ptr(mask), ptr(noise), ptr(idx), ptr(batch_cols),
             ptr(cov), 0.0, ptr(klw), 1.0, ptr(gr), ptr(gk), ptr(nbt[0:]), ptr(nbt[1:]), st)
        torch.cuda.synchronize()
        return mask, noise, cov, gr, gk

    mask, noise, cov, gr, gk = draw(7, 0, 0)
    assert set(mask.unique().tolist()) <= {0, 1}
    n = B * H
    assert abs(float(mask.float().mean()) - (1 - p)) < 5 * np.sqrt(p * (1 - p) / n)
    # rows and columns are not correlated copies of each other
    assert abs(float(mask[:, 0].float().mean()) - (1 - p)) < 0.03 and abs(float(mask[0].float().mean()) - (1 - p)) < 0.15
    z = noise.double().flatten()
    m = z.numel()
    assert abs(float(z.mean())) < 5 / np.sqrt(m)
    assert abs(float(z.var()) - 1.0) < 5 * np.sqrt(2.0 / m)
    assert abs(float((z**3).mean())) < 5 * np.sqrt(15.0 / m)
    assert abs(float((z**4).mean()) - 3.0) < 5 * np.sqrt(96.0 / m)
    assert float(z.abs().max()) < 6.5
    assert torch.equal(cov, batch_cols[idx])
    assert torch.allclose(gr, torch.full_like(gr, 1.0 / B)) and torch.allclose(gk, torch.full_like(gk, 0.25 / B))
    assert nbt.tolist() == [1, 1]
    # reproducible, and fresh per step / rank / seed
    mask2, noise2, *_ = draw(7, 0, 0)
    assert torch.equal(mask, mask2) and torch.equal(noise, noise2)
    for args in ((7, 1, 0), (7, 0, 1), (8, 0, 0)):
        mask3, noise3, *_ = draw(*args)
        assert float((mask3 != mask).float().mean()) > 0.1
        assert abs(float((noise3 * noise).mean())) < 5 / np.sqrt(m)
    # ragged sizes (tail paths of the 16-wide mask / 4-wide noise stores)
    Bs, Hs, Ls = 37, 7, 3
    mk = torch.full((Bs * Hs + 16,), 9, dtype=torch.uint8, device=dev)
    nz = torch.full((Bs * Ls + 4,), 9.0, device=dev)
    call("scvi_step_prologue", 1, ptr(counter), 0, Bs, Hs, Ls, 0.5, ptr(mk), ptr(nz), None, None, None, 0.0, None, 1.0,
         None, None, None, None, st)
    torch.cuda.synchronize()
    assert int(mk[: Bs * Hs].max()) <= 1 and mk[Bs * Hs :].eq(9).all()
    assert nz[Bs * Ls :].eq(9.0).all() and float(nz[: Bs * Ls].abs().max()) < 6.5


def test_adam_graph_counter_advances_without_extra_kernel():
    import ctypes as C

    from scvi_tools_b200._lib import call

    dev = "cuda:0"
    n = 300_001
    torch.manual_seed(0)
    p = torch.randn(n + 3, device=dev)[:n]
    p = torch.randn(((n + 3) // 4) * 4, device=dev)
    g = torch.randn_like(p)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    ref = torch.nn.Parameter(p.clone())
    opt = torch.optim.Adam([ref], lr=1e-3, eps=0.01, weight_decay=1e-6)
    counter = torch.zeros(2, dtype=torch.int64, device=dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for step in range(3):
        call("scvi_adam_step_graph", C.c_void_p(p.data_ptr()), C.c_void_p(g.data_ptr()), C.c_void_p(m.data_ptr()),
             C.c_void_p(v.data_ptr()), p.numel(), C.c_void_p(counter.data_ptr()), 1e-3, 0.9, 0.999, 0.01, 1e-6, 1.0, st)
        ref.grad = g.clone()
        opt.step()
        torch.cuda.synchronize()
        assert counter.tolist() == [step + 1, 0]
    assert rel_err(p.cpu(), ref.detach().cpu()) < 1e-6


@pytest.mark.parametrize("device_gather", [True, False], ids=["device-gather", "host-gather"])
def test_streamed_uint8_stage_trains_like_resident_counts(device_gather):
    """`HostStreamedCounts(compact_stage=True)`: uint8 host counts land unexpanded in uint8 staging rows (rows gathered
    by the GPU straight from the pinned host matrix, or by the host threads) and the tcgen05 kernels read them as
    they are -- identical weights to training on the HBM-resident matrix, ragged last minibatch included."""
    import scvi_tools_b200 as S
    from scvi_tools_b200.data import DeviceResidentCounts, HostStreamedCounts
    from scvi_tools_b200.train import FusedTrainingPlan

    X, batch = O.synthetic_iid(700, 200, 3, seed=2)  # 200 % 16 != 0: padded staging rows
    out = []
    for streamed in (False, True):
        torch.manual_seed(0)
        m = S.VAE(n_input=200, n_batch=3, n_hidden=128, precision=1).train()
        if streamed:
            host = HostStreamedCounts(X, batch=batch, batch_size=256, device="cuda:0", compact_stage=True,
                                      device_gather=device_gather)
            assert host.dtype == torch.uint8 and host.stage.X.dtype == torch.uint8 and host.stage.X.stride(0) % 16 == 0
            assert host.device_gather == device_gather
            plan = FusedTrainingPlan(m, host.stage, batch_size=256, seed=3)
            hist = plan.fit_streamed(host, max_epochs=2)
        else:
            plan = FusedTrainingPlan(m, DeviceResidentCounts(X, batch=batch, device="cuda:0"), batch_size=256, seed=3)
            hist = plan.fit(max_epochs=2)
        assert hist[-1]["n_cells"] == X.shape[0]
        out.append(({k: v.detach().clone() for k, v in m.state_dict().items()}, hist))
    for k in out[0][0]:
        assert torch.equal(out[0][0][k], out[1][0][k]), k
    assert out[0][1][-1]["train_loss"] == out[1][1][-1]["train_loss"]


def test_streamed_hybrid_gather_trains_like_resident_counts():
    """`HostStreamedCounts(hybrid_fraction=0.4)`: part of every minibatch pulled by the GPU from the mapped pinned
    matrix, the rest gathered by the host thread pool and copied -- same rows, same order, identical weights."""
    import scvi_tools_b200 as S
    from scvi_tools_b200.data import DeviceResidentCounts, HostStreamedCounts
    from scvi_tools_b200.train import FusedTrainingPlan

    X, batch = O.synthetic_iid(700, 208, 3, seed=2)
    out = []
    for streamed in (False, True):
        torch.manual_seed(0)
        m = S.VAE(n_input=208, n_batch=3, n_hidden=128, precision=1).train()
        if streamed:
            host = HostStreamedCounts(X, batch=batch, batch_size=256, device="cuda:0", compact_stage=True,
                                      hybrid_fraction=0.4)
            if not host.device_gather:
                pytest.skip("pinned host memory is not mapped into the device address space")
            plan = FusedTrainingPlan(m, host.stage, batch_size=256, seed=3)
            hist = plan.fit_streamed(host, max_epochs=2)
        else:
            plan = FusedTrainingPlan(m, DeviceResidentCounts(X, batch=batch, device="cuda:0"), batch_size=256, seed=3)
            hist = plan.fit(max_epochs=2)
        out.append(({k: v.detach().clone() for k, v in m.state_dict().items()}, hist))
    for k in out[0][0]:
        assert torch.equal(out[0][0][k], out[1][0][k]), k
