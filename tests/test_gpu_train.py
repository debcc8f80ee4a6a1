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
