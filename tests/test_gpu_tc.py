"""GPU: the tcgen05 tensor-core decoder path (SCVI_PREC_TC, error-compensated fp16 split) against the
reference golden vectors, the oracle and the FFMA path -- same fp32-path tolerances."""
import numpy as np
import pytest
import torch

from oracle import scvi_oracle as O
from tests._golden import grad_close, grad_scale, load_vae_case, max_rel

pytestmark = pytest.mark.gpu
PREC_TC = 1


def _run(G, B, nb, lik, precision, seed):
    import scvi_tools_b200 as S

    X, batch = O.synthetic_iid(max(B, nb * 40) + nb, G, nb, seed=seed)  # synthetic_iid rounds down to a multiple of nb
    torch.manual_seed(seed)
    m = S.VAE(n_input=G, n_batch=nb, gene_likelihood=lik, precision=precision)
    P = {k: v.detach().clone() for k, v in m.state_dict().items()}
    m = m.cuda().train()
    x = torch.from_numpy(X[:B])
    bi = torch.from_numpy(batch[:B]).view(-1, 1)
    g = torch.Generator().manual_seed(seed + 1)
    mask = (torch.rand(B, 128, generator=g) >= 0.1).to(torch.uint8)
    noise = torch.randn(B, 10, generator=g)
    t = {"X": x.cuda(), "batch": bi.cuda(), "labels": torch.zeros(B, 1, dtype=torch.long, device="cuda")}
    with S.rng_feed(drop_masks=[mask], noise=[noise]):
        _, _, lo = m(t, loss_kwargs={"kl_weight": 0.3})
    lo.loss.backward()
    grads = {k: p.grad.detach().cpu() for k, p in m.named_parameters() if p.grad is not None}
    return lo, grads, (P, x, bi, mask, noise)


@pytest.mark.parametrize(("B", "G", "nb", "lik"), [(300, 2000, 8, "zinb"), (129, 4000, 32, "nb"), (77, 333, 2, "zinb"),
                                                   (1000, 200, 3, "nb")])
def test_tc_forward_matches_oracle(B, G, nb, lik):
    lo, grads, (P, x, bi, mask, noise) = _run(G, B, nb, lik, PREC_TC, seed=B)
    torch.set_num_threads(8)
    o, og = O.loss_and_grads(P, x, bi, n_batch=nb, gene_likelihood=lik, kl_weight=0.3, drop_mask=mask, noise=noise)
    assert abs(float(lo.loss.detach()) - float(o["loss"].detach())) <= 1e-5 * abs(float(o["loss"].detach()))
    assert max_rel(lo.reconstruction_loss["reconstruction_loss"].detach().cpu(), o["reconst"].detach()) < 1e-5
    gs = max(float(v.abs().max()) for v in og.values())
    for k in O.trained_keys(lik):
        assert grad_close(grads[k], og[k], 1e-4, gs), k


@pytest.mark.parametrize("name", ["c1_zinb", "ragged_zinb", "onebatch_nb"])
def test_tc_vs_reference_golden(name):
    import scvi_tools_b200 as S

    P, inp, out = load_vae_case(name)
    G = P["px_r"].shape[0]
    m = S.VAE(n_input=G, n_batch=inp["n_batch"], gene_likelihood=inp["likelihood"], precision=PREC_TC)
    sd = m.state_dict()
    sd.update({k: v.clone() for k, v in P.items()})
    m.load_state_dict(sd)
    m = m.cuda().train()
    B = inp["x"].shape[0]
    t = {"X": torch.from_numpy(inp["x"]).cuda(), "batch": torch.from_numpy(inp["batch"]).cuda(),
         "labels": torch.zeros(B, 1, dtype=torch.long, device="cuda")}
    with S.rng_feed(drop_masks=[torch.from_numpy(inp["drop_mask"])], noise=[torch.from_numpy(inp["noise"])]):
        _, _, lo = m(t, loss_kwargs={"kl_weight": inp["kl_weight"]})
    lo.loss.backward()
    assert abs(float(lo.loss.detach()) - float(out["loss"])) <= 1e-5 * abs(float(out["loss"]))
    assert max_rel(lo.reconstruction_loss["reconstruction_loss"].detach().cpu(), out["reconst"]) < 1e-5
    gs = grad_scale(out)
    for k, p in m.named_parameters():
        if k in O.trained_keys(inp["likelihood"]):
            assert grad_close(p.grad.cpu(), out["grad::" + k], 1e-4, gs), k


def test_tc_large_batch_matches_ffma_path():
    """B = 4096 x G = 2000 (several CTAs per split, full tiles): TC and FFMA agree row by row."""
    lo_tc, g_tc, _ = _run(2000, 4096, 8, "zinb", PREC_TC, seed=7)
    lo_32, g_32, _ = _run(2000, 4096, 8, "zinb", 0, seed=7)
    a = lo_tc.reconstruction_loss["reconstruction_loss"].detach().cpu()
    b = lo_32.reconstruction_loss["reconstruction_loss"].detach().cpu()
    assert max_rel(a, b) < 5e-6
    gs = max(float(v.abs().max()) for v in g_32.values())
    for k in g_32:
        assert grad_close(g_tc[k], g_32[k], 1e-4, gs), k
