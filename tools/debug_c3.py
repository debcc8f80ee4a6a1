"""Debug aid (GPU): per-gradient error of the fused plan vs the oracle for a given shape, both precisions."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import scvi_oracle as O
from tests._golden import rel_err
from tests.test_gpu_benchconfig import _build, _resident_counts, N_ROWS, B

dev = torch.device("cuda:0")
for (G, nb, lik) in [(4000, 32, "nb"), (4000, 8, "zinb"), (2000, 32, "nb"), (4000, 32, "zinb")]:
    X = _resident_counts(dev, G)
    g = torch.Generator().manual_seed(11)
    idx = torch.randperm(N_ROWS, generator=g)[:B].sort().values
    mask = (torch.rand(B, 128, generator=g) >= 0.1).to(torch.uint8)
    noise = torch.randn(B, 10, generator=g)
    ref = None
    for prec in (0, 1):
        import scvi_tools_b200 as S
        from scvi_tools_b200.data import DeviceResidentCounts
        from scvi_tools_b200.train import FusedTrainingPlan
        batch = torch.arange(N_ROWS, device=dev) * nb // N_ROWS
        torch.manual_seed(0)
        m = S.VAE(n_input=G, n_batch=nb, gene_likelihood=lik, precision=prec).train()
        P0 = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
        plan = FusedTrainingPlan(m, DeviceResidentCounts(X, batch=batch, device=dev), batch_size=B,
                                 n_epochs_kl_warmup=None, max_kl_weight=0.37, use_cuda_graph=False)
        plan.set_epoch(0)
        plan.training_step(idx.to(dev), drop_mask=mask, noise=noise)
        torch.cuda.synchronize()
        if ref is None:
            x = X[idx.to(dev)].cpu(); bi = batch[idx.to(dev)].cpu().view(-1, 1)
            out, grads = O.loss_and_grads(P0, x, bi, n_batch=nb, gene_likelihood=lik, kl_weight=0.37, drop_mask=mask, noise=noise)
            P64 = {k: v.double() for k, v in P0.items()}
            out64, grads64 = O.loss_and_grads(P64, x.double(), bi, n_batch=nb, gene_likelihood=lik, kl_weight=0.37, drop_mask=mask, noise=noise.double())
            ref = True
        print(f"--- G={G} nb={nb} {lik} precision={prec}")
        for k in O.trained_keys(lik):
            print(f"  {k:55s} vs f32 oracle {rel_err(plan.Gd[k].cpu(), grads[k]):.2e}  vs f64 {rel_err(plan.Gd[k].cpu(), grads64[k]):.2e}   oracle32 vs 64 {rel_err(grads[k], grads64[k]):.2e}")
        print("  dy1 amax", float(plan.dy1.abs().max()), "dy1 median abs", float(plan.dy1.abs().median()), "dh amax", float(plan.dh.abs().max()))
