"""A few eager fused-plan steps at the bench shape on a small resident matrix: the target of `ncu` captures
(python tools/profile_step.py [genes] [likelihood] [batches] [u8])."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import scvi_tools_b200 as S
from scvi_tools_b200.data import DeviceResidentCounts
from scvi_tools_b200.train import FusedTrainingPlan

G = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
lik = sys.argv[2] if len(sys.argv) > 2 else "zinb"
nb = int(sys.argv[3]) if len(sys.argv) > 3 else 8
compress = "u8" if (len(sys.argv) > 4 and sys.argv[4] == "u8") else None
N, B = 131072, 8192
dev = torch.device("cuda:0")
gen = torch.Generator(device=dev).manual_seed(0)
lam = torch._standard_gamma(torch.full((1,), 5.0, device=dev).expand(N, G), generator=gen) * (0.7 / 0.3)
X = torch.poisson(lam, generator=gen) * (torch.rand(N, G, device=dev, generator=gen) >= 0.3)
data = DeviceResidentCounts(X, batch=torch.arange(N, device=dev) * nb // N, device=dev, compress=compress)
torch.manual_seed(0)
m = S.VAE(n_input=G, n_batch=nb, gene_likelihood=lik, precision=1).train()
plan = FusedTrainingPlan(m, data, batch_size=B, use_cuda_graph=False)
plan.set_epoch(0)
for i in range(3):
    idx = torch.randperm(N, device=dev)[:B].sort().values
    plan.training_step(idx)
torch.cuda.synchronize()
print("ok", float(plan.out3[0]))
