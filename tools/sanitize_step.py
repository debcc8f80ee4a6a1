"""One full fused step (tcgen05 path, uint8-resident counts) at B = 8192 and at a ragged B, eager launches, for
`compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitize_step.py [B ...]`."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import scvi_tools_b200 as S
from scvi_tools_b200.data import DeviceResidentCounts
from scvi_tools_b200.train import FusedTrainingPlan

dev = torch.device("cuda:0")
G, nb = 2000, 8
sizes = [int(v) for v in sys.argv[1:]] or [8192, 1000]
n_rows = 4 * max(sizes)
gen = torch.Generator(device=dev).manual_seed(0)
X = torch.poisson(torch.rand(n_rows, G, device=dev, generator=gen) * 8.0, generator=gen)
batch = torch.arange(n_rows, device=dev) * nb // n_rows
for B in sizes:
    for compress in ("u8", None):
        torch.manual_seed(0)
        m = S.VAE(n_input=G, n_batch=nb, gene_likelihood="zinb", n_hidden=128, n_latent=10, precision=1).train()
        plan = FusedTrainingPlan(m, DeviceResidentCounts(X, batch=batch, device=dev, compress=compress), batch_size=B,
                                 use_cuda_graph=False)
        plan.set_epoch(0)
        idx = torch.randperm(n_rows, device=dev)[:B].sort().values
        out = plan.training_step(idx)
        torch.cuda.synchronize()
        print(f"B={B} resident={'uint8' if compress else 'float32'} loss={float(out[0]):.4f}", flush=True)
print("done")
