"""Throughput lines for the widened workloads of BASELINE.json (configs[3] SCANVI, configs[4] TOTALVI) -- not the
headline bench (bench.py), same timing rules: warm-up, CUDA events on the launching stream, median of blocks,
resident counts gathered at random from a matrix larger than L2.

    python tools/bench_workloads.py scanvi  [--cells 500000 --genes 2000 --labels 20 --batch-size 1024]
    python tools/bench_workloads.py totalvi [--cells 200000 --genes 4000 --proteins 100 --batch-size 256]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch


def counts(n, g, dev, seed=0, chunk=65536):
    gen = torch.Generator(device=dev).manual_seed(seed)
    X = torch.empty(n, g, dtype=torch.float32, device=dev)
    conc = torch.full((1,), 5.0, device=dev)
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        lam = torch._standard_gamma(conc.expand(e - s, g), generator=gen) * (0.7 / 0.3)
        X[s:e] = torch.poisson(lam, generator=gen) * (torch.rand(e - s, g, device=dev, generator=gen) >= 0.3)
    return X


WORLD = int(os.environ.get("WORLD_SIZE", "1"))
RANK = int(os.environ.get("RANK", "0"))
PG = None


def barrier():
    if WORLD > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()


def timed(step, K, R, W):
    for i in range(W):
        step(i)
    barrier()
    out = []
    n = W
    for _ in range(R):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _i in range(K):
            step(n)
            n += 1
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if WORLD > 1:  # max over ranks
            t = torch.tensor([ms], device="cuda")
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            ms = float(t)
        out.append(ms)
    return float(np.median(out)), out


def scanvi(a, dev):
    from scvi_tools_b200.data import DeviceResidentCounts
    from scvi_tools_b200.scanvi import SCANVAE
    from scvi_tools_b200.semisupervised import SemiSupervisedTrainingPlan

    X = counts(a.cells, a.genes, dev)
    g = torch.Generator().manual_seed(1)
    labels = torch.randint(0, a.labels, (a.cells,), generator=g)
    labels[torch.rand(a.cells, generator=g) < 0.5] = a.labels  # half of the cells unlabelled
    data = DeviceResidentCounts(X, batch=torch.arange(a.cells) * 4 // a.cells, labels=labels, device=dev)
    torch.manual_seed(0)
    m = SCANVAE(n_input=a.genes, n_batch=4, n_labels=a.labels, gene_likelihood="zinb").to(dev)
    plan = SemiSupervisedTrainingPlan(m, n_classes=a.labels, use_cuda_graph=not a.eager, process_group=PG)
    B = a.batch_size
    lab_idx = torch.nonzero(labels != a.labels).view(-1).to(dev)

    def step(i):
        gi = torch.Generator(device=dev).manual_seed(1000 * RANK + i)
        f = torch.randint(0, a.cells, (B,), device=dev, generator=gi)
        l = lab_idx[torch.randint(0, lab_idx.numel(), (B,), device=dev, generator=gi)]
        plan.training_step((data.minibatch(f), data.minibatch(l)))

    ms, blocks = timed(step, a.steps, 3, 5)
    return {"workload": f"SCANVI zinb semi-supervised step, synthetic {a.cells} cells x {a.genes} genes, {a.labels} labels, "
                        f"full + labelled minibatch of {B} cells each, classification_ratio 50",
            "metric": "SCANVI train cells/sec (full-minibatch cells)", "value": WORLD * a.steps * B / (ms / 1e3), "unit": "cells/s",
            "ms_per_step": ms / a.steps, "steps": a.steps, "blocks_ms": blocks, "path": "autograd over the library's kernels "
            "(SCANVAE.forward) + multi-tensor Adam; the [B,G] side is the fused decoder + likelihood kernel"}


def totalvi(a, dev):
    from scvi_tools_b200.adversarial import AdversarialTrainingPlan
    from scvi_tools_b200.totalvi import TOTALVAE

    X = counts(a.cells, a.genes, dev)
    Y = counts(a.cells, a.proteins, dev, seed=3) * 4
    nb = 4
    batch = (torch.arange(a.cells, device=dev) * nb // a.cells).view(-1, 1)
    torch.manual_seed(0)
    m = TOTALVAE(n_input_genes=a.genes, n_input_proteins=a.proteins, n_batch=nb).to(dev)
    plan = AdversarialTrainingPlan(m, lr=4e-3, adversarial_classifier=True, n_batch=nb, use_cuda_graph=not a.eager,
                                   process_group=PG)
    B = a.batch_size

    def step(i):
        gi = torch.Generator(device=dev).manual_seed(1000 * RANK + i)
        f = torch.randint(0, a.cells, (B,), device=dev, generator=gi)
        t = {"X": X[f], "proteins": Y[f], "batch": batch[f], "labels": torch.zeros(B, 1, dtype=torch.long, device=dev)}
        plan.training_step(t)

    ms, blocks = timed(step, a.steps, 3, 5)
    return {"workload": f"TOTALVI step (adversarial plan), synthetic {a.cells} cells x {a.genes} genes + {a.proteins} proteins, "
                        f"batch_size {B}", "metric": "TOTALVI train cells/sec", "value": WORLD * a.steps * B / (ms / 1e3),
            "unit": "cells/s", "ms_per_step": ms / a.steps, "steps": a.steps, "blocks_ms": blocks}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("which", choices=["scanvi", "totalvi"])
    ap.add_argument("--cells", type=int, default=None)
    ap.add_argument("--genes", type=int, default=None)
    ap.add_argument("--labels", type=int, default=20)
    ap.add_argument("--proteins", type=int, default=100)
    ap.add_argument("--batch-size", type=int, default=None)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--eager", action="store_true", help="no whole-step CUDA graph")
    a = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if WORLD > 1:
        torch.distributed.init_process_group("nccl", device_id=dev)
        PG = torch.distributed.group.WORLD
    if a.which == "scanvi":
        a.cells, a.genes, a.batch_size = a.cells or 500_000, a.genes or 2000, a.batch_size or 1024
        line = scanvi(a, dev)
    else:
        a.cells, a.genes, a.batch_size = a.cells or 200_000, a.genes or 4000, a.batch_size or 256
        line = totalvi(a, dev)
    line.update({"n_gpus": WORLD, "data": "synthetic", "dtype": "f32", "whole_step_cuda_graph": not a.eager,
                 "scaling": "weak", "parallelism": f"dp{WORLD}: one all-reduce of the flat gradient arena per optimizer"})
    if RANK == 0:
        print(json.dumps(line))
    if WORLD > 1:
        torch.distributed.barrier()
        os._exit(0)
