"""GPU: the reference algorithm as EAGER PyTorch fp32 on the same B200, timed beside this library's fused step
("Reference timed beside it" item 2 -- the honest comparator, since no reference Blackwell
kernel exists).

The reference package cannot travel to the GPU box, so the eager arm is the oracle restatement
(`oracle/scvi_oracle.py`, pinned against the unmodified reference by `tests/test_oracle_pins.py`) moved to
`cuda:0`: the same torch ops the reference module issues (`Linear`, `BatchNorm1d` math, `softmax`, `lgamma`,
`softplus`, autograd backward) followed by a per-tensor Adam.  Two data routes, as in the reference:
  (a) `AnnDataLoader`: numpy fancy-index on the host matrix + H2D copy every step
      (src/scvi/data/_anntorchdataset.py:143-152);
  (b) `DeviceBackedDataSplitter`: the matrix resident on the device, `tensor[idx]`
      (src/scvi/dataloaders/_data_splitting.py:589-732).
The oracle is only the thing compared AGAINST here; the product arm is `FusedTrainingPlan.training_step`.
The line is printed (and written to gpurun_out/ when that directory exists); the test asserts only that the
fused step is faster than eager PyTorch on the same GPU.
"""
import functools
import json
import os

import numpy as np
import pytest
import torch

from oracle import scvi_oracle as O

pytestmark = pytest.mark.gpu

N_CELLS, G, NB, H, L = 32768, 2000, 8, 128, 10


@functools.lru_cache(maxsize=1)
def _counts():
    X, batch = O.synthetic_iid(N_CELLS, G, NB, seed=0)
    return np.ascontiguousarray(X, dtype=np.float32), np.asarray(batch).reshape(-1).astype(np.int64)


def _time_steps(fn, warmup, steps):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(steps):
        fn()
    t1.record()
    torch.cuda.synchronize()
    return t0.elapsed_time(t1) / steps


class _EagerStep:
    """zero_grad -> forward -> ELBO -> backward -> Adam, plain torch ops on cuda:0 (oracle restatement)."""

    def __init__(self, P, X_host, batch_host, B, route):
        dev = torch.device("cuda:0")
        self.P = {k: v.to(dev) for k, v in P.items()}
        self.keys = O.trained_keys("zinb")
        self.M = {k: torch.zeros_like(self.P[k]) for k in self.keys}
        self.V = {k: torch.zeros_like(self.P[k]) for k in self.keys}
        self.X_host, self.batch_host = X_host, batch_host
        self.route = route
        if route == "device":
            self.X_dev = torch.from_numpy(X_host).to(dev)
            self.batch_dev = torch.from_numpy(batch_host).to(dev)
        self.B, self.t, self.dev = B, 0, dev
        self.rng = np.random.default_rng(1)

    def __call__(self):
        B = self.B
        idx = self.rng.integers(0, self.X_host.shape[0], size=B)
        if self.route == "host":
            x = torch.from_numpy(O.gather_rows(self.X_host, idx)).to(self.dev)
            bi = torch.from_numpy(self.batch_host[idx]).to(self.dev).view(-1, 1)
        else:
            ii = torch.from_numpy(idx).to(self.dev)
            x = self.X_dev[ii]
            bi = self.batch_dev[ii].view(-1, 1)
        mask = torch.rand(B, H, device=self.dev) >= 0.1
        noise = torch.randn(B, L, device=self.dev)
        out, grads = O.loss_and_grads(self.P, x, bi, n_batch=NB, gene_likelihood="zinb", kl_weight=0.0,
                                      drop_mask=mask, noise=noise)
        self.t += 1
        for k in self.keys:
            self.P[k], self.M[k], self.V[k] = O.adam_step(self.P[k], grads[k], self.M[k], self.V[k], self.t)
        self.P.update(out["bn_buffers"])
        return out["loss"]


@pytest.mark.parametrize("B", [128, 8192])
def test_fused_step_vs_eager_pytorch_on_the_same_gpu(B):
    import scvi_tools_b200 as S
    from scvi_tools_b200.data import DeviceResidentCounts
    from scvi_tools_b200.train import FusedTrainingPlan

    X, batch = _counts()
    torch.manual_seed(0)
    m = S.VAE(n_input=G, n_batch=NB, gene_likelihood="zinb", n_hidden=H, n_latent=L, precision=1).to("cuda:0")
    m.train()
    P = {k: v.detach().clone().cpu() for k, v in m.state_dict().items()}

    steps = 30 if B == 8192 else 100
    eager_dev = _time_steps(_EagerStep(P, X, batch, B, "device"), 5, steps)
    eager_host = _time_steps(_EagerStep(P, X, batch, B, "host"), 3, steps)

    data = DeviceResidentCounts(X, batch=batch, device="cuda:0")
    plan = FusedTrainingPlan(m, data, batch_size=B, use_cuda_graph=True)
    plan.set_epoch(0)
    g = torch.Generator(device="cuda:0").manual_seed(1)
    idxs = [torch.randint(0, N_CELLS, (B,), device="cuda:0", generator=g).sort().values for _ in range(8)]
    it = iter(range(10**9))
    fused = _time_steps(lambda: plan.training_step(idxs[next(it) % 8]), 10, 10 * steps)
    loss = float(plan.out3[0])
    assert np.isfinite(loss)

    line = {
        "what": "SCVI zinb train step on one B200, same GPU for every arm", "cells": N_CELLS, "genes": G, "batch_size": B,
        "fused_ms": fused, "eager_torch_device_resident_ms": eager_dev, "eager_torch_host_gather_ms": eager_host,
        "fused_cells_per_s": B / fused * 1e3, "eager_device_cells_per_s": B / eager_dev * 1e3,
        "eager_host_cells_per_s": B / eager_host * 1e3,
        "speedup_vs_eager_device": eager_dev / fused, "speedup_vs_eager_host": eager_host / fused,
        "eager_arm": "oracle restatement of the reference step as plain torch ops on cuda:0 (fp32, autograd, per-tensor Adam)",
    }
    print("\nSAME_GPU_COMPARATOR " + json.dumps(line))
    if os.path.isdir("gpurun_out"):
        with open(os.path.join("gpurun_out", f"same_gpu_comparator_B{B}.json"), "w") as f:
            f.write(json.dumps(line) + "\n")
    assert fused < eager_dev, line


# ---------------------------------------------------------------------------------------------------------------
# The same eager arm under data parallelism (SURVEY 8d "Reference timed beside it"): what torch
# DDP does for the reference's `Trainer(strategy="ddp")` -- every rank runs the eager step on its own minibatch, the
# gradients are mean-all-reduced over NCCL (DDP's single 25 MB bucket: one flattened all-reduce; the arena is 3 MB),
# every rank applies Adam.  Weak scaling, 8192 cells per rank and step, counts resident on each device.
# ---------------------------------------------------------------------------------------------------------------
def _ddp_eager_worker(rank, world, port, q):
    try:
        import torch.distributed as dist

        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
        B = 8192
        X, batch = _counts()
        import scvi_tools_b200 as S

        torch.manual_seed(0)
        m = S.VAE(n_input=G, n_batch=NB, gene_likelihood="zinb", n_hidden=H, n_latent=L, precision=1)
        P = {k: v.detach().clone().to(dev) for k, v in m.state_dict().items()}
        keys = O.trained_keys("zinb")
        M = {k: torch.zeros_like(P[k]) for k in keys}
        V = {k: torch.zeros_like(P[k]) for k in keys}
        X_dev = torch.from_numpy(X).to(dev)
        b_dev = torch.from_numpy(batch).to(dev)
        gen = torch.Generator(device=dev).manual_seed(100 + rank)
        t = [0]

        def step():
            ii = torch.randint(0, N_CELLS, (B,), device=dev, generator=gen)
            mask = torch.rand(B, H, device=dev, generator=gen) >= 0.1
            noise = torch.randn(B, L, device=dev, generator=gen)
            out, grads = O.loss_and_grads(P, X_dev[ii], b_dev[ii].view(-1, 1), n_batch=NB, gene_likelihood="zinb",
                                          kl_weight=0.0, drop_mask=mask, noise=noise)
            flat = torch.cat([grads[k].reshape(-1) for k in keys])
            dist.all_reduce(flat)
            flat /= world
            t[0] += 1
            off = 0
            for k in keys:
                n = P[k].numel()
                P[k], M[k], V[k] = O.adam_step(P[k], flat[off : off + n].view_as(P[k]), M[k], V[k], t[0])
                off += n
            P.update(out["bn_buffers"])

        for _ in range(5):
            step()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        steps = 20
        e0.record()
        for _ in range(steps):
            step()
        e1.record()
        dist.barrier()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        chk = torch.stack([P[k].double().sum() for k in keys]).sum().view(1)
        both = [torch.empty_like(chk) for _ in range(world)]
        dist.all_gather(both, chk)
        same = all(torch.equal(both[0], b) for b in both)
        q.put((rank, "ok", float(ms), same))
    except Exception:  # noqa: BLE001
        import traceback

        q.put((rank, traceback.format_exc(), 0.0, False))
    finally:
        os._exit(0)


@pytest.mark.timeout(900)
@pytest.mark.parametrize("world", [2, 4, 8])
def test_eager_ddp_comparator_weak_scaling(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs >= {world} GPUs")
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = 29900 + (os.getpid() % 1000) + world
    procs = [ctx.Process(target=_ddp_eager_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=600)
        if p.is_alive():
            p.kill()
            pytest.fail("eager DDP worker hung")
    res = [q.get() for _ in range(world)]
    for r in res:
        assert r[1] == "ok", r[1]
    ms = max(r[2] for r in res)
    assert all(r[3] for r in res), "replicas diverged"
    line = {"what": "reference algorithm as eager PyTorch fp32 + gradient mean all-reduce (DDP semantics), weak scaling",
            "n_gpus": world, "batch_size_per_gpu": 8192, "genes": G, "ms_per_step": ms,
            "cells_per_s": world * 8192 / ms * 1e3, "replicas_identical": True}
    print("\nEAGER_DDP_COMPARATOR " + json.dumps(line))
    if os.path.isdir("gpurun_out"):
        with open(os.path.join("gpurun_out", f"eager_ddp_comparator_n{world}.json"), "w") as f:
            f.write(json.dumps(line) + "\n")
