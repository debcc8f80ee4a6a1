"""GPU (>= 2 devices; skipped on one): the data-parallel step, numerically.

What torch DDP guarantees the reference (`Trainer(strategy="ddp")`, reference tests/model/test_multigpu.py:29-36):
replicas start equal, every step applies the MEAN of the per-replica gradients, replicas stay equal.  Checked for
every `ddp_mode` of `FusedTrainingPlan` (the fused peer-memory all-reduce + Adam "p2p" and the three NCCL
routes):

* replicas built from different seeds hold rank 0's parameters after construction (initial broadcast);
* one step with fed dropout masks / noise: parameters and Adam moments equal -- bit for bit -- Adam applied to
  (g_0 + g_1) / 2, with g_r the gradient a single-process plan computes on rank r's minibatch (BatchNorm
  statistics per replica, as under DDP);
* CUDA-graph replay == eager sequence over several steps with the library's rank-keyed random streams, and the
  parameter arenas are identical on all ranks afterwards.
"""
import os
import traceback

import pytest
import torch

pytestmark = pytest.mark.gpu

# (ddp_mode, p2p_buckets): sharded / one-shot / one-shot in two buckets over peer memory, then the NCCL routes
MODES = [("p2p", 0), ("p2p", 1), ("p2p", 2), ("split", 0), ("overlap", 0), ("graph", 0)]


def _worker(rank, world, port, q):
    try:
        import ctypes as C

        import torch.distributed as dist

        import scvi_tools_b200 as S
        from oracle import scvi_oracle as O
        from scvi_tools_b200._lib import call
        from scvi_tools_b200.data import DeviceResidentCounts
        from scvi_tools_b200.train import FusedTrainingPlan

        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
        pg = dist.group.WORLD
        N, G, nb, B = 4096, 512, 3, 256
        X, batch = O.synthetic_iid(N + 2, G, nb, seed=0)
        data = DeviceResidentCounts(X, batch=batch, device=dev)
        n_obs = data.n_obs

        def build(pgroup, graph, mode="p2p", lik="zinb", seed=0, buckets=0):
            torch.manual_seed(seed)
            m = S.VAE(n_input=G, n_batch=nb, gene_likelihood=lik, precision=1).train()
            kw = {} if pgroup is None else {"process_group": pgroup, "ddp_mode": mode, "p2p_buckets": buckets}
            plan = FusedTrainingPlan(m, data, batch_size=B, n_epochs_kl_warmup=None, max_kl_weight=0.5,
                                     use_cuda_graph=graph, seed=3, **kw)
            plan.set_epoch(0)
            return plan

        def inputs(r, step):
            g = torch.Generator().manual_seed(1000 * step + r)
            idx = torch.randperm(n_obs, generator=g)[:B].sort().values
            mask = (torch.rand(B, 128, generator=g) >= 0.1).to(torch.uint8)
            noise = torch.randn(B, 10, generator=g)
            return idx, mask, noise

        for mode, buckets in MODES:
            # ---- initial broadcast + one fed step against single-process arithmetic
            plan = build(pg, False, mode, seed=11 * rank, buckets=buckets)  # deliberately different initial weights per rank
            if mode == "p2p":
                assert plan.exchange is not None and plan.ddp_mode == "p2p", "peer memory could not be mapped"
            p0 = plan.p.clone()
            both = [torch.empty_like(p0) for _ in range(world)]
            dist.all_gather(both, p0)
            assert all(torch.equal(both[0], b) for b in both), f"{mode}: replicas differ after construction"
            idx, mask, noise = inputs(rank, 0)
            plan.training_step(idx.to(dev), drop_mask=mask, noise=noise)
            torch.cuda.synchronize()
            gsum = torch.zeros_like(p0)
            for r in range(world):
                single = build(None, False)
                single.p.copy_(p0)
                ix, mk, nz = inputs(r, 0)
                single.training_step(ix.to(dev), drop_mask=mk, noise=nz)
                gsum += single.g  # world == 2: a + b is order-independent
            exp_p, exp_m, exp_v = p0.clone(), torch.zeros_like(p0), torch.zeros_like(p0)
            st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
            call("scvi_adam_step", C.c_void_p(exp_p.data_ptr()), C.c_void_p(gsum.data_ptr()), C.c_void_p(exp_m.data_ptr()),
                 C.c_void_p(exp_v.data_ptr()), exp_p.numel(), 1, 1e-3, 0.9, 0.999, 0.01, 1e-6, 1.0 / world, st)
            torch.cuda.synchronize()
            assert float((exp_p - p0).abs().max()) > 0
            lo, hi = 0, p0.numel()
            if mode == "p2p" and buckets == 0:  # sharded: this rank keeps the Adam moments of its own slice only
                lo, hi = plan.exchange.shard_bounds()
                assert 0 <= lo < hi <= p0.numel()
            assert torch.equal(plan.p, exp_p), (mode, buckets, "p", float((plan.p - exp_p).abs().max()))
            for name, got, want in (("m", plan.m, exp_m), ("v", plan.v, exp_v)):
                assert torch.equal(got[lo:hi], want[lo:hi]), (mode, buckets, name, float((got - want).abs().max()))
            assert int(plan.step_counter[0]) == 1
            if hasattr(plan, "close"):
                plan.close()
            # ---- graph replay == eager sequence; replicas stay identical
            end = []
            for graph in (True, False):
                plan = build(pg, graph, mode, buckets=buckets)
                for step in range(5):
                    ix, _, _ = inputs(rank, 10 + step)
                    plan.training_step(ix.to(dev))
                torch.cuda.synchronize()
                lo, hi = plan.exchange.shard_bounds() if (mode == "p2p" and buckets == 0) else (0, plan.p.numel())
                end.append((plan.p.clone(), plan.m[lo:hi].clone(), plan.v[lo:hi].clone(), plan.bn1_rm.clone(),
                            plan.bn2_rv.clone()))
                assert int(plan.step_counter[0]) == 5
                plan.close()
            for a, b in zip(end[0], end[1], strict=True):
                assert torch.equal(a, b), f"{mode}/{buckets}: graph replay differs from the eager sequence"
            both = [torch.empty_like(end[0][0]) for _ in range(world)]
            dist.all_gather(both, end[0][0])
            assert all(torch.equal(both[0], b) for b in both), f"{mode}: replicas diverged"
            assert torch.isfinite(end[0][0]).all()
        dist.barrier()
        q.put((rank, "ok"))
    except Exception:  # noqa: BLE001
        q.put((rank, traceback.format_exc()))
    finally:
        os._exit(0)  # graph-captured NCCL kernels can block communicator teardown


@pytest.mark.timeout(900)
def test_two_rank_step_matches_single_process_arithmetic():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = 29700 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=800)
        if p.is_alive():
            p.kill()
            pytest.fail("data-parallel worker hung")
    res = dict(q.get() for _ in range(2))
    for r in range(2):
        assert res[r] == "ok", f"rank {r}:\n{res[r]}"


def _adv_worker(rank, world, port, q):
    """Data-parallel `AdversarialTrainingPlan` (one all-reduce of each optimizer's flat gradient arena per step):
    replicas built from different seeds start from rank 0's weights; after one step the parameters equal Adam applied
    to the mean of the two single-process gradients, bit for bit; replicas stay identical over further steps, with
    and without the whole-step CUDA graph."""
    try:
        import ctypes as C

        import numpy as np
        import torch.distributed as dist

        import scvi_tools_b200 as S
        from scvi_tools_b200 import nn as N
        from scvi_tools_b200._lib import call
        from scvi_tools_b200.adversarial import AdversarialTrainingPlan
        from scvi_tools_b200.totalvi import TOTALVAE

        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
        pg = dist.group.WORLD

        def mask(shape, p, device):  # fixed, capture-safe stand-ins for the random draws
            n = int(np.prod(shape))
            return ((torch.arange(n, device=device) * 7919) % 10 != 0).to(torch.uint8).view(shape)

        def noise(shape, device):
            n = int(np.prod(shape))
            return torch.sin(torch.arange(n, device=device, dtype=torch.float32) * 0.37).view(shape)

        import scvi_tools_b200.module as M_
        import scvi_tools_b200.totalvi as T_

        N._next_mask, N._next_noise = mask, noise
        M_._next_noise = T_._next_noise = noise  # the modules bound the draw function by name at import
        rng = np.random.default_rng(1)
        n, G, P, nb, B = 1024, 80, 10, 3, 64
        X = torch.from_numpy(rng.poisson(3.0, size=(n, G)).astype(np.float32)).to(dev)
        Y = torch.from_numpy(rng.poisson(30.0, size=(n, P)).astype(np.float32)).to(dev)
        bat = torch.from_numpy(rng.integers(0, nb, size=(n, 1))).to(dev)
        lab = torch.zeros(n, 1, dtype=torch.long, device=dev)

        def mb(r, t):
            s = slice((2 * t + r) * B, (2 * t + r + 1) * B)
            return {"X": X[s], "proteins": Y[s], "batch": bat[s], "labels": lab[s]}

        def build(seed, group, graph=False):
            torch.manual_seed(seed)
            m = TOTALVAE(n_input_genes=G, n_input_proteins=P, n_batch=nb, n_panel=nb, n_hidden=64).to(dev)
            plan = AdversarialTrainingPlan(m, lr=4e-3, adversarial_classifier=True, n_epochs_kl_warmup=16,
                                           use_cuda_graph=graph, process_group=group)
            plan.current_epoch = 3
            return plan

        ddp = build(11 * rank + 5, pg)  # different initial weights per rank before the broadcast
        sd0 = {k: v.detach().clone() for k, v in ddp.module.state_dict().items()}
        cd0 = {k: v.detach().clone() for k, v in ddp.adversarial_classifier.state_dict().items()}
        both = [torch.empty_like(sd0["px_r"]) for _ in range(world)]
        dist.all_gather(both, sd0["px_r"].contiguous())
        assert torch.equal(both[0], both[1]), "replicas differ after construction"
        ddp.training_step(mb(rank, 0))
        torch.cuda.synchronize()
        # expectation: single-process plans from the same state on each rank's minibatch -> mean gradient -> Adam
        exp = []
        for r in range(world):
            sp = build(0, None)
            sp.module.load_state_dict(sd0)
            sp.adversarial_classifier.load_state_dict(cd0)
            sp.training_step(mb(r, 0))
            torch.cuda.synchronize()
            exp.append(sp)
        assert exp[0].opt1.names == ddp.opt1.names
        for which in ("opt1", "opt2"):
            g = sum(getattr(e, which).arena["g"] for e in exp)
            ref = build(0, None)
            ref.module.load_state_dict(sd0)
            ref.adversarial_classifier.load_state_dict(cd0)
            ref.training_step(mb(0, 0))  # lays the arenas out (values overwritten below)
            a = getattr(ref, which).arena
            named = dict((ref.module if which == "opt1" else ref.adversarial_classifier).named_parameters())
            src = sd0 if which == "opt1" else cd0
            for k, off in zip(getattr(ref, which).names, a["offs"], strict=True):
                a["p"][off : off + named[k].numel()].copy_(src[k].reshape(-1))
            a["m"].zero_()
            a["v"].zero_()
            o = getattr(ref, which)
            st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
            call("scvi_adam_step", C.c_void_p(a["p"].data_ptr()), C.c_void_p(g.data_ptr()), C.c_void_p(a["m"].data_ptr()),
                 C.c_void_p(a["v"].data_ptr()), a["p"].numel(), 1, o.lr, o.betas[0], o.betas[1], o.eps, o.wd, 1.0 / world, st)
            torch.cuda.synchronize()
            got = getattr(ddp, which).arena["p"]
            assert torch.equal(got, a["p"]), (which, float((got - a["p"]).abs().max()))
        # replicas stay identical, eager and as a whole-step graph (NCCL all-reduce captured with the step)
        for graph in (False, True):
            plan = build(3, pg, graph)
            for t in range(6):
                plan.current_epoch = t
                plan.training_step(mb(rank, t))
            torch.cuda.synchronize()
            if graph:
                assert plan._graph.n_replays == 3
            p = plan.opt1.arena["p"]
            both = [torch.empty_like(p) for _ in range(world)]
            dist.all_gather(both, p)
            assert torch.equal(both[0], both[1]) and torch.isfinite(p).all(), f"replicas diverged (graph={graph})"
        dist.barrier()
        q.put((rank, "ok"))
    except Exception:  # noqa: BLE001
        q.put((rank, traceback.format_exc()))
    finally:
        os._exit(0)


@pytest.mark.timeout(900)
def test_two_rank_adversarial_plan_matches_single_process_arithmetic():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = 29800 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_adv_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=800)
        if p.is_alive():
            p.kill()
            pytest.fail("data-parallel worker hung")
    res = dict(q.get() for _ in range(2))
    for r in range(2):
        assert res[r] == "ok", f"rank {r}:\n{res[r]}"
