#!/usr/bin/env python
"""bench.py -- SCVI train cells/s on synthetic counts (BASELINE.json metric), one JSON line.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference algorithm on host cores

Workload (config.workload): SCVI ZINB VAE, 1,000,000 cells x 2,000 genes, 8 batches, n_hidden 128,
n_latent 10, counts NB(5, .3) x Bernoulli(.7) (the synthetic_iid recipe), fp32.  One "step" is one
minibatch training step (gather -> forward -> ELBO -> backward -> Adam) on `--batch-size` cells.
`value` times K steps with the count matrix resident in HBM (8 GB, random row gather: inputs far
larger than L2); `e2e` times the same step fed from pinned HOST minibatches (H2D inside the timed
region, loss read back every step).  Under torchrun each rank holds a replica of the matrix and
trains its own shard of every global batch (weak scaling, gradient allreduce over NCCL).
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

# NCCL's version banner (NCCL_DEBUG=VERSION/INFO on some boxes) must not land on stdout next to the JSON line
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")

import numpy as np  # noqa: E402
import torch  # noqa: E402

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_CELLS, N_GENES, N_BATCH, N_HIDDEN, N_LATENT = 1_000_000, 2000, 8, 128, 10


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch-size", type=int, default=8192)
    ap.add_argument("--cells", type=int, default=N_CELLS)
    ap.add_argument("--genes", type=int, default=N_GENES)
    ap.add_argument("--likelihood", default="zinb", choices=["zinb", "nb"])
    ap.add_argument("--batches", type=int, default=N_BATCH, help="number of batch-covariate categories")
    ap.add_argument("--precision", default="tc", choices=["tc", "fp32"],
                    help="GEMM arithmetic of the fused decoder kernels: tc = tcgen05 error-compensated fp16 split "
                         "(fp32-grade accuracy, parity-tested), fp32 = FFMA")
    ap.add_argument("--unsorted-batches", action="store_true",
                    help="feed each minibatch's row ids in shuffled order (default: ascending within the minibatch, "
                         "as FusedTrainingPlan.fit does -- same cells, better DRAM page / TLB locality)")
    ap.add_argument("--ddp-mode", default="p2p", choices=["p2p", "graph", "overlap", "split"],
                    help="multi-GPU gradient exchange (FusedTrainingPlan.ddp_mode): p2p = this library's fused "
                         "all-reduce + Adam over NVLink peer memory, the others = NCCL routes")
    ap.add_argument("--p2p-buckets", type=int, default=None,
                    help="FusedTrainingPlan.p2p_buckets (ddp-mode p2p): default = one-shot up to 2 GPUs, sharded beyond")
    ap.add_argument("--blocks", type=int, default=5,
                    help="the timed region is repeated this many times (each EXACTLY --steps steps between "
                         "barrier + synchronize); ms_per_step is the median block")
    ap.add_argument("--zero-frac", type=float, default=0.3,
                    help="fraction of structural zeros in the synthetic counts (0.3 = synthetic_iid; 0.9 ~ real scRNA)")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the extra workloads of the default line (B=128, 90%%-zeros counts, config C3)")
    ap.add_argument("--e2e-dtype", default="auto", choices=["auto", "u16", "f32"],
                    help="host count type of the e2e leg: auto = narrowest lossless (uint8 / uint16 / float32)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--hybrid-fraction", type=float, default=0.45,
                    help="e2e hybrid route: fraction of every minibatch the GPU pulls from the pinned host matrix "
                         "itself while the host threads gather the rest (0 = skip the route)")
    ap.add_argument("--resident", default="auto", choices=["auto", "u8", "f32"],
                    help="type of the HBM-resident count matrix: uint8 when every count is < 256 (auto), or fp32")
    ap.add_argument("--value-only", action="store_true",
                    help="tuning aid: time the resident-counts step only (no e2e / roofline / baseline legs)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    return ap.parse_args()


def workload_name(a):
    return (f"SCVI {a.likelihood} VAE train step, synthetic {a.cells} cells x {a.genes} genes, {a.batches} batches, "
            f"n_hidden={N_HIDDEN}, n_latent={N_LATENT}, batch_size={a.batch_size} per GPU")


# ----------------------------------------------------------------------------------------------
# synthetic data
# ----------------------------------------------------------------------------------------------
def synth_counts_device(n, g, device, seed=0, chunk=65536, zero_frac=0.3):
    """NB(5, 0.3) x Bernoulli(1 - zero_frac) (reference src/scvi/data/_built_in_data/_synthetic.py:54-58 with
    zero_frac = 0.3), generated on the device in chunks: Gamma(5, scale=(1-p)/p) -> Poisson."""
    gen = torch.Generator(device=device).manual_seed(seed)
    X = torch.empty(n, g, dtype=torch.float32, device=device)
    conc = torch.full((1,), 5.0, device=device)
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        lam = torch._standard_gamma(conc.expand(e - s, g), generator=gen) * (0.7 / 0.3)
        cnt = torch.poisson(lam, generator=gen)
        keep = torch.rand(e - s, g, device=device, generator=gen) >= zero_frac
        X[s:e] = cnt * keep
    return X


def config_dict(a, world):
    """The `config` object both arms print (the reference arm times the same workload on the host cores)."""
    return {"workload": workload_name(a), "parallelism": f"dp{world}", "global_batch": a.batch_size * world,
            "cells": a.cells, "genes": a.genes, "batches": a.batches, "likelihood": a.likelihood,
            "batch_size_per_gpu": a.batch_size, "n_hidden": N_HIDDEN, "n_latent": N_LATENT}


def synth_counts_host(n, g, seed=0):
    rng = np.random.default_rng(seed)
    X = rng.negative_binomial(5, 0.3, size=(n, g)) * rng.binomial(1, 0.7, size=(n, g))
    return X.astype(np.float32)


# ----------------------------------------------------------------------------------------------
# CPU arm: the reference algorithm (oracle port) on host cores
# ----------------------------------------------------------------------------------------------
class CpuReferenceStep:
    """`zero_grad -> module(batch, kl_weight) -> backward -> Adam(lr=1e-3, eps=.01, wd=1e-6)` restated by
    oracle/scvi_oracle.py (the reference itself cannot travel to the GPU box)."""

    def __init__(self, a, n_rows):
        from oracle import scvi_oracle as O

        self.O = O
        torch.manual_seed(0)
        import scvi_tools_b200 as S  # only to obtain identically-initialised parameters (host tensors)

        m = S.VAE(n_input=a.genes, n_batch=a.batches, gene_likelihood=a.likelihood, n_hidden=N_HIDDEN, n_latent=N_LATENT)
        self.P = {k: v.detach().clone() for k, v in m.state_dict().items()}
        self.keys = O.trained_keys(a.likelihood)
        self.M = {k: torch.zeros_like(self.P[k]) for k in self.keys}
        self.V = {k: torch.zeros_like(self.P[k]) for k in self.keys}
        self.X = synth_counts_host(n_rows, a.genes, seed=0)
        self.batch = (np.arange(n_rows) * a.batches // n_rows).astype(np.int64)
        self.a = a
        self.t = 0
        self.rng = np.random.default_rng(1)
        self.g = torch.Generator().manual_seed(2)

    def step(self):
        a, O = self.a, self.O
        B = a.batch_size
        idx = self.rng.integers(0, self.X.shape[0], size=B)
        x = torch.from_numpy(O.gather_rows(self.X, idx))  # AnnTorchDataset.__getitem__
        bi = torch.from_numpy(self.batch[idx]).view(-1, 1)
        mask = torch.rand(B, N_HIDDEN, generator=self.g) >= 0.1
        noise = torch.randn(B, N_LATENT, generator=self.g)
        out, grads = O.loss_and_grads(self.P, x, bi, n_batch=a.batches, gene_likelihood=a.likelihood, kl_weight=0.0,
                                      drop_mask=mask, noise=noise)
        self.t += 1
        for k in self.keys:
            self.P[k], self.M[k], self.V[k] = O.adam_step(self.P[k], grads[k], self.M[k], self.V[k], self.t)
        self.P.update(out["bn_buffers"])
        return float(out["loss"].detach())


def cpu_baseline(a, seconds):
    """Bounded sample: as many full-size steps as fit in ~`seconds` (after 1 warm-up step)."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ref = CpuReferenceStep(a, n_rows=max(4 * a.batch_size, 20000))
    ref.step()
    n, t0 = 0, time.perf_counter()
    while True:
        ref.step()
        n += 1
        el = time.perf_counter() - t0
        if el >= seconds or n >= 50:
            break
    return {"value": n * a.batch_size / el, "unit": "cells/s", "cores": cores, "kind": "port",
            "sample": f"{n} full steps of {a.batch_size} cells on a {ref.X.shape[0]}-cell host matrix ({el:.1f} s)"}


def run_reference_arm(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ref = CpuReferenceStep(a, n_rows=max(4 * a.batch_size, 20000))
    steps = a.steps
    for _ in range(a.warmup):
        ref.step()
    # bound the run: the CPU step at B=8192 takes O(1 s)
    t0 = time.perf_counter()
    done = 0
    for _ in range(steps):
        ref.step()
        done += 1
        if time.perf_counter() - t0 > 150:
            break
    el = time.perf_counter() - t0
    v = done * a.batch_size / el
    line = {
        "impl": "reference", "metric": "SCVI train cells/sec", "value": v, "unit": "cells/s", "n_gpus": a.gpus,
        "steps": done, "warmup": a.warmup, "ms_per_step": 1e3 * el / done, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(a, int(os.environ.get("WORLD_SIZE", "1"))),
        "details": {"timing": "host perf_counter; CPU arm (rank 0 only)"},
        "cpu_baseline": {"value": v, "unit": "cells/s", "cores": cores, "kind": "port",
                         "sample": f"{done} steps of {a.batch_size} cells (oracle port of the reference step, torch CPU, {cores} threads)"},
        "e2e": {"value": v, "unit": "cells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """Samples SM clocks / throttle reasons during the timed region (NVML every 20 ms; nvidia-smi
    subprocess every 100 ms if NVML cannot be loaded)."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.sm, self.mx, self.reasons = index, [], 0.0, set()
        self._stop_evt = threading.Event()
        self.nv = self.h = None
        try:
            import pynvml as nv

            nv.nvmlInit()
            self.nv, self.h = nv, nv.nvmlDeviceGetHandleByIndex(index)
            self.mx = float(nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM))
            self.bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown,
                         "hw_thermal_slowdown": nv.nvmlClocksEventReasonHwThermalSlowdown,
                         "sw_thermal_slowdown": nv.nvmlClocksEventReasonSwThermalSlowdown,
                         "sw_power_cap": nv.nvmlClocksEventReasonSwPowerCap}
        except Exception:
            self.nv = None

    def _sample_nvml(self):
        nv = self.nv
        self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
        r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        for name, bit in self.bits.items():
            if r & bit:
                self.reasons.add(name)

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                              "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
        r = [c.strip() for c in out.strip().split(",")]
        self.sm.append(float(r[0]))
        self.mx = max(self.mx, float(r[1]))
        for name, v in zip(self.NAMES, r[2:6]):
            if v.lower().startswith("active"):
                self.reasons.add(name)

    def run(self):
        while not self._stop_evt.is_set():
            try:
                if self.nv is not None:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                pass
            self._stop_evt.wait(0.02 if self.nv is not None else 0.1)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=5)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx or None,
                "reasons": sorted(self.reasons), "samples": len(self.sm),
                "source": "nvml" if self.nv is not None else "nvidia-smi"}


def algorithmic_bytes(a, B, n_arena, xbytes=4):
    """SURVEY.md section 8(d) / DESIGN.md section 4: algorithmic HBM bytes per launch of the main kernels (every input
    read once, every output written once), keyed by the launch label.  `xbytes` = bytes per count: 4 is SURVEY's
    figure (fp32 operands, what `roofline.achieved` is defined on); 1 is what the kernels move with uint8-resident
    counts (reported beside it)."""
    G, H = a.genes, N_HIDDEN
    heads = 2 if a.likelihood == "zinb" else 1
    W = heads * 4 * G * H  # head weights, fp32
    xb = xbytes
    return {
        # K2: x rows + h + (lse, library) in; dh + (reconst, S) out; head weights + 5 per-gene constants in
        "scvi_decoder_nll_fwd(tc nll+dh)": B * (xb * G + 4 * H + 16) + B * (4 * H + 8) + W + 20 * G,
        # K1: h + scale-head weights + bias in; (max, sumexp) out
        "scvi_decoder_nll_fwd(tc lse)": B * (4 * H) + 4 * G * H + 4 * G + 8 * B,
        # K3: x rows + h + 4 row constants in; head weights in; dW, db, dpx_r out
        "scvi_decoder_nll_bwd(tc dW)": B * (xb * G + 4 * H + 16) + W + 20 * G + W + heads * 4 * G + 4 * G,
        # E1: x rows + W in; y + library out
        "scvi_linear_fwd(tc)": B * (xb * G + 8) + 4 * G * H + B * (4 * H + 4),
        # E2: x rows + dy in; dW out
        "scvi_linear_bwd(tc dW)": B * (xb * G + 8 + 4 * H) + 4 * G * H,
        "scvi_gather_rows": 2 * 4 * B * G + 8 * B,
        "scvi_adam_step_graph": 28 * n_arena,
    }


def profile_step(plan, a, reps=20):
    """Per-kernel durations of the eager step, measured with CUDA events recorded by the library behind
    every launch (scvi_profile_begin/_end) on the launching stream.  Rows are gathered at random from the
    resident matrix (> L2) and a delay kernel keeps the launch queue full, so consecutive events bracket
    one kernel.  Returns {label: {ms, launches_per_step}}."""
    import ctypes as C

    from scvi_tools_b200._lib import call, load

    lib = load()
    X = plan.data.X
    state = (plan.p, plan.m, plan.v, plan.step_counter, plan.bn1_rm, plan.bn1_rv, plan.bn2_rm, plan.bn2_rv,
             plan.bn1_nbt, plan.bn2_nbt)
    snap = [t.clone() for t in state]
    out = torch.empty(plan.B, plan.G, device=plan.dev)
    world, plan.world = plan.world, 1  # this leg runs on rank 0 only: no collectives
    tot, cnt = {}, {}
    buf = C.create_string_buffer(1 << 16)
    for r in range(reps + 2):
        plan.idx.copy_(torch.randint(0, X.shape[0], (plan.B,), device=plan.dev))
        torch.cuda.synchronize()
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        call("scvi_profile_begin", st, 3000)
        plan._prologue(st)  # the step without the data-parallel all-reduce (this leg runs on rank 0 only)
        plan._forward_backward(st)
        plan._adam(st)
        call("scvi_gather_rows_u8" if X.dtype == torch.uint8 else "scvi_gather_rows", X.data_ptr(), X.shape[0], plan.G,
             X.stride(0), plan.idx.data_ptr(), plan.B, out.data_ptr(), plan.G, st)
        n = lib.scvi_profile_end(buf, len(buf))
        if n < 0:
            raise RuntimeError("scvi_profile_end: " + lib.scvi_last_error_string().decode())
        if r < 2:
            continue
        for ln in buf.value.decode().splitlines():
            name, c, ms = ln.split("\t")
            tot[name] = tot.get(name, 0.0) + float(ms)
            cnt[name] = cnt.get(name, 0) + int(c)
    for dst, src in zip(state, snap, strict=True):
        dst.copy_(src)
    plan.world = world
    return {k: {"ms": tot[k] / cnt[k], "launches_per_step": cnt[k] / reps} for k in tot}


def time_logprob_kernels(a, dev, peak, reps=10):
    """BASELINE.json's second metric: achieved HBM GB/s of the standalone NB / ZINB log-likelihood kernels
    (`NegativeBinomial.log_prob` on materialised mu/theta/pi) on [32768, G] operands (> L2), CUDA events."""
    import ctypes as C

    from scvi_tools_b200._lib import call

    B, G = 32768, a.genes
    gen = torch.Generator(device=dev).manual_seed(5)
    x = torch.poisson(torch.full((B, G), 3.0, device=dev), generator=gen)
    mu = torch.rand(B, G, device=dev, generator=gen) * 5 + 0.01
    pi = torch.randn(B, G, device=dev, generator=gen)
    theta = torch.rand(G, device=dev, generator=gen) * 3 + 0.1
    out, go = torch.empty(B, G, device=dev), torch.ones(B, G, device=dev)
    dmu, dpi = torch.empty(B, G, device=dev), torch.empty(B, G, device=dev)
    dth = torch.empty((B + 63) // 64, G, device=dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    legs = {
        "nb_logprob_fwd": (12 * B * G + 4 * G, lambda: call("scvi_nb_logprob_fwd", p(x), G, p(mu), G, p(theta), 0, B, G, 1e-8,
                                                           p(out), G, None, st)),
        "zinb_logprob_fwd": (16 * B * G + 4 * G, lambda: call("scvi_zinb_logprob_fwd", p(x), G, p(mu), G, p(theta), 0, p(pi), G,
                                                             B, G, 1e-8, p(out), G, None, st)),
        "nb_logprob_bwd": (20 * B * G, lambda: call("scvi_nb_logprob_bwd", p(x), G, p(mu), G, p(theta), 0, p(go), G, B, G, 1e-8,
                                                     p(dmu), p(dth), st)),
        "zinb_logprob_bwd": (28 * B * G, lambda: call("scvi_zinb_logprob_bwd", p(x), G, p(mu), G, p(theta), 0, p(pi), G, p(go), G,
                                                       B, G, 1e-8, p(dmu), p(dth), p(dpi), st)),
    }
    res = {}
    for name, (nbytes, fn) in legs.items():
        ts = []
        for r in range(reps + 3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            e1.synchronize()
            if r >= 3:
                ts.append(e0.elapsed_time(e1))
        ms = float(np.median(ts))
        res[name] = {"ms": ms, "algorithmic_bytes": nbytes, "GBps": nbytes / ms / 1e6, "frac_of_hbm_peak": nbytes / ms / 1e6 / peak,
                     "shape": [B, G]}
    return res


def timed_blocks(step, K, R, barrier, world, dev):
    """R repetitions of the contract's timed region: barrier + synchronize, EXACTLY K steps between two CUDA events
    on the launching stream, barrier + synchronize; each block's time is the max over ranks.  Returns the list (ms)."""
    out, n = [], 0
    for _ in range(R):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _i in range(K):
            step(n)
            n += 1
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            ms = float(t)
        out.append(ms)
    return out


def resident_type(a, X):
    """"u8" when the resident matrix can be kept as uint8 counts (tcgen05 kernels, every count < 256), else None."""
    if a.resident == "f32" or a.precision != "tc":
        return None
    ok = float(X.max()) < 256.0
    if a.resident == "u8" and not ok:
        raise SystemExit("--resident u8: counts >= 256 present")
    return "u8" if ok else None


def make_plan(a, data, B, pg, genes=None, likelihood=None, batches=None, seed=0):
    import scvi_tools_b200 as S
    from scvi_tools_b200.train import FusedTrainingPlan

    torch.manual_seed(seed)
    m = S.VAE(n_input=genes or a.genes, n_batch=batches or a.batches, gene_likelihood=likelihood or a.likelihood,
              n_hidden=N_HIDDEN, n_latent=N_LATENT, precision=1 if a.precision == "tc" else 0).train()
    plan = FusedTrainingPlan(m, data, batch_size=B, process_group=pg, use_cuda_graph=True, ddp_mode=a.ddp_mode,
                             p2p_buckets=a.p2p_buckets)
    plan.set_epoch(0)
    return plan


def minibatch_order(n_cells, B, dev, sort=True, seed=0):
    g = torch.Generator(device=dev).manual_seed(seed)
    order = torch.randperm(n_cells, device=dev, generator=g)
    n_chunks = n_cells // B
    if sort:  # a minibatch is a set: ascending row ids inside each one (as FusedTrainingPlan.fit does)
        order[: n_chunks * B] = order[: n_chunks * B].view(n_chunks, B).sort(dim=1).values.reshape(-1)
    return order, n_chunks


def quick_rate(plan, n_cells, B, dev, K, R, W, barrier, world):
    """cells/s of `plan` on resident counts: W warm-up steps, median of R blocks of K steps."""
    order, n_chunks = minibatch_order(n_cells, B, dev, seed=1)
    for i in range(W):
        plan.training_step(order[(i % n_chunks) * B : (i % n_chunks + 1) * B])
    ms = timed_blocks(lambda i: plan.training_step(order[((W + i) % n_chunks) * B : ((W + i) % n_chunks + 1) * B]),
                      K, R, barrier, world, dev)
    med = float(np.median(ms))
    return {"cells_per_s": K * B * world / (med / 1e3), "ms_per_step": med / K, "steps": K, "blocks_ms": ms,
            "final_loss": float(plan.out3[0])}


def run_b200_arm(a):
    from scvi_tools_b200.data import DeviceResidentCounts, HostStreamedCounts, narrowest_count_dtype

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (this arm has no CPU fallback)")
    # stdout carries ONE JSON line: anything a library prints on fd 1 meanwhile (NCCL's version banner ignores
    # NCCL_DEBUG_FILE on some builds) goes to stderr; fd 1 is restored right before the line is printed
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    pg = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)
        pg = dist.group.WORLD
    # under data parallelism the first steps also warm NCCL / peer mappings / clocks on every rank: >= 20
    B, K, R = a.batch_size, a.steps, max(a.blocks, 1)
    W = max(a.warmup, 3) if world == 1 else max(a.warmup, 20)

    X = synth_counts_device(a.cells, a.genes, dev, seed=rank, zero_frac=a.zero_frac)
    batch = (torch.arange(a.cells, device=dev) * a.batches // a.cells)
    compress = resident_type(a, X)
    data = DeviceResidentCounts(X, batch=batch, device=dev, compress=compress)
    if compress:
        del X
        X = data.X
    plan = make_plan(a, data, B, pg)
    order, n_chunks = minibatch_order(a.cells, B, dev, sort=not a.unsorted_batches)

    def chunk(i):
        j = i % n_chunks
        return order[j * B : (j + 1) * B]

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    # ---- resident timing (value)
    for i in range(W):
        plan.training_step(chunk(i))
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    blocks_ms = timed_blocks(lambda i: plan.training_step(chunk(W + i)), K, R, barrier, world, dev)
    clocks = sampler.stop()
    ms = float(np.median(blocks_ms))
    loss_end = float(plan.out3[0])
    launches_per_step = plan.kernel_launches_per_step  # this library's kernels recorded in the step graph
    value = K * B * world / (ms / 1e3)

    # ---- data-parallel checks: replicas identical after the timed steps; cost of the exchange
    ddp = None
    if world > 1:
        torch.cuda.synchronize()
        bits = plan.p.view(torch.int32).to(torch.int64)
        sig = torch.stack([bits.sum(), (bits * torch.arange(1, bits.numel() + 1, device=dev)).sum()])
        sigs = [torch.empty_like(sig) for _ in range(world)]
        torch.distributed.all_gather(sigs, sig)
        same = all(torch.equal(sigs[0], t) for t in sigs)
        solo = make_plan(a, data, B, None)  # the same step without any exchange, on this rank alone
        solo_rate = quick_rate(solo, a.cells, B, dev, K, min(R, 3), 5, barrier, 1)
        t = torch.tensor([solo_rate["ms_per_step"]], device=dev)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ddp = {"mode": plan.ddp_mode, "params_identical_across_ranks": bool(same),
               "param_checksum": [int(v) for v in sigs[0].tolist()],
               "single_rank_ms_per_step": float(t), "exchange_exposed_us": 1e3 * (ms / K - float(t)),
               "numerical_test": "tests/test_gpu_multigpu.py (bit-equality with single-process arithmetic, N=2)"}
        del solo
        if not same:
            print("bench.py: replicas diverged", file=sys.stderr)

    if a.value_only:
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        if rank == 0:
            print(json.dumps({"metric": "SCVI train cells/sec", "value": value, "unit": "cells/s", "n_gpus": world,
                              "steps": K, "warmup": W, "ms_per_step": ms / K, "config": config_dict(a, world),
                              "blocks_ms": blocks_ms, "ddp_check": ddp, "clocks": clocks, "partial": "value-only"}),
                  flush=True)
        if world > 1:
            plan._graphs = None
            torch.cuda.synchronize()
            torch.distributed.barrier()
            sys.stdout.flush()
            os._exit(0)
        return

    # ---- end to end (A): counts stay in pinned HOST memory (the whole matrix, narrowest lossless integer type);
    # every step submits its random row ids, the rows cross PCIe inside the timed region (gathered by the GPU
    # straight from the mapped host pages, or by the host threads + a copy-engine H2D), are expanded to fp32
    # rows on the device, the fused step runs, and the loss is copied back and read on the host every step.
    n_host_rows = a.cells
    hdt = (torch.float32 if a.e2e_dtype == "f32" else torch.uint8 if X.dtype == torch.uint8
           else narrowest_count_dtype(X[: min(a.cells, 262144)]))
    if a.e2e_dtype == "u16" and hdt == torch.uint8:
        hdt = torch.uint16
    if hdt != torch.float32 and float(X.max()) >= (256 if hdt == torch.uint8 else 65536):
        hdt = torch.uint16 if float(X.max()) < 65536 else torch.float32
    host_mat = torch.empty(n_host_rows, a.genes, dtype=hdt, pin_memory=True)
    for s0 in range(0, n_host_rows, 65536):
        host_mat[s0 : s0 + 65536].copy_(X[s0 : s0 + 65536].to(hdt))
    host_batch = batch.cpu()
    order_host = order.cpu()
    loss_host = [torch.zeros(3).pin_memory() for _ in range(2)]
    loss_done = [torch.cuda.Event(), torch.cuda.Event()]
    losses = []

    def run_e2e(feeder, submit):
        plan.attach_streamed(feeder)

        def loop(n, base):
            # every step's loss is copied to the host and read there; the read of step i happens after step i+1
            # has been enqueued (one step of lookahead, as a logger thread would), so the GPU never idles on it
            losses.clear()
            submit(0, base)
            for i in range(n):
                s = i % 2
                plan.training_step(feeder.acquire(s))
                feeder.release(s)
                loss_host[s].copy_(plan.out3, non_blocking=True)
                loss_done[s].record()
                if i + 1 < n:  # the next minibatch crosses PCIe while this step computes
                    submit((i + 1) % 2, base + i + 1)
                if i > 0:
                    loss_done[1 - s].synchronize()
                    losses.append(float(loss_host[1 - s][0]))
            if n > 0:
                loss_done[(n - 1) % 2].synchronize()
                losses.append(float(loss_host[(n - 1) % 2][0]))

        loop(W, 0)
        out, base = [], W
        for _ in range(R):
            barrier()
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record()
            loop(K, base)
            t1.record()
            barrier()
            base += K
            msb = t0.elapsed_time(t1)
            if world > 1:
                t = torch.tensor([msb], device=dev)
                torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
                msb = float(t)
            out.append(msb)
        return float(np.median(out)), out

    def rows_of(i):
        j = i % n_chunks
        return order_host[j * B : (j + 1) * B]

    legs = {}
    # uint8 host counts + tcgen05 kernels: the H2D copy lands in the uint8 rows the step reads (no expansion kernel)
    compact = hdt == torch.uint8 and a.precision == "tc" and a.resident != "f32"
    for name, devg, frac in (("device_gather", True, 0.0), ("host_gather", False, 0.0),
                             ("hybrid_gather", True, a.hybrid_fraction)):
        if name == "hybrid_gather" and frac <= 0.0:
            continue
        feeder = HostStreamedCounts(host_mat, batch=host_batch, batch_size=B, device=dev, dtype=hdt, device_gather=devg,
                                    compact_stage=compact, hybrid_fraction=frac,
                                    host_threads=max(1, (os.cpu_count() or 1) // world))  # the ranks share the host cores
        if devg and not feeder.device_gather:
            continue
        med, allb = run_e2e(feeder, lambda slot, i, f=feeder: f.submit(slot, rows=rows_of(i)))
        legs[name] = {"cells_per_s": K * B * world / (med / 1e3), "ms_per_step": med / K, "blocks_ms": allb}
        del feeder
    # (B) round-1 figure for continuity: 4 pre-formed pinned minibatches cycled (no per-step row gather)
    n_pre = 4
    pre_x = [host_mat[i * B : (i + 1) * B].clone().pin_memory() for i in range(n_pre)]
    pre_b = [host_batch[i * B : (i + 1) * B].clone().pin_memory() for i in range(n_pre)]
    feeder = HostStreamedCounts(torch.zeros(1, a.genes), batch=None, batch_size=B, device=dev, dtype=hdt,
                                device_gather=False, compact_stage=compact)
    med, allb = run_e2e(feeder, lambda slot, i: feeder.submit(slot, host_x=pre_x[i % n_pre], host_b=pre_b[i % n_pre]))
    legs["preformed_minibatches"] = {"cells_per_s": K * B * world / (med / 1e3), "ms_per_step": med / K, "blocks_ms": allb}
    best = max((k for k in legs if k != "preformed_minibatches"), key=lambda k: legs[k]["cells_per_s"])
    e2e = {"value": legs[best]["cells_per_s"], "unit": "cells/s",
           "h2d_bytes_per_step": B * a.genes * host_mat.element_size() + B * 8 + B * 8, "d2h_bytes_per_step": 12,
           "host_dtype": str(hdt).replace("torch.", ""), "losses_read_on_host": len(losses), "route": best,
           "routes": legs, "host_matrix_rows": n_host_rows,
           "how": "whole count matrix in pinned host memory (narrowest lossless count type); per step: random row ids "
                  "-> rows over PCIe (device_gather: a gather kernel reads the mapped host pages; hybrid_gather: the GPU "
                  "pulls a fraction of the rows that way WHILE the host threads gather the rest; host_gather: "
                  "index_select on the host threads + cudaMemcpyAsync; uint8 counts land in the uint8 staging rows the "
                  "tcgen05 kernels read), fused step, "
                  "loss D2H each step read on the host one step behind the launch; preformed_minibatches = the "
                  "round-1 figure (no per-step row gather), kept for continuity"}
    del host_mat, pre_x

    # ---- per-kernel roofline (rank 0)
    line = None
    extras = {}
    plan.data = data
    plan.batch_cols = (data.batch + plan.L).to(torch.int32)
    plan._graphs = None
    if rank == 0:
        kern = profile_step(plan, a)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
        ab = algorithmic_bytes(a, B, plan.n_arena)
        ab1 = algorithmic_bytes(a, B, plan.n_arena, xbytes=1) if compress else None
        for k, v in kern.items():
            v["algorithmic_bytes"] = ab.get(k)
            if ab.get(k):
                v["GBps"] = ab[k] / v["ms"] / 1e6
                v["frac_of_hbm_peak"] = v["GBps"] / peak
                if ab1:  # what the kernel actually has to move with uint8-resident counts
                    v["bytes_uint8_counts"] = ab1[k]
                    v["GBps_uint8_counts"] = ab1[k] / v["ms"] / 1e6
        step_kernels = {k: v for k, v in kern.items() if k != "scvi_gather_rows"}
        dom = max(step_kernels, key=lambda k: step_kernels[k]["ms"] * step_kernels[k]["launches_per_step"])
        eager_ms = sum(v["ms"] * v["launches_per_step"] for v in step_kernels.values())
        traffic = None
        try:  # DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture
            tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            ent = tr.get(dom)
            if (ent and ent["batch_size"] == B and ent["genes"] == a.genes and ent["likelihood"] == a.likelihood
                    and ent.get("resident_counts", "float32") == ("uint8" if compress else "float32")):
                traffic = ent["dram_bytes_per_launch"]
        except Exception:
            pass
        roof = {"bound": "hbm", "kernel": dom, "achieved": kern[dom].get("GBps"), "peak": peak, "unit": "GB/s",
                "frac": (kern[dom]["GBps"] / peak) if kern[dom].get("GBps") else None, "traffic": traffic,
                "algorithmic_bytes": kern[dom]["algorithmic_bytes"], "kernel_ms": kern[dom]["ms"],
                "algorithmic_bytes_uint8_counts": kern[dom].get("bytes_uint8_counts"),
                "achieved_uint8_counts": kern[dom].get("GBps_uint8_counts"),
                "peak_source": peak_src, "share_of_step": kern[dom]["ms"] * kern[dom]["launches_per_step"] / eager_ms,
                "note": "HBM is the nearest roofline of the step's streaming kernels; the dominant kernel is limited "
                        "by the issue rate of the per-element likelihood math, not by bytes (DESIGN.md section 4). "
                        "`achieved` / `frac` are on SURVEY 8(d)'s algorithmic bytes (fp32 counts, 4 B each); with the "
                        "uint8-resident counts this run uses the kernel moves `algorithmic_bytes_uint8_counts` "
                        "(`achieved_uint8_counts` GB/s) and `traffic` is the ncu DRAM figure of that uint8 kernel"}
    # ---- the other workloads SURVEY 8(d) asks for, as extra keys of the same line (every rank runs them: the
    # plans are data-parallel); each is W' warm-up steps + the median of 3 blocks of K steps
    if not a.no_extras:
        Kx, Rx = K, 3
        # reference default minibatch (src/scvi/model/base/_training_mixin.py:78)
        p128 = make_plan(a, data, 128, pg, seed=1)
        extras["batch_size_128"] = quick_rate(p128, a.cells, 128, dev, max(Kx, 50), Rx, W, barrier, world)
        del p128
        # ~90 % zeros (real scRNA-seq sparsity): thin the resident matrix in place, same shape
        if a.zero_frac < 0.85:
            keep = (1.0 - 0.9) / (1.0 - a.zero_frac)
            g = torch.Generator(device=dev).manual_seed(77 + rank)
            for s0 in range(0, a.cells, 65536):
                X[s0 : s0 + 65536].mul_(torch.rand(X[s0 : s0 + 65536].shape, device=dev, generator=g) < keep)
            psp = make_plan(a, data, B, pg, seed=2)
            extras["zeros_90pct"] = quick_rate(psp, a.cells, B, dev, Kx, Rx, W, barrier, world)
            del psp
    del plan.data, data, X  # free the resident matrix before the standalone-kernel leg
    if not a.no_extras and not (a.genes == 4000 and a.likelihood == "nb" and a.batches == 32):
        # BASELINE.json configs[2]: 1M x 4000 genes, NB, 32 batches
        import copy

        a3 = copy.copy(a)
        a3.genes, a3.likelihood, a3.batches = 4000, "nb", 32
        X3 = synth_counts_device(a.cells, 4000, dev, seed=50 + rank, zero_frac=a.zero_frac)
        d3 = DeviceResidentCounts(X3, batch=torch.arange(a.cells, device=dev) * 32 // a.cells, device=dev,
                                  compress=resident_type(a, X3))
        p3 = make_plan(a3, d3, B, pg, seed=3)
        extras["c3_1Mx4000_nb_32batches"] = quick_rate(p3, a.cells, B, dev, K, 3, W, barrier, world)
        extras["c3_1Mx4000_nb_32batches"]["workload"] = workload_name(a3)
        del p3, d3, X3
    if rank == 0:
        logprob = time_logprob_kernels(a, dev, peak)
        cpu = None if a.no_cpu_baseline else cpu_baseline(a, a.cpu_seconds)
        cfg = config_dict(a, world)
        line = {
            "metric": "SCVI train cells/sec", "value": value, "unit": "cells/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": cfg,
            "details": {"cache": "inputs larger than L2: rows gathered at random from the HBM-resident count matrix "
                                 f"({a.cells * a.genes * (1 if compress else 4) / 1e9:.1f} GB)",
                        "resident_counts": ("uint8 (every count < 256; the kernels' arithmetic is the fp32 path's, "
                                            "bit for bit)" if compress else "float32"),
                        "precision": ("tcgen05 fp16 hi/lo split GEMMs, fp32 accumulate (fp32-grade, parity-tested)"
                                      if a.precision == "tc" else "fp32 FFMA GEMMs"), "cuda_graph": True,
                        "timing": f"median of {R} blocks of exactly {K} steps, each between barrier+synchronize, "
                                  "CUDA events on the launching stream, max over ranks",
                        "blocks_ms": blocks_ms, "zero_frac": a.zero_frac},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches_per_step * K,
            "launches_per_step": launches_per_step, "ddp_check": ddp,
            "roofline": roof, "kernels": kern, "nb_loss_kernel_hbm": logprob, "cpu_baseline": cpu,
            "final_loss": loss_end, "also": extras,
        }
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    if line is not None:
        print(json.dumps(line), flush=True)
    if world > 1:
        # the step graphs hold captured collectives: release them before the communicator goes away, and do not
        # wait on the communicator teardown (it can block on graph-captured collectives)
        plan._graphs = None
        torch.cuda.synchronize()
        torch.distributed.barrier()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def main():
    a = parse()
    if os.environ.get("SCVI_BENCH_FAULT"):  # debugging aid: dump all Python stacks if the run hangs
        import faulthandler

        faulthandler.dump_traceback_later(float(os.environ["SCVI_BENCH_FAULT"]), exit=True)
    if a.impl == "reference":
        run_reference_arm(a)
    else:
        run_b200_arm(a)


if __name__ == "__main__":
    main()
