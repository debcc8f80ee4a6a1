"""`ArenaAdam`: torch.optim.Adam semantics (reference src/scvi/train/_trainingplans.py:475-482: lr 1e-3, eps 0.01,
weight decay 1e-6 added to the gradient) for the autograd-driven plans, as ONE launch of the library's
multi-tensor Adam (`scvi_adam_step`) over a flat parameter / gradient / moment arena.

The arena is laid out at the first `step()` over the parameters that hold a gradient by then (torch's Adam skips
parameters whose `.grad` is None; so does this one -- and it refuses, loudly, a parameter that starts receiving
gradient later, because its moments would have no slot).  Parameters become views of the parameter arena.

Gradients: `zero_grad()` sets every `.grad` to None, so autograd's first accumulation into a parameter KEEPS the
tensor the backward kernel wrote (no `grad += new` launch per parameter, no arena memset); `step()` moves all of
them into the flat gradient arena with ONE launch (`scvi_gather_segments`, pointer table by value in the launch
parameters -- capturable), which is also the buffer the data-parallel all-reduce runs on.
"""

from __future__ import annotations

import ctypes as C

import torch

from ._lib import call


def broadcast_module_state(module, process_group):
    """Rank 0's parameters and buffers to every rank, once, at plan construction (what torch DDP does when it wraps
    a module: replicas must start equal for the gradient mean to keep them equal)."""
    if process_group is None:
        return
    import torch.distributed as dist

    src = dist.get_global_rank(process_group, 0)
    with torch.no_grad():
        for t in list(module.parameters()) + list(module.buffers()):
            dist.broadcast(t.data, src=src, group=process_group)


class ArenaAdam:
    def __init__(self, named_parameters, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 0.01,
                 weight_decay: float = 1e-6, capturable: bool = False, process_group=None):
        """`named_parameters`: a callable returning (name, parameter) pairs, e.g. `module.named_parameters`.
        `capturable`: keep the step count on the device (`scvi_adam_step_graph`), so that `step()` can be captured
        into a CUDA graph; a learning-rate change then needs a re-capture (the rate is a launch argument)."""
        self._named = named_parameters
        self.capturable = capturable
        # data parallelism (torch DDP's contract, reference `Trainer(strategy="ddp")`): parameters start as rank 0's,
        # every step applies the MEAN of the ranks' gradients -- one all-reduce of the flat gradient arena
        self.pg = process_group
        self.world = 1
        if process_group is not None:
            import torch.distributed as dist

            self.world = dist.get_world_size(process_group)
        self.lr, self.betas, self.eps, self.wd = lr, betas, eps, weight_decay
        self.arena = None
        self.names: list[str] = []
        self._frozen: list[str] = []

    def _build(self):
        params = [(k, p) for k, p in self._named() if p.requires_grad and p.grad is not None]
        if not params:
            raise RuntimeError("no parameter received a gradient")
        dev = params[0][1].device
        if dev.type != "cuda":
            raise RuntimeError("ArenaAdam: parameters must live on a CUDA device (there is no CPU path)")
        offs, off = [], 0
        for _, p in params:
            offs.append(off)
            off += (p.numel() + 3) // 4 * 4
        P = torch.zeros(off, dtype=torch.float32, device=dev)
        G = torch.zeros_like(P)
        for (_, p), o in zip(params, offs, strict=True):
            n = p.numel()
            P[o : o + n].copy_(p.data.reshape(-1))
            p.data = P[o : o + n].view(p.shape)
        self.arena = {"p": P, "g": G, "m": torch.zeros_like(P), "v": torch.zeros_like(P), "step": 0, "offs": offs,
                      "counter": torch.zeros(2, dtype=torch.int64, device=dev)}
        self.names = [k for k, _ in params]
        taken = set(self.names)
        self._frozen = [k for k, p in self._named() if p.requires_grad and k not in taken]
        n = len(params)
        self._offs_c = (C.c_int64 * n)(*offs)
        self._ns_c = (C.c_int64 * n)(*[p.numel() for _, p in params])

    def zero_grad(self):
        for _, p in self._named():
            p.grad = None

    def step(self):
        if self.arena is None:
            self._build()
        a = self.arena
        named = dict(self._named())
        for k in self._frozen:
            if named[k].grad is not None:
                raise RuntimeError(f"parameter {k} received a gradient after the optimizer arena was laid out")
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        # every gradient tensor -> its slot of the flat arena, one launch (a parameter without gradient: zeros)
        n = len(self.names)
        srcs = (C.c_void_p * n)()
        keep = []
        for i, k in enumerate(self.names):
            g = named[k].grad
            if g is None:
                continue
            if g.dtype != torch.float32 or not g.is_contiguous():
                g = g.float().contiguous()
                keep.append(g)
            srcs[i] = g.data_ptr()
        call("scvi_gather_segments", srcs, self._offs_c, self._ns_c, n, C.c_void_p(a["g"].data_ptr()), st)
        a["step"] += 1
        scale = 1.0
        if self.world > 1:
            import torch.distributed as dist

            dist.all_reduce(a["g"], group=self.pg)  # captured with the step when the plan runs as a CUDA graph
            scale = 1.0 / self.world
        if self.capturable:
            call("scvi_adam_step_graph", C.c_void_p(a["p"].data_ptr()), C.c_void_p(a["g"].data_ptr()),
                 C.c_void_p(a["m"].data_ptr()), C.c_void_p(a["v"].data_ptr()), a["p"].numel(),
                 C.c_void_p(a["counter"].data_ptr()), self.lr, self.betas[0], self.betas[1], self.eps, self.wd, scale, st)
            return
        call("scvi_adam_step", C.c_void_p(a["p"].data_ptr()), C.c_void_p(a["g"].data_ptr()),
             C.c_void_p(a["m"].data_ptr()), C.c_void_p(a["v"].data_ptr()), a["p"].numel(), a["step"], self.lr,
             self.betas[0], self.betas[1], self.eps, self.wd, scale, st)

    @property
    def n_steps(self) -> int:
        """Steps applied so far (device counter when capturable: this read synchronises)."""
        if self.arena is None:
            return 0
        return int(self.arena["counter"][0]) if self.capturable else self.arena["step"]
