"""Whole-step CUDA graphs for the autograd-driven plans (semi-supervised, adversarial).

Those steps are ~150-300 small launches (this library's kernels through `torch.autograd.Function`s plus torch's
element-wise glue): launch- and interpreter-bound at the reference's minibatch sizes (128-1024 cells).  A step
whose shapes repeat is captured ONCE -- zero_grad, forward, backward, both optimizers -- and replayed from static
input buffers; everything that changes between steps without a shape change lives on the device (the KL weight,
the Adam step counters, torch's graph-safe Philox offsets for dropout / reparameterisation noise).

`WholeStepGraph(step_fn)`: `step_fn(inputs)` runs one full step on a pytree of tensors and returns any object
holding output tensors.  Calls are eager for the first `warmup` occurrences of a shape signature (which also builds
the optimizer arenas and opts kernels into their shared-memory sizes), then captured; a ragged last minibatch gets
its own graph after it has been seen `warmup` times, until then it runs eagerly -- on the same kernels either way.
Outputs of a replayed step are STATIC tensors: they are overwritten by the next replay of the same graph.

Streams.  Autograd's gradient-accumulation nodes remember the stream they were created on, and a node that outlives
its step (anything still referencing the previous step's autograd graph keeps it) would tie a capture to that
stream -- fatally if it is the legacy default stream.  Every step of a graphed plan, eager or captured or replayed,
therefore runs on ONE side stream owned by this object, fenced against the caller's current stream on entry and
exit; `step_fn` should also return detached tensors only.
"""

from __future__ import annotations

import gc

import torch
from torch.utils import _pytree as pytree


class WholeStepGraph:
    def __init__(self, step_fn, warmup: int = 3, max_graphs: int = 4):
        self.step_fn = step_fn
        self.warmup = warmup
        self.max_graphs = max_graphs
        self._seen: dict = {}
        self._graphs: dict = {}
        self._pool = None
        self._stream = None
        self.n_replays = 0
        self.n_eager = 0

    @staticmethod
    def _signature(leaves):
        return tuple((tuple(t.shape), t.dtype, t.device.index) if isinstance(t, torch.Tensor) else ("const", repr(t))
                     for t in leaves)

    def invalidate(self):
        """Drop every captured graph (a launch argument baked into them changed, e.g. a learning rate)."""
        self._graphs.clear()

    def __call__(self, inputs):
        leaves, spec = pytree.tree_flatten(inputs)
        if self._stream is None:
            dev = next(t.device for t in leaves if isinstance(t, torch.Tensor))
            self._stream = torch.cuda.Stream(device=dev)
        caller = torch.cuda.current_stream(self._stream.device)
        self._stream.wait_stream(caller)
        with torch.cuda.stream(self._stream):
            out = self._run(inputs, leaves, spec)
        caller.wait_stream(self._stream)
        return out

    def _run(self, inputs, leaves, spec):
        sig = self._signature(leaves)
        entry = self._graphs.get(sig)
        if entry is None:
            n = self._seen.get(sig, 0)
            self._seen[sig] = n + 1
            if n < self.warmup or len(self._graphs) >= self.max_graphs:
                self.n_eager += 1
                return self.step_fn(inputs)
            entry = self._capture(leaves, spec, sig)
        graph, static_leaves, out = entry
        for s, t in zip(static_leaves, leaves, strict=True):
            if isinstance(s, torch.Tensor):
                s.copy_(t, non_blocking=True)
        graph.replay()
        self.n_replays += 1
        return out

    def _capture(self, leaves, spec, sig):
        static_leaves = [t.clone() if isinstance(t, torch.Tensor) else t for t in leaves]
        static_in = pytree.tree_unflatten(static_leaves, spec)
        graph = torch.cuda.CUDAGraph()
        gc.collect()  # autograd graphs of earlier steps that only reference cycles keep alive
        torch.cuda.synchronize()
        # torch.distributions' argument validation reads its verdict on the host: not capturable.  The eager warm-up
        # steps validated the same expressions; inside the graph the checks are off.
        from torch.distributions import Distribution

        prev = Distribution._validate_args
        Distribution.set_default_validate_args(False)
        try:
            kw = {"stream": self._stream}
            if self._pool is not None:
                kw["pool"] = self._pool
            with torch.cuda.graph(graph, **kw):
                out = self.step_fn(static_in)
        finally:
            Distribution.set_default_validate_args(prev)
        if self._pool is None:
            self._pool = graph.pool()
        entry = (graph, static_leaves, out)
        self._graphs[sig] = entry
        # the capture itself executed nothing: the caller's replay below performs this step
        return entry
