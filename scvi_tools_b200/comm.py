"""Peer-memory gradient exchange for the data-parallel training plan (single node, NVLink / NVSwitch).

Host side of `scvi_allreduce_adam` (csrc/comm.cu): every rank allocates one exchange segment through the
C-ABI, the 64-byte IPC handles travel over the process group's host channel (`all_gather_object`), every rank
opens every other rank's segment, and the gradient arena of the plan lives inside the segment (two copies,
alternating by step parity) so the fused all-reduce + Adam kernel reads the peers' gradients directly.

Stands in for `DistributedDataParallel`'s reducer under the reference's `Trainer(strategy="ddp")`
(reference tests/model/test_multigpu.py:29-36, src/scvi/model/_utils.py:25-34).
"""

from __future__ import annotations

import ctypes as C

import torch

from ._lib import call, query


class _RawCuda:
    """`__cuda_array_interface__` carrier: lets torch view a raw device pointer without copying."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}


class PeerGradientExchange:
    """One exchange segment per rank: `g[parity]` (fp32 views of `n_floats`) + the control block.

    `available` is False (and nothing is allocated) when peer access cannot be set up on every rank; the
    caller then falls back to NCCL."""

    def __init__(self, n_floats: int, process_group, device):
        dist = torch.distributed
        self.pg = process_group
        self.world = dist.get_world_size(process_group)
        self.rank = dist.get_rank(process_group)
        self.dev = torch.device(device)
        self.n = int(n_floats)
        self.arena_bytes = (self.n * 4 + 255) // 256 * 256
        self.ctrl_off = 3 * self.arena_bytes  # [g parity 0][g parity 1][parameters][control block]
        self.nbytes = self.ctrl_off + int(query("scvi_comm_ctrl_bytes"))
        self.base = None
        self.peers: list[int] = []
        self.available = False
        ok, handle = 1, bytes(64)
        try:
            ptr, hbuf = C.c_void_p(), C.create_string_buffer(64)
            with torch.cuda.device(self.dev):
                call("scvi_comm_alloc", self.nbytes, C.byref(ptr), hbuf)
            self.base, handle = int(ptr.value), hbuf.raw
        except RuntimeError:
            ok = 0
        handles = [None] * self.world
        dist.all_gather_object(handles, (ok, handle), group=process_group)
        if all(h[0] for h in handles):
            try:
                for r, (_, h) in enumerate(handles):
                    if r == self.rank:
                        self.peers.append(self.base)
                        continue
                    p = C.c_void_p()
                    with torch.cuda.device(self.dev):
                        call("scvi_comm_open", C.create_string_buffer(h, 64), C.byref(p))
                    self.peers.append(int(p.value))
            except RuntimeError:
                ok = 0
        else:
            ok = 0
        flag = torch.tensor([ok], device=self.dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=process_group)
        if int(flag.item()) != 1:
            self.close()
            return
        self.available = True
        raw = torch.as_tensor(_RawCuda(self.base, self.nbytes), device=self.dev)
        self._raw = raw
        self.g = [raw[i * self.arena_bytes : i * self.arena_bytes + self.n * 4].view(torch.float32) for i in (0, 1)]
        # the parameter arena of the plan lives here as well: the sharded exchange stores updated slices into it
        self.p = raw[2 * self.arena_bytes : 2 * self.arena_bytes + self.n * 4].view(torch.float32)
        PV = C.c_void_p * self.world
        self._peer_g = [PV(*[p + i * self.arena_bytes for p in self.peers]) for i in (0, 1)]
        self._peer_p = PV(*[p + 2 * self.arena_bytes for p in self.peers])
        self._peer_ctrl = PV(*[p + self.ctrl_off for p in self.peers])
        dist.barrier(group=process_group)  # nobody launches before every rank has opened every segment

    def allreduce_adam(self, parity: int, slot: int, p, m, v, begin: int, end: int, step_counter, advance: bool,
                       lr, beta1, beta2, eps, wd, stream, blocks: int = 0, g_sum=None):
        call("scvi_allreduce_adam", self._peer_g[parity], self._peer_ctrl, self.world, self.rank, slot,
             C.c_void_p(p.data_ptr()), C.c_void_p(m.data_ptr()), C.c_void_p(v.data_ptr()),
             None if g_sum is None else C.c_void_p(g_sum.data_ptr()), begin, end, C.c_void_p(step_counter.data_ptr()),
             1 if advance else 0, lr, beta1, beta2, eps, wd, blocks, stream)

    def allreduce_adam_sharded(self, parity: int, slot: int, m, v, step_counter, lr, beta1, beta2, eps, wd, stream,
                               blocks: int = 0):
        call("scvi_allreduce_adam_sharded", self._peer_g[parity], self._peer_p, self._peer_ctrl, self.world, self.rank,
             slot, C.c_void_p(m.data_ptr()), C.c_void_p(v.data_ptr()), self.n, C.c_void_p(step_counter.data_ptr()),
             lr, beta1, beta2, eps, wd, blocks, stream)

    def shard_bounds(self):
        """[lo, hi) in floats of the slice whose parameters / Adam moments this rank maintains (sharded exchange)."""
        n4 = self.n // 4
        chunk = (n4 + self.world - 1) // self.world
        lo = min(self.rank * chunk, n4)
        return 4 * lo, 4 * min(lo + chunk, n4)

    def close(self):
        """Unmap the peers' segments and free this rank's (call after a barrier + synchronize)."""
        for r, p in enumerate(self.peers):
            if r != self.rank and p:
                try:
                    call("scvi_comm_close", C.c_void_p(p))
                except RuntimeError:
                    pass
        self.peers = []
        self.g = None
        self.p = None
        self._raw = None
        if self.base:
            call("scvi_comm_free", C.c_void_p(self.base))
            self.base = None
        self.available = False
