"""HBM-resident count matrix + on-device minibatch index stream.

Replaces the reference's per-step host path -- `AnnDataLoader` / `AnnTorchDataset.__getitem__`
(src/scvi/dataloaders/_ann_dataloader.py:21-138, src/scvi/data/_anntorchdataset.py:115-169) and
the Lightning host->device copy -- the way `DeviceBackedDataSplitter`
(src/scvi/dataloaders/_data_splitting.py:589-732) does, but with the index stream of
`BatchDistributedSampler` (src/scvi/dataloaders/_samplers.py:6-89): every epoch a global
permutation seeded `seed + epoch`, rank-strided shard, padded (or tail-dropped) to equal
length, chunks of the per-replica batch size.  The matrix is uploaded once; a step moves no
count data over PCIe.
"""

from __future__ import annotations

import ctypes as C
import math

import numpy as np
import torch

from . import ops
from ._lib import call


def epoch_indices(n_obs: int, rank: int, world: int, epoch: int, seed: int = 0, shuffle: bool = True,
                  drop_dataset_tail: bool = False) -> torch.Tensor:
    """This rank's sample order for `epoch` (host int64 tensor) -- torch `DistributedSampler.__iter__`
    semantics, which `BatchDistributedSampler` inherits."""
    if shuffle:
        g = torch.Generator()
        g.manual_seed(seed + epoch)
        idx = torch.randperm(n_obs, generator=g)
    else:
        idx = torch.arange(n_obs)
    if drop_dataset_tail and n_obs % world != 0:
        num = math.ceil((n_obs - world) / world)
    else:
        num = math.ceil(n_obs / world)
    total = num * world
    if not drop_dataset_tail:
        pad = total - n_obs
        if pad > 0:
            reps = math.ceil(pad / n_obs)
            idx = torch.cat([idx, idx.repeat(reps)[:pad]])
    else:
        idx = idx[:total]
    return idx[rank:total:world].contiguous()


def narrowest_count_dtype(X) -> torch.dtype:
    """uint8 / uint16 when every entry of `X` is an integer in range (the expansion to fp32 is then exact),
    else float32."""
    X = torch.as_tensor(X)
    if X.numel() == 0:
        return torch.float32
    lo, hi = float(X.min()), float(X.max())
    if lo < 0 or hi >= 65536:
        return torch.float32
    if X.dtype.is_floating_point and not bool((X == X.round()).all()):
        return torch.float32
    return torch.uint8 if hi < 256 else torch.uint16


class HostStreamedCounts:
    """Count matrix kept in pinned HOST memory (for data that should not, or does not, live in HBM) and fed to
    the training step minibatch by minibatch -- the reference's `AnnDataLoader` route
    (src/scvi/dataloaders/_ann_dataloader.py:21-138 + the Lightning host->device copy), made PCIe-frugal: the
    matrix is stored once in the narrowest lossless integer type (`narrowest_count_dtype`), a minibatch crosses
    PCIe in that type on a copy stream (double buffered) and `scvi_gather_rows_u8/_u16` expands it to the fp32
    rows the kernels read -- bit-identical to `data[indexes].astype(float32)`.

    `stage` is a `DeviceResidentCounts` of 2 x batch_size rows; the plan trains on rows `slot*B .. slot*B+n`.

    `submit(rows=...)` gathers the minibatch rows ON THE DEVICE, straight out of the pinned host matrix (the
    pinned pages are mapped into the GPU's address space): only the 8-byte row ids are copied first, then one
    gather kernel on the copy stream pulls the compact rows over PCIe and expands them into the staging slot --
    no host-side copy, no host thread touching count data.  `device_gather=False` keeps the host-side
    `index_select` into a pinned buffer followed by a plain H2D copy.
    """

    def __init__(self, X, batch=None, batch_size: int = 128, device="cuda", dtype: torch.dtype | None = None,
                 device_gather: bool = True, compact_stage: bool = False, host_threads: int = 0,
                 hybrid_fraction: float = 0.0):
        """`compact_stage`: keep the staging rows in uint8 (host matrix of uint8 counts only) -- the tensor-core
        training kernels read uint8 counts directly (SCVI_X_U8 / SCVI_LIN_X_U8), so the H2D copy lands in the rows
        the step reads and nothing is expanded."""
        self.dev = torch.device(device)
        self.host_threads = int(host_threads)  # threads of the host-side row gather (0 = all hardware threads)
        # > 0: that fraction of every minibatch is pulled by the GPU straight from the mapped pinned matrix (no host
        # core involved, ~25 GB/s of PCIe reads) WHILE the host threads gather the rest for a copy-engine H2D: the two
        # routes add up when the host gather is the bottleneck
        self.hybrid_fraction = float(hybrid_fraction)
        self._gather_stream = None
        X = torch.as_tensor(np.asarray(X.toarray()) if hasattr(X, "toarray") else X)
        self.dtype = narrowest_count_dtype(X) if dtype is None else dtype
        self.host = X.to(self.dtype).contiguous().pin_memory()
        n = X.shape[0]
        b = torch.zeros(n, dtype=torch.int64) if batch is None else torch.as_tensor(np.asarray(batch)).reshape(-1).to(torch.int64)
        self.host_batch = b.contiguous().pin_memory()
        self.B = int(batch_size)
        G = X.shape[1]
        self.compact_stage = bool(compact_stage) and self.dtype == torch.uint8
        if self.compact_stage:
            self.stage = DeviceResidentCounts(torch.zeros(2 * self.B, G, dtype=torch.uint8), compress="u8",
                                              batch=torch.zeros(2 * self.B, dtype=torch.int64), device=self.dev)
        else:
            self.stage = DeviceResidentCounts(torch.zeros(2 * self.B, G), batch=torch.zeros(2 * self.B, dtype=torch.int64),
                                              device=self.dev)
        # the staging rows hold whatever the host matrix holds: integer counts unless it had to stay float32
        self.stage.integer_counts = self.dtype != torch.float32 or DeviceResidentCounts._all_integer(self.host)
        self._compact = (None if (self.dtype == torch.float32 or self.compact_stage)
                         else torch.empty(2 * self.B, G, dtype=self.dtype, device=self.dev))
        self._pin = [torch.empty(self.B, G, dtype=self.dtype).pin_memory() for _ in range(2)]
        self._pin_b = [torch.empty(self.B, dtype=torch.int64).pin_memory() for _ in range(2)]
        self.copy_stream = torch.cuda.Stream(device=self.dev)
        self.ready = [torch.cuda.Event(), torch.cuda.Event()]
        self.freed = [torch.cuda.Event(), torch.cuda.Event()]
        for e in self.freed:
            e.record()
        self.slot_rows = [torch.arange(self.B, device=self.dev), torch.arange(self.B, 2 * self.B, device=self.dev)]
        self.on_batch = None  # callable(lo, n) run on the copy stream after a slot's batch codes landed
        self.device_gather = False
        if device_gather and n > 0:
            try:  # device addresses of the pinned matrix / batch codes (UVA: normally the same numbers)
                dx, db = C.c_void_p(), C.c_void_p()
                with torch.cuda.device(self.dev):
                    call("scvi_host_device_pointer", C.c_void_p(self.host.data_ptr()), C.byref(dx))
                    call("scvi_host_device_pointer", C.c_void_p(self.host_batch.data_ptr()), C.byref(db))
                self._dev_x, self._dev_b = int(dx.value), int(db.value)
                self._pin_idx = [torch.empty(self.B, dtype=torch.int64).pin_memory() for _ in range(2)]
                self._dev_idx = torch.empty(2 * self.B, dtype=torch.int64, device=self.dev)
                self.device_gather = True
            except RuntimeError:
                self.device_gather = False

    @property
    def n_obs(self) -> int:
        return self.host.shape[0]

    @property
    def n_vars(self) -> int:
        return self.host.shape[1]

    @property
    def h2d_bytes_per_row(self) -> int:
        return self.n_vars * self.host.element_size() + 8

    def submit(self, slot: int, rows=None, host_x: torch.Tensor | None = None, host_b: torch.Tensor | None = None) -> int:
        """Enqueue the copy of one minibatch into staging slot `slot` (0 / 1): either rows `rows` (host int64
        indices) of the host matrix -- gathered into a pinned buffer first -- or an already formed pinned
        minibatch (`host_x` in this object's dtype, `host_b`).  Returns the number of cells."""
        if host_x is None and self.device_gather and self.hybrid_fraction > 0.0:
            return self._submit_hybrid(slot, torch.as_tensor(rows, dtype=torch.int64))
        if host_x is None and self.device_gather:
            return self._submit_device_gather(slot, torch.as_tensor(rows, dtype=torch.int64))
        if host_x is None:
            rows = torch.as_tensor(rows, dtype=torch.int64)
            n = rows.numel()
            self.freed[slot].synchronize()  # the pinned buffer of this slot is reused
            # the library's host thread pool (rows are 1-4 KB and random: latency-bound, prefetched per thread)
            pin = self._pin[slot]
            call("scvi_host_gather_rows", C.c_void_p(self.host.data_ptr()), self.n_obs,
                 self.n_vars * self.host.element_size(), self.host.stride(0) * self.host.element_size(),
                 C.c_void_p(rows.data_ptr()), n, C.c_void_p(pin.data_ptr()), pin.stride(0) * pin.element_size(),
                 self.host_threads)
            torch.index_select(self.host_batch, 0, rows, out=self._pin_b[slot][:n])
            host_x, host_b = self._pin[slot][:n], self._pin_b[slot][:n]
        n = host_x.shape[0]
        lo = slot * self.B
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self.freed[slot])
            if self._compact is None:
                self.stage.X[lo : lo + n].copy_(host_x, non_blocking=True)
            else:
                self._compact[lo : lo + n].copy_(host_x, non_blocking=True)
                ops.gather_rows(self._compact[lo : lo + n], None, out=self.stage.X[lo : lo + n])
            self.stage.batch[lo : lo + n].copy_(host_b, non_blocking=True)
            if self.on_batch is not None:
                self.on_batch(lo, n)
            self.ready[slot].record(self.copy_stream)
        return n

    def _enqueue_device_gather(self, slot: int, rows: torch.Tensor, lo: int, st_obj) -> None:
        """Device-side gather of `rows` (host int64) into staging rows lo .. lo + len(rows) on stream `st_obj`."""
        n = rows.numel()
        G = self.n_vars
        off = lo - slot * self.B
        self._pin_idx[slot][off : off + n].copy_(rows)
        fn = ("scvi_gather_rows_u8_raw" if self.compact_stage else
              {torch.uint8: "scvi_gather_rows_u8", torch.uint16: "scvi_gather_rows_u16",
               torch.float32: "scvi_gather_rows"}[self.dtype])
        with torch.cuda.stream(st_obj):
            st = C.c_void_p(st_obj.cuda_stream)
            idx = self._dev_idx[lo : lo + n]
            idx.copy_(self._pin_idx[slot][off : off + n], non_blocking=True)
            out = self.stage.X[lo : lo + n]
            call(fn, C.c_void_p(self._dev_x), self.n_obs, G, self.host.stride(0), C.c_void_p(idx.data_ptr()), n,
                 C.c_void_p(out.data_ptr()), out.stride(0), st)
            call("scvi_gather_i64", C.c_void_p(self._dev_b), self.n_obs, C.c_void_p(idx.data_ptr()), n,
                 C.c_void_p(self.stage.batch[lo : lo + n].data_ptr()), st)

    def _submit_device_gather(self, slot: int, rows: torch.Tensor) -> int:
        n = rows.numel()
        lo = slot * self.B
        self.freed[slot].synchronize()  # the pinned index buffer of this slot is reused
        self.copy_stream.wait_event(self.freed[slot])
        self._enqueue_device_gather(slot, rows, lo, self.copy_stream)
        with torch.cuda.stream(self.copy_stream):
            if self.on_batch is not None:
                self.on_batch(lo, n)
            self.ready[slot].record(self.copy_stream)
        return n

    def _submit_hybrid(self, slot: int, rows: torch.Tensor) -> int:
        """First `hybrid_fraction` of the rows by the GPU over PCIe (zero-copy reads of the mapped matrix), the rest by
        the host thread pool + a copy-engine H2D; the two proceed concurrently."""
        n = rows.numel()
        lo = slot * self.B
        nd = max(0, min(n, int(round(n * self.hybrid_fraction)) // 16 * 16))
        if self._gather_stream is None:
            self._gather_stream = torch.cuda.Stream(device=self.dev)
        gs = self._gather_stream
        self.freed[slot].synchronize()  # pinned buffers of this slot are reused
        gs.wait_event(self.freed[slot])
        if nd > 0:
            self._enqueue_device_gather(slot, rows[:nd], lo, gs)  # asynchronous: runs during the host gather below
        nh = n - nd
        if nh > 0:
            pin = self._pin[slot]
            hrows = rows[nd:].contiguous()
            call("scvi_host_gather_rows", C.c_void_p(self.host.data_ptr()), self.n_obs,
                 self.n_vars * self.host.element_size(), self.host.stride(0) * self.host.element_size(),
                 C.c_void_p(hrows.data_ptr()), nh, C.c_void_p(pin.data_ptr()), pin.stride(0) * pin.element_size(),
                 self.host_threads)
            torch.index_select(self.host_batch, 0, hrows, out=self._pin_b[slot][:nh])
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self.freed[slot])
            if nh > 0:
                dst = self.stage.X[lo + nd : lo + n] if self._compact is None else self._compact[lo + nd : lo + n]
                dst.copy_(self._pin[slot][:nh], non_blocking=True)
                if self._compact is not None:
                    ops.gather_rows(self._compact[lo + nd : lo + n], None, out=self.stage.X[lo + nd : lo + n])
                self.stage.batch[lo + nd : lo + n].copy_(self._pin_b[slot][:nh], non_blocking=True)
            self.copy_stream.wait_stream(gs)
            if self.on_batch is not None:
                self.on_batch(lo, n)
            self.ready[slot].record(self.copy_stream)
        return n

    def minibatch(self, idx) -> dict:
        """Synchronous `{X, batch, labels}` dict of the cells `idx` (getters; same contract as
        `DeviceResidentCounts.minibatch`)."""
        rows = torch.as_tensor(idx).to("cpu", torch.int64).reshape(-1)
        x = self.host.index_select(0, rows).to(self.dev)
        X = x if self.dtype == torch.float32 else ops.gather_rows(x, None)
        b = self.host_batch.index_select(0, rows).to(self.dev).view(-1, 1)
        return {"X": X, "batch": b, "labels": torch.zeros_like(b)}

    def acquire(self, slot: int, n: int | None = None) -> torch.Tensor:
        """Make the current stream wait for slot `slot`; returns the stage row ids of its cells."""
        torch.cuda.current_stream().wait_event(self.ready[slot])
        return self.slot_rows[slot] if n is None or n == self.B else self.slot_rows[slot][:n]

    def release(self, slot: int):
        """Call after the step that read slot `slot` has been enqueued."""
        self.freed[slot].record()


class DeviceResidentCSR:
    """Sparse count matrix resident in HBM in CSR form -- int64 row pointers, uint16 column ids (int32 when
    n_vars > 65536) and values in the narrowest lossless count type -- with the same minibatch contract as
    `DeviceResidentCounts`: `minibatch(idx)` densifies the rows on the device (`scvi_gather_rows_csr`), the
    reference's `data[indexes].toarray().astype(float32)` (src/scvi/data/_anntorchdataset.py:143-152) without
    the host round trip.  At ~10 % density a 1M x 2000 matrix takes 0.6 GB instead of 8 GB dense fp32.

    `X`: scipy.sparse matrix (any format), or a `(indptr, indices, data, shape)` tuple of arrays."""

    def __init__(self, X, batch=None, labels=None, device="cuda"):
        dev = torch.device(device)
        if hasattr(X, "tocsr"):
            X = X.tocsr()
            X.sum_duplicates()
            X.sort_indices()
            indptr, indices, values, shape = X.indptr, X.indices, X.data, X.shape
        else:
            indptr, indices, values, shape = X
        indptr = torch.as_tensor(np.asarray(indptr)).to(torch.int64)
        indices = torch.as_tensor(np.asarray(indices))
        values = torch.as_tensor(np.asarray(values))
        self.shape = (int(shape[0]), int(shape[1]))
        if indptr.numel() != self.shape[0] + 1 or int(indptr[-1]) != indices.numel() or indices.numel() != values.numel():
            raise ValueError("inconsistent CSR arrays")
        if indices.numel() and (int(indices.min()) < 0 or int(indices.max()) >= self.shape[1]):
            raise ValueError("CSR column index out of range")
        vdt = narrowest_count_dtype(values)
        idt = torch.uint16 if self.shape[1] <= 65536 else torch.int32
        self.indptr = indptr.to(dev).contiguous()
        np_idt = np.uint16 if idt == torch.uint16 else np.int32
        self.indices = torch.from_numpy(indices.numpy().astype(np_idt)).to(dev).contiguous()
        self.values = values.to(vdt).to(dev).contiguous()
        self.integer_counts = vdt != torch.float32 or DeviceResidentCounts._all_integer(self.values.view(1, -1))
        n = self.shape[0]

        def col(v):
            if v is None:
                return torch.zeros(n, dtype=torch.int64, device=dev)
            v = torch.as_tensor(np.asarray(v) if not isinstance(v, torch.Tensor) else v)
            return v.reshape(-1).to(torch.int64).to(dev).contiguous()

        self.batch = col(batch)
        self.labels = col(labels)

    @property
    def n_obs(self) -> int:
        return self.shape[0]

    @property
    def n_vars(self) -> int:
        return self.shape[1]

    @property
    def nbytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in (self.indptr, self.indices, self.values))

    def densify(self, idx: torch.Tensor | None, out: torch.Tensor | None = None) -> torch.Tensor:
        return ops.gather_rows_csr(self.indptr, self.indices, self.values, self.n_vars, idx, out=out)

    def minibatch(self, idx: torch.Tensor) -> dict:
        idx = idx.to(self.indptr.device, dtype=torch.int64)
        return {
            "X": self.densify(idx),
            "batch": ops.gather_i64(self.batch, idx).view(-1, 1),
            "labels": ops.gather_i64(self.labels, idx).view(-1, 1),
        }


class DeviceResidentCounts:
    """Count matrix [N, G] (fp32, or uint16-compressed) + batch / label columns resident in HBM."""

    def __init__(self, X, batch=None, labels=None, device="cuda", compress_u16: bool = False,
                 compress: str | None = None):
        dev = torch.device(device)
        if isinstance(X, np.ndarray):
            X = torch.from_numpy(X)
        if hasattr(X, "toarray"):  # scipy sparse: densify once on the host (reference :152 does it per batch)
            X = torch.from_numpy(np.asarray(X.toarray()))
        if compress == "auto":
            X = X.to(narrowest_count_dtype(X))
        elif compress_u16 or compress == "u16":
            if float(X.max()) >= 65536 or float(X.min()) < 0:
                raise ValueError("uint16 residency needs counts in [0, 65535]")
            X = X.to(torch.uint16)
        elif compress == "u8":
            if float(X.max()) >= 256 or float(X.min()) < 0:
                raise ValueError("uint8 residency needs counts in [0, 255]")
            X = X.to(torch.uint8)
        elif X.dtype != torch.float32:
            X = X.to(torch.float32)
        if X.dtype == torch.uint8 and X.shape[1] % 16:
            # the tensor-core kernels copy uint8 rows in 16-byte pieces: rows start 16-byte aligned
            G = X.shape[1]
            store = torch.zeros(X.shape[0], (G + 15) // 16 * 16, dtype=torch.uint8, device=dev)
            store[:, :G].copy_(X)
            self.X = store[:, :G]
        else:
            self.X = X.to(dev).contiguous()
        n = self.X.shape[0]
        # non-negative integer counts (what a counts layer holds): lets the training kernels take their
        # straight-line per-element path (include/scvi_b200.h, SCVI_X_INTEGER); verified once, here
        self.integer_counts = self.X.dtype != torch.float32 or self._all_integer(self.X)

        def col(v):
            if v is None:
                return torch.zeros(n, dtype=torch.int64, device=dev)
            v = torch.as_tensor(np.asarray(v) if not isinstance(v, torch.Tensor) else v)
            return v.reshape(-1).to(torch.int64).to(dev).contiguous()

        self.batch = col(batch)
        self.labels = col(labels)

    @staticmethod
    def _all_integer(X: torch.Tensor, chunk: int = 1 << 16) -> bool:
        for s in range(0, X.shape[0], chunk):
            x = X[s : s + chunk]
            if not bool(((x >= 0) & (x == x.floor()) & (x < 8388608.0)).all()):
                return False
        return True

    @property
    def n_obs(self) -> int:
        return self.X.shape[0]

    @property
    def n_vars(self) -> int:
        return self.X.shape[1]

    def minibatch(self, idx: torch.Tensor) -> dict:
        """The dict `AnnTorchDataset.__getitem__` + collate produce, built on the device:
        X float32 (n, G), batch / labels int64 (n, 1)."""
        idx = idx.to(self.X.device, dtype=torch.int64)
        return {
            "X": ops.gather_rows(self.X, idx),
            "batch": ops.gather_i64(self.batch, idx).view(-1, 1),
            "labels": ops.gather_i64(self.labels, idx).view(-1, 1),
        }

    def iter_epoch(self, batch_size: int, epoch: int = 0, seed: int = 0, rank: int = 0, world: int = 1,
                   shuffle: bool = True, drop_last: bool = False, drop_dataset_tail: bool = False):
        """Yield device index tensors, one per minibatch (the sampler's batches)."""
        order = epoch_indices(self.n_obs, rank, world, epoch, seed, shuffle, drop_dataset_tail).to(self.X.device)
        for s in range(0, order.numel(), batch_size):
            chunk = order[s : s + batch_size]
            if chunk.numel() < batch_size and drop_last:
                break
            yield chunk
