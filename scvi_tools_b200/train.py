"""Fused SCVI training plan: the whole minibatch step as one fixed sequence of C-ABI kernels.

Stands in for `TrainingPlan.training_step` + Lightning's `loss.backward()` + `Adam.step()`
(reference src/scvi/train/_trainingplans.py:427-454,475-532): KL warm-up weight `epoch/400`
(:31-72), `torch.optim.Adam(lr=1e-3, eps=0.01, weight_decay=1e-6)` over the parameters that
receive gradient, mean-reduced loss.  No autograd graph is built: forward and backward are
the explicit kernel sequence of DESIGN.md section 3, reading minibatch rows straight from the
HBM-resident count matrix (`row_idx`), writing gradients into one flat arena that a single
multi-tensor Adam launch consumes.  The sequence is captured once in a CUDA graph.

Supported configuration = the reference default SCVI module: n_layers=1, one-hot batch
covariate, observed library, dispersion="gene", gene_likelihood in {"zinb","nb"},
BatchNorm in encoder and decoder, no LayerNorm.  Anything else: use the autograd path
(`VAE.forward` + torch optimizer), which runs on the same kernels.
"""

from __future__ import annotations

import ctypes as C
import os

import torch

from . import _lib
from ._lib import call, query
from .data import DeviceResidentCounts, epoch_indices
from .ops import BN_EPS, BN_MOMENTUM
from .ops import gather_i64 as ops_gather_i64


def compute_kl_weight(epoch, step, n_epochs_kl_warmup=400, n_steps_kl_warmup=None, max_kl_weight=1.0,
                      min_kl_weight=0.0):
    """KL warm-up schedule (reference src/scvi/train/_trainingplans.py:31-72)."""
    if min_kl_weight > max_kl_weight:
        raise ValueError(f"min_kl_weight={min_kl_weight} is larger than max_kl_weight={max_kl_weight}.")
    slope = max_kl_weight - min_kl_weight
    if n_epochs_kl_warmup:
        if epoch < n_epochs_kl_warmup:
            return slope * (epoch / n_epochs_kl_warmup) + min_kl_weight
    elif n_steps_kl_warmup:
        if step < n_steps_kl_warmup:
            return slope * (step / n_steps_kl_warmup) + min_kl_weight
    return max_kl_weight


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


ENC = "z_encoder.encoder.fc_layers.Layer 0."
DEC = "decoder.px_decoder.fc_layers.Layer 0."
# arena order: decoder-head gradients first (they are ready first in backward and are 2/3 of the
# bytes), so the data-parallel allreduce of that chunk overlaps the rest of backward
ARENA_ORDER = [
    "px_r",
    "decoder.px_scale_decoder.0.weight", "decoder.px_scale_decoder.0.bias",
    "decoder.px_dropout_decoder.weight", "decoder.px_dropout_decoder.bias",
    DEC + "0.weight", DEC + "0.bias", DEC + "1.weight", DEC + "1.bias",
    "z_encoder.mean_encoder.weight", "z_encoder.mean_encoder.bias",
    "z_encoder.var_encoder.weight", "z_encoder.var_encoder.bias",
    ENC + "0.weight", ENC + "0.bias", ENC + "1.weight", ENC + "1.bias",
]
N_HEAD_TENSORS = 5


class FusedTrainingPlan:
    """One object = module + optimizer state + preallocated step buffers for a fixed batch size."""

    def __init__(self, module, data: DeviceResidentCounts, batch_size: int = 128, lr: float = 1e-3,
                 eps: float = 0.01, weight_decay: float = 1e-6, betas=(0.9, 0.999),
                 n_epochs_kl_warmup: int | None = 400, n_steps_kl_warmup: int | None = None,
                 max_kl_weight: float = 1.0, min_kl_weight: float = 0.0, seed: int = 0,
                 use_cuda_graph: bool = True, process_group=None, ddp_mode: str = "p2p", p2p_buckets: int | None = None):
        m = module
        self._check_supported(m)
        self.module, self.data = m, data
        self.B = int(batch_size)
        self.lr, self.eps, self.wd, self.betas = lr, eps, weight_decay, betas
        self.n_epochs_kl_warmup, self.n_steps_kl_warmup = n_epochs_kl_warmup, n_steps_kl_warmup
        self.max_kl_weight, self.min_kl_weight = max_kl_weight, min_kl_weight
        self.seed = seed
        self.use_cuda_graph = use_cuda_graph
        self.pg = process_group
        # data parallel gradient exchange (world > 1):
        #   "p2p"     (default) this library's fused all-reduce + Adam over NVLink peer memory (csrc/comm.cu), one
        #             captured graph per step parity; the decoder-head bucket (2/3 of the arena, ready after K3) is
        #             exchanged and updated on a side stream under the rest of the backward, the remainder after it.
        #             Falls back to "split" when peer memory cannot be mapped on every rank.
        #   "split"   forward+backward graph | eager ncclAllReduce | Adam graph
        #   "graph"   the ncclAllReduce captured inside one step graph
        #   "overlap" as "graph", the decoder-head bucket reduced on a side stream
        if ddp_mode not in ("p2p", "graph", "overlap", "split"):
            raise ValueError(f"ddp_mode must be 'p2p', 'graph', 'overlap' or 'split' (got {ddp_mode!r})")
        self.world = 1 if process_group is None else torch.distributed.get_world_size(process_group)
        self.rank = 0 if process_group is None else torch.distributed.get_rank(process_group)
        self.exchange = None
        if self.world > 1 and ddp_mode == "p2p":
            from .comm import PeerGradientExchange

            self.exchange = PeerGradientExchange(self._arena_layout(module)[1], process_group, data.X.device)
            if not self.exchange.available:
                self.exchange, ddp_mode = None, "split"
        self.ddp_mode = ddp_mode
        self.ddp_overlap = ddp_mode == "overlap"
        # "p2p" forms: 0 = sharded: one launch at the end of the step in which every rank reduces + updates 1/world
        # of the arena and stores the new parameters into all replicas (csrc/comm.cu); 1 = one launch, every rank
        # reduces the whole arena (one-shot); 2 = one-shot in two buckets, the decoder-head bucket on a side
        # stream under the rest of the backward.  Measured (tools/ddp_modes.sh, profiles/r02_ddp_modes.txt), ms
        # per step against 0.452-0.455 on one GPU:   2 GPUs: 0.473 / 0.467 / 0.476, NCCL routes 0.494-0.500;
        # 8 GPUs: 0.491 / 0.506 / 0.509.  Default (None): one-shot up to 2 ranks, sharded beyond.
        if p2p_buckets is None:
            p2p_buckets = 1 if self.world <= 2 else 0
        if p2p_buckets not in (0, 1, 2):
            raise ValueError("p2p_buckets must be None, 0, 1 or 2")
        self.p2p_buckets = int(p2p_buckets)
        self._n_exec = 0  # whole steps executed so far: its low bit selects the gradient buffer (p2p double buffer)
        self.zinb = m.gene_likelihood == "zinb"
        self.dev = data.X.device
        self.G, self.H, self.L, self.nb = m.n_input, m.n_hidden, m.n_latent, m.n_batch
        self.current_epoch = 0
        self.global_step = 0
        self.kernel_launches_per_step = 0
        self._build_arena()
        self._alloc_step_buffers()
        self._graphs = None
        self._side = None
        self._ev_prologue = None
        self.overlap = True  # independent launches on a second stream / parallel graph branches
        if self.world > 1:
            self._comm = torch.cuda.Stream(device=self.dev)
            self._ev_heads_ready, self._ev_heads_reduced = torch.cuda.Event(), torch.cuda.Event()
            # the head-bucket kernel runs beside the rest of the backward: a grid small enough to co-reside
            self.head_blocks = 48
            self._sync_initial_state()

    def close(self):
        """Release the peer-memory exchange segment (data-parallel "p2p" mode); the plan is unusable afterwards."""
        self._graphs = None
        if self.exchange is not None:
            torch.cuda.synchronize()
            torch.distributed.barrier(group=self.pg)
            # the module's parameters move out of the segment before it is freed
            new_p = self.p.clone()
            named = dict(self.module.named_parameters())
            offs = self._arena_layout(self.module)[2]
            for k in self.keys:
                view = new_p[offs[k] : offs[k] + named[k].numel()].view(named[k].shape)
                named[k].data = view
                self.P[k] = view
            self.p = new_p
            self._g2, self._Gd2 = [], []
            self.exchange.close()
            self.exchange = None

    def _sync_initial_state(self):
        """Every replica starts from rank 0's parameters and BatchNorm buffers, as torch DDP does at construction
        (`DistributedDataParallel.__init__` -> `_sync_module_states`); the Adam moments and the step counter are
        zero on every rank."""
        for t in (self.p, self.bn1_rm, self.bn1_rv, self.bn2_rm, self.bn2_rv, self.bn1_nbt, self.bn2_nbt):
            torch.distributed.broadcast(t, src=0, group=self.pg)

    def _broadcast_buffers(self):
        """BN running buffers follow rank 0 (DDP `broadcast_buffers` semantics), once per epoch."""
        if self.world > 1:
            for b in (self.bn1_rm, self.bn1_rv, self.bn2_rm, self.bn2_rv):
                torch.distributed.broadcast(b, src=0, group=self.pg)

    @staticmethod
    def _check_supported(m):
        ok = (
            m.n_layers == 1 and m.dispersion == "gene" and m.use_observed_lib_size and m.log_variational
            and not m.encode_covariates and m.latent_distribution == "normal" and not m.use_size_factor_key
            and m.batch_representation == "one-hot"
            and m.n_hidden in (64, 128, 256) and m.n_latent <= 32
            and m.z_encoder.encoder.use_batch_norm and m.decoder.px_decoder.use_batch_norm
            and not m.z_encoder.encoder.use_layer_norm and not m.decoder.px_decoder.use_layer_norm
            and len(m.decoder.px_decoder.n_cat_list) == 1 and m.decoder.px_decoder.n_cont == 0
        )
        if not ok:
            raise NotImplementedError(
                "FusedTrainingPlan covers the default SCVI module configuration; use VAE.forward + a torch "
                "optimizer (same kernels through autograd) for other options"
            )

    # ------------------------------------------------------------------ arena
    @staticmethod
    def _arena_layout(module):
        """(keys, n_floats, offsets, n_head) of the flat arena for `module`."""
        named = dict(module.named_parameters())
        zinb = module.gene_likelihood == "zinb"
        keys = [k for k in ARENA_ORDER if not (("px_dropout" in k) and not zinb)]
        offs, off = {}, 0
        for k in keys:
            offs[k] = off
            off += (named[k].numel() + 3) // 4 * 4  # keep every slice 16-byte aligned
        return keys, off, offs, offs[keys[N_HEAD_TENSORS if zinb else 3]]

    def _build_arena(self):
        named = dict(self.module.named_parameters())
        keys, off, offs, self.n_head = self._arena_layout(self.module)
        self.keys = keys
        self.n_arena = off
        dev = self.dev
        self.p = self.exchange.p.zero_() if self.exchange is not None else torch.zeros(off, dtype=torch.float32, device=dev)
        # gradient arena(s): inside the peer-mapped exchange segment, one per step parity, under "p2p"
        self._g2 = self.exchange.g if self.exchange is not None else [torch.zeros(off, dtype=torch.float32, device=dev)]
        self.m = torch.zeros(off, dtype=torch.float32, device=dev)
        self.v = torch.zeros(off, dtype=torch.float32, device=dev)
        self.P, self._Gd2 = {}, [{} for _ in self._g2]
        for k in keys:
            prm = named[k]
            n = prm.numel()
            view = self.p[offs[k] : offs[k] + n].view(prm.shape)
            view.copy_(prm.data.to(dev))
            prm.data = view  # the module's parameter now lives in the arena
            self.P[k] = view
            for g, Gd in zip(self._g2, self._Gd2, strict=True):
                Gd[k] = g[offs[k] : offs[k] + n].view(prm.shape)
        self.step_counter = torch.zeros(2, dtype=torch.int64, device=dev)  # {steps done, Adam's arrival ticket}
        self.klw = torch.zeros(1, dtype=torch.float32, device=dev)
        # make sure the BN buffers are on the device
        self.module.to(dev)
        sd = dict(self.module.named_buffers())
        self.bn1_rm, self.bn1_rv = sd[ENC + "1.running_mean"], sd[ENC + "1.running_var"]
        self.bn2_rm, self.bn2_rv = sd[DEC + "1.running_mean"], sd[DEC + "1.running_var"]
        self.bn1_nbt, self.bn2_nbt = sd[ENC + "1.num_batches_tracked"], sd[DEC + "1.num_batches_tracked"]

    @property
    def _x_integer(self):
        """The matrix the step reads holds non-negative integer counts only (its data class verified it)."""
        return bool(getattr(self.data, "integer_counts", False))

    def _parity(self, n_exec=None):
        return ((self._n_exec if n_exec is None else n_exec) & 1) if len(self._g2) == 2 else 0

    @property
    def g(self):
        """The gradient arena the most recent step wrote (this rank's gradients; under NCCL modes: their sum)."""
        return self._g2[self._parity(max(self._n_exec - 1, 0))]

    @property
    def Gd(self):
        return self._Gd2[self._parity(max(self._n_exec - 1, 0))]

    # ------------------------------------------------------------------ buffers
    def _alloc_step_buffers(self):
        B, G, H, L, dev = self.B, self.G, self.H, self.L, self.dev
        f = lambda *s: torch.empty(*s, dtype=torch.float32, device=dev)  # noqa: E731
        T = (B + _lib.BN_TILE - 1) // _lib.BN_TILE
        self.idx = torch.zeros(B, dtype=torch.int64, device=dev)
        self.cov_cols = torch.zeros(B, 1, dtype=torch.int32, device=dev)
        self.batch_cols = (self.data.batch + L).to(torch.int32)  # one-hot column of Wd per cell
        self.y1, self.a1, self.dy1, self.da1 = f(B, H), f(B, H), f(B, H), f(B, H)
        self.y2, self.h, self.dy2, self.dh = f(B, H), f(B, H), f(B, H), f(B, H)
        self.part1, self.part2 = f(T, 2, H), f(T, 2, H)
        self.mean1, self.rstd1, self.mean2, self.rstd2 = f(H), f(H), f(H), f(H)
        self.library = f(B)
        self.maskb = torch.ones(B, H, dtype=torch.bool, device=dev)
        self.mask = self.maskb.view(torch.uint8)  # keep-mask, 1 byte per element
        self.B_cur = B
        self.noise = f(B, L)
        self.q_m, self.q_v, self.z, self.dz = f(B, L), f(B, L), f(B, L), f(B, L)
        self.kl, self.reconst, self.stats = f(B), f(B), f(B, 4)
        self.g_rec, self.g_kl = f(B), f(B)
        self.out3 = torch.zeros(3, dtype=torch.float32, device=dev)
        enc = self.module.z_encoder.encoder.fc_layers[0]
        self.p_drop = float(enc[4].p) if enc[4] is not None else 0.0
        lik = _lib.LIK_ZINB if self.zinb else _lib.LIK_NB
        nbytes = max(
            query("scvi_linear_bwd_ws_bytes", B, G, H, G),
            query("scvi_linear_bwd_ws_bytes", B, L, H, L + max(self.nb, 0)),
            query("scvi_bn_bwd_ws_bytes", B, H),
            query("scvi_latent_bwd_ws_bytes", B, H, L),
            query("scvi_decoder_nll_bwd_ws_bytes", B, H, G, lik),
        )
        self.ws = torch.empty(int(nbytes), dtype=torch.uint8, device=dev)
        ncols = L + (self.nb if self.nb > 1 else 0)
        # fused middle (BN/ReLU/dropout, latent heads + KL, decoder hidden layer and their backward)
        self.use_mid = H == 128 and L <= 16 and ncols <= 80
        self.ws_mid = (torch.empty(int(query("scvi_mid_ws_bytes", B, L, ncols)), dtype=torch.uint8, device=dev)
                       if self.use_mid else None)
        self.dy1_amax = f(int(query("scvi_mid_n_tiles", B)))  # max |dy1| per row tile (mid_bwd -> encoder dW)
        # SCVI_AOPS=1 (integer counts + tcgen05 kernels): the fused likelihood kernel keeps its per-element operands in
        # the forward workspace and the head-weight gradients are a pure tensor-core GEMM over them (SCVI_PREC_AOPS).
        # Opt-in: parity-green, but keeping the operands costs the fused kernel ~15 us and the dispersion derivative
        # has to be re-derived per element inside the GEMM kernel: 0.427 ms per step against 0.379 ms (DESIGN.md
        # section 9, experiments/README.md)
        self._aops = (_lib.PREC_AOPS if (self.module.precision == _lib.PREC_TC and H == 128 and self._x_integer
                                         and os.environ.get("SCVI_AOPS", "0") == "1") else 0)
        nf = query("scvi_decoder_nll_fwd_ws_bytes", B, H, G, lik,
                   self.module.precision | (_lib.X_INTEGER | self._aops if self._aops else 0))
        self.ws_fwd = torch.empty(int(nf), dtype=torch.uint8, device=dev) if nf > 0 else None
        self.enc_tc = _lib.LIN_TC if (self.module.precision == _lib.PREC_TC and H == 128) else 0
        ne = query("scvi_linear_fwd_ws_bytes", B, G, H, _lib.LIN_LOG1P | self.enc_tc)
        self.ws_enc = torch.empty(int(ne), dtype=torch.uint8, device=dev) if ne > 0 else None

    # ------------------------------------------------------------------ the step
    def _forward_backward(self, st, overlap: bool = False, parity: int | None = None, prologue: dict | None = None):
        """Kernel sequence of one minibatch: forward, ELBO, backward into the gradient arena.
        Rows are read from the resident matrix through `self.idx`.  `st` = cudaStream_t of the current stream.

        `overlap`: the launches that depend on nothing the main chain is computing run on a second stream
        (fork / join with events, captured as parallel branches of the step graph): the decoder-head packing
        (weights only) next to the encoder + middle forward, the ELBO scalar reduction (not needed by the
        backward, whose seeds come from the step prologue) next to the decoder backward."""
        B, G, H, L = self.B_cur, self.G, self.H, self.L
        parity = self._parity() if parity is None else parity
        P, Gd, X = self.P, self._Gd2[parity], self.data.X
        if X.dtype == torch.uint8:
            if not (self.enc_tc and self.module.precision == _lib.PREC_TC):
                raise NotImplementedError("uint8-resident counts need the tcgen05 kernels (precision=1, n_hidden=128)")
            if X.stride(0) % 16 or X.data_ptr() % 16:
                raise ValueError("uint8-resident counts need 16-byte aligned rows (DeviceResidentCounts pads them)")
        elif X.dtype != torch.float32:
            raise NotImplementedError("fused training reads fp32 or uint8 resident counts (uint16 residency: gather path)")
        ldx = X.stride(0)
        x_lin = _lib.LIN_X_U8 if X.dtype == torch.uint8 else 0
        lik = _lib.LIK_ZINB if self.zinb else _lib.LIK_NB
        prec = self.module.precision
        # how the training kernels may read the counts (uint8 bytes / guaranteed-integer fp32 / any fp32)
        xkind = 0
        if prec == _lib.PREC_TC:
            xkind = _lib.X_U8 if X.dtype == torch.uint8 else (_lib.X_INTEGER if self._x_integer else 0)
        aops = self._aops if xkind else 0
        Ws, bs = P["decoder.px_scale_decoder.0.weight"], P["decoder.px_scale_decoder.0.bias"]
        Wp = P.get("decoder.px_dropout_decoder.weight")
        bp = P.get("decoder.px_dropout_decoder.bias")
        overlap = overlap and prec == _lib.PREC_TC
        main = torch.cuda.current_stream()
        if overlap:
            if self._side is None:
                self._side = torch.cuda.Stream(device=self.dev)
            side = self._side
            side.wait_stream(main)
            if prologue is not None:
                # the step prologue (random buffers, covariate columns, backward seeds) feeds the middle of the
                # step, not the encoder's input layer: it runs beside E1
                self._prologue(C.c_void_p(side.cuda_stream), **prologue)
                if self._ev_prologue is None:
                    self._ev_prologue = torch.cuda.Event()
                self._ev_prologue.record(side)
            call("scvi_decoder_nll_prepack", B, H, G, _p(Ws), _p(bs), _p(Wp), _p(bp), _p(P["px_r"]), lik, prec, 1,
                 _p(self.ws_fwd), C.c_void_p(side.cuda_stream))
        if prologue is not None and not overlap:
            self._prologue(st, **prologue)
        # ---- encoder layer 0: log1p + library + GEMM + BN tile stats (reads X rows via idx)
        call("scvi_linear_fwd", _p(X), ldx, _p(self.idx), B, G, H, _p(P[ENC + "0.weight"]), G, _p(P[ENC + "0.bias"]),
             None, 0, _lib.LIN_LOG1P | _lib.LIN_LIBRARY | self.enc_tc | x_lin, _p(self.y1), H, _p(self.library), _p(self.part1),
             _p(self.ws_enc), st)
        drop = self.p_drop > 0
        dscale = 1.0 / (1.0 - self.p_drop) if drop else 1.0
        inj = self.nb > 1
        ncols = L + (self.nb if inj else 0)
        if overlap and prologue is not None:
            main.wait_event(self._ev_prologue)
        if self.use_mid:
            call("scvi_mid_fwd", _p(self.y1), _p(self.part1), B, H, L, _p(P[ENC + "1.weight"]), _p(P[ENC + "1.bias"]),
                 _p(self.bn1_rm), _p(self.bn1_rv), _p(self.mask) if drop else None, dscale, _p(self.noise),
                 _p(P["z_encoder.mean_encoder.weight"]), _p(P["z_encoder.mean_encoder.bias"]),
                 _p(P["z_encoder.var_encoder.weight"]), _p(P["z_encoder.var_encoder.bias"]),
                 self.module.z_encoder.var_eps, _p(P[DEC + "0.weight"]), ncols, _p(P[DEC + "0.bias"]),
                 _p(self.cov_cols) if inj else None, _p(P[DEC + "1.weight"]), _p(P[DEC + "1.bias"]), _p(self.bn2_rm),
                 _p(self.bn2_rv), BN_EPS, BN_MOMENTUM, _p(self.a1), _p(self.q_m), _p(self.q_v), _p(self.z), _p(self.kl),
                 _p(self.y2), _p(self.h), _p(self.mean1), _p(self.rstd1), _p(self.mean2), _p(self.rstd2),
                 _p(self.ws_mid), st)
        else:
            self._middle_forward_unfused(st, B, H, L, drop, dscale, inj, ncols)
        # ---- fused heads + softmax + likelihood
        # the ELBO is a mean over the minibatch, so d(loss)/d(reconst) = 1/B is known before the forward
        # (seeded by the step prologue): the fused kernel emits dh together with the likelihood
        if overlap:
            main.wait_stream(side)  # packed weight tiles ready
        call("scvi_decoder_nll_fwd", _p(self.h), H, B, H, G, _p(Ws), _p(bs), _p(Wp), _p(bp), _p(P["px_r"]), _p(X), ldx,
             _p(self.idx), _p(self.library), lik, prec | xkind | aops | (_lib.PREC_PREPACKED if overlap else 0), _p(self.reconst),
             _p(self.stats), _p(self.g_rec), _p(self.dh), H, _p(self.ws_fwd), st)
        if overlap:
            side.wait_stream(main)
            call("scvi_elbo_reduce", _p(self.reconst), _p(self.kl), B, 0.0, _p(self.klw), _p(self.out3),
                 C.c_void_p(side.cuda_stream))
        else:
            call("scvi_elbo_reduce", _p(self.reconst), _p(self.kl), B, 0.0, _p(self.klw), _p(self.out3), st)
        # ---- backward
        call("scvi_decoder_nll_bwd", _p(self.h), H, B, H, G, _p(Ws), _p(bs), _p(Wp), _p(bp), _p(P["px_r"]), _p(X), ldx,
             _p(self.idx), _p(self.library), _p(self.stats), _p(self.g_rec), lik, prec | xkind | aops, None, H,
             _p(Gd["decoder.px_scale_decoder.0.weight"]), _p(Gd["decoder.px_scale_decoder.0.bias"]),
             _p(Gd.get("decoder.px_dropout_decoder.weight")), _p(Gd.get("decoder.px_dropout_decoder.bias")),
             _p(Gd["px_r"]), _p(self.ws), _p(self.ws_fwd) if prec == _lib.PREC_TC else None, st)
        self._after_head_grads(st, parity)
        if self.use_mid:
            call("scvi_mid_bwd", _p(self.dh), _p(self.y2), _p(self.mean2), _p(self.rstd2), _p(P[DEC + "1.weight"]),
                 _p(P[DEC + "1.bias"]), _p(self.z), _p(P[DEC + "0.weight"]), ncols, _p(self.cov_cols) if inj else None,
                 _p(self.g_kl), _p(self.q_m), _p(self.q_v), _p(self.noise), self.module.z_encoder.var_eps,
                 _p(P["z_encoder.mean_encoder.weight"]), _p(P["z_encoder.var_encoder.weight"]), _p(self.a1), _p(self.y1),
                 _p(self.mean1), _p(self.rstd1), _p(P[ENC + "1.weight"]), _p(P[ENC + "1.bias"]),
                 _p(self.mask) if drop else None, dscale, B, H, L, _p(self.dy1), _p(Gd[DEC + "1.weight"]),
                 _p(Gd[DEC + "1.bias"]), _p(Gd[DEC + "0.weight"]), _p(Gd[DEC + "0.bias"]),
                 _p(Gd["z_encoder.mean_encoder.weight"]), _p(Gd["z_encoder.mean_encoder.bias"]),
                 _p(Gd["z_encoder.var_encoder.weight"]), _p(Gd["z_encoder.var_encoder.bias"]), _p(Gd[ENC + "1.weight"]),
                 _p(Gd[ENC + "1.bias"]), _p(Gd[ENC + "0.bias"]), _p(self.dy1_amax), _p(self.ws_mid), st)
        else:
            self._middle_backward_unfused(st, B, H, L, drop, dscale, inj, ncols, Gd)
        call("scvi_linear_bwd", _p(X), ldx, _p(self.idx), B, G, H, _p(P[ENC + "0.weight"]), G, G, None, 0,
             _lib.LIN_LOG1P | self.enc_tc | x_lin, _p(self.dy1), H, _p(Gd[ENC + "0.weight"]),
             None if self.use_mid else _p(Gd[ENC + "0.bias"]), None, 0,
             _p(self.dy1_amax) if self.use_mid else None, self.dy1_amax.numel() if self.use_mid else 0, _p(self.ws), st)
        if overlap:
            main.wait_stream(side)  # join: the ELBO scalars are part of the step's result

    def _middle_forward_unfused(self, st, B, H, L, drop, dscale, inj, ncols):
        P = self.P
        call("scvi_bn_finalize", _p(self.part1), B, H, 1, BN_EPS, BN_MOMENTUM, _p(self.bn1_rm), _p(self.bn1_rv),
             _p(self.mean1), _p(self.rstd1), st)
        drop = self.p_drop > 0
        call("scvi_bn_act_drop_fwd", _p(self.y1), H, B, H, _p(self.mean1), _p(self.rstd1), _p(P[ENC + "1.weight"]),
             _p(P[ENC + "1.bias"]), 1, _p(self.mask) if drop else None, 1.0 / (1.0 - self.p_drop) if drop else 1.0,
             _p(self.a1), H, st)
        # ---- latent heads + rsample + KL
        call("scvi_latent_fwd", _p(self.a1), H, B, H, L, _p(P["z_encoder.mean_encoder.weight"]),
             _p(P["z_encoder.mean_encoder.bias"]), _p(P["z_encoder.var_encoder.weight"]),
             _p(P["z_encoder.var_encoder.bias"]), _p(self.noise), self.module.z_encoder.var_eps, _p(self.q_m),
             _p(self.q_v), _p(self.z), _p(self.kl), st)
        # ---- decoder layer 0 with the batch covariate as a weight-column gather
        inj = self.nb > 1
        ncols = L + (self.nb if inj else 0)
        call("scvi_linear_fwd", _p(self.z), L, None, B, L, H, _p(P[DEC + "0.weight"]), ncols, _p(P[DEC + "0.bias"]),
             _p(self.cov_cols) if inj else None, 1 if inj else 0, 0, _p(self.y2), H, None, _p(self.part2), None, st)
        call("scvi_bn_finalize", _p(self.part2), B, H, 1, BN_EPS, BN_MOMENTUM, _p(self.bn2_rm), _p(self.bn2_rv),
             _p(self.mean2), _p(self.rstd2), st)
        call("scvi_bn_act_drop_fwd", _p(self.y2), H, B, H, _p(self.mean2), _p(self.rstd2), _p(P[DEC + "1.weight"]),
             _p(P[DEC + "1.bias"]), 1, None, 1.0, _p(self.h), H, st)

    def _middle_backward_unfused(self, st, B, H, L, drop, dscale, inj, ncols, Gd):
        P = self.P
        call("scvi_bn_act_drop_bwd", _p(self.dh), H, _p(self.y2), H, B, H, _p(self.mean2), _p(self.rstd2),
             _p(P[DEC + "1.weight"]), _p(P[DEC + "1.bias"]), 1, None, 1.0, 1, _p(self.dy2), H, _p(Gd[DEC + "1.weight"]),
             _p(Gd[DEC + "1.bias"]), _p(self.ws), st)
        call("scvi_linear_bwd", _p(self.z), L, None, B, L, H, _p(P[DEC + "0.weight"]), ncols, ncols,
             _p(self.cov_cols) if inj else None, 1 if inj else 0, 0, _p(self.dy2), H, _p(Gd[DEC + "0.weight"]),
             _p(Gd[DEC + "0.bias"]), _p(self.dz), L, None, 0, _p(self.ws), st)
        call("scvi_latent_bwd", _p(self.a1), H, B, H, L, _p(P["z_encoder.mean_encoder.weight"]),
             _p(P["z_encoder.var_encoder.weight"]), _p(self.noise), self.module.z_encoder.var_eps, _p(self.q_m),
             _p(self.q_v), None, None, _p(self.dz), _p(self.g_kl), _p(self.da1), H,
             _p(Gd["z_encoder.mean_encoder.weight"]), _p(Gd["z_encoder.mean_encoder.bias"]),
             _p(Gd["z_encoder.var_encoder.weight"]), _p(Gd["z_encoder.var_encoder.bias"]), _p(self.ws), st)
        call("scvi_bn_act_drop_bwd", _p(self.da1), H, _p(self.y1), H, B, H, _p(self.mean1), _p(self.rstd1),
             _p(P[ENC + "1.weight"]), _p(P[ENC + "1.bias"]), 1, _p(self.mask) if drop else None,
             1.0 / (1.0 - self.p_drop) if drop else 1.0, 1, _p(self.dy1), H, _p(Gd[ENC + "1.weight"]),
             _p(Gd[ENC + "1.bias"]), _p(self.ws), st)

    def _after_head_grads(self, st, parity):
        """Between the decoder-head backward and the rest: under data parallelism the head gradients (the first
        `n_head` floats of the arena) are exchanged here, on a side stream, under the rest of the backward --
        "p2p": fused peer-memory all-reduce + Adam of that bucket; "overlap": its ncclAllReduce."""
        if self.world == 1 or not ((self.exchange is not None and self.p2p_buckets > 1) or self.ddp_overlap):
            return
        main = torch.cuda.current_stream()
        self._ev_heads_ready.record(main)
        self._comm.wait_event(self._ev_heads_ready)
        with torch.cuda.stream(self._comm):
            if self.exchange is not None:
                self.exchange.allreduce_adam(parity, 0, self.p, self.m, self.v, 0, self.n_head, self.step_counter,
                                             False, self.lr, self.betas[0], self.betas[1], self.eps, self.wd,
                                             C.c_void_p(self._comm.cuda_stream), blocks=self.head_blocks)
            else:
                torch.distributed.all_reduce(self._g2[parity][: self.n_head], group=self.pg)
            self._ev_heads_reduced.record(self._comm)

    def _reduce_gradients(self, parity=0):
        """NCCL modes: sum the gradient arena over the ranks (Adam applies 1/world)."""
        if self.world == 1 or self.exchange is not None:
            return
        g = self._g2[parity]
        if self.ddp_overlap:
            torch.distributed.all_reduce(g[self.n_head :], group=self.pg)
            torch.cuda.current_stream().wait_event(self._ev_heads_reduced)
        else:
            torch.distributed.all_reduce(g, group=self.pg)

    def _adam(self, st, parity=None):
        """The optimizer step; under "p2p" the remaining bucket's fused all-reduce + Adam (which also advances the
        step counter), after the head bucket's kernel on the side stream has finished."""
        parity = self._parity() if parity is None else parity
        if self.exchange is not None and self.world > 1 and self.p2p_buckets == 0:
            self.exchange.allreduce_adam_sharded(parity, 2, self.m, self.v, self.step_counter, self.lr, self.betas[0],
                                                 self.betas[1], self.eps, self.wd, st)
            return
        if self.exchange is not None and self.world > 1:
            begin = 0
            if self.p2p_buckets > 1:
                torch.cuda.current_stream().wait_event(self._ev_heads_reduced)
                begin = self.n_head
            self.exchange.allreduce_adam(parity, 1, self.p, self.m, self.v, begin, self.n_arena,
                                         self.step_counter, True, self.lr, self.betas[0], self.betas[1], self.eps,
                                         self.wd, st)
            return
        call("scvi_adam_step_graph", _p(self.p), _p(self._g2[parity]), _p(self.m), _p(self.v), self.n_arena,
             _p(self.step_counter), self.lr, self.betas[0], self.betas[1], self.eps, self.wd, 1.0 / self.world, st)

    # ------------------------------------------------------------------ driving the step
    def _prologue(self, st, feed_rng: bool = True, gen_mask: bool | None = None, gen_noise: bool | None = None):
        """Dropout keep-mask, reparameterisation noise, batch-covariate columns, ELBO backward seeds and the
        BatchNorm step counters in one launch (library Philox streams keyed by seed / step / rank).
        `gen_mask` / `gen_noise` (default: `feed_rng`) select which of the two random buffers are drawn; a
        buffer the caller filled itself is left alone."""
        B = self.B_cur
        gen_mask = feed_rng if gen_mask is None else gen_mask
        gen_noise = feed_rng if gen_noise is None else gen_noise
        drop = gen_mask and self.p_drop > 0
        call("scvi_step_prologue", self.seed, _p(self.step_counter), self.rank, B, self.H, self.L, self.p_drop,
             _p(self.mask) if drop else None, _p(self.noise) if gen_noise else None, _p(self.idx),
             _p(self.batch_cols) if self.nb > 1 else None, _p(self.cov_cols) if self.nb > 1 else None, 0.0,
             _p(self.klw), 1.0, _p(self.g_rec), _p(self.g_kl), _p(self.bn1_nbt), _p(self.bn2_nbt), st)

    def _whole_step(self, gen_mask: bool = True, gen_noise: bool = True, parity: int | None = None):
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        parity = self._parity() if parity is None else parity
        self._forward_backward(st, overlap=self.overlap, parity=parity,
                               prologue={"gen_mask": gen_mask, "gen_noise": gen_noise})
        self._reduce_gradients(parity)
        self._adam(st, parity)

    def _capture(self):
        """Capture the full-batch step once.  With data parallelism the graph is cut at the gradient
        allreduce (fwd+bwd graph -> NCCL -> Adam graph)."""
        self.B_cur = self.B
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):  # warm-up outside capture (lazy module loading, cudaFuncSetAttribute)
            state = (self.p, self.m, self.v, self.step_counter, self.bn1_rm, self.bn1_rv, self.bn2_rm, self.bn2_rv,
                     self.bn1_nbt, self.bn2_nbt)
            snap = tuple(t.clone() for t in state)
            n0 = query("scvi_launch_count")
            self._whole_step()
            self._n_exec += 1  # the warm-up is a real exchange on every rank: the parity moves with it
            self.kernel_launches_per_step = query("scvi_launch_count") - n0  # this library's kernels only
            for dst, src in zip(state, snap, strict=True):
                dst.copy_(src)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        self._graphs = []
        for parity in range(len(self._g2)):  # "p2p": one graph per gradient buffer, replayed alternately
            if self.world == 1 or self.ddp_mode != "split":  # collectives are captured with the kernels
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    self._whole_step(parity=parity)
                self._graphs.append((g,))
            else:
                g1, g2 = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
                with torch.cuda.graph(g1):
                    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
                    self._forward_backward(st, overlap=self.overlap, parity=parity, prologue={})
                with torch.cuda.graph(g2):
                    self._adam(C.c_void_p(torch.cuda.current_stream().cuda_stream), parity)
                self._graphs.append((g1, g2))

    def set_epoch(self, epoch: int):
        self.current_epoch = epoch
        klw = compute_kl_weight(epoch, self.global_step, self.n_epochs_kl_warmup, self.n_steps_kl_warmup,
                                self.max_kl_weight, self.min_kl_weight)
        self.klw.fill_(float(klw))

    def training_step(self, idx: torch.Tensor, drop_mask=None, noise=None):
        """One optimisation step on the cells `idx` (device int64).  Returns the device tensor
        `out3` = (loss, sum reconst, sum kl) without synchronising.  Explicit `drop_mask` /
        `noise` (parity tests) force the eager path."""
        n = idx.numel()
        self.idx[:n].copy_(idx, non_blocking=True)
        if not self.n_epochs_kl_warmup and self.n_steps_kl_warmup:
            # step-based warm-up: the weight follows `global_step` (reference _trainingplans.py:31-72,248-258)
            self.klw.fill_(float(compute_kl_weight(self.current_epoch, self.global_step, self.n_epochs_kl_warmup,
                                                   self.n_steps_kl_warmup, self.max_kl_weight, self.min_kl_weight)))
        fed = drop_mask is not None or noise is not None
        if fed or n != self.B or not self.use_cuda_graph:
            self.B_cur = n
            if drop_mask is not None:
                self.mask[:n].copy_(drop_mask.to(self.dev, dtype=torch.uint8))
            if noise is not None:
                self.noise[:n].copy_(noise.to(self.dev))
            # whichever of the two was not supplied is drawn from the library streams
            self._whole_step(gen_mask=drop_mask is None, gen_noise=noise is None)
            self.B_cur = self.B
        else:
            if self._graphs is None:
                self._capture()
            parity = self._parity()
            graphs = self._graphs[parity]
            if len(graphs) == 1:
                graphs[0].replay()
            else:  # forward + backward | all-reduce | Adam
                graphs[0].replay()
                torch.distributed.all_reduce(self._g2[parity], group=self.pg)
                graphs[1].replay()
        self._n_exec += 1
        self.global_step += 1
        return self.out3

    def attach_streamed(self, host):
        """Train from pinned-host minibatches (`HostStreamedCounts`): the step reads rows of `host.stage`, which
        the copy stream fills one slot ahead of the compute stream."""
        if host.B != self.B or host.n_vars != self.G:
            raise ValueError("HostStreamedCounts batch size / gene count differ from the plan's")
        self.data = host.stage
        self.batch_cols = torch.zeros(2 * self.B, dtype=torch.int32, device=self.dev)
        L = self.L

        def batch_codes_to_columns(lo, n):  # runs on the copy stream, right after the batch codes landed
            self.batch_cols[lo : lo + n].copy_(host.stage.batch[lo : lo + n] + L)

        host.on_batch = batch_codes_to_columns
        self._graphs = None  # re-capture against the staging matrix
        self.host = host

    def fit_streamed(self, host, max_epochs: int, shuffle: bool = True, drop_last: bool = False):
        """`fit` for host-resident counts: minibatch i+1 crosses PCIe (compact integer rows) while step i runs."""
        if getattr(self, "host", None) is not host:
            self.attach_streamed(host)
        history = []
        first = self.current_epoch
        for epoch in range(first, first + max_epochs):
            self.set_epoch(epoch)
            order = epoch_indices(host.n_obs, self.rank, self.world, epoch, self.seed, shuffle)
            chunks = [order[s : s + self.B] for s in range(0, order.numel(), self.B)]
            chunks = [c.sort().values for c in chunks if not (c.numel() < self.B and (drop_last or c.numel() < 2))]
            acc = torch.zeros(3, dtype=torch.float64, device=self.dev)
            n_cells = 0
            if chunks:
                host.submit(0, chunks[0])
            for i, c in enumerate(chunks):
                rows = host.acquire(i % 2, c.numel())
                out = self.training_step(rows)
                host.release(i % 2)
                if i + 1 < len(chunks):  # host gather + H2D of the next minibatch overlap this step
                    host.submit((i + 1) % 2, chunks[i + 1])
                acc[0] += out[0].double() * c.numel()
                acc[1:] += out[1:].double()
                n_cells += c.numel()
            self._broadcast_buffers()
            a = acc.cpu()
            history.append({"epoch": epoch, "train_loss": float(a[0]) / max(n_cells, 1),
                            "reconstruction_loss": float(a[1]) / max(n_cells, 1),
                            "kl_local": float(a[2]) / max(n_cells, 1), "n_cells": n_cells,
                            "kl_weight": float(self.klw.item())})
        self.current_epoch = first + max_epochs
        return history

    def fit_csr(self, csr, max_epochs: int, shuffle: bool = True, drop_last: bool = False):
        """`fit` for CSR-resident counts (`DeviceResidentCSR`): each minibatch is densified on the device into a
        fixed fp32 staging matrix of `batch_size` rows (one `scvi_gather_rows_csr` launch ahead of the step graph);
        the step itself is unchanged and reads staging rows 0..n-1."""
        if getattr(self, "csr", None) is not csr:
            if csr.n_vars != self.G:
                raise ValueError("DeviceResidentCSR gene count differs from the plan's")
            if self.data.n_obs != self.B:  # the plan was built on something else: swap in a staging matrix
                self.data = DeviceResidentCounts(torch.zeros(self.B, self.G),
                                                 batch=torch.zeros(self.B, dtype=torch.int64), device=self.dev)
            self.data.integer_counts = csr.integer_counts  # what densify() will write into the staging rows
            self.batch_cols = torch.zeros(self.B, dtype=torch.int32, device=self.dev)
            self._graphs = None
            self.csr = csr
            self._stage_rows = torch.arange(self.B, device=self.dev)
        history = []
        first = self.current_epoch
        for epoch in range(first, first + max_epochs):
            self.set_epoch(epoch)
            order = epoch_indices(csr.n_obs, self.rank, self.world, epoch, self.seed, shuffle).to(self.dev)
            acc = torch.zeros(3, dtype=torch.float64, device=self.dev)
            n_cells = 0
            for s in range(0, order.numel(), self.B):
                chunk = order[s : s + self.B].sort().values
                n = chunk.numel()
                if n < self.B and (drop_last or n < 2):
                    break
                csr.densify(chunk, out=self.data.X[:n])
                self.batch_cols[:n].copy_(ops_gather_i64(csr.batch, chunk) + self.L)
                out = self.training_step(self._stage_rows[:n])
                acc[0] += out[0].double() * n
                acc[1:] += out[1:].double()
                n_cells += n
            self._broadcast_buffers()
            a = acc.cpu()
            history.append({"epoch": epoch, "train_loss": float(a[0]) / max(n_cells, 1),
                            "reconstruction_loss": float(a[1]) / max(n_cells, 1),
                            "kl_local": float(a[2]) / max(n_cells, 1), "n_cells": n_cells,
                            "kl_weight": float(self.klw.item())})
        self.current_epoch = first + max_epochs
        return history

    def fit(self, max_epochs: int, shuffle: bool = True, drop_last: bool = False, log_every_epoch: bool = True):
        """Train for `max_epochs` over this rank's shard; returns the per-epoch history
        (train loss, ELBO pieces per cell) -- the numbers `TrainingPlan` logs."""
        history = []
        first = self.current_epoch
        for epoch in range(first, first + max_epochs):
            self.set_epoch(epoch)
            order = epoch_indices(self.data.n_obs, self.rank, self.world, epoch, self.seed, shuffle).to(self.dev)
            # a minibatch is a set of cells: ascending row ids inside each one cost nothing (one sort per epoch)
            # and give the row gathers better DRAM page / TLB locality
            nfull = order.numel() // self.B
            if nfull:
                order[: nfull * self.B] = order[: nfull * self.B].view(nfull, self.B).sort(dim=1).values.reshape(-1)
            if order.numel() > nfull * self.B:
                order[nfull * self.B :] = order[nfull * self.B :].sort().values
            acc = torch.zeros(3, dtype=torch.float64, device=self.dev)
            n_cells = 0
            for s in range(0, order.numel(), self.B):
                chunk = order[s : s + self.B]
                if chunk.numel() < self.B and (drop_last or chunk.numel() < 2):
                    break
                out = self.training_step(chunk)
                acc[0] += out[0].double() * chunk.numel()
                acc[1:] += out[1:].double()
                n_cells += chunk.numel()
            self._broadcast_buffers()
            if log_every_epoch:
                a = acc.cpu()
                history.append({"epoch": epoch, "train_loss": float(a[0]) / max(n_cells, 1),
                                "reconstruction_loss": float(a[1]) / max(n_cells, 1),
                                "kl_local": float(a[2]) / max(n_cells, 1), "n_cells": n_cells,
                                "kl_weight": float(self.klw.item())})
        self.current_epoch = first + max_epochs
        return history
