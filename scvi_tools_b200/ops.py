"""torch.autograd wrappers over the C-ABI kernels (PyTorch is plumbing: memory, streams, autograd graph).

Each Function names the reference code it stands in for; all arithmetic happens in
``libscvi_b200.so``.  Inputs must be CUDA float32; there is no CPU path.
"""

from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import call, query

BN_EPS = 1e-3  # reference src/scvi/nn/_base_components.py:111
BN_MOMENTUM = 0.01


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _check(t, name, dtype=torch.float32):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"scvi_tools_b200: `{name}` must be a CUDA tensor (there is no CPU fallback)")
    if t.dtype != dtype:
        raise TypeError(f"scvi_tools_b200: `{name}` must be {dtype}, got {t.dtype}")
    return t


def _rowmajor(t, name, dtype=torch.float32):
    """2-D tensor with unit inner stride (row stride free); copies only if needed."""
    _check(t, name, dtype)
    if t.dim() != 2:
        raise ValueError(f"`{name}` must be 2-D, got shape {tuple(t.shape)}")
    if t.stride(1) != 1 or t.stride(0) < t.shape[1]:
        t = t.contiguous()
    return t


def _ws(nbytes, device):
    return torch.empty(int(nbytes), dtype=torch.uint8, device=device)


# --------------------------------------------------------------------------------------
# a1: gather
# --------------------------------------------------------------------------------------
def gather_rows(X: torch.Tensor, idx: torch.Tensor | None, out: torch.Tensor | None = None) -> torch.Tensor:
    """``X[idx].astype(float32)`` on the device (reference src/scvi/data/_anntorchdataset.py:143-147).
    ``X`` is float32, or a compact uint16 / uint8 count matrix (expanded exactly); ``idx=None`` with a
    compact ``X`` expands all its rows in order (the staged-minibatch case)."""
    if idx is None:
        if X.dtype == torch.float32:
            raise TypeError("gather_rows: idx=None (expand in place order) is for compact integer matrices")
        B = X.shape[0]
    else:
        idx = _check(idx, "idx", torch.int64).contiguous()
        B = idx.numel()
    G = X.shape[1]
    if out is None:
        out = torch.empty(B, G, dtype=torch.float32, device=X.device)
    if B == 0:
        return out
    ip = _ptr(idx) if idx is not None else None
    if X.dtype == torch.float32:
        X = _rowmajor(X, "X")
        call("scvi_gather_rows", _ptr(X), X.shape[0], G, X.stride(0), ip, B, _ptr(out), out.stride(0), _stream())
    elif X.dtype in (torch.uint16, torch.uint8):
        X = _rowmajor(X, "X", X.dtype)
        fn = "scvi_gather_rows_u16" if X.dtype == torch.uint16 else "scvi_gather_rows_u8"
        call(fn, _ptr(X), X.shape[0], G, X.stride(0), ip, B, _ptr(out), out.stride(0), _stream())
    else:
        raise TypeError(f"count matrix must be float32, uint16 or uint8, got {X.dtype}")
    return out


def gather_rows_csr(indptr, indices, values, n_vars: int, idx: torch.Tensor | None, out: torch.Tensor | None = None,
                    n_bad: torch.Tensor | None = None) -> torch.Tensor:
    """Dense fp32 rows ``idx`` (None = all rows in order) of a CSR matrix resident on the device
    (reference ``data[indexes].toarray().astype(float32)``, src/scvi/data/_anntorchdataset.py:143-152)."""
    indptr = _check(indptr, "indptr", torch.int64).contiguous()
    n_rows = indptr.numel() - 1
    if indices.dtype not in (torch.uint16, torch.int32) or values.dtype not in (torch.float32, torch.uint16, torch.uint8):
        raise TypeError("CSR indices must be uint16 / int32 and values float32 / uint16 / uint8")
    _check(indices, "indices", indices.dtype)
    _check(values, "values", values.dtype)
    if idx is not None:
        idx = _check(idx, "idx", torch.int64).contiguous()
    B = n_rows if idx is None else idx.numel()
    if out is None:
        out = torch.empty(B, n_vars, dtype=torch.float32, device=indptr.device)
    if B == 0:
        return out
    kind = {torch.float32: 0, torch.uint16: 1, torch.uint8: 2}[values.dtype]
    call("scvi_gather_rows_csr", _ptr(indptr), _ptr(indices.contiguous()), indices.element_size(),
         _ptr(values.contiguous()), kind, n_rows, n_vars, _ptr(idx), B, _ptr(out), out.stride(0), _ptr(n_bad), _stream())
    return out


def gather_i64(src: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    src = _check(src, "src", torch.int64).contiguous().view(-1)
    idx = _check(idx, "idx", torch.int64).contiguous()
    out = torch.empty(idx.numel(), dtype=torch.int64, device=src.device)
    if idx.numel() == 0:
        return out
    call("scvi_gather_i64", _ptr(src), src.numel(), _ptr(idx), idx.numel(), _ptr(out), _stream())
    return out


# --------------------------------------------------------------------------------------
# a7 / a8: standalone likelihoods
# --------------------------------------------------------------------------------------
class _LogProb(torch.autograd.Function):
    """log_nb_positive / log_zinb_positive (reference _negative_binomial.py:36-135)."""

    @staticmethod
    def forward(ctx, x, mu, theta, pi, eps, bad_flag):
        zinb = pi is not None
        shape = torch.broadcast_shapes(x.shape, mu.shape, theta.shape, *([pi.shape] if zinb else []))
        G = shape[-1]
        per_gene = theta.numel() == G and (theta.dim() == 1 or all(s == 1 for s in theta.shape[:-1]))
        xe = _check(x, "x").expand(shape).reshape(-1, G).contiguous()
        mue = _check(mu, "mu").expand(shape).reshape(-1, G).contiguous()
        the = _check(theta, "theta").reshape(-1).contiguous() if per_gene else theta.expand(shape).reshape(-1, G).contiguous()
        pie = _check(pi, "pi").expand(shape).reshape(-1, G).contiguous() if zinb else None
        B = xe.shape[0]
        out = torch.empty(B, G, dtype=torch.float32, device=x.device)
        tld = 0 if per_gene else G
        if zinb:
            call("scvi_zinb_logprob_fwd", _ptr(xe), G, _ptr(mue), G, _ptr(the), tld, _ptr(pie), G, B, G, eps,
                 _ptr(out), G, _ptr(bad_flag), _stream())
        else:
            call("scvi_nb_logprob_fwd", _ptr(xe), G, _ptr(mue), G, _ptr(the), tld, B, G, eps, _ptr(out), G,
                 _ptr(bad_flag), _stream())
        ctx.save_for_backward(xe, mue, the, pie)
        ctx.meta = (shape, mu.shape, theta.shape, pi.shape if zinb else None, per_gene, eps, zinb)
        return out.view(shape)

    @staticmethod
    def backward(ctx, gout):
        xe, mue, the, pie = ctx.saved_tensors
        shape, mu_shape, theta_shape, pi_shape, per_gene, eps, zinb = ctx.meta
        B, G = xe.shape
        g = gout.expand(shape).reshape(B, G).contiguous().float()
        need_mu, need_th, need_pi = ctx.needs_input_grad[1], ctx.needs_input_grad[2], zinb and ctx.needs_input_grad[3]
        dmu = torch.empty_like(mue) if need_mu else None
        dpi = torch.empty_like(mue) if need_pi else None
        T = (B + _lib.BN_TILE - 1) // _lib.BN_TILE
        dth = None
        if need_th:
            dth = torch.empty((T, G) if per_gene else (B, G), dtype=torch.float32, device=xe.device)
        tld = 0 if per_gene else G
        if zinb:
            call("scvi_zinb_logprob_bwd", _ptr(xe), G, _ptr(mue), G, _ptr(the), tld, _ptr(pie), G, _ptr(g), G, B, G,
                 eps, _ptr(dmu), _ptr(dth), _ptr(dpi), _stream())
        else:
            call("scvi_nb_logprob_bwd", _ptr(xe), G, _ptr(mue), G, _ptr(the), tld, _ptr(g), G, B, G, eps, _ptr(dmu),
                 _ptr(dth), _stream())
        if need_th and per_gene:
            red = torch.empty(G, dtype=torch.float32, device=xe.device)
            call("scvi_colsum_partials", _ptr(dth), T, G, _ptr(red), 0, _stream())
            dth = red.view(theta_shape)
        elif need_th:
            dth = dth.view(shape).sum_to_size(theta_shape)
        if need_mu:
            dmu = dmu.view(shape).sum_to_size(mu_shape)
        if need_pi:
            dpi = dpi.view(shape).sum_to_size(pi_shape)
        return None, dmu, dth, dpi, None, None


def nb_log_prob(x, mu, theta, eps=1e-8, bad_flag=None):
    return _LogProb.apply(x, mu, theta, None, eps, bad_flag)


def zinb_log_prob(x, mu, theta, pi, eps=1e-8, bad_flag=None):
    return _LogProb.apply(x, mu, theta, pi, eps, bad_flag)


class _MixtureLogProb(torch.autograd.Function):
    """log_mixture_nb with one shared inverse dispersion (reference _negative_binomial.py:138-216)."""

    @staticmethod
    def forward(ctx, x, mu1, mu2, theta, pi, eps, bad_flag):
        shape = torch.broadcast_shapes(x.shape, mu1.shape, mu2.shape, theta.shape, pi.shape)
        Pn = shape[-1]
        per_col = theta.numel() == Pn and (theta.dim() == 1 or all(s == 1 for s in theta.shape[:-1]))
        ex = lambda t, n: _check(t, n).expand(shape).reshape(-1, Pn).contiguous()  # noqa: E731
        xe, m1, m2, pie = ex(x, "x"), ex(mu1, "mu1"), ex(mu2, "mu2"), ex(pi, "pi_logits")
        the = _check(theta, "theta").reshape(-1).contiguous() if per_col else ex(theta, "theta")
        B = xe.shape[0]
        out = torch.empty(B, Pn, dtype=torch.float32, device=x.device)
        if B > 0 and Pn > 0:
            call("scvi_mixture_nb_logprob_fwd", _ptr(xe), Pn, _ptr(m1), Pn, _ptr(m2), Pn, _ptr(the),
                 0 if per_col else Pn, _ptr(pie), Pn, B, Pn, eps, _ptr(out), Pn, _ptr(bad_flag), _stream())
        ctx.save_for_backward(xe, m1, m2, the, pie)
        ctx.meta = (shape, mu1.shape, mu2.shape, theta.shape, pi.shape, per_col, eps)
        return out.view(shape)

    @staticmethod
    def backward(ctx, gout):
        xe, m1, m2, the, pie = ctx.saved_tensors
        shape, s1, s2, sth, spi, per_col, eps = ctx.meta
        B, Pn = xe.shape
        g = gout.expand(shape).reshape(B, Pn).contiguous().float()
        n1, n2, nth, npi = ctx.needs_input_grad[1:5]
        d1 = torch.empty_like(m1) if n1 else None
        d2 = torch.empty_like(m1) if n2 else None
        dpi = torch.empty_like(m1) if npi else None
        T = (B + _lib.BN_TILE - 1) // _lib.BN_TILE
        dth = torch.empty((T, Pn) if per_col else (B, Pn), dtype=torch.float32, device=xe.device) if nth else None
        if B > 0 and Pn > 0:
            call("scvi_mixture_nb_logprob_bwd", _ptr(xe), Pn, _ptr(m1), Pn, _ptr(m2), Pn, _ptr(the),
                 0 if per_col else Pn, _ptr(pie), Pn, _ptr(g), Pn, B, Pn, eps, _ptr(d1), _ptr(d2), _ptr(dth),
                 _ptr(dpi), _stream())
        elif nth:
            dth.zero_()
        if nth and per_col and T > 0:
            red = torch.empty(Pn, dtype=torch.float32, device=xe.device)
            call("scvi_colsum_partials", _ptr(dth), T, Pn, _ptr(red), 0, _stream())
            dth = red.view(sth)
        elif nth and per_col:
            dth = torch.zeros(sth, dtype=torch.float32, device=xe.device)
        elif nth:
            dth = dth.view(shape).sum_to_size(sth)
        if n1:
            d1 = d1.view(shape).sum_to_size(s1)
        if n2:
            d2 = d2.view(shape).sum_to_size(s2)
        if npi:
            dpi = dpi.view(shape).sum_to_size(spi)
        return None, d1, d2, dth, dpi, None, None


def mixture_nb_log_prob(x, mu1, mu2, theta, pi_logits, eps=1e-8, bad_flag=None):
    return _MixtureLogProb.apply(x, mu1, mu2, theta, pi_logits, eps, bad_flag)


# --------------------------------------------------------------------------------------
# a2 / a3: FCLayers linear layer with covariate injection
# --------------------------------------------------------------------------------------
class FusedLinear(torch.autograd.Function):
    """cat(x, one_hot(cat)...) + nn.Linear (reference _base_components.py:148-188,198-215);
    optional log1p / library pre-ops (reference _vae.py:370-374).  Returns
    ``(y, bn_partials, library)``; the last two carry no gradient."""

    @staticmethod
    def forward(ctx, x, W, bias, cov_cols, n_in, flags, row_idx, want_bn):
        x = _rowmajor(x, "x")
        W = _rowmajor(W, "W")
        B = x.shape[0] if row_idx is None else row_idx.numel()
        N = W.shape[0]
        n_cov = 0 if cov_cols is None else cov_cols.shape[1]
        if cov_cols is not None:
            cov_cols = _check(cov_cols, "cov_cols", torch.int32).contiguous()
        if row_idx is not None:
            row_idx = _check(row_idx, "row_idx", torch.int64).contiguous()
        dev = x.device
        y = torch.empty(B, N, dtype=torch.float32, device=dev)
        T = (B + _lib.BN_TILE - 1) // _lib.BN_TILE
        part = torch.empty(T, 2, N, dtype=torch.float32, device=dev) if want_bn else None
        lib = torch.empty(B, dtype=torch.float32, device=dev) if (flags & _lib.LIN_LIBRARY) else None
        nws = query("scvi_linear_fwd_ws_bytes", B, n_in, N, flags)
        ws = _ws(nws, dev) if nws > 0 else None
        call("scvi_linear_fwd", _ptr(x), x.stride(0), _ptr(row_idx), B, n_in, N, _ptr(W), W.stride(0), _ptr(bias),
             _ptr(cov_cols), n_cov, flags, _ptr(y), N, _ptr(lib), _ptr(part), _ptr(ws), _stream())
        ctx.save_for_backward(x, W, cov_cols, row_idx)
        ctx.meta = (n_in, flags, bias is not None)
        outs = (y, part if want_bn else y.new_empty(0), lib.view(B, 1) if lib is not None else y.new_empty(0))
        ctx.mark_non_differentiable(outs[1], outs[2])
        return outs

    @staticmethod
    def backward(ctx, dy, _gpart, _glib):
        x, W, cov_cols, row_idx = ctx.saved_tensors
        n_in, flags, has_bias = ctx.meta
        dy = _rowmajor(dy.float(), "dy")
        B, N = dy.shape
        ncols = W.shape[1]
        n_cov = 0 if cov_cols is None else cov_cols.shape[1]
        need_x = ctx.needs_input_grad[0]
        if need_x and (flags & _lib.LIN_LOG1P):
            raise RuntimeError("gradient w.r.t. the count matrix is not defined on this path")
        dW = torch.empty_like(W, memory_format=torch.contiguous_format) if ctx.needs_input_grad[1] else None
        db = torch.empty(N, dtype=torch.float32, device=dy.device) if (has_bias and ctx.needs_input_grad[2]) else None
        dx = torch.empty(B, n_in, dtype=torch.float32, device=dy.device) if need_x else None
        ws = _ws(query("scvi_linear_bwd_ws_bytes", B, n_in, N, ncols), dy.device)
        call("scvi_linear_bwd", _ptr(x), x.stride(0), _ptr(row_idx), B, n_in, N, _ptr(W), W.stride(0), ncols,
             _ptr(cov_cols), n_cov, flags, _ptr(dy), dy.stride(0), _ptr(dW), _ptr(db), _ptr(dx), n_in, None, 0,
             _ptr(ws), _stream())
        return dx, dW, db, None, None, None, None, None


# --------------------------------------------------------------------------------------
# a3: BatchNorm + ReLU + Dropout
# --------------------------------------------------------------------------------------
class BNActDrop(torch.autograd.Function):
    """nn.BatchNorm1d(momentum=.01, eps=1e-3) -> ReLU -> Dropout of FCLayers._build_layer
    (reference _base_components.py:102-115).  `bn_partials` comes from FusedLinear (or is
    computed here).  Updates the running buffers in place when training."""

    @staticmethod
    def forward(ctx, y, part, gamma, beta, run_mean, run_var, mask, drop_scale, relu, training, use_bn):
        y = _rowmajor(y, "y")
        B, N = y.shape
        dev = y.device
        mean = rstd = None
        if use_bn:
            if training and B < 2:
                raise ValueError("Expected more than 1 value per channel when training")  # torch's message
            if training and (part is None or part.numel() == 0):
                T = (B + _lib.BN_TILE - 1) // _lib.BN_TILE
                part = torch.empty(T, 2, N, dtype=torch.float32, device=dev)
                call("scvi_bn_partials", _ptr(y), y.stride(0), B, N, _ptr(part), _stream())
            mean = torch.empty(N, dtype=torch.float32, device=dev)
            rstd = torch.empty(N, dtype=torch.float32, device=dev)
            call("scvi_bn_finalize", _ptr(part) if training else None, B, N, int(training), BN_EPS, BN_MOMENTUM,
                 _ptr(run_mean), _ptr(run_var), _ptr(mean), _ptr(rstd), _stream())
        if mask is not None:
            mask = _check(mask, "drop_mask", torch.uint8).contiguous()
        a = torch.empty(B, N, dtype=torch.float32, device=dev)
        call("scvi_bn_act_drop_fwd", _ptr(y), y.stride(0), B, N, _ptr(mean), _ptr(rstd), _ptr(gamma) if use_bn else None,
             _ptr(beta) if use_bn else None, int(relu), _ptr(mask), float(drop_scale), _ptr(a), N, _stream())
        ctx.save_for_backward(y, mean, rstd, gamma if use_bn else None, beta if use_bn else None, mask)
        ctx.meta = (float(drop_scale), bool(relu), bool(training), bool(use_bn))
        return a

    @staticmethod
    def backward(ctx, da):
        y, mean, rstd, gamma, beta, mask = ctx.saved_tensors
        drop_scale, relu, training, use_bn = ctx.meta
        da = _rowmajor(da.float(), "da")
        B, N = y.shape
        dev = y.device
        dy = torch.empty(B, N, dtype=torch.float32, device=dev)
        want_g = use_bn and gamma is not None and ctx.needs_input_grad[2]
        want_b = use_bn and beta is not None and ctx.needs_input_grad[3]
        dgamma = torch.empty(N, dtype=torch.float32, device=dev) if want_g else None
        dbeta = torch.empty(N, dtype=torch.float32, device=dev) if want_b else None
        ws = _ws(query("scvi_bn_bwd_ws_bytes", B, N), dev)
        call("scvi_bn_act_drop_bwd", _ptr(da), da.stride(0), _ptr(y), y.stride(0), B, N, _ptr(mean), _ptr(rstd),
             _ptr(gamma), _ptr(beta), int(relu), _ptr(mask), drop_scale, int(training), _ptr(dy), N, _ptr(dgamma),
             _ptr(dbeta), _ptr(ws), _stream())
        return dy, None, dgamma, dbeta, None, None, None, None, None, None, None


# --------------------------------------------------------------------------------------
# a4 / a9: latent heads + reparameterisation + KL
# --------------------------------------------------------------------------------------
class LatentHeads(torch.autograd.Function):
    """mean/var heads, `exp + var_eps`, rsample and KL(q || N(0,1)) (reference
    _base_components.py:311-317, _vae.py:563-565).  Returns (q_m, q_v, z, kl)."""

    @staticmethod
    def forward(ctx, a, Wm, bm, Wv, bv, noise, var_eps):
        a = _rowmajor(a, "a")
        B, H = a.shape
        L = Wm.shape[0]
        Wm, Wv = Wm.contiguous(), Wv.contiguous()
        noise = _check(noise, "noise").contiguous()
        dev = a.device
        q_m = torch.empty(B, L, dtype=torch.float32, device=dev)
        q_v = torch.empty_like(q_m)
        z = torch.empty_like(q_m)
        kl = torch.empty(B, dtype=torch.float32, device=dev)
        call("scvi_latent_fwd", _ptr(a), a.stride(0), B, H, L, _ptr(Wm), _ptr(bm), _ptr(Wv), _ptr(bv), _ptr(noise),
             float(var_eps), _ptr(q_m), _ptr(q_v), _ptr(z), _ptr(kl), _stream())
        ctx.save_for_backward(a, Wm, Wv, noise, q_m, q_v)
        ctx.var_eps = float(var_eps)
        return q_m, q_v, z, kl

    @staticmethod
    def backward(ctx, g_qm, g_qv, g_z, g_kl):
        a, Wm, Wv, noise, q_m, q_v = ctx.saved_tensors
        B, H = a.shape
        L = Wm.shape[0]
        dev = a.device

        def prep(g):
            return None if g is None else g.contiguous().float()

        g_qm, g_qv, g_z, g_kl = prep(g_qm), prep(g_qv), prep(g_z), prep(g_kl)
        da = torch.empty(B, H, dtype=torch.float32, device=dev)
        dWm, dWv = torch.empty_like(Wm), torch.empty_like(Wv)
        dbm = torch.empty(L, dtype=torch.float32, device=dev)
        dbv = torch.empty(L, dtype=torch.float32, device=dev)
        ws = _ws(query("scvi_latent_bwd_ws_bytes", B, H, L), dev)
        call("scvi_latent_bwd", _ptr(a), a.stride(0), B, H, L, _ptr(Wm), _ptr(Wv), _ptr(noise), ctx.var_eps, _ptr(q_m),
             _ptr(q_v), _ptr(g_qm), _ptr(g_qv), _ptr(g_z), _ptr(g_kl), _ptr(da), H, _ptr(dWm), _ptr(dbm), _ptr(dWv),
             _ptr(dbv), _ptr(ws), _stream())
        return da, dWm, dbm, dWv, dbv, None, None


class LNActDrop(torch.autograd.Function):
    """nn.LayerNorm(n_out, elementwise_affine=False) -> ReLU -> Dropout of FCLayers._build_layer
    (reference _base_components.py:102-115, `use_layer_norm`)."""

    @staticmethod
    def forward(ctx, y, mask, drop_scale, relu, eps):
        y = _rowmajor(y, "y")
        B, N = y.shape
        a = torch.empty(B, N, dtype=torch.float32, device=y.device)
        stats = torch.empty(B, 2, dtype=torch.float32, device=y.device)
        if mask is not None:
            mask = _check(mask, "drop_mask", torch.uint8).contiguous()
        call("scvi_ln_act_drop_fwd", _ptr(y), y.stride(0), B, N, float(eps), int(relu), _ptr(mask), float(drop_scale),
             _ptr(a), N, _ptr(stats), _stream())
        ctx.save_for_backward(y, stats, mask)
        ctx.meta = (float(drop_scale), bool(relu))
        return a

    @staticmethod
    def backward(ctx, da):
        y, stats, mask = ctx.saved_tensors
        drop_scale, relu = ctx.meta
        da = _rowmajor(da.float(), "da")
        B, N = y.shape
        dy = torch.empty(B, N, dtype=torch.float32, device=y.device)
        call("scvi_ln_act_drop_bwd", _ptr(da), da.stride(0), _ptr(y), y.stride(0), B, N, _ptr(stats), int(relu),
             _ptr(mask), drop_scale, _ptr(dy), N, _stream())
        return dy, None, None, None, None


class EmbeddingRows(torch.autograd.Function):
    """`scvi.nn.Embedding.forward` (reference _embedding.py:29, nn.Embedding): rows `idx` of the table; the
    gradient is a per-row segmented sum in fixed order."""

    @staticmethod
    def forward(ctx, weight, idx):
        weight = _rowmajor(weight, "weight")
        idx = _check(idx, "idx", torch.int64).reshape(-1).contiguous()
        ctx.save_for_backward(idx)
        ctx.shape = tuple(weight.shape)
        return gather_rows(weight, idx)

    @staticmethod
    def backward(ctx, g):
        (idx,) = ctx.saved_tensors
        n_emb, D = ctx.shape
        g = _rowmajor(g.float(), "g")
        dW = torch.empty(n_emb, D, dtype=torch.float32, device=g.device)
        call("scvi_embedding_bwd", _ptr(g), g.stride(0), _ptr(idx), idx.numel(), D, n_emb, _ptr(dW), _stream())
        return dW, None


# --------------------------------------------------------------------------------------
# a5-a9: fused decoder heads + softmax + NB/ZINB NLL
# --------------------------------------------------------------------------------------
class DecoderNLL(torch.autograd.Function):
    """px_scale/px_dropout heads, softmax, exp(library) scaling, exp(px_r), NB/ZINB log_prob and
    the -sum over genes (reference _base_components.py:436-441, _vae.py:516-528,571,
    _negative_binomial.py:36-135) without materialising any [B,G] tensor.  Returns reconst [B]."""

    @staticmethod
    def forward(ctx, h, Ws, bs, Wp, bp, px_r, x, library, likelihood, row_idx, precision):
        h = _rowmajor(h, "h")
        x = _rowmajor(x, "x")
        Ws = _check(Ws, "Ws").contiguous()
        zinb = likelihood == _lib.LIK_ZINB
        if zinb:
            Wp = _check(Wp, "Wp").contiguous()
        if row_idx is not None:
            row_idx = _check(row_idx, "row_idx", torch.int64).contiguous()
        B, H = h.shape
        G = Ws.shape[0]
        ctx.lib_shape = library.shape
        library = _check(library, "library").reshape(-1).contiguous()
        dev = h.device
        reconst = torch.empty(B, dtype=torch.float32, device=dev)
        stats = torch.empty(B, 4, dtype=torch.float32, device=dev)
        # training fusion: d(reconst)/d(h) comes out of the same pass as the likelihood
        dh_unit = torch.empty(B, H, dtype=torch.float32, device=dev) if (ctx.needs_input_grad[0] and B > 0) else None
        nws = query("scvi_decoder_nll_fwd_ws_bytes", B, H, G, likelihood, precision)
        ws = _ws(nws, dev) if nws > 0 else None
        call("scvi_decoder_nll_fwd", _ptr(h), h.stride(0), B, H, G, _ptr(Ws), _ptr(bs), _ptr(Wp) if zinb else None,
             _ptr(bp) if zinb else None, _ptr(px_r), _ptr(x), x.stride(0), _ptr(row_idx), _ptr(library), likelihood,
             precision, _ptr(reconst), _ptr(stats), None, _ptr(dh_unit), H, _ptr(ws), _stream())
        ctx.save_for_backward(h, Ws, bs, Wp if zinb else None, bp if zinb else None, px_r, x, library, stats, row_idx,
                              dh_unit)
        ctx.meta = (likelihood, precision)
        return reconst

    @staticmethod
    def backward(ctx, g):
        h, Ws, bs, Wp, bp, px_r, x, library, stats, row_idx, dh_unit = ctx.saved_tensors
        likelihood, precision = ctx.meta
        zinb = likelihood == _lib.LIK_ZINB
        B, H = h.shape
        G = Ws.shape[0]
        dev = h.device
        g = g.contiguous().float()
        dh = None
        if ctx.needs_input_grad[0]:
            dh = dh_unit * g.unsqueeze(1) if dh_unit is not None else torch.empty(B, H, dtype=torch.float32, device=dev)
        dh_k = None if dh_unit is not None else dh  # already produced by the forward
        dWs = torch.empty_like(Ws)
        dbs = torch.empty(G, dtype=torch.float32, device=dev)
        dWp = torch.empty_like(Wp) if zinb else None
        dbp = torch.empty(G, dtype=torch.float32, device=dev) if zinb else None
        dr = torch.empty(G, dtype=torch.float32, device=dev)
        ws = _ws(query("scvi_decoder_nll_bwd_ws_bytes", B, H, G, likelihood), dev)
        call("scvi_decoder_nll_bwd", _ptr(h), h.stride(0), B, H, G, _ptr(Ws), _ptr(bs), _ptr(Wp), _ptr(bp), _ptr(px_r),
             _ptr(x), x.stride(0), _ptr(row_idx), _ptr(library), _ptr(stats), _ptr(g), likelihood, precision, _ptr(dh_k),
             H, _ptr(dWs), _ptr(dbs), _ptr(dWp), _ptr(dbp), _ptr(dr), _ptr(ws), None, _stream())
        # encoded library size (use_observed_lib_size=False): mu = exp(library) * px_scale, so
        # d(reconst)/d(library) = sum_g mu * d(reconst)/d(mu) -- the forward's fourth row statistic
        dlib = (g * stats[:, 3]).reshape(ctx.lib_shape) if ctx.needs_input_grad[7] else None
        return dh, dWs, dbs, dWp, dbp, dr, None, dlib, None, None, None


# --------------------------------------------------------------------------------------
# f1 / f2: per-row reductions of the semi-supervised / multi-modal losses
# --------------------------------------------------------------------------------------
class _NormalLogProbSum(torch.autograd.Function):
    """Normal(loc, scale).log_prob(value).sum(-1) (`second` = scale, or the variance when is_var): one forward and
    one backward launch where torch.distributions runs ~10 + ~20 element-wise kernels."""

    @staticmethod
    def forward(ctx, value, loc, second, is_var):
        D = value.shape[-1]
        lead = value.shape[:-1]
        v = _check(value, "value").reshape(-1, D).contiguous()
        m = _check(loc, "loc").reshape(-1, D).contiguous()
        s2 = _check(second, "scale").reshape(-1, D).contiguous()
        R = v.shape[0]
        out = torch.empty(R, dtype=torch.float32, device=v.device)
        call("scvi_normal_logprob_sum_fwd", _ptr(v), _ptr(m), _ptr(s2), R, D, int(is_var), _ptr(out), _stream())
        ctx.save_for_backward(v, m, s2)
        ctx.meta = (tuple(value.shape), bool(is_var))
        return out.view(lead)

    @staticmethod
    def backward(ctx, g):
        v, m, s2 = ctx.saved_tensors
        shape, is_var = ctx.meta
        R, D = v.shape
        g = g.reshape(-1).contiguous().float()
        nv, nm, ns = ctx.needs_input_grad[:3]
        dv = torch.empty_like(v) if nv else None
        dm = torch.empty_like(v) if nm else None
        ds = torch.empty_like(v) if ns else None
        if nv or nm or ns:
            call("scvi_normal_logprob_sum_bwd", _ptr(v), _ptr(m), _ptr(s2), _ptr(g), R, D, int(is_var), _ptr(dv), _ptr(dm),
                 _ptr(ds), _stream())
        rs = lambda t: None if t is None else t.view(shape)  # noqa: E731
        return rs(dv), rs(dm), rs(ds), None


def normal_log_prob_sum(value, loc, scale=None, var=None):
    """sum over the last axis of Normal(loc, scale or sqrt(var)).log_prob(value) (arguments broadcast together)."""
    second, is_var = (var, True) if var is not None else (scale, False)
    if not (value.shape == loc.shape == second.shape):
        value, loc, second = torch.broadcast_tensors(value, loc, second)
    return _NormalLogProbSum.apply(value, loc, second, is_var)


class _CategoricalKL(torch.autograd.Function):
    """kl_divergence(Categorical(probs=p), Categorical(probs=q)) over the last axis, q without gradient; the
    gradient is produced by a second launch of the same kernel."""

    @staticmethod
    def forward(ctx, p, q):
        C_ = p.shape[-1]
        pe = _check(p, "probs").reshape(-1, C_).contiguous()
        q = _check(q, "prior")
        shared = q.numel() == C_
        qe = q.reshape(-1).contiguous() if shared else q.expand(p.shape).reshape(-1, C_).contiguous()
        R = pe.shape[0]
        kl = torch.empty(R, dtype=torch.float32, device=pe.device)
        call("scvi_categorical_kl", _ptr(pe), _ptr(qe), 0 if shared else C_, R, C_, None, _ptr(kl), None, _stream())
        ctx.save_for_backward(pe, qe)
        ctx.meta = (tuple(p.shape), shared)
        return kl.view(p.shape[:-1])

    @staticmethod
    def backward(ctx, g):
        pe, qe = ctx.saved_tensors
        shape, shared = ctx.meta
        R, C_ = pe.shape
        if not ctx.needs_input_grad[0]:
            return None, None
        g = g.reshape(-1).contiguous().float()
        dp = torch.empty_like(pe)
        call("scvi_categorical_kl", _ptr(pe), _ptr(qe), 0 if shared else C_, R, C_, _ptr(g), None, _ptr(dp), _stream())
        return dp.view(shape), None


def categorical_kl(probs, prior_probs):
    """KL(Categorical(probs) || Categorical(prior_probs)) per row, torch's construction (see scvi_categorical_kl);
    `prior_probs`: one row of class probabilities or the shape of `probs`; it receives no gradient."""
    return _CategoricalKL.apply(probs, prior_probs.detach())


def decoder_params(h, Ws, bs, Wp, bp, library, want_scale=True, want_rate=True, want_dropout=True):
    """Materialise px_scale / px_rate / px_dropout for inference-time getters (no autograd)."""
    h = _rowmajor(h, "h")
    B, H = h.shape
    G = Ws.shape[0]
    dev = h.device
    mk = lambda: torch.empty(B, G, dtype=torch.float32, device=dev)  # noqa: E731
    scale = mk() if want_scale else None
    rate = mk() if want_rate else None
    drop = mk() if (want_dropout and Wp is not None) else None
    lib = library.reshape(-1).contiguous() if library is not None else None
    call("scvi_decoder_params_fwd", _ptr(h), h.stride(0), B, H, G, _ptr(Ws.contiguous()), _ptr(bs),
         _ptr(Wp.contiguous()) if Wp is not None else None, _ptr(bp), _ptr(lib), _ptr(scale), _ptr(rate), _ptr(drop),
         G, _stream())
    return scale, rate, drop
