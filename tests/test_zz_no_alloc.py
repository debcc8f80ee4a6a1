"""GPU: the allocation-free entry point `scvi_bn_act_drop_bwd_static` (scratch built into the library image)
against the existing `scvi_bn_act_drop_bwd` (caller-owned workspace), through the C-ABI.

This entry point has no counterpart in the reference; it was added for
callers that cannot allocate device memory.  Collected last on purpose (file name):
a failure here hides no other test.

Two calls in a row on different shapes -- the second one smaller and ragged, so the shared scratch still holds
the first call's partial sums when it starts -- each compared with the existing entry point on the same inputs
(same kernels, fixed-order sums: the tolerance below is the suite's 1e-6, the results are normally identical).
Also: shapes beyond the built-in scratch are rejected with a clear error, and the call itself moves the
device's free-memory figure by nothing.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _inputs(B, N, seed, dev):
    g = torch.Generator(device=dev).manual_seed(seed)
    y = torch.randn(B, N, device=dev, generator=g) * 1.7 + 0.3
    da = torch.randn(B, N, device=dev, generator=g)
    gamma = torch.rand(N, device=dev, generator=g) + 0.5
    beta = torch.randn(N, device=dev, generator=g) * 0.2
    mask = (torch.rand(B, N, device=dev, generator=g) >= 0.1).to(torch.uint8)
    mean = y.mean(0).contiguous()
    rstd = (1.0 / torch.sqrt(y.var(0, unbiased=False) + 1e-3)).contiguous()
    return y, da, gamma, beta, mask, mean, rstd


def _rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


def test_static_scratch_entry_point_matches_workspace_entry_point_twice():
    from scvi_tools_b200._lib import call, load, query

    dev = torch.device("cuda:0")
    lib = load()
    cap = query("scvi_bn_static_scratch_bytes")
    assert cap == 526592
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    # (B, N) of the two consecutive calls: both within the scratch, both on the same kernel variants
    for call_no, (B, N) in enumerate([(4096, 128), (1500, 96)]):
        assert query("scvi_bn_bwd_ws_bytes", B, N) <= cap
        y, da, gamma, beta, mask, mean, rstd = _inputs(B, N, seed=call_no, dev=dev)
        ws = torch.empty(query("scvi_bn_bwd_ws_bytes", B, N), dtype=torch.uint8, device=dev)
        ref = [torch.empty(B, N, device=dev), torch.empty(N, device=dev), torch.empty(N, device=dev)]
        got = [torch.full((B, N), float("nan"), device=dev), torch.full((N,), float("nan"), device=dev),
               torch.full((N,), float("nan"), device=dev)]
        common = (_ptr(da), N, _ptr(y), N, B, N, _ptr(mean), _ptr(rstd), _ptr(gamma), _ptr(beta), 1, _ptr(mask),
                  1.0 / 0.9, 1)
        call("scvi_bn_act_drop_bwd", *common, _ptr(ref[0]), N, _ptr(ref[1]), _ptr(ref[2]), _ptr(ws), st)
        torch.cuda.synchronize(dev)
        free_before = torch.cuda.mem_get_info(dev)[0]
        call("scvi_bn_act_drop_bwd_static", *common, _ptr(got[0]), N, _ptr(got[1]), _ptr(got[2]), st)
        torch.cuda.synchronize(dev)
        free_after = torch.cuda.mem_get_info(dev)[0]
        if call_no == 1:  # every kernel of the path is loaded by now: the call must not have taken any memory
            assert free_after >= free_before, (free_before, free_after)
        for name, a, b in zip(("dy", "dgamma", "dbeta"), got, ref, strict=True):
            assert torch.isfinite(a).all(), (call_no, name)
            assert _rel(a, b) <= 1e-6, (call_no, name, _rel(a, b))
    # beyond the built-in scratch: rejected, nothing launched, message names the limit
    B, N = 32768, 256
    assert query("scvi_bn_bwd_ws_bytes", B, N) > cap
    dummy = torch.zeros(8, device=dev)
    rc = lib.scvi_bn_act_drop_bwd_static(_ptr(dummy), N, _ptr(dummy), N, B, N, None, None, None, None, 1, None, 1.0, 1,
                                         _ptr(dummy), N, None, None, st)
    assert rc != 0
    msg = lib.scvi_last_error_string().decode()
    assert "526592" in msg and "scratch" in msg, msg
    torch.cuda.synchronize(dev)
