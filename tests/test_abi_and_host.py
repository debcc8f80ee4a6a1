"""CPU-side checks: the C-ABI library loads and exports every symbol include/scvi_b200.h declares
(with the argument counts the ctypes table uses), and the host modules mirror the reference's
state_dict / error behaviour.  No compute call is made (no GPU here)."""
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_decls():
    src = open(os.path.join(ROOT, "include", "scvi_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    decls = {}
    for m in re.finditer(r"\b(?:int|int64_t|const char\*)\s+(scvi_\w+)\s*\(([^;]*?)\)\s*;", src, flags=re.S):
        args = m.group(2).strip()
        n = 0 if args in ("", "void") else args.count(",") + 1
        decls[m.group(1)] = n
    return decls


def test_library_exports_every_declared_symbol():
    from scvi_tools_b200 import _lib

    decls = _header_decls()
    assert len(decls) >= 25
    lib = _lib.load()
    for name, nargs in decls.items():
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
        assert name in _lib.SIGNATURES, f"{name} missing from the ctypes table"
        assert len(_lib.SIGNATURES[name][1]) == nargs, f"{name}: header has {nargs} args"
    assert set(_lib.SIGNATURES) == set(decls)
    assert lib.scvi_b200_abi_version() == 3


def test_workspace_queries_are_pure_host_functions():
    from scvi_tools_b200 import _lib

    assert _lib.query("scvi_linear_bwd_ws_bytes", 128, 2000, 128, 2000) > 0
    assert _lib.query("scvi_bn_bwd_ws_bytes", 128, 128) > 0
    assert _lib.query("scvi_latent_bwd_ws_bytes", 128, 128, 10) > 0
    assert _lib.query("scvi_decoder_nll_bwd_ws_bytes", 8192, 128, 2000, 1) >= 4 * 2 * 2000 * 128


def test_linear_workspace_covers_the_split_reductions():
    """The FFMA linear layer splits its reduction for few-cell minibatches (csrc/linear.cu): the workspace queries
    must hand out room for the partial tiles / slabs exactly when a split can happen, and the dW slabs must cover the
    covariate and bias columns that ride along in the GEMM."""
    from scvi_tools_b200 import _lib

    q = _lib.query
    # forward: totalVI's encoder input layer, 256 cells x 4100 inputs -> 256 (16 tiles, 129 chunks): split-K partials
    ws = q("scvi_linear_fwd_ws_bytes", 256, 4100, 256, 0)
    assert ws >= 4 * 2 * 256 * 256 and ws % (4 * 256 * 256) == 0
    # a minibatch that fills the SMs with tiles, a short reduction, or the library row-sum flag: no split, no workspace
    assert q("scvi_linear_fwd_ws_bytes", 8192, 4100, 256, 0) == 0
    assert q("scvi_linear_fwd_ws_bytes", 256, 128, 256, 0) == 0
    assert q("scvi_linear_fwd_ws_bytes", 256, 4100, 256, _lib.LIN_LIBRARY) == 0
    # backward: dx slabs on top of the dW slabs when the N reduction is long and the output small
    base = q("scvi_linear_bwd_ws_bytes", 8192, 276, 4000, 276)
    few = q("scvi_linear_bwd_ws_bytes", 256, 276, 4000, 276)
    assert few > 0 and base > 0
    # dW slabs hold K + covariate columns + the ones column of the bias
    assert q("scvi_linear_bwd_ws_bytes", 1024, 10, 128, 30) >= 4 * 128 * 31
    # monotone in the number of weight columns
    assert q("scvi_linear_bwd_ws_bytes", 1024, 10, 128, 30) >= q("scvi_linear_bwd_ws_bytes", 1024, 10, 128, 10)


def test_state_dict_layout_matches_reference_golden():
    """Every parameter / buffer recorded from the reference module exists with the same shape."""
    import scvi_tools_b200 as S
    from tests._golden import load_vae_case

    for name in ("c1_zinb", "ragged_zinb", "onebatch_nb"):
        P, inp, _ = load_vae_case(name)
        m = S.VAE(n_input=P["px_r"].shape[0], n_batch=inp["n_batch"], gene_likelihood=inp["likelihood"])
        sd = m.state_dict()
        for k, v in P.items():
            assert k in sd and tuple(sd[k].shape) == tuple(v.shape), k
    m = S.VAE(n_input=100, n_batch=2)
    assert sum(p.numel() for p in m.parameters()) == 69926
    assert len(list(m.parameters())) == 27


def test_cpu_tensors_are_rejected_loudly():
    import scvi_tools_b200 as S

    m = S.VAE(n_input=20, n_batch=2).train()
    t = {"X": torch.ones(4, 20), "batch": torch.zeros(4, 1, dtype=torch.long), "labels": torch.zeros(4, 1, dtype=torch.long)}
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(t)
    with pytest.raises(RuntimeError):
        S.log_nb_positive(torch.ones(2, 3), torch.ones(2, 3), torch.ones(3))


def test_reference_error_behaviour():
    import scvi_tools_b200 as S

    with pytest.raises(ValueError):
        S.VAE(n_input=10, dispersion="bogus")
    with pytest.raises(ValueError):
        S.VAE(n_input=10, use_observed_lib_size=False)  # needs library_log_means / vars
    with pytest.raises(ValueError):
        S.NegativeBinomial(mu=None, total_count=None)
    with pytest.raises(ValueError):
        S.LossOutput(loss=torch.tensor(1.0))
    fc = S.FCLayers(10, 5, n_cat_list=[3, 1])
    assert fc.n_cat_list == [3, 0]  # a single-category covariate is dropped
    assert fc.fc_layers[0][0].in_features == 13
    lo = S.LossOutput(loss=torch.tensor(2.0), reconstruction_loss=torch.ones(3), kl_local={"a": torch.ones(3), "b": torch.ones(3)})
    assert lo.n_obs_minibatch == 3 and float(lo.kl_local_sum) == 6.0 and float(lo.reconstruction_loss_sum) == 3.0


def test_no_product_import_of_oracle():
    """The product package must never import the oracle."""
    pkg = os.path.join(ROOT, "scvi_tools_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(from|import)\s+\.*oracle", src, flags=re.M), fn
