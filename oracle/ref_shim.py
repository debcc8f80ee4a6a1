"""TEST INFRASTRUCTURE ONLY -- loader for the *unmodified* reference hot-path files.

The reference (scverse/scvi-tools, mounted read-only at ``/root/reference``) cannot be
imported as a package in the build container (``lightning``, ``anndata``, ``pyro`` ... are
absent).  Its hot-path files, however, depend only on torch once a ~40 line stub ``scvi``
package is in ``sys.modules``.  This module builds that stub and
loads, with ``importlib``, exactly these reference files where they lie (nothing is copied):

    src/scvi/_constants.py                 src/scvi/data/_constants.py
    src/scvi/utils/_decorators.py          src/scvi/distributions/_constraints.py
    src/scvi/distributions/_negative_binomial.py   src/scvi/distributions/_normal.py
    src/scvi/nn/_utils.py  src/scvi/nn/_embedding.py  src/scvi/nn/_base_components.py
    src/scvi/module/_constants.py          src/scvi/module/base/_decorators.py
    src/scvi/module/base/_embedding_mixin.py   src/scvi/module/base/_base_module.py
    src/scvi/module/_vae.py
    src/scvi/module/_utils.py  src/scvi/module/_classifier.py  src/scvi/module/_scanvae.py   (SCANVI, 8f rank 1)
    src/scvi/module/_totalvae.py                                                             (TOTALVI, 8f rank 2)

It only works where ``/root/reference`` exists (the build container).  It is used by
``oracle/make_golden.py`` (to generate ``tests/golden/*.npz``) and by the
``tests/test_oracle_pins.py`` cross-check of ``oracle/scvi_oracle.py``.  Nothing under
``scvi_tools_b200/`` may import it.
"""

from __future__ import annotations

import importlib.util
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("SCVI_REFERENCE_ROOT", "/root/reference")
_SRC = os.path.join(REFERENCE_ROOT, "src", "scvi")

_LOADED = None


def available() -> bool:
    return os.path.isfile(os.path.join(_SRC, "module", "_vae.py"))


def _pkg(name: str) -> types.ModuleType:
    mod = types.ModuleType(name)
    mod.__path__ = []  # mark as package
    sys.modules[name] = mod
    return mod


def _load(name: str, relpath: str) -> types.ModuleType:
    spec = importlib.util.spec_from_file_location(name, os.path.join(_SRC, relpath))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def load():
    """Return a namespace holding the reference's own classes/functions for the hot path."""
    global _LOADED
    if _LOADED is not None:
        return _LOADED
    if not available():
        raise RuntimeError(f"reference sources not found under {REFERENCE_ROOT}")
    if "scvi" in sys.modules and not getattr(sys.modules["scvi"], "_b200_shim", False):
        raise RuntimeError("a real `scvi` package is already imported; shim refuses to shadow it")

    scvi = _pkg("scvi")
    scvi._b200_shim = True
    scvi.settings = types.SimpleNamespace(warnings_stacklevel=2, seed=0, batch_size=128)
    scvi.REGISTRY_KEYS = _load("scvi._constants", "_constants.py").REGISTRY_KEYS

    data = _pkg("scvi.data")
    data._constants = _load("scvi.data._constants", "data/_constants.py")
    scvi.data = data

    utils = _pkg("scvi.utils")
    dec = _load("scvi.utils._decorators", "utils/_decorators.py")
    utils.unsupported_if_adata_minified = dec.unsupported_if_adata_minified
    utils.error_on_missing_dependencies = lambda *a, **k: None
    scvi.utils = utils

    # empty pyro: _base_module.py does `import pyro` and uses it only inside Pyro classes
    pyro = types.ModuleType("pyro")
    sys.modules.setdefault("pyro", pyro)

    dist = _pkg("scvi.distributions")
    _load("scvi.distributions._constraints", "distributions/_constraints.py")
    nbm = _load("scvi.distributions._negative_binomial", "distributions/_negative_binomial.py")
    nrm = _load("scvi.distributions._normal", "distributions/_normal.py")
    dist.NegativeBinomial = nbm.NegativeBinomial
    dist.ZeroInflatedNegativeBinomial = nbm.ZeroInflatedNegativeBinomial
    dist.NegativeBinomialMixture = nbm.NegativeBinomialMixture
    dist.Poisson = nbm.Poisson
    dist.Normal = nrm.Normal
    scvi.distributions = dist

    nn = _pkg("scvi.nn")
    from torch.nn.functional import one_hot

    nn.one_hot = one_hot
    _load("scvi.nn._utils", "nn/_utils.py")
    emb = _load("scvi.nn._embedding", "nn/_embedding.py")
    nn.Embedding = emb.Embedding
    bc = _load("scvi.nn._base_components", "nn/_base_components.py")
    for k in ("FCLayers", "Encoder", "DecoderSCVI", "Decoder", "DecoderTOTALVI", "EncoderTOTALVI"):
        setattr(nn, k, getattr(bc, k))
    scvi.nn = nn

    module = _pkg("scvi.module")
    mconst = _load("scvi.module._constants", "module/_constants.py")
    base = _pkg("scvi.module.base")
    mdec = _load("scvi.module.base._decorators", "module/base/_decorators.py")
    base.auto_move_data = mdec.auto_move_data
    pyro_stub = types.ModuleType("scvi.module.base._pyro")
    pyro_stub.AutoMoveDataPredictive = object
    sys.modules["scvi.module.base._pyro"] = pyro_stub
    bm = _load("scvi.module.base._base_module", "module/base/_base_module.py")
    base.LossOutput = bm.LossOutput
    base.BaseModuleClass = bm.BaseModuleClass
    base.BaseMinifiedModeModuleClass = bm.BaseMinifiedModeModuleClass
    base.SupervisedModuleClass = bm.SupervisedModuleClass
    em = _load("scvi.module.base._embedding_mixin", "module/base/_embedding_mixin.py")
    base.EmbeddingModuleMixin = em.EmbeddingModuleMixin
    module.base = base
    vae = _load("scvi.module._vae", "module/_vae.py")
    module.VAE = vae.VAE
    _load("scvi.module._utils", "module/_utils.py")
    cls = _load("scvi.module._classifier", "module/_classifier.py")
    module.Classifier = cls.Classifier
    scanvae = _load("scvi.module._scanvae", "module/_scanvae.py")
    module.SCANVAE = scanvae.SCANVAE
    # _totalvae.py imports BaseModelClass for a type annotation only
    model_pkg = _pkg("scvi.model")
    model_base = _pkg("scvi.model.base")
    model_base.BaseModelClass = object
    model_pkg.base = model_base
    scvi.model = model_pkg
    totalvae = _load("scvi.module._totalvae", "module/_totalvae.py")
    module.TOTALVAE = totalvae.TOTALVAE
    scvi.module = module

    _LOADED = types.SimpleNamespace(
        VAE=vae.VAE,
        SCANVAE=scanvae.SCANVAE,
        TOTALVAE=totalvae.TOTALVAE,
        Classifier=cls.Classifier,
        Decoder=bc.Decoder,
        FCLayers=bc.FCLayers,
        Encoder=bc.Encoder,
        DecoderSCVI=bc.DecoderSCVI,
        NegativeBinomial=nbm.NegativeBinomial,
        ZeroInflatedNegativeBinomial=nbm.ZeroInflatedNegativeBinomial,
        log_nb_positive=nbm.log_nb_positive,
        log_zinb_positive=nbm.log_zinb_positive,
        log_mixture_nb=nbm.log_mixture_nb,
        NegativeBinomialMixture=nbm.NegativeBinomialMixture,
        LossOutput=bm.LossOutput,
        MODULE_KEYS=mconst.MODULE_KEYS,
        REGISTRY_KEYS=scvi.REGISTRY_KEYS,
    )
    return _LOADED
