"""scvi_tools_b200 -- B200-native (sm_100a) implementation of the SCVI `VAE` minibatch training hot path.

Only what that path needs lives here: the CUDA kernels + C ABI (`csrc/`, `include/scvi_b200.h`)
and the host-side mirror of the reference's nn / distributions / module surface.
Importing the package does not need a GPU; running any op does (there is no CPU fallback).
"""

from . import _lib
from .distributions import (NegativeBinomial, NegativeBinomialMixture, ZeroInflatedNegativeBinomial, log_mixture_nb,
                            log_nb_positive, log_zinb_positive)
from .module import MODULE_KEYS, REGISTRY_KEYS, VAE, LossOutput
from .nn import DecoderSCVI, Encoder, FCLayers, rng_feed
from .scanvi import SCANVAE, Classifier, Decoder
from .totalvi import TOTALVAE, DecoderTOTALVI, EncoderTOTALVI
from .model import SCVI  # noqa: E402  (after the module classes it builds on)

__all__ = [
    "SCVI", "VAE", "SCANVAE", "Classifier", "Decoder", "TOTALVAE", "EncoderTOTALVI", "DecoderTOTALVI", "LossOutput", "MODULE_KEYS", "REGISTRY_KEYS", "FCLayers", "Encoder", "DecoderSCVI",
    "NegativeBinomial", "ZeroInflatedNegativeBinomial", "NegativeBinomialMixture", "log_mixture_nb", "log_nb_positive", "log_zinb_positive", "rng_feed",
]
__version__ = "0.1.0"
