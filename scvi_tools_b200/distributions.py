"""`NegativeBinomial` / `ZeroInflatedNegativeBinomial` and the free log-likelihood functions.

Surface of the reference's src/scvi/distributions/_negative_binomial.py:36-135,331-586 with
`log_prob` computed by the standalone B200 likelihood kernels.  The ZINB `log_prob` keeps
the reference's always-on support check (:575-582) but evaluates it inside the kernel and
raises the `UserWarning` lazily (on `.check_support()` or interpreter exit of the call when
`SCVI_B200_SYNC_VALIDATE=1`), so the training step has no host sync.
"""

from __future__ import annotations

import os
import warnings

import torch
from torch.distributions import Distribution, Gamma, constraints
from torch.distributions import Poisson as PoissonTorch
from torch.distributions.utils import broadcast_all, lazy_property, logits_to_probs, probs_to_logits

from . import ops

_SUPPORT_MSG = "The value argument must be within the support of the distribution"


def log_nb_positive(x, mu, theta, eps: float = 1e-8, log_fn=None, lgamma_fn=None):
    """NB log-likelihood per element (reference :99-135).  `log_fn`/`lgamma_fn` are accepted
    for signature compatibility; the kernel uses its own fp32 log / lgamma."""
    return ops.nb_log_prob(x, mu, theta, eps)


def log_zinb_positive(x, mu, theta, pi, eps: float = 1e-8, log_fn=None, lgamma_fn=None):
    """ZINB log-likelihood per element (reference :36-96); `pi` is the zero-inflation logit."""
    return ops.zinb_log_prob(x, mu, theta, pi, eps)


def _convert_counts_logits_to_mean_disp(total_count, logits):
    """(total_count, logits) -> (mu, theta) as the reference's helper (:245-262)."""
    theta = total_count
    mu = logits.exp() * theta
    return mu, theta


class NegativeBinomial(Distribution):
    """Negative binomial in the (mu, theta) or (total_count, probs|logits) parameterisation."""

    arg_constraints = {}
    support = constraints.nonnegative_integer

    def __init__(self, total_count=None, probs=None, logits=None, mu=None, theta=None, scale=None,
                 validate_args: bool = False):
        self._eps = 1e-8
        if (mu is None) == (total_count is None):
            raise ValueError(
                "Please use one of the two possible parameterizations. Refer to the documentation "
                "for more information."
            )
        if total_count is not None and (logits is not None or probs is not None):
            logits = logits if logits is not None else probs_to_logits(probs)
            total_count = total_count.type_as(logits)
            total_count, logits = broadcast_all(total_count, logits)
            mu, theta = _convert_counts_logits_to_mean_disp(total_count, logits)
            scale = mu / torch.sum(mu, dim=-1, keepdim=True)
            self._theta_compact = None
        else:
            self._theta_compact = theta  # [G] stays [G] for the kernel (per-gene constants hoisted)
            mu, theta = broadcast_all(mu, theta)
        self.mu, self.theta, self.scale = mu, theta, scale
        self._bad = None
        super().__init__(batch_shape=mu.shape, validate_args=False)
        self._validate_args = bool(validate_args)

    @property
    def mean(self):
        return self.mu

    @property
    def variance(self):
        return self.mean + (self.mean**2) / self.theta

    def get_normalized(self, key):
        if key == "mu":
            return self.mu
        if key == "scale":
            return self.scale
        raise ValueError(f"normalized key {key} not recognized")

    def _gamma(self):
        return Gamma(concentration=self.theta, rate=self.theta / self.mu)

    @torch.inference_mode()
    def sample(self, sample_shape=None):
        sample_shape = sample_shape or torch.Size()
        rate = torch.clamp(self._gamma().sample(sample_shape), max=1e8)
        return PoissonTorch(rate).sample()

    def _theta_arg(self):
        return self._theta_compact if self._theta_compact is not None else self.theta

    def _flag(self, value):
        self._bad = torch.zeros(1, dtype=torch.int32, device=value.device)
        return self._bad

    def check_support(self):
        """Raise the reference's UserWarning if the last `log_prob` saw out-of-support values."""
        if self._bad is not None and int(self._bad.item()) != 0:
            warnings.warn(_SUPPORT_MSG, UserWarning, stacklevel=2)
            return False
        return True

    def log_prob(self, value):
        flag = self._flag(value) if self._validate_args else None
        out = ops.nb_log_prob(value, self.mu, self._theta_arg(), self._eps, flag)
        if flag is not None and os.environ.get("SCVI_B200_SYNC_VALIDATE", "1") == "1":
            self.check_support()
        return out

    def __repr__(self):
        return f"{self.__class__.__name__}(mu: {tuple(self.mu.shape)}, theta: {tuple(self.theta.shape)})"


class ZeroInflatedNegativeBinomial(NegativeBinomial):
    """Zero-inflated negative binomial; `zi_logits` are the zero-inflation logits."""

    def __init__(self, total_count=None, probs=None, logits=None, mu=None, theta=None, zi_logits=None, scale=None,
                 validate_args: bool = False):
        super().__init__(total_count=total_count, probs=probs, logits=logits, mu=mu, theta=theta, scale=scale,
                         validate_args=validate_args)
        self.zi_logits, self.mu, self.theta = broadcast_all(zi_logits, self.mu, self.theta)

    @property
    def mean(self):
        return (1 - self.zi_probs) * self.mu

    @property
    def variance(self):
        pi = self.zi_probs
        return (1 - pi) * self.mu * (self.mu + self.theta + pi * self.mu * self.theta) / self.theta

    @lazy_property
    def zi_probs(self):
        return logits_to_probs(self.zi_logits, is_binary=True)

    @torch.inference_mode()
    def sample(self, sample_shape=None):
        sample_shape = sample_shape or torch.Size()
        samp = super().sample(sample_shape=sample_shape)
        is_zero = torch.rand_like(samp) <= self.zi_probs
        return torch.where(is_zero, torch.zeros_like(samp), samp)

    def log_prob(self, value):
        # the reference validates here regardless of validate_args (:575-582); the kernel does it,
        # and the warning is raised synchronously only outside training-style use
        flag = self._flag(value)
        out = ops.zinb_log_prob(value, self.mu, self._theta_arg(), self.zi_logits, 1e-8, flag)
        if os.environ.get("SCVI_B200_SYNC_VALIDATE", "1") == "1" and not torch.is_grad_enabled():
            self.check_support()
        elif os.environ.get("SCVI_B200_SYNC_VALIDATE", "1") == "1" and not out.requires_grad:
            self.check_support()
        return out


def log_mixture_nb(x, mu_1, mu_2, theta_1, theta_2, pi_logits, eps: float = 1e-8, log_fn=None, lgamma_fn=None):
    """Log-likelihood of a two-component NB mixture (reference :138-216); `pi_logits` is the logit of
    belonging to the first (background, for totalVI) component.  One shared inverse dispersion
    (`theta_2 is None`, the totalVI case) runs as one fused kernel; two dispersions combine the NB kernel
    outputs."""
    if theta_2 is None:
        return ops.mixture_nb_log_prob(x, mu_1, mu_2, theta_1, pi_logits, eps)
    log_nb_1 = ops.nb_log_prob(x, mu_1, theta_1, eps)
    log_nb_2 = ops.nb_log_prob(x, mu_2, theta_2, eps)
    return torch.logsumexp(torch.stack((log_nb_1, log_nb_2 - pi_logits)), dim=0) - torch.nn.functional.softplus(-pi_logits)


class NegativeBinomialMixture(Distribution):
    """Negative binomial mixture (reference :589-717): components (`mu1`, `theta1`) and (`mu2`, `theta2` or
    `theta1`), `mixture_logits` = logit of component 1."""

    arg_constraints = {}
    support = constraints.nonnegative_integer

    def __init__(self, mu1, mu2, theta1, mixture_logits, theta2=None, validate_args: bool = False):
        self._theta1_compact = theta1
        self.mu1, self.theta1, self.mu2, self.mixture_logits = broadcast_all(mu1, theta1, mu2, mixture_logits)
        super().__init__(batch_shape=self.mu1.shape, validate_args=False)
        self._validate_args = bool(validate_args)
        self.theta2 = broadcast_all(mu1, theta2) if theta2 is not None else None
        self._theta2_arg = theta2
        self._bad = None

    @property
    def mean(self):
        pi = self.mixture_probs
        return pi * self.mu1 + (1 - pi) * self.mu2

    def get_normalized(self, key):
        if key == "mu":
            return self.rate
        if key == "scale":
            return self.scale
        raise ValueError(f"normalized key {key} not recognized")

    @lazy_property
    def mixture_probs(self):
        return logits_to_probs(self.mixture_logits, is_binary=True)

    @torch.inference_mode()
    def sample(self, sample_shape=None):
        sample_shape = sample_shape or torch.Size()
        mixing = torch.distributions.Bernoulli(self.mixture_probs).sample()
        mu = self.mu1 * mixing + self.mu2 * (1 - mixing)
        theta = self.theta1 if self.theta2 is None else self.theta1 * mixing + self.theta2[1] * (1 - mixing)
        rate = torch.clamp(Gamma(concentration=theta, rate=theta / mu).sample(sample_shape), max=1e8)
        return PoissonTorch(rate).sample()

    def check_support(self):
        if self._bad is not None and int(self._bad.item()) != 0:
            warnings.warn(_SUPPORT_MSG, UserWarning, stacklevel=2)
            return False
        return True

    def log_prob(self, value):
        if self._theta2_arg is not None:
            return log_mixture_nb(value, self.mu1, self.mu2, self.theta1, self._theta2_arg, self.mixture_logits)
        self._bad = torch.zeros(1, dtype=torch.int32, device=value.device)
        out = ops.mixture_nb_log_prob(value, self.mu1, self.mu2, self._theta1_compact, self.mixture_logits, 1e-8, self._bad)
        if os.environ.get("SCVI_B200_SYNC_VALIDATE", "1") == "1" and not out.requires_grad:
            self.check_support()  # the reference warns on out-of-support values (:690-697)
        return out

    def __repr__(self):
        return f"{self.__class__.__name__}(mu1: {tuple(self.mu1.shape)}, theta1: {tuple(self.theta1.shape)})"
