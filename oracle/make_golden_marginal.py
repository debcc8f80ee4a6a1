"""TEST INFRASTRUCTURE ONLY -- tests/golden/marginal_*.npz: `VAE.marginal_ll` (importance-sampled log p(x),
src/scvi/module/_vae.py:665-747, behind `SCVI.get_marginal_ll`) from the UNMODIFIED reference module in eval
mode on the inputs and weights of an existing fixture.  Every normal draw (`Normal.rsample` in the encoder,
`Normal.sample((n,))` for the extra importance samples) goes through a recording feed so the CUDA path can be
handed the same noise.

    python oracle/make_golden_marginal.py      (build container only: needs /root/reference)
"""

from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import ref_shim  # noqa: E402
from oracle.make_golden_options import GOLD, RecordingFeed  # noqa: E402


class SampleFeed(RecordingFeed):
    """RecordingFeed that also routes `Normal.sample(shape)` through the recorded generator."""

    def __enter__(self):
        super().__enter__()
        from torch.distributions import Normal

        self._old_sample = Normal.sample
        feed = self

        def fed_sample(dist, sample_shape=torch.Size()):
            shape = torch.Size(sample_shape) + dist.loc.shape
            noise = torch.randn(shape, generator=feed.g)
            feed.noises.append(noise)
            return (dist.loc + noise * dist.scale).detach()

        Normal.sample = fed_sample
        return self

    def __exit__(self, *exc):
        from torch.distributions import Normal

        Normal.sample = self._old_sample
        super().__exit__(*exc)


def marginal_case(src, n_mc, n_per, seed):
    R = ref_shim.load()
    z = np.load(os.path.join(GOLD, src + ".npz"))
    P0 = {k[len("param::"):]: torch.from_numpy(z[k]) for k in z.files if k.startswith("param::")}
    x, bi = torch.from_numpy(z["in::x"]), torch.from_numpy(z["in::batch"])
    n_batch, lik = int(z["in::n_batch"]), str(z["in::likelihood"])
    m = R.VAE(n_input=x.shape[1], n_batch=n_batch, gene_likelihood=lik)
    sd = m.state_dict()
    sd.update(P0)
    m.load_state_dict(sd)
    m.eval()
    tensors = {"X": x, "batch": bi, "labels": torch.zeros_like(bi)}
    with torch.no_grad(), SampleFeed(seed) as feed:
        ll = m.marginal_ll(tensors, n_mc_samples=n_mc, return_mean=False, n_mc_samples_per_pass=n_per)
    save = {"marginal_ll": ll.numpy(), "n_mc": np.int64(n_mc), "n_per": np.int64(n_per), "src": np.array(src)}
    for i, n in enumerate(feed.noises):
        save[f"noise::{i}"] = n.numpy()
    name = f"marginal_{src}_{n_mc}x{n_per}"
    path = os.path.join(GOLD, name + ".npz")
    np.savez_compressed(path, **save)
    print(f"{name}: mean log p(x) = {float(ll.mean()):.6f}, {len(feed.noises)} draws -> {os.path.getsize(path)/1e3:.0f} kB")


if __name__ == "__main__":
    torch.set_num_threads(1)
    marginal_case("c1_zinb", 6, 3, 31)
    marginal_case("c1_zinb", 4, 1, 32)
    marginal_case("onebatch_nb", 5, 5, 33)
