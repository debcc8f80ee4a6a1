"""Debug aid (GPU): which gradients are NaN for the encoded-library option under precision=1."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import scvi_tools_b200 as S
from tests.test_gpu_options import load_option_case
from tests._golden import rel_err

c = load_option_case("opt_libenc_nb")
dev = torch.device("cuda:0")
for prec in (0, 1):
    kw = dict(c["kwargs"])
    for k in ("library_log_means", "library_log_vars"):
        if k in c["extra"]:
            kw[k] = c["extra"][k]
    m = S.VAE(**kw, precision=prec)
    m.load_state_dict(c["params"], strict=True)
    m = m.to(dev).train()
    tensors = {k: v.to(dev) for k, v in c["tensors"].items()}
    print("kwargs", kw.keys(), "X", tensors["X"].shape)
    with S.rng_feed(drop_masks=c["masks"], noise=c["noises"]):
        inf, gen, lo = m(tensors, loss_kwargs={"kl_weight": float(c["extra"]["kl_weight"])})
    hooks = {}
    lo.loss.backward()
    torch.cuda.synchronize()
    print("precision", prec, "loss", float(lo.loss))
    for k, p in m.named_parameters():
        if p.grad is None:
            continue
        ref = c["out"].get("grad::" + k)
        nan = bool(torch.isnan(p.grad).any())
        print(f"  {k:55s} nan={nan} rel={rel_err(p.grad.cpu(), ref) if ref is not None and not nan else None}")
