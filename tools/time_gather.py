"""Time scvi_gather_rows alone: random rows of a resident fp32 matrix larger than L2 ([B, G] out of [n, G]).

    python tools/time_gather.py
"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from scvi_tools_b200._lib import call

dev = torch.device("cuda:0")
n, G, B = 262144, 2000, 8192
X = torch.rand(n, G, device=dev)
out = torch.empty(B, G, device=dev)
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
ts = []
for r in range(25):
    idx = torch.randint(0, n, (B,), device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    call("scvi_gather_rows", C.c_void_p(X.data_ptr()), n, G, G, C.c_void_p(idx.data_ptr()), B, C.c_void_p(out.data_ptr()), G, st)
    e1.record()
    e1.synchronize()
    if r >= 5:
        ts.append(e0.elapsed_time(e1))
    assert r or torch.equal(out, X[idx])
ms = float(np.median(ts))
nbytes = 2 * 4 * B * G + 8 * B
print(f"scvi_gather_rows [{B} of {n}, {G}] {ms * 1000:.1f} us  {nbytes / ms / 1e6:.0f} GB/s")
