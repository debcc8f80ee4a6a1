"""Debug aid (GPU, library built with -DSCVI_K2_TRACE): per-tile phase durations of K2's CTA (0,0) in SM clocks."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import tools.profile_step  # noqa: F401  (runs 3 eager steps at the bench shape)
from scvi_tools_b200._lib import load

lib = load()
buf = (C.c_longlong * (3 * 64 * 8))()
rc = lib.scvi_debug_k2_trace(buf, 3 * 64 * 8)
assert rc == 0, rc
t = np.array(buf, dtype=np.int64).reshape(3, 64, 8)
nt = 31  # tiles per CTA at G = 2000 with 2 gene splits
for who, name in ((0, "epilogue warp 4"), (1, "epilogue warp 19")):
    a = t[who, 1:nt]
    print(name)
    print("  end of tile -> top ", np.median(a[1:, 6] - a[:-1, 5]))
    print("  issue next x loads ", np.median(a[:, 7] - a[:, 6]))
    print("  prep x (uses loads)", np.median(a[:, 0] - a[:, 7]))
    print("  wait tfull+wfull   ", np.median(a[:, 1] - a[:, 0]))
    print("  tcgen05.ld         ", np.median(a[:, 2] - a[:, 1]))
    print("  math + split       ", np.median(a[:, 3] - a[:, 2]))
    print("  wait sempty        ", np.median(a[:, 4] - a[:, 3]))
    print("  store + proxy fence", np.median(a[:, 5] - a[:, 4]))
    print("  tile period        ", np.median(np.diff(t[who, 1:nt, 0])))
a = t[2, 1:nt]
print("MMA issuer")
print("  logits issue (before wait sfull)", np.median(a[:, 0][1:] - a[:, 2][:-1]))
print("  wait sfull                      ", np.median(a[:, 1] - a[:, 0]))
print("  dh issue + commits              ", np.median(a[:, 2] - a[:, 1]))
print("  tile period                     ", np.median(np.diff(a[:, 0])))
print("first tiles (epilogue warp 4):")
print(t[0, :4] - t[0, 0, 0])
