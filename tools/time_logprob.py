"""Time the stand-alone NB / ZINB log-prob kernels alone (BASELINE.json's second metric), as bench.py does it.

    python tools/time_logprob.py
"""
import os
import sys
import types

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench

peak = bench.measured_peaks()["hbm_GBps"] if hasattr(bench, "measured_peaks") else 6424.7
r = bench.time_logprob_kernels(types.SimpleNamespace(genes=2000), torch.device("cuda:0"), peak, reps=20)
for k, v in r.items():
    print(k, round(v["ms"] * 1000, 1), "us", round(v["GBps"]), "GB/s", round(v["frac_of_hbm_peak"], 3))
