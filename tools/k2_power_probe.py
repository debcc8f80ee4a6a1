"""Loop the fused decoder forward (LSE + NLL + dh) for ~1 s per batch size and sample SM clock / power."""
import ctypes as C
import sys
import threading
import time

import torch

sys.path.insert(0, ".")
import bench  # noqa: E402
import scvi_tools_b200 as S  # noqa: E402
from scvi_tools_b200 import _lib  # noqa: E402
from scvi_tools_b200._lib import call  # noqa: E402
from scvi_tools_b200.data import DeviceResidentCounts  # noqa: E402
from scvi_tools_b200.train import FusedTrainingPlan, _p  # noqa: E402

import pynvml as nv  # noqa: E402

nv.nvmlInit()
h = nv.nvmlDeviceGetHandleByIndex(0)
dev = torch.device("cuda:0")
X = bench.synth_counts_device(262144, 2000, dev)
batch = torch.arange(262144, device=dev) * 8 // 262144
for B in (7168, 8192, 9472):
    data = DeviceResidentCounts(X, batch=batch, device=dev)
    torch.manual_seed(0)
    m = S.VAE(n_input=2000, n_batch=8, gene_likelihood="zinb", precision=1).train()
    plan = FusedTrainingPlan(m, data, batch_size=B, use_cuda_graph=False)
    plan.set_epoch(0)
    plan.training_step(torch.randperm(262144, device=dev)[:B])
    P = plan.P
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    Ws, bs = P["decoder.px_scale_decoder.0.weight"], P["decoder.px_scale_decoder.0.bias"]
    Wp, bp = P["decoder.px_dropout_decoder.weight"], P["decoder.px_dropout_decoder.bias"]

    def k2():
        call("scvi_decoder_nll_fwd", _p(plan.h), 128, B, 128, 2000, _p(Ws), _p(bs), _p(Wp), _p(bp), _p(P["px_r"]), _p(X),
             X.stride(0), _p(plan.idx), _p(plan.library), _lib.LIK_ZINB, 1, _p(plan.reconst), _p(plan.stats), _p(plan.g_rec),
             _p(plan.dh), 128, _p(plan.ws_fwd), st)

    rows, stop = [], threading.Event()

    def sample():
        while not stop.is_set():
            rows.append((nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM), nv.nvmlDeviceGetPowerUsage(h) / 1000.0))
            time.sleep(0.01)

    for _ in range(20):
        k2()
    torch.cuda.synchronize()
    th = threading.Thread(target=sample)
    th.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 3000
    e0.record()
    for _ in range(n):
        k2()
    e1.record()
    e1.synchronize()
    stop.set()
    th.join()
    clk = sorted(r[0] for r in rows)
    pw = sorted(r[1] for r in rows)
    print(f"B={B}: {e0.elapsed_time(e1) / n * 1000:.1f} us per fwd call; SM clock min/med/max {clk[0]}/{clk[len(clk)//2]}/{clk[-1]} MHz; "
          f"power med/max {pw[len(pw)//2]:.0f}/{pw[-1]:.0f} W; {len(rows)} samples")
