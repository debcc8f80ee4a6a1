#!/bin/bash
# usage: tools/ddp_modes.sh N  -- step time of every data-parallel exchange mode at N GPUs (value leg only)
N=$1
if [ -n "$2" ]; then MODES=("p2p --p2p-buckets 0" "p2p --p2p-buckets 1" "p2p --p2p-buckets 2"); else MODES=("p2p --p2p-buckets 0" "p2p --p2p-buckets 2" "p2p --p2p-buckets 1" "split" "overlap" "graph"); fi
for mode in "${MODES[@]}"; do
  tag=$(echo $mode | tr -d ' -')
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29520 \
    bench.py --gpus $N --steps 50 --warmup 20 --no-extras --no-cpu-baseline --value-only --ddp-mode $mode \
    > gpurun_out/ddp_${N}_${tag}.json 2> gpurun_out/ddp_${N}_${tag}.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/ddp_${N}_${tag}.json") if l.startswith("{")][-1])
    print("$mode", "N=$N", round(d["ms_per_step"],4), "ms/step", d["ddp_check"])
except Exception as e:
    print("$mode failed", e)
PY
done
