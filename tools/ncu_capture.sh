#!/bin/bash
# One round's ncu evidence (run on the GPU box through gpurun; a number printed under ncu is never a bench value):
#   tools/ncu_capture.sh r02       -> gpurun_out/<tag>_{bench.json,launches_ncu.csv,full.ncu-rep,mid.ncu-rep}
set -u
TAG=${1:-r02}
OUT=gpurun_out
ARGS="--cells 262144 --steps 2 --warmup 3 --no-cpu-baseline --no-extras --blocks 1"
python bench.py --steps 200 --warmup 20 --no-cpu-baseline --no-extras > $OUT/${TAG}_bench_quick.json 2> $OUT/${TAG}_bench_quick.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/${TAG}_launches_ncu.csv \
    python bench.py $ARGS > $OUT/${TAG}_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k regex:"tc_nll_dh_kernel|tc_dw_kernel|tc_decoder_fwd_kernel|tc_enc_fwd_kernel|tc_enc_dw_kernel" -c 10 \
    -f -o $OUT/${TAG}_full python bench.py $ARGS > $OUT/${TAG}_ncu_full.log 2>&1
ncu --set full --clock-control none --kernel-name-base demangled \
    -k regex:"mid_|combine|reduce|colsum|pack_|adam|prologue|gather" -c 40 \
    -f -o $OUT/${TAG}_mid python bench.py $ARGS > $OUT/${TAG}_ncu_mid.log 2>&1
# the reports stay on the box unless small: the raw pages travel (gpurun_out is capped at 64 MiB)
ncu -i $OUT/${TAG}_full.ncu-rep --page raw --csv > $OUT/${TAG}_full_raw.csv 2>/dev/null
ncu -i $OUT/${TAG}_mid.ncu-rep --page raw --csv > $OUT/${TAG}_mid_raw.csv 2>/dev/null
rm -f $OUT/${TAG}_mid.ncu-rep
ls -la $OUT
