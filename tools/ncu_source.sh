#!/bin/bash
# Source-level capture (SASS + stall samples) of ONE fused kernel, run on the GPU box:  bash tools/ncu_source.sh <tag> <d> [fftn]
# Writes gpurun_out/<tag>_d<d>.src.csv (source page) and .raw.csv; the .ncu-rep stays on the box (64 MiB return limit).
set -e
TAG=$1; D=$2; FFTN=${3:-2048}
O=gpurun_out
python tools/prof_run.py 8192 $D $FFTN > /dev/null
ncu --set full --clock-control none --import-source on -k regex:r2iq -s 1 -c 1 -o $O/${TAG}_d$D -f python tools/prof_run.py 8192 $D $FFTN > $O/ncu_src.log 2>&1
ncu -i $O/${TAG}_d$D.ncu-rep --page source --csv > $O/${TAG}_d$D.src.csv
ncu -i $O/${TAG}_d$D.ncu-rep --page raw --csv > $O/${TAG}_d$D.raw.csv
rm -f $O/${TAG}_d$D.ncu-rep
