#!/bin/bash
# ncu --set full captures of the channelizer kernels that profile_round.sh does not cover, run on the GPU box (one GPU):
#   bash tools/chan_profile.sh r02e
# FFTN 2048 d = 2, 3, 4 (r2iq_channels_pair_kernel) and FFTN 8192 d = 4, 5 (r2iq_wide_channels_kernel), 64 channels x 2048 blocks.
# Every ncu command follows a plain run of the same command that exited 0; the raw page is exported on the box and the
# report deleted (gpurun returns at most 64 MiB).  Summary: python tools/summarize_chan.py <tag>  ->  profiles/<tag>_chan_kernels.json
set -e
TAG=${1:-r02e}
O=gpurun_out
mkdir -p $O
: > $O/${TAG}_chan_plain.txt
cap() {   # <stem> <command...>
    local stem=$1; shift
    "$@" >> $O/${TAG}_chan_plain.txt 2>&1
    ncu --set full --clock-control none -k regex:channels -s 2 -c 1 -o $O/$stem -f "$@" > $O/$stem.log 2>&1
    ncu -i $O/$stem.ncu-rep --page raw --csv > $O/$stem.raw.csv
    rm -f $O/$stem.ncu-rep
}
for d in 2 3 4; do cap ${TAG}_chan2048_d$d python tools/chan_bench.py 64 $d 2048 3; done
for d in 4 5; do cap ${TAG}_chan8192_d$d python tools/chan_bench8.py 64 $d 2048; done
cat $O/${TAG}_chan_plain.txt
