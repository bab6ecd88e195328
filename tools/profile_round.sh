#!/bin/bash
# Evidence pass of a round, run on the GPU box (one GPU):  bash tools/profile_round.sh r02
# Every ncu command follows a plain run of the same command that exited 0.  gpurun brings back at most 64 MiB, so the
# .ncu-rep files are turned into their raw page (all metrics per launch) and, for d = 0 and d = 5 of each family, their
# source page (SASS with stall samples) ON THE BOX and then deleted.  Outputs land in gpurun_out/ and become the tracked
# summaries through `python tools/summarize_profiles.py <tag> 8192` back in the container.
set -e
TAG=${1:-r02}
O=gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-channelizer"
$B > $O/${TAG}_bench_short.json
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:r2iq -c 120 --csv --log-file $O/${TAG}_launches.csv $B > /dev/null 2>&1
export_rep() {   # <rep stem> <kernel regex for source pages...>
    local stem=$1; shift
    ncu -i $O/$stem.ncu-rep --page raw --csv > $O/$stem.raw.csv
    for k in "$@"; do
        ncu -i $O/$stem.ncu-rep --page source --csv --kernel-name "regex:$k" > "$O/$stem.src.$(echo $k | tr -c 'A-Za-z0-9_\n' '_').csv" 2>/dev/null || true
    done
    rm -f $O/$stem.ncu-rep
}
python tools/prof_run.py 8192
ncu --set full --clock-control none --import-source on -k regex:r2iq_pair -s 7 -c 7 -o $O/${TAG}_full -f python tools/prof_run.py 8192 > $O/ncu_full.log 2>&1
export_rep ${TAG}_full 'r2iq_pair_kernel<0' 'r2iq_pair_kernel<3' 'r2iq_pair_kernel<5'
python tools/prof_run.py 8192 0,1,2,3,4,5,6 8192
ncu --set full --clock-control none --import-source on -k regex:r2iq_wide -s 7 -c 7 -o $O/${TAG}_wide_full -f python tools/prof_run.py 8192 0,1,2,3,4,5,6 8192 > $O/ncu_wide.log 2>&1
export_rep ${TAG}_wide_full 'r2iq_wide_pair_kernel<0' 'r2iq_wide_pair_kernel<3' 'r2iq_wide_pair_kernel<5'
python tools/chan_bench.py > $O/${TAG}_chan_bench.txt 2>&1 || true
ncu --set full --clock-control none --import-source on -k regex:r2iq_channels -c 2 -o $O/${TAG}_chan_full -f python tools/chan_bench.py > $O/ncu_chan.log 2>&1 || true
[ -f $O/${TAG}_chan_full.ncu-rep ] && export_rep ${TAG}_chan_full 'r2iq_channels'
python tools/mixer_bench.py 8192 > $O/${TAG}_mixer_bench.txt 2>&1 || true
du -sh $O
