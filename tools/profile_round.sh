#!/bin/bash
# Evidence pass of a round, run on the GPU box (one GPU):  bash tools/profile_round.sh r02
# Every ncu command follows a plain run of the same command that exited 0.  Outputs land in gpurun_out/ and are turned
# into the tracked summaries by `python tools/summarize_profiles.py <tag> 8192` back in the container.
set -e
TAG=${1:-r02}
B="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-channelizer"
$B > gpurun_out/${TAG}_bench_short.json
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:r2iq -c 120 --csv --log-file gpurun_out/${TAG}_launches.csv $B > /dev/null 2>&1
python tools/prof_run.py 8192
ncu --set full --clock-control none --import-source on -k regex:r2iq_pair -s 7 -c 7 -o gpurun_out/${TAG}_full -f python tools/prof_run.py 8192 > gpurun_out/ncu_full.log 2>&1
python tools/prof_run.py 8192 0,1,2,3,4,5,6 8192
ncu --set full --clock-control none --import-source on -k regex:r2iq_wide -s 7 -c 7 -o gpurun_out/${TAG}_wide_full -f python tools/prof_run.py 8192 0,1,2,3,4,5,6 8192 > gpurun_out/ncu_wide.log 2>&1
python tools/mixer_bench.py 8192
ls -la gpurun_out/*.ncu-rep
