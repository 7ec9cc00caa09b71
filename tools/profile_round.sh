#!/bin/bash
# Run on the GPU box (gpurun): plain bench first, then the ncu passes of the same commands.
# Outputs under gpurun_out/ (scratch); tools/ncu_summarize.py turns them into profiles/.
set -u
R=${1:-r02}
O=gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --sweep 65536 --no-cpu --no-configs"
timeout 300 $CMD > $O/${R}_profcmd.json 2> $O/${R}_profcmd.err || { echo "plain run failed"; tail -5 $O/${R}_profcmd.err; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${R}_launches_bench_jet256.csv $CMD > $O/ncu1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:jet2_kernel -s 4 -c 1 -o $O/${R}_jet2_kernel -f $CMD > $O/ncu2.log 2>&1
for cfg in "0 1 2" "2 1 2"; do
  tag=$(echo $cfg | tr -d ' ')
  timeout 120 python tools/run_survey.py 16 2000 $cfg > $O/ncu3.log 2>&1 || { echo "survey run failed"; exit 1; }
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:tile_kernel -s 1 -c 1 -o $O/${R}_survey_tile_$tag -f python tools/run_survey.py 16 2000 $cfg > $O/ncu4.log 2>&1
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/${R}_launches_survey_212.csv python tools/run_survey.py 16 2000 2 1 2 > $O/ncu5.log 2>&1
ls -la $O/*.ncu-rep
