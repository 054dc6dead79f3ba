#!/bin/bash
# ncu --set full captures of five factor launches of one cfg1 step (400 systems); the plain run goes first.
# The launch indices are those of `-k regex:factor_` in the third step of bench.py --profile (67 factor
# launches per step, profiles/r2_launches_cfg1_F400.csv).  One GPU, run under gpurun.
set -x
python bench.py --profile --freqs 400 --no-north-star > gpurun_out/plain.log 2>&1 || exit 1
for spec in 139:panel64 151:panel128 162:level128 200:level256top 198:level256l2; do
  skip=${spec%%:*}; name=${spec##*:}
  ncu --set full --clock-control none --import-source on -k regex:factor_ --launch-skip $skip --launch-count 1 \
      -f -o gpurun_out/r2r_$name python bench.py --profile --freqs 400 --no-north-star > gpurun_out/ncu_$name.log 2>&1
  echo "$name rc=$?"
done
ls -la gpurun_out/*.ncu-rep
