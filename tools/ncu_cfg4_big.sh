#!/bin/bash
# ncu --set full capture of one large-front factor launch of the north-star config (cfg4, 148 systems):
# the 11th factor_level_kernel<256> launch of the third step = level 12 (4 fronts per system, s = 660-870).
set -x
export JPB_SCHUR_SPLIT=${JPB_SCHUR_SPLIT:-0}
python bench.py --profile --workload cfg4 --freqs 148 --no-north-star > gpurun_out/plain4.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"factor_level_kernel.*int.256" --launch-skip 34 --launch-count 1 \
    -f -o gpurun_out/r3b_cfg4_level12 python bench.py --profile --workload cfg4 --freqs 148 --no-north-star > gpurun_out/ncu4b.log 2>&1
echo "rc=$?"
ncu -i gpurun_out/r3b_cfg4_level12.ncu-rep --page raw --csv > gpurun_out/r3b_cfg4_level12_raw.csv 2>/dev/null
ncu -i gpurun_out/r3b_cfg4_level12.ncu-rep --page source --csv > gpurun_out/r3b_cfg4_level12_source.csv 2>/dev/null
ls -la gpurun_out/r3b_*
