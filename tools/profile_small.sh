#!/bin/bash
# Reduced refresh of tools/profile_round.sh (run on the GPU box through gpurun): the bench line, the ncu launch list of
# the timed step, and `--set full` captures of the kernels named in $2 (a regex) only.
tag=${1:-r2g}
kernels=${2:-"cdef_kernel|picture_copy_kernel|lr_kernel"}
out=gpurun_out
set -x
python bench.py --steps 5 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err || exit 1
EAV1_BENCH_PROFILE=1 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file $out/${tag}_launches.csv python bench.py --only none --steps 2 --warmup 3 > $out/${tag}_ncu_launches.log 2>&1
EAV1_BENCH_PROFILE=1 EAV1_BENCH_NO_GRAPH=1 ncu --profile-from-start off --set full --clock-control none --import-source on \
    -k regex:"$kernels" --launch-count 90 \
    -o $out/${tag}_full python bench.py --only none --steps 2 --warmup 3 > $out/${tag}_ncu_full.log 2>&1
ncu -i $out/${tag}_full.ncu-rep --page raw --csv > $out/${tag}_full_raw.csv 2>/dev/null
rm -f $out/${tag}_full.ncu-rep
ls -la $out/${tag}_*
