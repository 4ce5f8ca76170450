#!/bin/bash
# Run on the GPU box (through gpurun): bench without a profiler first, then the ncu launch list of
# the same command and `--set full` captures of each kernel. Outputs land in gpurun_out/<tag>_*.
tag=${1:-r1}
out=gpurun_out
set -x
python bench.py --steps 5 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err || exit 1
EAV1_BENCH_PROFILE=1 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file $out/${tag}_launches.csv python bench.py --only none --steps 2 --warmup 3 > $out/${tag}_ncu_launches.log 2>&1
# --set full over the first 40 frames of the timed step (key frame, an ARF, ~38 inter frames: >= 10 launches of every
# kernel, spread over the clip instead of its first three frames). Direct launches (EAV1_BENCH_NO_GRAPH) so that the
# capture can stop after a launch count; the kernels and their arguments are those of the graph.
EAV1_BENCH_PROFILE=1 EAV1_BENCH_NO_GRAPH=1 ncu --profile-from-start off --set full --clock-control none --import-source on \
    -k regex:"recon_frame_kernel|resid_kernel|mc_kernel|cdef_kernel|lr_kernel|deblock_kernel" --launch-count 210 \
    -o $out/${tag}_full python bench.py --only none --steps 2 --warmup 3 > $out/${tag}_ncu_full.log 2>&1
python tools/vorbis_bench.py > $out/${tag}_vorbis.json 2> $out/${tag}_vorbis.err
ncu --set full --clock-control none --import-source on -k regex:vorbis_kernel --launch-skip 6 --launch-count 1 \
    -o $out/${tag}_vorbis_full python tools/vorbis_bench.py > $out/${tag}_ncu_vorbis.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"fg_apply_kernel|fg_prepare_kernel" --launch-skip 120 --launch-count 4 \
    -o $out/${tag}_grain_full python -m pytest tests/test_filmgrain.py -m gpu -x -q > $out/${tag}_ncu_grain.log 2>&1
ls -la $out/${tag}_*
# the reports themselves are large (source import): keep their raw pages, drop the .ncu-rep files so
# that gpurun_out stays under the 64 MiB that travel back
for r in full vorbis_full grain_full; do
  ncu -i $out/${tag}_$r.ncu-rep --page raw --csv > $out/${tag}_${r}_raw.csv 2>/dev/null
  rm -f $out/${tag}_$r.ncu-rep
done
ls -la $out/${tag}_*
