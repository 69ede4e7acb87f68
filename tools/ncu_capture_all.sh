#!/bin/bash
# Capture one `ncu --set full` report per stage kernel at a workload's shapes, digest every report to text on
# the box (tools/ncu_summary.py), and keep gpurun_out/ below the 64 MiB pull limit.
# Usage (on the GPU box): tools/ncu_capture_all.sh [workload] [stages...]
WL=${1:-nyu_b8}; shift
STAGES=${@:-head 2x2 4x4 8x8}
OUT=gpurun_out; mkdir -p $OUT
for s in $STAGES; do
  for d in bwd fwd; do
    if [ $d = bwd ]; then K='regex:chain_bwd_fused|chain_tc_dgrad|head_bwd_mma'; else K='regex:chain_tc_fwd|head_fwd_mma'; fi
    ncu --set full --clock-control none --import-source on -k "$K" --launch-skip 3 -c 1 -f -o $OUT/${d}_$s \
        python tools/stage_times.py $WL 5 $s > $OUT/ncu_${d}_$s.log 2>&1
    python tools/ncu_summary.py $OUT/${d}_$s.ncu-rep > $OUT/sum_${d}_$s.txt 2>&1
    python tools/ncu_hot.py $OUT/${d}_$s.ncu-rep 40 > $OUT/hot_${d}_$s.txt 2>&1
  done
done
gzip -f $OUT/*.ncu-rep
# drop the largest reports until the directory fits
while [ $(du -sm $OUT | cut -f1) -gt 58 ]; do rm -f $(ls -S $OUT/*.ncu-rep.gz | head -1); done
ls -la $OUT
