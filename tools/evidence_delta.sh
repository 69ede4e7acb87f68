#!/bin/bash
# Short refresh after a change to the wide-chain backward kernels: GPU tests, the two bench lines, the ncu launch
# list and `ncu --set full` digests of the dgrad and TMA weight-gradient kernels. Output: gpurun_out/ (text only).
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > $OUT/r01_pytest_gpu.log
python bench.py > $OUT/bench.json 2> $OUT/bench.err
python bench.py --workload kitti_b16 --no-cpu-baseline > $OUT/bench_kitti.json 2>> $OUT/bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/r01_launches.csv \
    python bench.py --steps 2 --warmup 3 --kernel-only --no-graph > $OUT/ncu_launches.log 2>&1
for s in 4x4 8x8; do
  ncu --set full --clock-control none -k regex:chain_wgrad_tma --launch-skip 1 -c 1 -f -o /tmp/wgrad_$s \
      python tools/stage_times.py nyu_b8 5 $s > /dev/null 2>&1
  python tools/ncu_summary.py /tmp/wgrad_$s.ncu-rep "$s wgrad" > $OUT/r01_ncu_wgrad_$s.txt 2>&1
  ncu --set full --clock-control none -k regex:chain_tc_dgrad --launch-skip 3 -c 1 -f -o /tmp/dgrad_$s \
      python tools/stage_times.py nyu_b8 5 $s > /dev/null 2>&1
  python tools/ncu_summary.py /tmp/dgrad_$s.ncu-rep "$s bwd" > $OUT/r01_ncu_bwd_$s.txt 2>&1
done
