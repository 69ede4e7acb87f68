#!/bin/bash
# Regenerate the round's evidence on a B200 box (one gpurun call): bench lines, reference arm, ncu launch list,
# one `ncu --set full` capture per stage kernel (digested on the box), batch sweep, eager-PyTorch baseline.
# Everything lands in gpurun_out/ev/ (text only, a few hundred KB); copy what should be judged into profiles/.
# Usage: tools/evidence.sh [round-tag]   e.g. tools/evidence.sh r01
TAG=${1:-r02}
OUT=gpurun_out/ev; mkdir -p $OUT
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > $OUT/${TAG}_pytest_gpu.log
python bench.py > $OUT/${TAG}_bench.json 2> $OUT/bench.err
python bench.py --workload kitti_b16 --no-cpu-baseline > $OUT/${TAG}_bench_kitti_b16.json 2>> $OUT/bench.err
python bench.py --impl reference --steps 5 --warmup 3 > $OUT/${TAG}_bench_reference.json 2>> $OUT/bench.err
python tools/stage_times.py nyu_b8 > $OUT/${TAG}_stage_times_nyu_b8.txt 2>&1
python tools/stage_times.py kitti_b16 > $OUT/${TAG}_stage_times_kitti_b16.txt 2>&1
python tests/perf_eager_gpu_baseline.py nyu_b8 > $OUT/${TAG}_eager_gpu_baseline.json 2>> $OUT/bench.err
[ -n "$SWEEP" ] && python tools/sweep.py > $OUT/${TAG}_sweep_batch_1_256_f32_bf16_n1.md 2>> $OUT/bench.err
python bench.py --dtype bf16 --no-cpu-baseline > $OUT/${TAG}_bench_bf16.json 2>> $OUT/bench.err
# launch list of the same command as the bench's timed loop (eager launches so that ncu sees every kernel)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --kernel-only --no-graph > $OUT/ncu_launches.log 2>&1
# full captures, digested here; the reports themselves stay on the box
for s in head 2x2 4x4 8x8; do
  for d in bwd fwd; do
    if [ $d = bwd ]; then K='regex:chain_bwd_fused|chain_tc_dgrad|head_bwd_mma'; else K='regex:chain_tc_fwd|head_fwd_mma'; fi
    ncu --set full --clock-control none --import-source on -k "$K" --launch-skip 3 -c 1 -f -o /tmp/${d}_$s \
        python tools/stage_times.py nyu_b8 5 $s > $OUT/ncu_${d}_$s.log 2>&1
    python tools/ncu_summary.py /tmp/${d}_$s.ncu-rep "$s $d" > $OUT/${TAG}_ncu_${d}_$s.txt 2>&1
    python tools/ncu_hot.py /tmp/${d}_$s.ncu-rep 30 > $OUT/${TAG}_hot_${d}_$s.txt 2>&1
  done
done
python tools/ncu_summary.py /tmp/bwd_head.ncu-rep --json > $OUT/traffic.json 2>/dev/null
ncu --set full --clock-control none -k regex:chain_wgrad_tma --launch-skip 1 -c 1 -f -o /tmp/wgrad_4x4 \
    python tools/stage_times.py nyu_b8 5 4x4 > $OUT/ncu_wgrad.log 2>&1
python tools/ncu_summary.py /tmp/wgrad_4x4.ncu-rep "4x4 wgrad" > $OUT/${TAG}_ncu_wgrad_4x4.txt 2>&1
ls -la $OUT
