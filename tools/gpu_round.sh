#!/bin/bash
# GPU rounds, one ncu invocation per gpurun call:
#   main : GPU tests, smoke, bench (+ reference arm), op benches, then the ncu launch list of the bench command
#   full : ncu --set full of the hot kernels of one bench step (digit prologue, tcgen05 keyswitch, blind rotation)
#   wide : ncu --set full of the small-level (wide) blind-rotation kernel at 148 ciphertexts
#   pair : ncu --set full of the 2-CTA-cluster (pair) blind-rotation kernel at 74 ciphertexts
set -x
MODE=${1:-main}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-op-latency"
if [ "$MODE" = main ]; then
  nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap --format=csv -lms 500 > gpurun_out/clocks.csv &
  SMI=$!
  timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
  tail -5 gpurun_out/pytest_gpu.log
  timeout 600 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
  timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; cat gpurun_out/bench.json
  timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err; cat gpurun_out/bench_ref.json
  kill $SMI
  timeout 600 python bench_ops.py > gpurun_out/bench_ops.json 2> gpurun_out/bench_ops.err; cat gpurun_out/bench_ops.json
  timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
elif [ "$MODE" = full ]; then
  # 3 matching launches per step (ks_digits, keyswitch_umma, blind_rotate_v2); skip the 3 warm-up steps
  timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'blind_rotate|keyswitch|ks_digits' -s 9 -c 3 -f -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
  tail -3 gpurun_out/ncu_full.log
elif [ "$MODE" = wide ]; then
  timeout 300 python tools/dev_br.py 148 > gpurun_out/plain_w.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:blind_rotate_wide -s 1 -c 1 -f -o gpurun_out/prof_wide python tools/dev_br.py 148 > gpurun_out/ncu_wide.log 2>&1
elif [ "$MODE" = pair ]; then
  timeout 300 python tools/dev_br.py 74 > gpurun_out/plain_p.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:blind_rotate_pair -s 1 -c 1 -f -o gpurun_out/prof_pair python tools/dev_br.py 74 > gpurun_out/ncu_pair.log 2>&1
fi
