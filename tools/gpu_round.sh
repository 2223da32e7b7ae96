#!/bin/bash
# GPU evidence passes, one ncu invocation per gpurun call (B200_PROFILING.md):
#   launch : ncu launch list (gpu__time_duration.sum) of one short bench command
#   full   : ncu --set full of the hot kernels of one bench step (digit prologue, tcgen05 keyswitch, blind rotation)
#   wide   : ncu --set full of the small-level (wide) blind-rotation kernel at 148 ciphertexts
#   v2     : ncu --set full of the throughput blind-rotation kernel at one round of 592 ciphertexts
#   pair   : ncu --set full of the 2-CTA-cluster (pair) blind-rotation kernel at 74 ciphertexts
set -x
MODE=${1:-launch}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-op-latency"
if [ "$MODE" = launch ]; then
  timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
  tail -3 gpurun_out/ncu_launch.log
elif [ "$MODE" = full ]; then
  # 3 matching launches per step (ks_digits, keyswitch_umma, blind_rotate_v2); skip the 3 warm-up steps
  timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'blind_rotate|keyswitch|ks_digits' -s 9 -c 3 -f -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
  tail -3 gpurun_out/ncu_full.log
elif [ "$MODE" = wide ]; then
  timeout 300 python tools/dev_br.py 148 > gpurun_out/plain_w.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:blind_rotate_wide -s 1 -c 1 -f -o gpurun_out/prof_wide python tools/dev_br.py 148 > gpurun_out/ncu_wide.log 2>&1
elif [ "$MODE" = v2 ]; then
  timeout 300 python tools/dev_br.py 592 > gpurun_out/plain_v2.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:blind_rotate_v2 -s 1 -c 1 -f -o gpurun_out/prof_v2 python tools/dev_br.py 592 > gpurun_out/ncu_v2.log 2>&1
elif [ "$MODE" = pair ]; then
  timeout 300 python tools/dev_br.py 74 > gpurun_out/plain_p.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:blind_rotate_pair -s 1 -c 1 -f -o gpurun_out/prof_pair python tools/dev_br.py 74 > gpurun_out/ncu_pair.log 2>&1
fi
