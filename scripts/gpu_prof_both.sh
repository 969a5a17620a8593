set -x
mkdir -p gpurun_out
CMD="python bench.py --mode train --steps 2 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain_both.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:wkv6_.*_chunk -s 28 -c 24 -o gpurun_out/prof_both $CMD > gpurun_out/ncu_both.log 2>&1
tail -3 gpurun_out/ncu_both.log
