set -x
mkdir -p gpurun_out
CMD="python bench.py --mode train --steps 2 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain_bwd.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:wkv6_bwd_chunk -s 14 -c 2 -o gpurun_out/prof_bwd_chunk $CMD > gpurun_out/ncu_bwd.log 2>&1
tail -3 gpurun_out/ncu_bwd.log
