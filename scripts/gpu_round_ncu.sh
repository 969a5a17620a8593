# second call of the round deliverables: one ncu --set full capture of forward and backward launches of the training bench
set -x
mkdir -p gpurun_out
CMD="python bench.py --mode train --steps 2 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:wkv6_.*_chunk -s 32 -c 8 -f -o gpurun_out/prof_both $CMD > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
