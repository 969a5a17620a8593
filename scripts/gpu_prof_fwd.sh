# ncu full capture of two launches of the forward chunk kernel (variant chosen by the environment) inside the bench workload
set -x
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-train"
$CMD > gpurun_out/plain_fwd.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:wkv6_fwd_chunk -s 14 -c 2 -f -o gpurun_out/prof_fwd $CMD > gpurun_out/ncu_fwd.log 2>&1
tail -3 gpurun_out/ncu_fwd.log
