# one ncu --set full capture with source of each chunk kernel (inside bench.py's train_timed range)
set -x
mkdir -p gpurun_out
CMD="python bench.py --legs none --steps 2 --warmup 1"
timeout 300 $CMD > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err || exit 1
timeout 600 ncu --set full --clock-control none --import-source on --nvtx -c 1 -f --nvtx-include "train_timed/" -k regex:wkv6_bwd_chunk -o gpurun_out/prof_bwd $CMD > gpurun_out/ncu_b.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on --nvtx -c 1 -f --nvtx-include "train_timed/" -k regex:wkv6_fwd_chunk -o gpurun_out/prof_fwd_train $CMD > gpurun_out/ncu_a.log 2>&1
tail -2 gpurun_out/ncu_b.log
ls -la gpurun_out/*.ncu-rep
