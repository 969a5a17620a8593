# backward kernel: per-chunk trace + one ncu --set full capture with source (inside bench.py's train_timed range)
set -x
mkdir -p gpurun_out
timeout 200 python scripts/trace_phases.py > gpurun_out/trace.txt 2>&1; grep "^bwd" gpurun_out/trace.txt | cut -c1-900
CMD="python bench.py --legs none --steps 2 --warmup 1"
timeout 300 $CMD > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err || exit 1
timeout 600 ncu --set full --clock-control none --import-source on --nvtx -c 1 -f --nvtx-include "train_timed/" -k regex:wkv6_bwd_chunk -o gpurun_out/prof_bwd $CMD > gpurun_out/ncu_b.log 2>&1
tail -2 gpurun_out/ncu_b.log
ls -la gpurun_out/*.ncu-rep
