# Round-2 profile pass (GPU box): launch list of the default bench command + one ncu --set full capture of each chunk kernel
# inside bench.py's own timed regions (NVTX ranges train_timed / infer_timed).  Outputs under gpurun_out/.
set -x
mkdir -p gpurun_out
CMD="python bench.py --legs infer --steps 2 --warmup 1"
$CMD > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
NCU="ncu --set full --clock-control none --import-source on --nvtx -c 1 -f"
$NCU --nvtx-include "train_timed/" -k regex:wkv6_fwd_chunk -o gpurun_out/prof_fwd_train $CMD > gpurun_out/ncu_a.log 2>&1
$NCU --nvtx-include "train_timed/" -k regex:wkv6_bwd_chunk -o gpurun_out/prof_bwd $CMD > gpurun_out/ncu_b.log 2>&1
$NCU --nvtx-include "infer_timed/" -k regex:wkv6_fwd_chunk -o gpurun_out/prof_fwd $CMD > gpurun_out/ncu_c.log 2>&1
tail -2 gpurun_out/ncu_a.log gpurun_out/ncu_b.log gpurun_out/ncu_c.log
ls -la gpurun_out/*.ncu-rep
