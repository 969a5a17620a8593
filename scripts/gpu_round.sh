# round deliverables: full GPU test suite, default bench, reference arm, ncu launch list (the --set full capture is scripts/gpu_round_ncu.sh, a call of its own)
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench_infer.json 2> gpurun_out/bench_infer.err; echo "bench rc=$?"
cat gpurun_out/bench_infer.json
timeout 300 python bench.py --mode train --steps 30 --no-cpu-baseline > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err; echo "train rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>&1; cat gpurun_out/bench_ref.json
CMD="python bench.py --mode train --steps 2 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_train.csv $CMD > gpurun_out/ncu1.log 2>&1
