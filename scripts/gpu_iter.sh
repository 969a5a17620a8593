# one GPU iteration: parity tests, default bench (inference + training figure), phase trace
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_infer.json 2> gpurun_out/bench_infer.err; echo "bench rc=$?"
cat gpurun_out/bench_infer.json
tail -3 gpurun_out/bench_infer.err
python scripts/trace_phases.py > gpurun_out/trace.txt 2>&1; cat gpurun_out/trace.txt | tail -6
