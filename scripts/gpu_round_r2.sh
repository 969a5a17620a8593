# Round deliverables (one B200): full GPU test suite, default bench (all legs), reference arm, ncu launch list and one
# ncu --set full capture of each chunk kernel inside bench.py's timed regions, per-CTA trace.  Outputs under gpurun_out/.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; echo "bench rc=$?"
cat gpurun_out/bench_1gpu.json
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cat gpurun_out/bench_ref.json
timeout 900 bash scripts/gpu_prof_r2.sh > gpurun_out/prof_r2.log 2>&1; echo "prof rc=$?"
timeout 300 python -m paper_accurate_fast_cheap_b200.build --trace > gpurun_out/build_trace.log 2>&1
timeout 200 python scripts/trace_phases.py > gpurun_out/trace.txt 2>&1; grep "^bwd" gpurun_out/trace.txt | cut -c1-300
