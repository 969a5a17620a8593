# backward-kernel iteration: balanced-grid / chunk / parity / long tests, short bench (headline only), phase + CTA trace
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_balanced.py -x -q -m gpu > gpurun_out/pytest_bal.log 2>&1; echo "pytest bal rc=$?"
tail -15 gpurun_out/pytest_bal.log
timeout 600 python -m pytest tests/test_gpu_chunk.py tests/test_gpu_parity.py tests/test_gpu_long.py -x -q -m gpu > gpurun_out/pytest_bwd.log 2>&1; echo "pytest rc=$?"
tail -6 gpurun_out/pytest_bwd.log
timeout 300 python bench.py --steps 50 --legs none > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_quick.json'))
print('BWD us', d['roofline']['kernel_ms']*1e3, 'frac', d['roofline']['frac'], 'FWD us', d['roofline_fwd']['kernel_ms']*1e3, 'step ms', d['ms_per_step'], 'Mfps', d['value']/1e6)
"
timeout 120 python scripts/trace_phases.py > gpurun_out/trace.txt 2>&1; grep "^bwd" gpurun_out/trace.txt | cut -c1-600
