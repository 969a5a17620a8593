# quick GPU iteration: chunk parity tests, short bench (no CPU baseline / e2e), phase + CTA trace
set -x
mkdir -p gpurun_out
timeout 180 python -m pytest tests/test_gpu_chunk.py -x -q > gpurun_out/pytest_chunk.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/pytest_chunk.log
WKV6_DEBUG_OCC=1 timeout 180 python bench.py --steps 50 --no-cpu-baseline --no-e2e $BENCH_EXTRA > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_quick.json'))
print('FWD us', d['roofline']['kernel_ms']*1e3, 'frac', d['roofline']['frac'], 'Mfps', d['value']/1e6)
t=d.get('train')
if t: print('BWD us', t['roofline_bwd']['kernel_ms']*1e3, 'frac', t['roofline_bwd']['frac'], 'train Mfps', t['value']/1e6)
"
grep wkv6 gpurun_out/bench_quick.err
timeout 120 python scripts/trace_phases.py > gpurun_out/trace.txt 2>&1; tail -6 gpurun_out/trace.txt
