# A/B of the programmatic-dependent-launch attribute on the bench workload
for v in 0 1; do
  if [ $v = 1 ]; then export WKV6_NO_PDL=1; fi
  timeout 200 python bench.py --steps 50 --no-cpu-baseline --no-e2e > gpurun_out/bench_ab$v.json
  python -c "
import json; d=json.load(open('gpurun_out/bench_ab$v.json'))
print('NO_PDL=$v FWD us', d['roofline']['kernel_ms']*1e3, d['value']/1e6, 'BWD us', d['train']['roofline_bwd']['kernel_ms']*1e3, d['train']['value']/1e6)"
done
