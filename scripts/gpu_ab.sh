for v in "" "WKV6_FWD_CHUNK_WS=1"; do
  echo "== $v"
  env $v timeout 180 python bench.py --steps 50 --no-cpu-baseline --no-e2e --no-train > gpurun_out/ab.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/ab.json'))
print('FWD us', d['roofline']['kernel_ms']*1e3, 'frac', d['roofline']['frac'])"
  env $v timeout 120 python scripts/trace_phases.py 2>&1 | grep -E "fwd t32|solo|loop tops"
done
