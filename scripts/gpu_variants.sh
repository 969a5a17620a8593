# A/B of kernel variants built into variants/*.so (same box; loaded through WKV6_B200_LIB): backward errors vs the fp64 oracle,
# per-CTA timeline, short bench
mkdir -p gpurun_out
for v in variants/*.so; do
  echo "=== $v"
  export WKV6_B200_LIB=$PWD/$v
  timeout 300 python scripts/dbg_bwd.py 2>&1 | grep "chunk" | grep "stress\|T750" | awk '{print $1,$2,$3,$4,$5,$6,$8,$9,$10,$11,$12,$13}'
  timeout 200 python scripts/trace_phases.py 2>&1 | grep "bwd loop tops\|kernel span" | cut -c1-260
  timeout 300 python bench.py --steps 50 --legs none 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('BWD us', round(d['roofline']['kernel_ms']*1e3,2), 'FWD us', round(d['roofline_fwd']['kernel_ms']*1e3,2), 'step ms', round(d['ms_per_step'],4))"
done
