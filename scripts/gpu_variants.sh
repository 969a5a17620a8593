# A/B of backward kernel variants built into variants/*.so (same box): errors vs fp64 oracle + per-chunk cycles + bench
mkdir -p gpurun_out
cp paper_accurate_fast_cheap_b200/libwkv6_b200.so /tmp/lib_orig.so
for v in variants/*.so; do
  echo "=== $v"
  cp $v paper_accurate_fast_cheap_b200/libwkv6_b200.so
  timeout 300 python scripts/dbg_bwd.py 2>&1 | grep chunk | awk '{print $1,$2,$3,$4,$5,$6,$8,$9,$10,$11,$12,$13}'
  timeout 200 python scripts/trace_phases.py 2>&1 | grep "bwd loop tops\|kernel span" | cut -c1-260
  timeout 300 python bench.py --steps 50 --legs none 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('BWD us', round(d['roofline']['kernel_ms']*1e3,2), 'step ms', round(d['ms_per_step'],4))"
done
cp /tmp/lib_orig.so paper_accurate_fast_cheap_b200/libwkv6_b200.so
