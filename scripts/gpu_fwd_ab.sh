for v in variants/*.so; do
  echo "=== $v"
  export WKV6_B200_LIB=$PWD/$v
  timeout 300 python bench.py --steps 50 --legs infer 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('BWD us', round(d['roofline']['kernel_ms']*1e3,2), 'FWDtrain us', round(d['roofline_fwd']['kernel_ms']*1e3,2), 'FWDinfer us', round(d['infer']['roofline']['kernel_ms']*1e3,2), 'step ms', round(d['ms_per_step'],4), 'infer ms', round(d['infer']['ms_per_step'],4))"
done
unset WKV6_B200_LIB
timeout 600 python -m pytest tests/test_gpu_chunk.py tests/test_gpu_parity.py tests/test_gpu_long.py -x -q -m gpu 2>&1 | tail -2
