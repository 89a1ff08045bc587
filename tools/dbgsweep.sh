for b in 296 512; do for d in 0 7 256 263; do
  HM_TRUNK_DBG=$d python bench.py --no-cpu --batch $b --steps 50 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('batch $b dbg $d trunk_ms', d['per_kernel_ms']['trunk.b0-14'], 'single', d['single_stream']['ms_per_step'], 'mhz', d['clocks']['sm_mhz'])"
done; done
