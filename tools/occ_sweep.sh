for x in 0 1500 4000 7000 12000 20000; do
  for w in ball16384 ball4096; do
    URGYM_EXTRA_SMEM=$x python bench.py --workload $w --steps 30 --warmup 5 --no-cpu --no-e2e 2>/dev/null | python -c "
import sys, json
d=json.loads(sys.stdin.read()); print('extra', $x, '$w', 'env-steps/s %.0f' % d['value'], 'ticks/s %.3g' % d['physics_ticks_per_s'], 'ms %.2f' % d['ms_per_step'])"
  done
done
