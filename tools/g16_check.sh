for g in 32 16; do for w in ball16384 ball4096 grasp16384; do
URGYM_GROUP=$g python bench.py --workload $w --steps 20 --warmup 5 --no-cpu --no-e2e 2>/dev/null | python -c "
import sys, json
d=json.loads(sys.stdin.read()); print('G', $g, '$w', 'env-steps/s %.0f' % d['value'], 'ticks/s %.3g' % d['physics_ticks_per_s'], 'ms %.2f' % d['ms_per_step'])"
done; done
