timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "resynchronised or single_tick or reset_ticks or env_logic" 2>&1 | tail -n 6
for w in ball4096 ball16384 grasp16384 golf16384; do
  python bench.py --workload $w --steps 20 --warmup 5 --no-cpu --no-e2e 2>/dev/null | python -c "
import sys, json
d=json.loads(sys.stdin.read()); print('$w', 'env-steps/s %.0f' % d['value'], 'ticks/s %.3g' % d['physics_ticks_per_s'], 'ms %.2f' % d['ms_per_step'])"
done
