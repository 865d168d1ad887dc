for w in ball4096 ball16384 two16384 push16384 grasp16384 golf16384 pushdiff16384 golfjoints16384; do
  python bench.py --workload $w --steps 20 --warmup 5 --cpu-seconds 4 2>/dev/null | python -c "
import sys, json
d=json.loads(sys.stdin.read()); print('$w', 'env-steps/s %.0f' % d['value'], 'e2e %.0f' % d['e2e']['value'], 'ticks/step %.1f' % d['ticks_per_env_step'], 'ticks/s %.3g' % d['physics_ticks_per_s'], 'ms %.2f' % d['ms_per_step'], 'cpu %.0f (%d cores)' % (d['cpu_baseline']['value'], d['cpu_baseline']['cores']))"
done
