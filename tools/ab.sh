# A/B of prebuilt libraries with the same ABI: tools/ab.sh A.so B.so ... (files under gpurun_ab/)
for v in "$@"; do for w in ball4096 ball16384 grasp16384; do
  URGYM_SO=$PWD/gpurun_ab/$v python bench.py --workload $w --steps 20 --warmup 5 --no-cpu --no-e2e 2>/dev/null | python -c "
import sys, json
d=json.loads(sys.stdin.read()); print('$v', '$w', 'env-steps/s %.0f' % d['value'], 'ticks/s %.3g' % d['physics_ticks_per_s'], 'ms %.2f' % d['ms_per_step'])"
done; done
