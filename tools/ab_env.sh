# Throughput A/B of (library, environment) variants: tools/ab_env.sh "lib.so VAR=val ..." ... -- workload ...
# (libraries under gpurun_ab/, same ABI). One line per variant and workload.
variants=(); wls=(); seen=0
for a in "$@"; do if [ "$a" = "--" ]; then seen=1; elif [ $seen = 0 ]; then variants+=("$a"); else wls+=("$a"); fi; done
[ ${#wls[@]} = 0 ] && wls=(ball4096 ball16384 two16384)
for v in "${variants[@]}"; do set -- $v; so=$1; shift; for w in "${wls[@]}"; do
  env URGYM_SO=$PWD/gpurun_ab/$so "$@" python bench.py --workload $w --steps 20 --warmup 5 --no-cpu --no-e2e 2>/dev/null | python -c "
import sys, json
d=json.loads(sys.stdin.read()); print('$v', '$w', 'env-steps/s %.0f' % d['value'], 'ticks/s %.3g' % d['physics_ticks_per_s'], 'ms %.2f' % d['ms_per_step'])"
done; done
