python tools/ab_bits.py gpurun_ab/coop.so urgym_b200/liburgym_b200.so > gpurun_out/ab_bits_ptr.txt 2>&1; cat gpurun_out/ab_bits_ptr.txt
for w in ball4096 two16384 push16384 grasp16384 golf16384; do for so in coop.so new; do
  if [ $so = new ]; then export URGYM_SO=$PWD/urgym_b200/liburgym_b200.so; else export URGYM_SO=$PWD/gpurun_ab/$so; fi
  python bench.py --workload $w --steps 12 --warmup 4 --no-cpu --no-e2e --no-side 2>/dev/null | python -c "
import sys, json
d=json.loads(sys.stdin.read()); print('$so', '$w', 'env-steps/s %.0f' % d['value'], 'ticks/s %.3g' % d['physics_ticks_per_s'], 'ms %.2f' % d['ms_per_step'])"
done; done 2>&1 | tee gpurun_out/ab_perf_ptr.txt
unset URGYM_SO
timeout 900 python -m pytest tests -m gpu -x -q -k "shipped_actors or single_tick or parity_build" 2>&1 | tail -4
