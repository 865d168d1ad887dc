for w in 1 2 3 4 6 13; do URGYM_WARPS=$w python tools/lockstep_probe.py BallBalance 16384; done
for w in 1 4; do URGYM_WARPS=$w python tools/lockstep_probe.py CubesGrasp 16384; done
