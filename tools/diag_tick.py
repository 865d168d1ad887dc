"""One tick from an identical state, fp32 and fp64 kernels against the CPU checker: per-env error and where it sits."""
import sys
sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import numpy as np, torch
import oracle_lib as O
from urgym_b200.batch import EnvBatch
np.set_printoptions(precision=2, linewidth=220)
name = sys.argv[1] if len(sys.argv) > 1 else 'Golf'
n = 16
for prec in ('f32', 'f64'):
    ob, gb = O.OracleBatch(name, n, seed=5), EnvBatch(name, n, seed=5, precision=prec)
    ob.reset()
    rng = np.random.default_rng(1)
    ob.step(rng.uniform(-0.1, 0.1, (n, ob.ad)).astype(np.float32).astype(np.float64), u=rng.uniform(0, 1, n))
    s0 = ob.get_state().astype(np.float32)
    gb.set_state(torch.tensor(s0.astype(np.float64) if prec == 'f64' else s0)); ob.set_state(s0.astype(np.float64))
    ob.tick(1); gb.tick(1)
    po, co = ob.contacts()
    so, sg = ob.get_state(), gb.get_state().cpu().numpy().astype(np.float64)
    d = np.abs(sg - so)
    print(prec, 'per-env max err over q[0:12]', d[:, :12].max(1))
    print(prec, 'per-env max err over qd[12:24]', d[:, 12:24].max(1))
    print(prec, 'per-env max err over free bodies', d[:, 38:77].max(1))
    print(prec, 'oracle contact counts', co)
    i = int(d[:, :24].max(1).argmax())
    print(prec, 'worst env', i, 'q err', d[i, :12], 'qd err', d[i, 12:24])
    print(prec, 'stats', gb.stats())
