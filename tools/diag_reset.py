"""Per-element difference of the first observation after reset(): parity build (fp64) vs the CPU checker."""
import sys
sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import numpy as np
import oracle_lib as O
from urgym_b200.batch import EnvBatch
np.set_printoptions(precision=3, linewidth=200)
for name in sys.argv[1:] or ['CubesGrasp']:
    n = 32
    ob, gb = O.OracleBatch(name, n, seed=3), EnvBatch(name, n, seed=3, precision='f64')
    oo, ont = ob.reset()
    go = gb.reset().cpu().numpy().astype(np.float64)
    d = np.abs(go - oo)
    print(name, 'ticks', ont[:4], gb.reset_ticks.cpu().numpy()[:4], 'max err per obs column:')
    print(d.max(0))
    i = np.unravel_index(d.argmax(), d.shape)
    print('worst env', i, 'oracle', oo[i[0]], 'gpu', go[i[0]])
    so = ob.get_state(); sg = gb.get_state().cpu().numpy()
    ds = np.abs(so - sg)
    print('state err per slot (max over envs):'); print(ds.max(0))
