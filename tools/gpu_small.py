import sys
sys.path.insert(0, '.')
import torch
from urgym_b200.batch import EnvBatch
name = sys.argv[1] if len(sys.argv) > 1 else 'BallBalance'
gb = EnvBatch(name, 2, seed=3)
o = gb.reset(); torch.cuda.synchronize()
print(o, gb.reset_ticks)
