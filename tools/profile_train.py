"""One training step (batch 32, 448x448) bracketed by cudaProfilerStart/Stop for `ncu --profile-from-start off`.
usage: python tools/profile_train.py [fp32|bf16] [batch]"""
import sys

import numpy as np
import torch

sys.path.insert(0, '.')
import bench_train
from fcn_object_detector_b200.trainer import Solver

prec = sys.argv[1] if len(sys.argv) > 1 else 'bf16'
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
proto, cm = bench_train.train_files()
s = Solver(proto, cm, batch=B, height=448, width=448, precision=prec)
g = 28
rs = np.random.RandomState(0)
tops = [torch.from_numpy(rs.uniform(0, 255, (B, 3, 448, 448)).astype(np.float32)).cuda()] + \
       [torch.from_numpy(rs.uniform(0, 1, (B, c, g, g)).astype(np.float32)).cuda() for c in (1, 4, 4, 4, 4)]
for _ in range(2):
    s.forward_backward_device([t.data_ptr() for t in tops])
    s.apply_update()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
s.forward_backward_device([t.data_ptr() for t in tops])
s.apply_update()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print('losses', s.losses())
s.close()
