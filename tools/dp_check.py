"""Under torchrun (one process per GPU): the data-parallel update over peer memory against the NCCL all-reduce path.
Same snapshot, same per-rank minibatches, two iterations each; compares the reduced gradients of this rank's shard, the
parameters, and that all replicas hold identical parameters.  usage: torchrun --nproc-per-node N tools/dp_check.py [fp32|bf16]"""
import ctypes as C
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, '.')
import bench_train
from fcn_object_detector_b200 import _lib
from fcn_object_detector_b200.trainer import DataParallelSolver, Solver

prec = sys.argv[1] if len(sys.argv) > 1 else 'bf16'
rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dist.init_process_group('nccl', device_id=torch.device('cuda', local))
proto, cm = bench_train.train_files() if rank == 0 else (None, None)
files = [proto, cm]
dist.broadcast_object_list(files, src=0)
proto, cm = files
B, H, W = 4, 224, 224
g = (H // 16, W // 16)
rs = np.random.RandomState(100 + rank)
tops = [torch.from_numpy(rs.uniform(0, 255, (B, 3, H, W)).astype(np.float32)).cuda()] + \
       [torch.from_numpy((rs.uniform(0, 1, (B, c) + g) > 0.7).astype(np.float32)).cuda() for c in (1, 4, 4, 4, 4)]
ptrs = [t.data_ptr() for t in tops]
res = {}
for mode in ('nccl', 'peer'):
    s = Solver(proto, cm, batch=B, height=H, width=W, device=local, seed=5 + rank, precision=prec, dropout_ratio_override=0.0)
    dp = DataParallelSolver(s, collective=mode)
    for _ in range(2):
        dp.step_device(ptrs)
    torch.cuda.synchronize()
    dist.barrier()
    params = torch.as_tensor(type('A', (), {'__cuda_array_interface__': {'shape': (s.n_params,), 'typestr': '<f4', 'data': (s.param_ptr, False), 'version': 2, 'strides': None}})(), device='cuda:%d' % local).clone()
    grads = dp.grads.clone()
    # replicas identical?
    ref = params.clone()
    dist.broadcast(ref, src=0)
    same = bool(torch.equal(ref, params))
    # timing
    torch.cuda.synchronize(); dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        dp.step_device(ptrs)
    b.record(); torch.cuda.synchronize()
    res[mode] = dict(params=params, grads=grads, same=same, ms=a.elapsed_time(b) / 5, n=s.n_params)
    dist.barrier()
    del dp
    s.close()
n = res['nccl']['n']
per = ((n // 256 + world - 1) // world) * 256
lo, hi = min(n, per * rank), min(n, per * (rank + 1))
gn, gp = res['nccl']['grads'][lo:hi], res['peer']['grads'][lo:hi]
ge = float((gn - gp).abs().max() / gn.abs().max().clamp_min(1e-30))
pd = (res['nccl']['params'] - res['peer']['params']).abs()
frac = float((pd > 1e-6).float().mean())
print('rank %d: replicas identical nccl=%s peer=%s | reduced gradients of my shard: max rel diff %.2e | params differing by > 1e-6: %.2e of %d '
      '(max %.2e) | ms/step nccl %.3f peer %.3f' % (rank, res['nccl']['same'], res['peer']['same'], ge, frac, n, float(pd.max()), res['nccl']['ms'], res['peer']['ms']), flush=True)
ok = res['peer']['same'] and ge < 1e-4 and frac < 2e-3
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
