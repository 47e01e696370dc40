import os, sys
import numpy as np
sys.path.insert(0, '/root/repo')
from tests.util import write_stem_net, oracle_forward, gpu_forward, rel_err
d = '/tmp/diag_lrn'; os.makedirs(d, exist_ok=True)
tag, H, W, batch, k, pad, cout, shift, scale, c2 = ('deploy_like', 64, 96, 2, 7, 3, 64, -127.0, 1.0, 192)
proto, cm = write_stem_net(d, tag, H, W, batch=batch, seed=len(tag), k=k, pad=pad, cout=cout, shift=shift, scale=scale, c2=c2)
x = np.random.RandomState(4).uniform(0, 1, (batch, 3, H, W)).astype(np.float32)
ref = oracle_forward(proto, cm, x, want=['c2', 'n2', 'p2'])
fused, plan = gpu_forward(proto, cm, x, 'bf16', ['n2', 'p2'])
os.environ['FCNOD_NO_FUSION'] = '1'
unf, _ = gpu_forward(proto, cm, x, 'bf16', ['c2', 'n2', 'p2'])
del os.environ['FCNOD_NO_FUSION']
print('fused n2 vs oracle', rel_err(fused['n2'], ref['n2']), 'unfused n2 vs oracle', rel_err(unf['n2'], ref['n2']), 'unfused c2 vs oracle', rel_err(unf['c2'], ref['c2']))
print('fused vs unfused n2', rel_err(fused['n2'], unf['n2']))
# LRN of the unfused GPU c2 computed on the host in fp64: isolates the LRN arithmetic from the convolution error
c = unf['c2'].astype(np.float64)
sq = np.zeros((c.shape[0], c.shape[1] + 4) + c.shape[2:]); sq[:, 2:-2] = c * c
s = sum(sq[:, i:i + c.shape[1]] for i in range(5))
lrn = c * (1 + 1e-4 / 5 * s) ** -0.75
print('fused n2 vs host-LRN(unfused c2)', rel_err(fused['n2'], lrn), ' unfused n2 vs same', rel_err(unf['n2'], lrn))
e = np.abs(fused['n2'] - lrn); i = np.unravel_index(e.argmax(), e.shape); print('worst at', i, fused['n2'][i], lrn[i], 'channel', i[1])
per_ch = e.max(axis=(0, 2, 3)); print('max err per channel (top 8):', np.argsort(-per_ch)[:8], np.sort(-per_ch)[:8] * -1)
