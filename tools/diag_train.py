"""Diagnostic: per-blob / per-parameter errors of the training step against the oracle (no asserts)."""
import sys
import tempfile

import numpy as np
import torch

sys.path.insert(0, '.')
import tests.test_trainer_gpu as T
from fcn_object_detector_b200 import netspec
from fcn_object_detector_b200.caffemodel import write_synthetic_caffemodel
from fcn_object_detector_b200.trainer import Solver
from oracle import prototxt, train_net_ref
from oracle.caffemodel_io import read_caffemodel
from tests.util import rel_err

d = tempfile.mkdtemp()
proto = netspec.write_train_val_prototxt(d + '/tv.prototxt')
cm = write_synthetic_caffemodel(d + '/w.caffemodel', seed=3, num_cvg=1, num_bbox=4)
ratio = float(sys.argv[1]) if len(sys.argv) > 1 else -1.0
s = Solver(proto, cm, batch=T.N, height=T.H, width=T.W, seed=7, dropout_ratio_override=ratio)
layers = prototxt.layers(prototxt.parse(open(proto).read()))
W = {k: [np.asarray(a, np.float32) for a in v] for k, v in read_caffemodel(cm).items()}
data, grids = T._inputs(0)
dev = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (data,) + grids]
s.forward_backward_device([t.data_ptr() for t in dev])
torch.cuda.synchronize()
mask = s.dropout_mask()
tops = [L['top'][0] for L in train_net_ref.body_layers(layers) if L['type'] not in ('ReLU',)]
ref = train_net_ref.forward_backward(layers, W, data, grids, dropout_mask=mask, want=tops)
relu_tops = set(L['top'][0] for L in layers if L['type'] == 'ReLU')
for k in tops:
    if k not in ref['blob_grads']:
        continue
    e_d = rel_err(s.blob(k), ref['blobs'][k])
    e_g = rel_err(s.blob(k, 'diff'), ref['blob_grads'][k]) if k not in relu_tops else float('nan')
    print('%-28s data %.2e  diff %.2e  |ref diff| %.3e' % (k, e_d, e_g, np.abs(ref['blob_grads'][k]).max()))
for name in s.layers:
    gw, gb = ref['grads'][name]
    print('%-28s dW %.2e  db %.2e' % (name, rel_err(s.param(name, 0, 'grad').reshape(gw.shape), gw), rel_err(s.param(name, 1, 'grad'), gb)))


def where(tag, a, b):
    d = np.abs(a - b)
    tol = 1e-4 * np.abs(b).max()
    bad = np.argwhere(d > tol)
    print(tag, 'shape', a.shape, 'bad', len(bad), 'of', a.size, 'max at', np.unravel_index(d.argmax(), d.shape), 'got', a.flat[d.argmax()], 'want', b.flat[d.argmax()])
    if len(bad):
        for ax in range(a.ndim):
            u = np.unique(bad[:, ax])
            print('   axis', ax, 'bad indices', u[:20], '...' if len(u) > 20 else '', 'n =', len(u))


for name in ('inception_4e/pool_proj', 'inception_4d/pool_proj'):
    gw, gb = ref['grads'][name]
    where('dW ' + name, s.param(name, 0, 'grad').reshape(gw.shape), gw)
for k in ('inception_4e/pool', 'inception_5a/pool', 'inception_4d/pool'):
    where('diff ' + k, s.blob(k, 'diff'), ref['blob_grads'][k])
# the concat diff before the branch masks is not comparable; the pool_proj slice after the mask is: autograd's gradient of
# the ReLU output times (y > 0)
k = 'inception_4e/pool_proj'
y = ref['blobs'][k]
where('masked diff ' + k, s.blob(k, 'diff'), ref['blob_grads'][k] * (y > 0))
k = 'inception_4e/1x1'
y = ref['blobs'][k]
where('masked diff ' + k, s.blob(k, 'diff'), ref['blob_grads'][k] * (y > 0))
