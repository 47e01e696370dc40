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
prec = sys.argv[2] if len(sys.argv) > 2 else 'fp32'
s = Solver(proto, cm, batch=T.N, height=T.H, width=T.W, seed=7, dropout_ratio_override=ratio, precision=prec)
layers = prototxt.layers(prototxt.parse(open(proto).read()))
W = {k: [np.asarray(a, np.float32) for a in v] for k, v in read_caffemodel(cm).items()}
data, grids = T._inputs(0)
dev = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (data,) + grids]
s.forward_backward_device([t.data_ptr() for t in dev])
torch.cuda.synchronize()
mask = s.dropout_mask()
tops = [L['top'][0] for L in train_net_ref.body_layers(layers) if L['type'] not in ('ReLU',)]
argmax = {L['name']: s.pool_argmax(L['name']) for L in layers if L['type'] == 'Pooling'}
ref = train_net_ref.forward_backward(layers, W, data, grids, dropout_mask=mask, want=tops, pool_argmax=argmax)


def l2(a, b):
    return float(np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-30))

relu_tops = set(L['top'][0] for L in layers if L['type'] == 'ReLU')
for k in tops:
    if k not in ref['blob_grads']:
        continue
    e_d = rel_err(s.blob(k), ref['blobs'][k])
    e_g = rel_err(s.blob(k, 'diff'), ref['blob_grads'][k]) if k not in relu_tops else float('nan')
    print('%-28s data %.2e  diff %.2e  |ref diff| %.3e' % (k, e_d, e_g, np.abs(ref['blob_grads'][k]).max()))
for name in s.layers:
    gw, gb = ref['grads'][name]
    a = s.param(name, 0, 'grad').reshape(gw.shape)
    print('%-28s dW max %.2e l2 %.2e  db max %.2e l2 %.2e' % (name, rel_err(a, gw), l2(a, gw), rel_err(s.param(name, 1, 'grad'), gb), l2(s.param(name, 1, 'grad'), gb)))

