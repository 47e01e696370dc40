"""Training-side head of `models/train_val.prototxt` on the B200 path: the Slice / Eltwise mask-and-normalise chain and
the two losses (/root/reference/models/train_val.prototxt:36-72,2237-2281), consuming the network's `coverage` /
`bboxes` and the five label grids `DataArgumentationLayer` produces.  One fused kernel (fcnod_train_head_forward);
also returns the gradients that reach the two network outputs (the first step of the backward pass).
The rest of the training iteration (forward at TRAIN phase, backward of the body, ADAM, data-parallel all-reduce) is in
trainer.py / csrc/train_net.cu, which calls this head on the device.
"""
import ctypes as C

import numpy as np

from . import _lib

BBOX_LOSS_WEIGHT = 2.0          # train_val.prototxt:2273


def train_head_forward(coverage, bboxes, tops, bbox_loss_weight=BBOX_LOSS_WEIGHT, want_grads=True):
    """coverage (N,C,G,G), bboxes (N,4C,G,G): network outputs; tops = the layer's tops 1..5 (coverage-label, bbox-label,
    size-block, obj-block, coverage-block).  Returns dict(loss_bbox, loss_coverage, loss = loss_coverage + w * loss_bbox,
    d_coverage, d_bboxes)."""
    arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in (coverage, bboxes) + tuple(tops)]
    cov, bb, cl, bl, sz, ob, cb = arrs
    N, Cn, gh, gw = cov.shape
    if bb.shape != (N, 4 * Cn, gh, gw) or cl.shape != cov.shape or any(a.shape != bb.shape for a in (bl, sz, ob, cb)):
        raise ValueError('train head: blob shapes do not match the train_val.prototxt contract')
    losses = np.zeros(2, np.float32)
    dc = np.zeros_like(cov) if want_grads else None
    db = np.zeros_like(bb) if want_grads else None
    _lib.check(_lib.lib().fcnod_train_head_forward(_lib.ptr(cov), _lib.ptr(bb), _lib.ptr(cl), _lib.ptr(bl), _lib.ptr(sz), _lib.ptr(ob),
                                                   _lib.ptr(cb), N, Cn, gh, gw, C.c_float(bbox_loss_weight), _lib.ptr(losses),
                                                   _lib.ptr(dc), _lib.ptr(db)))
    return dict(loss_bbox=float(losses[0]), loss_coverage=float(losses[1]),
                loss=float(losses[1]) + bbox_loss_weight * float(losses[0]), d_coverage=dc, d_bboxes=db)
