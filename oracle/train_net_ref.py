"""CPU restatement of one training iteration of /root/reference/models/train_val.prototxt -- TEST INFRASTRUCTURE ONLY (oracle).

PARITY UNPINNED (no upstream vectors; Caffe / NVCaffe cannot run here).  What is restated, with its source:
  network body                       models/train_val.prototxt:27-35,76-2232 (the deploy body with 1 / 4 head channels)
  Dropout at TRAIN                   :2164-2172   y = x * mask / (1 - ratio), mask ~ Bernoulli(1 - ratio)   [ext: Caffe dropout_layer]
  label slicing, masks, the losses   :36-72,2237-2281   (oracle/train_ref.py)
  backward                           Caffe's layer-wise Backward [ext]: here torch autograd over the same fp32 forward graph
                                     (max-pool routes the gradient to the FIRST maximum of a window in row-major scan order,
                                     which is Caffe's `>` rule and torch's; Split layers sum the gradients of a blob's consumers)
  solver                             train/bounding_box/solver.prototxt:3-15: ADAM, base_lr 1e-4, lr_policy "step" (gamma 0.1,
                                     stepsize 10000), momentum 0.9, weight_decay 1e-7 (L2); momentum2 0.999 and delta 1e-8 are
                                     caffe.proto's defaults.  Per-blob lr_mult / decay_mult 1,1 (weights) and 2,0 (biases)
                                     (train_val.prototxt:82-89 and every other Convolution).
  Caffe's SGDSolver::ApplyUpdate [ext: sgd_solver.cpp, adam_solver.cpp]:
      rate = base_lr * gamma^floor(iter / stepsize)
      g   += weight_decay * decay_mult * w                                     (Regularize, L2)
      m    = b1 * m + (1 - b1) * g ;  v = b2 * v + (1 - b2) * g * g
      w   -= rate * lr_mult * sqrt(1 - b2^t) / (1 - b1^t) * m / (sqrt(v) + delta)        with t = iter + 1
  Multi-GPU (Caffe's NCCL solver [ext: parallel.cpp]): gradients are summed over the solvers and scaled by 1 / solver_count
  before ApplyUpdate -- `average_gradients` below.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

from oracle import net_ref

SOLVER = dict(base_lr=1e-4, gamma=0.1, stepsize=10000, momentum=0.9, momentum2=0.999, delta=1e-8, weight_decay=1e-7)
LR_MULT = (1.0, 2.0)
DECAY_MULT = (1.0, 0.0)


TRAIN_SIDE = ('Data', 'Slice', 'Eltwise', 'L1Loss', 'EuclideanLoss')


def body_layers(layers):
    """The network body of a train_val-form layer list (the train-side layers are restated in train_ref.py / below)."""
    return [L for L in layers if L['type'] not in TRAIN_SIDE]


def caffe_pool_argmax(x, k, s, p):
    """Caffe's MAX pooling forward on (N,C,H,W) numpy data [ext: pooling_layer.cpp]: windows [ph*s - p, min(ph*s - p + k, H + p))
    clipped to the blob, scanned row-major with `>` against -FLT_MAX, so the FIRST maximum wins.  Returns (y, idx) with idx =
    dh * k + dw relative to the unclipped window origin."""
    N, C, H, W = x.shape
    oh, ow = net_ref._caffe_pool_out(H, k, s, p), net_ref._caffe_pool_out(W, k, s, p)
    best = np.full((N, C, oh, ow), -np.finfo(np.float32).max, np.float32)
    idx = np.zeros((N, C, oh, ow), np.uint8)
    hh = np.arange(oh)[:, None] * s - p
    ww = np.arange(ow)[None, :] * s - p
    for dh in range(k):
        for dw in range(k):
            h, w = hh + dh, ww + dw
            ok = (h >= 0) & (h < min(H, H + p)) & (w >= 0) & (w < min(W, W + p)) & (h < hh + k) & (w < ww + k)
            v = x[:, :, np.clip(h, 0, H - 1), np.clip(w, 0, W - 1)]
            take = ok[None, None] & (v > best)
            best = np.where(take, v, best)
            idx = np.where(take, np.uint8(dh * k + dw), idx)
    return best, idx


def max_pool_routed(x, idx, k, s, p):
    """MAX pooling whose selection is GIVEN (idx as caffe_pool_argmax returns it): forward gathers, backward scatters to the
    same elements.  Used to compare gradients with the device's own routing: a near-tie inside a window may resolve to
    another element when the two forwards differ in the last bits, and the routing is then a different (equally valid)
    function of slightly different inputs."""
    N, C, H, W = x.shape
    idx = torch.as_tensor(np.asarray(idx).astype(np.int64))
    oh, ow = idx.shape[2:]
    h = torch.arange(oh).view(1, 1, oh, 1) * s - p + idx // k
    w = torch.arange(ow).view(1, 1, 1, ow) * s - p + idx % k
    flat = (h * W + w).reshape(N, C, oh * ow)
    return x.reshape(N, C, H * W).gather(2, flat).view(N, C, oh, ow)


def forward_backward(layers, weights, data, grids, dropout_mask=None, bbox_loss_weight=2.0, want=(), pool_argmax=None):
    """layers: oracle.prototxt.layers() of the network body; weights {name: [W, b]} float32; data (N,3,H,W) float32;
    grids = (cvg_label, bbox_label, size_block, obj_block, cvg_block) as the augmentation layer emits them;
    dropout_mask: (N,C,h,w) 0/1 array for the Dropout layer (None = no dropout, the TEST-phase identity).
    pool_argmax: optional {pooling layer name: idx} -- those layers use the given selection (max_pool_routed).
    Returns dict(loss_bbox, loss_coverage, grads {name: [dW, db]}, blobs {name: array} for `want`,
    blob_grads {name: array} for `want`)."""
    P = {}
    for name, wb in weights.items():
        P[name] = [torch.tensor(np.asarray(a, np.float32), requires_grad=True) for a in wb]
    blobs = {'data': torch.as_tensor(np.asarray(data, np.float32))}
    kept = {}
    for L in body_layers(layers):
        t = L['type']
        b = [blobs[n] for n in L['bottom']]
        if t == 'Power':
            y = L['shift'] + L['scale'] * b[0]
            assert L['power'] == 1.0
        elif t == 'Convolution':
            W, bias = P[L['name']]
            y = F.conv2d(b[0], W, bias, stride=L['stride'], padding=L['pad'])
        elif t == 'ReLU':
            y = torch.relu(b[0])
        elif t == 'Pooling':
            if pool_argmax is not None and L['name'] in pool_argmax:
                y = max_pool_routed(b[0], pool_argmax[L['name']], L['kernel'], L['stride'], L['pad'])
            else:
                y = net_ref.max_pool_caffe(b[0], L['kernel'], L['stride'], L['pad'])
        elif t == 'LRN':
            y = F.local_response_norm(b[0], L['local_size'], L['alpha'], L['beta'], L['k'])
        elif t == 'Concat':
            y = torch.cat(b, dim=1)
        elif t == 'Dropout':
            if dropout_mask is None:
                y = b[0]
            else:
                scale = np.float32(1.0) / (np.float32(1.0) - np.float32(L['dropout_ratio']))
                y = b[0] * torch.as_tensor(np.asarray(dropout_mask, np.float32)) * float(scale)
        elif t == 'Sigmoid':
            y = torch.sigmoid(b[0])
        else:
            raise NotImplementedError(t)
        blobs[L['top'][0]] = y
        if L['top'][0] in want and y.requires_grad:
            y.retain_grad()
            kept[L['top'][0]] = y
    cvg_label, bbox_label, size_block, obj_block, cvg_block = [torch.as_tensor(np.asarray(g, np.float32)) for g in grids]
    n = data.shape[0]
    target = (bbox_label * size_block) * obj_block
    pred = ((blobs['bboxes'] * cvg_block) * size_block) * obj_block
    loss_bbox = (pred - target).abs().double().sum() / n
    cdiff = blobs['coverage'] - cvg_label
    loss_cvg = (cdiff * cdiff).double().sum() / (2.0 * n)
    (loss_cvg + bbox_loss_weight * loss_bbox).backward()
    grads = {name: [p.grad.numpy().copy() if p.grad is not None else np.zeros(p.shape, np.float32) for p in wb]
             for name, wb in P.items()}
    return dict(loss_bbox=float(loss_bbox.detach()), loss_coverage=float(loss_cvg.detach()), grads=grads,
                blobs={k: blobs[k].detach().numpy() for k in want if k in blobs},
                blob_grads={k: v.grad.numpy().copy() for k, v in kept.items() if v.grad is not None})


def learning_rate(it, solver=SOLVER):
    return solver['base_lr'] * math.pow(solver['gamma'], it // solver['stepsize'])


def multipliers(layers):
    """{conv name: (lr_mult [w, b], decay_mult [w, b])} from the `param` records of the prototxt (one upstream layer,
    inception_5b/3x3_reduce, has a bias lr_mult of 1.0 instead of 2.0)."""
    return {L['name']: (L['lr_mult'], L['decay_mult']) for L in layers if L['type'] == 'Convolution'}


def adam_update(weights, grads, state, it, solver=SOLVER, mults=None):
    """One ApplyUpdate in float32, in Caffe's order of operations.  weights / grads {name: [W, b]}; state {name: [[m, v], [m, v]]}
    (created on first use); `it` = solver iteration BEFORE this update (0 for the first); mults = multipliers(layers)
    (default: 1,1 for weights and 2,0 for biases everywhere).  Returns the new weights."""
    f = np.float32
    b1, b2, eps = f(solver['momentum']), f(solver['momentum2']), f(solver['delta'])
    t = it + 1
    correction = f(math.sqrt(1.0 - math.pow(float(b2), t)) / (1.0 - math.pow(float(b1), t)))
    rate = f(learning_rate(it, solver))
    out = {}
    for name, wb in weights.items():
        st = state.setdefault(name, [[np.zeros_like(a, dtype=f), np.zeros_like(a, dtype=f)] for a in wb])
        new = []
        lr_mult, decay_mult = mults[name] if mults and name in mults else (LR_MULT, DECAY_MULT)
        for i, w in enumerate(wb):
            w = np.asarray(w, f)
            g = np.asarray(grads[name][i], f) + f(solver['weight_decay'] * decay_mult[i]) * w
            m, v = st[i]
            m[...] = b1 * m + (f(1) - b1) * g
            v[...] = b2 * v + (f(1) - b2) * (g * g)
            step = (rate * f(lr_mult[i]) * correction) * (m / (np.sqrt(v) + eps))
            new.append((w - step).astype(f))
        out[name] = new
    return out


def average_gradients(per_rank_grads):
    """Caffe's NCCL solver: all-reduce SUM, then scale by 1 / solver_count."""
    n = len(per_rank_grads)
    out = {}
    for name in per_rank_grads[0]:
        out[name] = [sum(np.asarray(g[name][i], np.float64) for g in per_rank_grads).astype(np.float32) / np.float32(n)
                     for i in range(len(per_rank_grads[0][name]))]
    return out
