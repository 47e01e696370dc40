"""CPU restatement of the network forward -- TEST INFRASTRUCTURE ONLY (oracle).

PARITY UNPINNED: the upstream repository holds no golden vectors for this path
and Caffe cannot run here; this restates Caffe's layer
semantics for the layers used by /root/reference/models/deploy.prototxt:1-2178:
  Power   (shift + scale*x)^power                       prototxt :8-16
  Convolution cross-correlation, out=floor((H+2p-k)/s)+1 e.g.     :20-47
  ReLU in place                                                    :48-53
  Pooling MAX, out=ceil((H+2p-k)/s)+1, windows clipped             :54-64
  LRN across channels, k=1, x*(1+alpha/n*sum x^2)^-beta            :65-75
  Concat along channels, Dropout=identity at TEST, Sigmoid         :2100-2149
Call site restated: net.forward(), /root/reference/scripts/fcn_object_detector.py:87.
"""
import math
import numpy as np
import torch
import torch.nn.functional as F


def _caffe_pool_out(h, k, s, p):
    o = int(math.ceil((h + 2 * p - k) / float(s))) + 1
    if p > 0 and (o - 1) * s >= h + p:
        o -= 1
    return o


def max_pool_caffe(x, k, s, p):
    H, W = x.shape[-2:]
    oh, ow = _caffe_pool_out(H, k, s, p), _caffe_pool_out(W, k, s, p)
    # pad right/bottom with -inf so clipped windows ignore the padding
    need_h = (oh - 1) * s + k - (H + p)
    need_w = (ow - 1) * s + k - (W + p)
    xp = F.pad(x, (p, max(need_w, 0), p, max(need_h, 0)), value=float('-inf'))
    y = F.max_pool2d(xp, k, s)
    return y[..., :oh, :ow]


def forward(layers, weights, data, want=None, act_round=None):
    """layers: oracle.prototxt.layers(); weights: {name: [W, b]}; data: (N,3,H,W) float32.

    act_round: optional callable applied to every conv/pool/LRN output (used only
    to *predict* bf16 storage error; the oracle proper is act_round=None).
    Returns dict of all blobs (or only those named in `want`).
    """
    blobs = {'data': torch.as_tensor(np.asarray(data), dtype=torch.float32)}
    rnd = act_round or (lambda t: t)
    with torch.no_grad():
        for L in layers:
            t = L['type']
            b = [blobs[n] for n in L['bottom']]
            if t == 'Power':
                y = (L['shift'] + L['scale'] * b[0])
                if L['power'] != 1.0:
                    y = y ** L['power']
            elif t == 'Convolution':
                W = torch.as_tensor(weights[L['name']][0])
                bias = torch.as_tensor(weights[L['name']][1]) if L.get('bias_term', True) else None
                y = rnd(F.conv2d(b[0], W, bias, stride=L['stride'], padding=L['pad']))
            elif t == 'ReLU':
                y = rnd(torch.relu(b[0]))
            elif t == 'Pooling':
                assert L['pool'] == 'MAX'
                y = max_pool_caffe(b[0], L['kernel'], L['stride'], L['pad'])
            elif t == 'LRN':
                y = rnd(F.local_response_norm(b[0], L['local_size'], L['alpha'], L['beta'], L['k']))
            elif t == 'Concat':
                y = torch.cat(b, dim=1)
            elif t == 'Dropout':
                y = b[0]
            elif t == 'Sigmoid':
                y = torch.sigmoid(b[0])
            else:
                raise NotImplementedError(t)
            blobs[L['top'][0]] = y
    if want is not None:
        return {k: blobs[k].numpy() for k in want}
    return {k: v.numpy() for k, v in blobs.items()}
