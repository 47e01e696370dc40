"""CPU: pins for the oracle that do not come from the oracle itself.

The reference holds no golden vectors for this path and cannot run here, so `oracle/` stays "parity
unpinned" at the upstream boundary.  What can be done is to stop it from being a single reading:

  1. KNOWN-ANSWER vectors derived by hand from Caffe's documented layer semantics (LRN's alpha/n scaling with a clipped
     channel window, ceil-mode pooling with clipped windows, the `pad > 0` last-window rule, cross-correlation) and from
     `JaccardCoeff.iou`'s text (argumentation_engine.py:24-55) at its edge cases -- the derivation is in each test;
  2. a SECOND, independent implementation of conv / pool / LRN in plain numpy, written from Caffe's loop structure
     (pooling_layer.cpp's hstart/hend clipping, lrn_layer.cpp's running window), cross-checked against
     `oracle.net_ref.forward` (torch) on the deploy net's stem;
  3. a FUZZ of `oracle.layer_ref.crop_window` / `flip_rect` against a line-for-line Python-3 transliteration of
     argumentation_engine.py:157-173,193-236,241-267 whose `/` is Python 2's (floor on two ints, decided at run time by
     the operand types) and whose `random.*` calls replay a script -- the oracle decides statically where `//` applies,
     the transliteration does not;
  4. the TRAINING HEAD (train_val.prototxt:36-72,2237-2281): a hand-derived known answer for both losses and their
     gradients, central differences of the loss against the analytic gradients, and the autograd oracle
     (train_net_ref) against the numpy head (train_ref) through a two-layer network with the weight gradients summed by hand;
  5. the BACKWARD rules the training oracle leans on: Caffe's first-maximum routing of the pooling gradient (known answer
     with ties, and the positions against a per-window loop), the LRN backward formula of lrn_layer.cpp in float64 loops.
"""
import math
import random

import numpy as np
import pytest

from oracle import layer_ref, net_ref


# ============================================================================ 1. known answers, derived by hand
def _run(layers, weights, x):
    return net_ref.forward(layers, weights, x)


def test_known_answer_lrn_alpha_over_n_and_clipped_window():
    """Caffe LRN, ACROSS_CHANNELS: scale_c = k + (alpha / n) * sum_{c' in [c - n/2, c + n/2] within the blob} x_c'^2,
    y_c = x_c * scale_c^-beta.  The divisor is always n (local_size), also where the window is clipped.
    One pixel, channels x = [1, 2, 3, 4, 5, 6], n = 5, alpha = 1, beta = 0.75, k = 1:
      c = 0: window {0,1,2}      sum = 1 + 4 + 9            = 14   scale = 1 + 14/5 = 3.8
      c = 1: window {0,1,2,3}    sum = 14 + 16              = 30   scale = 7.0
      c = 2: window {0..4}       sum = 30 + 25              = 55   scale = 12.0
      c = 3: window {1..5}       sum = 4+9+16+25+36         = 90   scale = 19.0
      c = 4: window {2..5}       sum = 9+16+25+36           = 86   scale = 18.2
      c = 5: window {3,4,5}      sum = 16+25+36             = 77   scale = 16.4"""
    x = np.arange(1, 7, dtype=np.float32).reshape(1, 6, 1, 1)
    L = [dict(name='n', type='LRN', bottom=['data'], top=['y'], local_size=5, alpha=1.0, beta=0.75, k=1.0)]
    y = _run(L, {}, x)['y'].reshape(-1)
    scale = [3.8, 7.0, 12.0, 19.0, 18.2, 16.4]
    want = [(c + 1) * math.pow(s, -0.75) for c, s in enumerate(scale)]
    assert np.allclose(y, want, rtol=2e-6, atol=0)
    # the deploy net's constants (alpha = 1e-4): 1 + 1e-4/5 * 55 = 1.0011
    L[0]['alpha'] = 1e-4
    y = _run(L, {}, x)['y'].reshape(-1)
    assert abs(y[2] - 3 * math.pow(1.0011, -0.75)) < 1e-6


def test_known_answer_pooling_ceil_mode_clipped_windows_and_pad_rule():
    """Caffe Pooling: pooled = ceil((H + 2p - k) / s) + 1; if p > 0 and (pooled - 1) * s >= H + p: pooled -= 1; a window is
    [ph*s - p, min(ph*s - p + k, H + p)) then clipped to [0, H): padding never wins a MAX.
      H = 6, k = 3, s = 2, p = 0: pooled = ceil(3/2) + 1 = 3; windows rows {0,1,2} {2,3,4} {4,5}   (the last one clipped)
      H = 5, k = 3, s = 2, p = 0: pooled = 2
      H = 3, k = 2, s = 2, p = 1: ceil(3/2) + 1 = 3, (3-1)*2 = 4 >= 3+1 -> 2; windows rows {0} {1,2}
      H = 4, k = 3, s = 1, p = 1 (the inception pools): pooled = 4; windows {0,1} {0,1,2} {1,2,3} {2,3}"""
    def pool(x, k, s, p):
        # the column vector repeated along the width (rows constant): the first output column is the 1-d answer
        L = [dict(name='p', type='Pooling', bottom=['data'], top=['y'], pool='MAX', kernel=k, stride=s, pad=p)]
        a = np.repeat(np.asarray(x, np.float32)[:, None], len(x), axis=1)
        return _run(L, {}, a.reshape(1, 1, len(x), len(x)))['y'][0, 0, :, 0]
    assert list(pool([5, 1, 2, 9, 3, 4], 3, 2, 0)) == [5, 9, 4]
    assert list(pool([5, 1, 2, 9, 3], 3, 2, 0)) == [5, 9]
    assert list(pool([-7, -2, -3], 2, 2, 1)) == [-7, -2]            # all negative: a zero pad would have won
    assert list(pool([1, -4, -6, -5], 3, 1, 1)) == [1, 1, -4, -5]
    # and the geometry the deploy net relies on: 224 -> 112 -> 56 -> 28 with 3x3/2 ceil-mode pools
    assert [net_ref._caffe_pool_out(h, 3, 2, 0) for h in (224, 112, 56)] == [112, 56, 28]


def test_known_answer_convolution_is_cross_correlation_with_floor_output():
    """Caffe Convolution: y[o, i, j] = b[o] + sum_c,r,s w[o, c, r, s] * x[c, i*st - p + r, j*st - p + s] (no kernel flip),
    out = floor((H + 2p - k) / st) + 1.  x = [[1,2,3],[4,5,6],[7,8,9]], w = [[1,2],[3,4]], b = 10:
      y[0,0] = 10 + 1*1 + 2*2 + 3*4 + 4*5 = 47     y[0,1] = 10 + 2 + 6 + 15 + 24 = 57
      y[1,0] = 10 + 4 + 10 + 21 + 32 = 77          y[1,1] = 10 + 5 + 12 + 24 + 36 = 87
    stride 2 on H = 4, k = 3, p = 0: floor(1/2) + 1 = 1 output (Caffe floors where pooling ceils)."""
    x = np.arange(1, 10, dtype=np.float32).reshape(1, 1, 3, 3)
    L = [dict(name='c', type='Convolution', bottom=['data'], top=['y'], stride=1, pad=0)]
    W = {'c': [np.asarray([[[[1, 2], [3, 4]]]], np.float32), np.asarray([10], np.float32)]}
    assert _run(L, W, x)['y'].reshape(-1).tolist() == [47, 57, 77, 87]
    L[0]['stride'] = 2
    W3 = {'c': [np.ones((1, 1, 3, 3), np.float32), np.zeros(1, np.float32)]}
    assert _run(L, W3, np.ones((1, 1, 4, 4), np.float32))['y'].shape == (1, 1, 1, 1)


def test_known_answer_jaccard_iou_edge_cases():
    """JaccardCoeff.iou (argumentation_engine.py:26-55): intersection [ix, iy, iw, ih] is rejected only when iw < 0 or
    ih < 0; the "union" is the BOUNDING BOX of both rects; score = (area(a^b) / area(aUb)) / (area(a) / area(b)) in fp32.
    cell a = [0, 0, 16, 16]:
      b = [16, 0, 16, 16] (touching on an edge): iw = min(16, 32) - 16 = 0 -> not rejected, area 0 -> score 0 -> label 0
      b = [0, 0, 16, 16]  (identical): (256/256) / (256/256) = 1 -> label 1
      b = [8, 8, 16, 16]: inter 8x8 = 64, bounding box 24x24 = 576, ratio 256/256: score = 64/576 = 0.111.. > 0.1 -> 1
      b = [9, 8, 16, 16]: inter 7x8 = 56, bounding box 25x24 = 600: score = 0.0933 -> 0
      b = [0, 0, 160, 160] (big rect over the cell): inter 256, box 25600, ratio 256/25600: score = 0.01/0.01 = 1 -> 1
      b = [500, 500, 10, 10] (outside): iw < 0 -> 0
      b = [4, 4, 0, 0] (zero area): inter 0x0 -> area(b) = 0 -> 0/0... the reference divides by area(a)/area(b) = inf:
          (0 / 256) / inf = 0 -> label 0 (no exception: numpy float32 division)"""
    a = [0, 0, 16, 16]
    assert layer_ref.iou(a, [16, 0, 16, 16]) == 0.0
    assert layer_ref.iou(a, [0, 0, 16, 16]) == 1.0
    s = layer_ref.iou(a, [8, 8, 16, 16])
    assert abs(float(s) - 64.0 / 576.0) < 1e-7 and float(s) > 0.1
    assert abs(float(layer_ref.iou(a, [9, 8, 16, 16])) - 56.0 / 600.0) < 1e-7
    assert abs(float(layer_ref.iou(a, [0, 0, 160, 160])) - 1.0) < 1e-6
    assert layer_ref.iou(a, [500, 500, 10, 10]) == 0.0
    with np.errstate(divide='ignore', invalid='ignore'):
        z = layer_ref.iou(a, [4, 4, 0, 0])
    assert float(z) == 0.0
    # and through the label grids: a rect outside the canvas labels nothing, a rect on the canvas edge labels its cells
    fg = layer_ref.bounding_box_parameterized_labels(448, 448, 16, 1, [[500, 500, 40, 40]], [0])[0]
    assert fg.sum() == 0
    fg = layer_ref.bounding_box_parameterized_labels(448, 448, 16, 1, [[432, 432, 16, 16]], [0])[0]
    assert fg[0, 27, 27] == 1 and fg.sum() == 1


# ============================================================================ 2. independent numpy conv / pool / LRN
def np_conv(x, w, b, stride, pad):
    """Direct sum over taps (Caffe's im2col + GEMM computes the same sums): float64 accumulate."""
    N, C, H, W = x.shape
    O, _, k, _ = w.shape
    oh, ow = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    xp = np.zeros((N, C, H + 2 * pad, W + 2 * pad), np.float64)
    xp[:, :, pad:pad + H, pad:pad + W] = x
    y = np.zeros((N, O, oh, ow), np.float64)
    for r in range(k):
        for s in range(k):
            patch = xp[:, :, r:r + stride * (oh - 1) + 1:stride, s:s + stride * (ow - 1) + 1:stride]
            y += np.einsum('nchw,oc->nohw', patch, w[:, :, r, s].astype(np.float64))
    return y + b.reshape(1, -1, 1, 1)


def np_pool_caffe(x, k, s, p):
    """pooling_layer.cpp, MAX: per output the window [hstart, hend) is built from the PADDED extent and then clipped."""
    N, C, H, W = x.shape

    def out(h):
        o = int(math.ceil((h + 2 * p - k) / float(s))) + 1
        if p > 0 and (o - 1) * s >= h + p:
            o -= 1
        return o
    oh, ow = out(H), out(W)
    y = np.empty((N, C, oh, ow), x.dtype)
    for ph in range(oh):
        hs = ph * s - p; he = min(hs + k, H + p); hs = max(hs, 0); he = min(he, H)
        for pw in range(ow):
            ws = pw * s - p; we = min(ws + k, W + p); ws = max(ws, 0); we = min(we, W)
            y[:, :, ph, pw] = x[:, :, hs:he, ws:we].max(axis=(2, 3))
    return y


def np_lrn(x, n, alpha, beta, k):
    """lrn_layer.cpp CrossChannelForward: scale = k + alpha / n * (sum of squares over the +-n/2 channel window)."""
    N, C, H, W = x.shape
    sq = np.zeros((N, C + n - 1, H, W), np.float64)
    sq[:, n // 2:n // 2 + C] = x.astype(np.float64) ** 2
    scale = np.full(x.shape, k, np.float64)
    for c in range(C):
        scale[:, c] += alpha / n * sq[:, c:c + n].sum(axis=1)
    return x * scale ** -beta


def test_independent_numpy_layers_agree_with_the_oracle_on_the_stem(oracle_net):
    """The deploy net's stem (deploy.prototxt:8-162: Power, conv1 7x7/2, ReLU, pool1, norm1, conv2 reduce, conv2 3x3, norm2,
    pool2) through the numpy layers above and through oracle.net_ref (torch): two implementations, one answer."""
    layers, weights = oracle_net
    stem = []
    for L in layers:
        stem.append(L)
        if L['name'] == 'pool2/3x3_s2':
            break
    x = np.random.RandomState(6).uniform(0, 1, (2, 3, 96, 80)).astype(np.float32)      # odd pooled sizes on purpose
    ref = net_ref.forward(stem, weights, x)
    blobs = {'data': x.astype(np.float64)}
    for L in stem:
        b = blobs[L['bottom'][0]]
        t = L['type']
        if t == 'Power':
            y = L['shift'] + L['scale'] * b
        elif t == 'Convolution':
            y = np_conv(b, weights[L['name']][0], weights[L['name']][1], L['stride'], L['pad'])
        elif t == 'ReLU':
            y = np.maximum(b, 0)
        elif t == 'Pooling':
            y = np_pool_caffe(b, L['kernel'], L['stride'], L['pad'])
        elif t == 'LRN':
            y = np_lrn(b, L['local_size'], L['alpha'], L['beta'], L['k'])
        else:
            raise AssertionError(t)
        blobs[L['top'][0]] = y
    for name in ('conv1/7x7_s2', 'pool1/3x3_s2', 'pool1/norm1', 'conv2/3x3', 'conv2/norm2', 'pool2/3x3_s2'):
        a, r = blobs[name], ref[name]
        assert a.shape == r.shape, name
        # conv1 sees x - 127: sums of ~147 terms of magnitude 127 * |w| cancel -> fp32 torch vs fp64 numpy differ by ~1e-5 rel
        assert np.abs(a - r).max() / np.abs(a).max() < 2e-5, (name, np.abs(a - r).max() / np.abs(a).max())
    assert blobs['pool2/3x3_s2'].shape == (2, 192, 12, 10)


# ============================================================================ 3. Python-2 transliteration fuzz
def py2div(a, b):
    """Python 2's `/`: floor division when both operands are integers, true division otherwise."""
    ints = (int, np.integer)
    if isinstance(a, ints) and isinstance(b, ints) and not isinstance(a, bool) and not isinstance(b, bool):
        return a // b
    return a / b


class Script(object):
    """Stands in for the `random` module: replays pre-drawn values in call order, checking the requested bounds."""

    def __init__(self, values):
        self.values = list(values)
        self.bounds = []

    def randint(self, a, b):
        v = self.values.pop(0)
        self.bounds.append((a, b))
        assert a <= v <= b, (a, v, b)
        return v


def t_flip_rect(image_shape, rect, flip_flag):
    """argumentation_engine.py:245-265, the rect part of flip_image (one rect)."""
    im_y, im_x = image_shape[0], image_shape[1]
    pt1 = (rect[0], rect[1])
    pt2 = (rect[0] + rect[2], rect[1] + rect[3])
    if flip_flag == -1:
        pt1 = (im_x - pt1[0] - 1, im_y - pt1[1] - 1)
        pt2 = (im_x - pt2[0] - 1, im_y - pt2[1] - 1)
    elif flip_flag == 0:
        pt1 = (pt1[0], im_y - pt1[1] - 1)
        pt2 = (pt2[0], im_y - pt2[1] - 1)
    elif flip_flag == 1:
        pt1 = (im_x - pt1[0] - 1, pt1[1])
        pt2 = (im_x - pt2[0] - 1, pt2[1])
    x = min(pt1[0], pt2[0])
    y = min(pt1[1], pt2[1])
    w = np.abs(pt2[0] - pt1[0])
    h = np.abs(pt2[1] - pt1[1])
    x = 0 if x < 0 else x
    y = 0 if y < 0 else y
    return [x, y, w, h]


def t_crop_image_dimension(image_shape, rect, widths, heights, rnd):
    """argumentation_engine.py:193-236, statement for statement, `/` -> py2div, `random` -> rnd; returns the slice the
    reference takes (`image[int(y):int(y+h), int(x):int(x+w)]`) and new_rect."""
    x = (rect[0] + py2div(rect[2], 2)) - widths[0]
    y = (rect[1] + py2div(rect[3], 2)) - heights[0]
    w = widths[1] + widths[0]
    h = heights[1] + heights[0]
    cx, cy = (rect[0] + py2div(rect[2], 2.0), rect[1] + py2div(rect[3], 2.0))
    shift_x, shift_y = (rnd.randint(0, int(py2div(w, 2))), rnd.randint(0, int(py2div(h, 2))))
    cx = (cx + shift_x) if rnd.randint(0, 1) else (cx - shift_x)
    cy = (cy + shift_y) if rnd.randint(0, 1) else (cy - shift_y)
    nx = int(cx - (py2div(w, 2)))
    ny = int(cy - (py2div(h, 2)))
    nw = int(w)
    nh = int(h)
    if nx > x:
        nx = x
        nw -= np.abs(nx - x)
    if ny > y:
        ny = y
        nh -= np.abs(ny - y)
    if nx + nw < x + w:
        nx += ((x + w) - (nx + nw))
    if ny + nh < y + h:
        ny += ((y + h) - (ny + nh))
    x = nx; y = ny; w = nw; h = nh
    x = 0 if x < 0 else x
    y = 0 if y < 0 else y
    w = ((w - (w + x) - image_shape[1])) if x > image_shape[1] else w
    h = ((h - (h + y) - image_shape[0])) if y > image_shape[0] else h
    sl = (slice(int(y), int(y + h)), slice(int(x), int(x + w)))
    new_rect = [int(rect[0] - x), int(rect[1] - y), rect[2], rect[3]]
    return sl, new_rect


def test_crop_and_flip_fuzz_against_python2_transliteration():
    rs = np.random.RandomState(12)
    checked = 0
    for it in range(4000):
        Hs, Ws = int(rs.randint(120, 700)), int(rs.randint(120, 900))
        w, h = int(rs.randint(8, min(Ws, 400))), int(rs.randint(8, min(Hs, 400)))
        rect = [int(rs.randint(0, Ws - w)), int(rs.randint(0, Hs - h)), w, h]
        if it % 3 == 0:      # numpy integer rects, as they arrive from the label blob
            rect = [np.int64(v) for v in rect]
        draw = layer_ref.draw_params(random.Random(it), Ws, Hs, [int(v) for v in rect])
        flag = draw['flip']
        # ---- transliteration: random_argumentation :146-173 with the scripted draws
        rect_flip = t_flip_rect((Hs, Ws, 3), rect, flag) if -2 < flag < 2 else list(rect)
        f1, f2 = draw['f1'], draw['f2']
        widths = (int(rect_flip[2] * f1), rect_flip[2] * f2)
        heights = (int(rect_flip[3] * f1), rect_flip[3] * f2)
        script = Script([draw['shift_x'], draw['shift_y'], draw['sign_x'], draw['sign_y']])
        sl, new_rect = t_crop_image_dimension((Hs, Ws, 3), rect_flip, widths, heights, script)
        # ---- oracle
        r = layer_ref.flip_rect([int(v) for v in rect], Ws, Hs, flag) if -2 < flag < 2 else [int(v) for v in rect]
        assert [int(v) for v in rect_flip] == [int(v) for v in r]
        (x, y, cw, ch), orect = layer_ref.crop_window(Ws, Hs, r, f1, f2, draw['shift_x'], draw['shift_y'], draw['sign_x'], draw['sign_y'])
        assert (int(y), int(y + ch), int(x), int(x + cw)) == (sl[0].start, sl[0].stop, sl[1].start, sl[1].stop), (it, rect, draw)
        assert [int(v) for v in orect] == [int(v) for v in new_rect]
        # the randint bounds the transliteration asked for are the ones the oracle's draw_params used
        bx, by = layer_ref.crop_bounds(r, f1, f2)
        assert script.bounds[0] == (0, bx) and script.bounds[1] == (0, by)
        checked += 1
    assert checked == 4000


# ============================================================================ 4. the training head (train_val.prototxt:36-72,2237-2281)
def _head_case():
    """Two frames, one grid cell, one class.  Frame 0 is a foreground cell, frame 1 a background cell (coverage-block 0:
    its bbox prediction is masked out and its labels are zero)."""
    f = np.float32
    coverage = np.array([0.8, 0.3], f).reshape(2, 1, 1, 1)
    cvg_label = np.array([1.0, 0.0], f).reshape(2, 1, 1, 1)
    bboxes = np.array([[10, -4, 6, 2], [7, 7, 7, 7]], f).reshape(2, 4, 1, 1)
    bbox_label = np.array([[8, -8, 6, 4], [0, 0, 0, 0]], f).reshape(2, 4, 1, 1)
    size_block = np.array([[0.1, 0.05, 0.1, 0.05], [0, 0, 0, 0]], f).reshape(2, 4, 1, 1)
    obj_block = np.array([[0.5] * 4, [0] * 4], f).reshape(2, 4, 1, 1)
    cvg_block = np.array([[1] * 4, [0] * 4], f).reshape(2, 4, 1, 1)
    return coverage, bboxes, cvg_label, bbox_label, size_block, obj_block, cvg_block


def test_known_answer_training_head_losses_and_gradients():
    """Derivation (Caffe's EuclideanLoss: sum (a-b)^2 / (2 num); NVCaffe's L1Loss: sum |a-b| / num; loss_weight 2 on the latter;
    Eltwise PROD chains of train_val.prototxt:53-72,2237-2266):
      coverage loss = ((0.8-1)^2 + (0.3-0)^2) / (2*2)                     = (0.04 + 0.09) / 4 = 0.0325
      frame 0 pred   = bboxes * 1 * size * obj = [10,-4,6,2] * [.1,.05,.1,.05] * .5 = [0.5, -0.1, 0.3, 0.05]
      frame 0 target = label * size * obj      = [8,-8,6,4]  * [.1,.05,.1,.05] * .5 = [0.4, -0.2, 0.3, 0.10]
      |pred - target| = [0.1, 0.1, 0, 0.05] -> 0.25; frame 1 is masked -> 0;  bbox loss = 0.25 / 2 = 0.125
      total = 0.0325 + 2 * 0.125 = 0.2825
      d coverage = (a - b) / num = [-0.1, 0.15]
      d bboxes (frame 0) = 2 * sign(pred - target) / num * obj * size * mask = [1, 1, 0, -1] * 0.5 * [.1,.05,.1,.05]
                         = [0.05, 0.025, 0, -0.025]   (sign(0) = 0, as caffe_cpu_sign gives);  frame 1: 0"""
    from oracle import train_ref
    out = train_ref.head_forward(*_head_case())
    assert abs(out['loss_coverage'] - 0.0325) < 1e-7 and abs(out['loss_bbox'] - 0.125) < 1e-7 and abs(out['loss'] - 0.2825) < 2e-7
    assert np.allclose(out['d_coverage'].reshape(-1), [-0.1, 0.15], atol=1e-7)
    assert np.allclose(out['d_bboxes'][0].reshape(-1), [0.05, 0.025, 0.0, -0.025], atol=1e-8)
    assert np.array_equal(out['d_bboxes'][1], np.zeros((4, 1, 1), np.float32))


def test_training_head_gradients_are_the_derivative_of_its_loss():
    """Central differences of the head's own loss (piecewise linear + quadratic: exact away from the kinks) against the
    analytic gradients it returns -- independent of the backward formulas of the module header."""
    from oracle import train_ref
    rs = np.random.RandomState(4)
    N, C, G = 3, 1, 4
    f = np.float32
    coverage = rs.uniform(0.05, 0.95, (N, C, G, G)).astype(f)
    bboxes = rs.uniform(-30, 30, (N, 4 * C, G, G)).astype(f)
    cvg_label = (rs.uniform(0, 1, (N, C, G, G)) > 0.5).astype(f)
    blk = np.repeat(cvg_label, 4, axis=1)
    bbox_label = (rs.uniform(-30, 30, (N, 4 * C, G, G)) * blk).astype(f)
    size_block = (blk * rs.uniform(0.01, 0.1, (N, 4 * C, G, G))).astype(f)
    obj_block = (blk * 0.25).astype(f)
    args = [coverage, bboxes, cvg_label, bbox_label, size_block, obj_block, blk]
    out = train_ref.head_forward(*args)

    def loss_with(i, arr):
        a = list(args)
        a[i] = arr
        return train_ref.head_forward(*a)['loss']

    for i, key, eps in ((0, 'd_coverage', 1e-2), (1, 'd_bboxes', 0.5)):
        base = args[i]
        for idx in [tuple(rs.randint(0, s) for s in base.shape) for _ in range(40)]:
            if i == 1:      # stay off the |.| kink: the step must not cross pred == target
                d = ((base[idx] - bbox_label[idx]) * size_block[idx] * obj_block[idx])
                if blk[idx] and abs(d) <= 2 * eps * size_block[idx] * obj_block[idx]:
                    continue
            p, m = base.astype(np.float64).copy(), base.astype(np.float64).copy()
            p[idx] += eps
            m[idx] -= eps
            num = (loss_with(i, p.astype(f)) - loss_with(i, m.astype(f))) / (float(p.astype(f)[idx]) - float(m.astype(f)[idx]))
            assert abs(num - out[key][idx]) < 2e-4 * max(1.0, abs(num)), (key, idx, num, out[key][idx])


def test_autograd_oracle_and_numpy_head_agree_through_a_two_layer_network():
    """oracle/train_net_ref.py (torch autograd over the restated graph) against oracle/train_ref.py (numpy, explicit backward)
    on a network small enough to follow by hand: data -> 1x1 conv -> Sigmoid = coverage, data -> 1x1 conv = bboxes.  Losses and
    the gradients reaching coverage / bboxes must agree, and the convolution's weight gradient must be the plain sum
    dW[o][c] = sum_{n,h,w} dY[n][o][h][w] * X[n][c][h][w] of that gradient (db = sum dY)."""
    from oracle import train_net_ref, train_ref
    rs = np.random.RandomState(6)
    N, G = 2, 3
    f = np.float32
    layers = [dict(name='cvg', type='Convolution', bottom=['data'], top=['cvg_raw'], stride=1, pad=0, kernel=1),
              dict(name='sig', type='Sigmoid', bottom=['cvg_raw'], top=['coverage']),
              dict(name='box', type='Convolution', bottom=['data'], top=['bboxes'], stride=1, pad=0, kernel=1)]
    weights = {'cvg': [rs.normal(0, 0.5, (1, 3, 1, 1)).astype(f), np.zeros(1, f)],
               'box': [rs.normal(0, 2.0, (4, 3, 1, 1)).astype(f), rs.normal(0, 1, 4).astype(f)]}
    data = rs.uniform(-1, 1, (N, 3, G, G)).astype(f)
    cvg_label = (rs.uniform(0, 1, (N, 1, G, G)) > 0.4).astype(f)
    blk = np.repeat(cvg_label, 4, axis=1)
    grids = (cvg_label, (rs.uniform(-3, 3, (N, 4, G, G)) * blk).astype(f), (blk * 0.125).astype(f), (blk * 0.5).astype(f), blk)
    ref = train_net_ref.forward_backward(layers, weights, data, grids, want=('coverage', 'bboxes'))
    head = train_ref.head_forward(ref['blobs']['coverage'], ref['blobs']['bboxes'], *grids)
    assert abs(ref['loss_bbox'] - head['loss_bbox']) < 1e-6 and abs(ref['loss_coverage'] - head['loss_coverage']) < 1e-6
    assert np.allclose(ref['blob_grads']['coverage'], head['d_coverage'], atol=1e-7)
    assert np.allclose(ref['blob_grads']['bboxes'], head['d_bboxes'], atol=1e-7)
    dW = np.einsum('nohw,nchw->oc', head['d_bboxes'].astype(np.float64), data.astype(np.float64))
    assert np.allclose(ref['grads']['box'][0].reshape(4, 3), dW, atol=1e-6)
    assert np.allclose(ref['grads']['box'][1], head['d_bboxes'].astype(np.float64).sum(axis=(0, 2, 3)), atol=1e-6)
    # through the sigmoid: d raw = d coverage * y * (1 - y)
    y = ref['blobs']['coverage'].astype(np.float64)
    d_raw = head['d_coverage'] * y * (1 - y)
    assert np.allclose(ref['grads']['cvg'][0].reshape(1, 3), np.einsum('nohw,nchw->oc', d_raw, data.astype(np.float64)), atol=1e-6)


# ============================================================================ 5. max-pool routing and LRN backward of the training oracle
def test_known_answer_first_maximum_routing_of_the_pooling_backward():
    """Caffe's MAX pooling (pooling_layer.cpp) scans a window row-major and replaces the running maximum only on `>`, so
    among equal values the FIRST one keeps the gradient.  5x5 input, 3x3 / 2 windows, no pad, ceil mode -> 2x2 outputs with
    window origins (0,0), (0,2), (2,0), (2,2).
      constant input: every window selects its own origin -> position 0; with top gradient [[1,2],[3,4]] the bottom gradient
      is 1 at (0,0), 2 at (0,2), 3 at (2,0), 4 at (2,2), zero elsewhere.
      a 9 planted at (2,2) (shared by all four windows): positions 8, 6, 2, 0 (= dh*3 + dw from each origin); its gradient is
      1 + 2 + 3 + 4 = 10."""
    import torch
    from oracle import train_net_ref
    x = np.ones((1, 1, 5, 5), np.float32)
    y, idx = train_net_ref.caffe_pool_argmax(x, 3, 2, 0)
    assert y.shape == (1, 1, 2, 2) and np.array_equal(idx.reshape(-1), [0, 0, 0, 0])
    top = torch.tensor([[1.0, 2.0], [3.0, 4.0]]).view(1, 1, 2, 2)
    xt = torch.tensor(x, requires_grad=True)
    (train_net_ref.max_pool_routed(xt, idx, 3, 2, 0) * top).sum().backward()
    want = np.zeros((5, 5), np.float32)
    want[0, 0], want[0, 2], want[2, 0], want[2, 2] = 1, 2, 3, 4
    assert np.array_equal(xt.grad.numpy()[0, 0], want)
    x[0, 0, 2, 2] = 9
    y, idx = train_net_ref.caffe_pool_argmax(x, 3, 2, 0)
    assert np.array_equal(y.reshape(-1), [9, 9, 9, 9]) and np.array_equal(idx.reshape(-1), [8, 6, 2, 0])
    xt = torch.tensor(x, requires_grad=True)
    out = train_net_ref.max_pool_routed(xt, idx, 3, 2, 0)
    assert np.array_equal(out.detach().numpy().reshape(-1), [9, 9, 9, 9])
    (out * top).sum().backward()
    want = np.zeros((5, 5), np.float32)
    want[2, 2] = 10
    assert np.array_equal(xt.grad.numpy()[0, 0], want)


def test_pool_argmax_agrees_with_the_forward_oracle_and_with_a_loop_restatement():
    """caffe_pool_argmax's values equal net_ref.max_pool_caffe's on the three pooling geometries of the network (3x3/2 pad 0
    ceil mode, 3x3/1 pad 1), and its positions equal a plain per-window loop with Caffe's clipping and `>` rule, ties included
    (inputs quantised to a few levels so that most windows hold ties)."""
    import torch
    from oracle import train_net_ref
    rs = np.random.RandomState(8)
    for (k, s, p, H, W) in ((3, 2, 0, 9, 12), (3, 2, 0, 8, 7), (3, 1, 1, 6, 5)):
        x = rs.randint(0, 4, (2, 3, H, W)).astype(np.float32)
        y, idx = train_net_ref.caffe_pool_argmax(x, k, s, p)
        assert np.array_equal(y, net_ref.max_pool_caffe(torch.tensor(x), k, s, p).numpy())
        oh, ow = y.shape[2:]
        for n in range(2):
            for c in range(3):
                for ph in range(oh):
                    for pw in range(ow):
                        hs, ws = ph * s - p, pw * s - p
                        he, we = min(hs + k, H + p), min(ws + k, W + p)
                        best, pos = -np.inf, 0
                        for h in range(max(hs, 0), min(he, H)):
                            for w in range(max(ws, 0), min(we, W)):
                                if x[n, c, h, w] > best:
                                    best, pos = x[n, c, h, w], (h - hs) * k + (w - ws)
                        assert (best, pos) == (y[n, c, ph, pw], idx[n, c, ph, pw]), (k, s, p, n, c, ph, pw)


def test_known_answer_lrn_backward():
    """Caffe's across-channel LRN, y_i = x_i * s_i^-beta with s_i = k + alpha/n * sum_{j in window(i)} x_j^2; by the chain rule
      dx_i = dy_i * s_i^-beta - (2 alpha beta / n) * x_i * sum_{j : i in window(j)} dy_j * y_j / s_j
    (lrn_layer.cpp's CrossChannelBackward).  Checked on 6 channels, n = 5 (windows clipped at the ends), against the formula
    evaluated in float64 with explicit loops -- the oracle gets this backward from autograd over
    torch.nn.functional.local_response_norm."""
    import torch
    import torch.nn.functional as F
    rs = np.random.RandomState(9)
    C, n, alpha, beta, k = 6, 5, 1e-2, 0.75, 1.0        # alpha large enough for the second term to matter
    x = rs.uniform(-3, 3, (1, C, 1, 2))
    dy = rs.uniform(-1, 1, (1, C, 1, 2))
    xt = torch.tensor(x, dtype=torch.float64, requires_grad=True)
    (F.local_response_norm(xt, n, alpha, beta, k) * torch.tensor(dy)).sum().backward()
    want = np.zeros_like(x)
    for w in range(2):
        xs, ds = x[0, :, 0, w], dy[0, :, 0, w]
        sc = np.array([k + alpha / n * sum(xs[j] ** 2 for j in range(max(0, i - 2), min(C, i + 3))) for i in range(C)])
        ys = xs * sc ** -beta
        for i in range(C):
            acc = sum(ds[j] * ys[j] / sc[j] for j in range(max(0, i - 2), min(C, i + 3)))
            want[0, i, 0, w] = ds[i] * sc[i] ** -beta - 2 * alpha * beta / n * xs[i] * acc
    assert np.allclose(xt.grad.numpy(), want, rtol=1e-10, atol=1e-12)
