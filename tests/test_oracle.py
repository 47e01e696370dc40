"""CPU: the oracle against the golden fixtures, cv2, and its own invariants."""
import os
import random

import numpy as np
import pytest

from tests.util import GOLDEN, REFERENCE, rel_err
from oracle import layer_ref, net_ref, postnet_ref, prototxt
from tests.golden import make_golden as G


def test_generated_prototxt_matches_reference():
    from fcn_object_detector_b200.netspec import deploy_prototxt
    path = os.path.join(REFERENCE, 'models', 'deploy.prototxt')
    if not os.path.exists(path):
        pytest.skip('upstream tree not present on this box')
    ref = prototxt.load(path)
    gen = prototxt.parse(deploy_prototxt())
    assert prototxt.input_spec(ref) == prototxt.input_spec(gen)
    assert prototxt.layers(ref) == prototxt.layers(gen)


def test_conv_shapes_match_prototxt():
    from fcn_object_detector_b200.netspec import conv_shapes, deploy_prototxt
    Ls = prototxt.layers(prototxt.parse(deploy_prototxt()))
    convs = [(L['name'], L['num_output'], L['kernel'], L['bias_value']) for L in Ls if L['type'] == 'Convolution']
    assert convs == [(n, co, k, b) for n, co, ci, k, b in conv_shapes()]
    assert sum(co * ci * k * k + co for n, co, ci, k, b in conv_shapes()) == 5994052


def test_group_rectangles_matches_cv2():
    cv2 = pytest.importorskip('cv2')
    rng = np.random.RandomState(11)
    for it in range(400):
        thr = int(rng.choice([1, 2, 3])); eps = float(rng.choice([0.1, 0.2, 0.3]))
        rects = []
        for c in range(rng.randint(0, 5)):
            x, y, w, h = rng.randint(0, 300), rng.randint(0, 300), rng.randint(20, 150), rng.randint(20, 150)
            for m in range(rng.randint(1, 12)):
                j = rng.randint(-int(0.15 * min(w, h)) - 1, int(0.15 * min(w, h)) + 2, size=4)
                rects.append([int(x + j[0]), int(y + j[1]), int(max(1, w + j[2])), int(max(1, h + j[3]))])
            if rng.rand() < 0.3:
                for m in range(rng.randint(1, 6)):
                    rects.append([int(x + w // 4 + rng.randint(-2, 3)), int(y + h // 4 + rng.randint(-2, 3)), w // 2, h // 2])
        rng.shuffle(rects)
        a_r, a_w = postnet_ref.group_rectangles(rects, thr, eps)
        b_r, b_w = cv2.groupRectangles([list(r) for r in rects], thr, eps)
        assert a_r == [list(map(int, r)) for r in b_r]
        assert a_w == [int(v) for v in np.ravel(b_w)]


def test_group_rectangles_edge_cases():
    assert postnet_ref.group_rectangles([], 3, 0.2) == ([], [])
    same = [[10, 10, 50, 50]] * 3
    assert postnet_ref.group_rectangles(same, 3, 0.2) == ([], [])          # needs > threshold members
    assert postnet_ref.group_rectangles(same + same[:1], 3, 0.2) == ([[10, 10, 50, 50]], [4])
    assert postnet_ref.group_rectangles(same, 0, 0.2) == ([list(r) for r in same], [1, 1, 1])


def test_postnet_golden():
    g = np.load(os.path.join(GOLDEN, 'postnet_planted.npz'))
    cov, bb = G.postnet_inputs()
    for tag, kw, shape in (('s16', dict(stride=16), (448, 448, 3)), ('s8', dict(stride=8), (448, 448, 3)),
                           ('vga', dict(), (480, 640, 3))):
        dets = [postnet_ref.postnet(cov[b], bb[b], shape, **kw) for b in range(cov.shape[0])]
        d, l, c = G.pack_dets(dets)
        assert np.array_equal(d, g['det_' + tag]) and np.array_equal(l, g['lab_' + tag]) and np.array_equal(c, g['cnt_' + tag])
    assert g['cnt_s16'].sum() > 0


def test_postnet_empty_and_all_zero():
    cov = np.zeros((4, 28, 28), np.float32); bb = np.zeros((16, 28, 28), np.float32)
    ob, ol = postnet_ref.postnet(cov, bb, (448, 448, 3))
    assert ob.shape == (0, 5) and ol.shape == (0,)


def test_layer_golden():
    g = np.load(os.path.join(GOLDEN, 'layer_small.npz'))
    data, lab, draws = G.layer_inputs()
    top, rects = layer_ref.layer_forward(data, lab, draws)
    for k, t in zip(('fg', 'bbox', 'size', 'obj', 'cvg'), top[1:]):
        assert np.array_equal(t.view(np.uint32), g[k].view(np.uint32)), k
    assert np.array_equal(np.asarray(rects, np.int32), g['rects'])
    assert np.allclose(top[0][:, :, ::37, ::41], g['data_probe'], atol=1e-6)
    assert g['cvg'].sum() > 0


def test_label_invariants():
    fg, bl, sl, ol, cl = layer_ref.bounding_box_parameterized_labels(448, 448, 16, 1, [(100, 120, 90, 70)], [0])
    assert fg.shape == (1, 28, 28) and bl.shape == (4, 28, 28)
    m = fg[0] == 1
    assert m.sum() > 0 and np.array_equal(cl[0] == 1, m) and np.array_equal(cl[0], cl[3])
    jj, ii = np.where(m)
    assert np.array_equal(bl[0][m], 100 - ii * 16) and np.array_equal(bl[3][m], 190 - jj * 16)
    assert np.all(sl[0][m] == np.float32(1.0 / 90)) and np.all(ol[0][m] == np.float32(256) / np.float32(6300))
    assert np.all(bl[:, ~m] == 0)


def test_iou_threshold_compared_in_double():
    # a score that is > 0.1 as float64(float32 value) but not after rounding the threshold to fp32 would differ;
    # the oracle must use the double compare
    s = np.float32(0.1)
    assert float(s) > 0.1 and not (s > np.float32(0.1))


def test_flip_rect_off_by_one():
    assert layer_ref.flip_rect([10, 20, 30, 40], 640, 480, 1) == [599, 20, 30, 40]     # W - x - w - 1
    assert layer_ref.flip_rect([10, 20, 30, 40], 640, 480, 0) == [10, 419, 30, 40]
    assert layer_ref.flip_rect([10, 20, 30, 40], 640, 480, -1) == [599, 419, 30, 40]


def test_resize_bilinear_matches_cv2():
    cv2 = pytest.importorskip('cv2')
    rs = np.random.RandomState(5)
    for shape in ((123, 211, 3), (480, 640, 3), (60, 33, 3)):
        roi = rs.rand(*shape).astype(np.float32)
        assert np.abs(postnet_ref.resize_bilinear(roi, 448, 448) - cv2.resize(roi, (448, 448))).max() < 1e-6
    # the reference passes INTER_CUBIC in the dst slot (argumentation_engine.py:120) => plain bilinear (A11)
    assert np.array_equal(cv2.resize(roi, (448, 448), cv2.INTER_CUBIC), cv2.resize(roi, (448, 448)))


def test_net_oracle_golden(oracle_net):
    Ls, W = oracle_net
    g = np.load(os.path.join(GOLDEN, 'net_seed2.npz'))
    o = net_ref.forward(Ls, W, G.net_inputs(), want=['coverage', 'bboxes', 'pool1/norm1'])
    assert o['coverage'].shape == (1, 4, 28, 28) and o['bboxes'].shape == (1, 16, 28, 28)
    assert rel_err(o['coverage'], g['coverage']) < 1e-5 and rel_err(o['bboxes'], g['bboxes']) < 1e-5
    assert rel_err(o['pool1/norm1'][:, ::8, ::16, ::16], g['norm1_probe']) < 1e-5


def test_caffe_pool_semantics():
    import torch
    x = torch.arange(25, dtype=torch.float32).reshape(1, 1, 5, 5)
    y = net_ref.max_pool_caffe(x, 3, 2, 0)            # ceil mode: 5 -> 2... (5-3)/2+1 = 2 exactly
    assert y.shape[-2:] == (2, 2)
    x = torch.arange(36, dtype=torch.float32).reshape(1, 1, 6, 6)
    y = net_ref.max_pool_caffe(x, 3, 2, 0)            # ceil((6-3)/2)+1 = 3, last window clipped
    assert y.shape[-2:] == (3, 3) and y[0, 0, 2, 2] == 35
    y = net_ref.max_pool_caffe(x, 3, 1, 1)
    assert y.shape[-2:] == (6, 6) and y[0, 0, 0, 0] == 7


def test_caffemodel_roundtrip(tmp_path):
    from fcn_object_detector_b200.caffemodel import write_caffemodel
    from oracle.caffemodel_io import read_caffemodel
    rs = np.random.RandomState(0)
    w = rs.randn(8, 3, 3, 3).astype(np.float32); b = rs.randn(8).astype(np.float32)
    for legacy in (False, True):
        p = write_caffemodel(str(tmp_path / ('m%d.caffemodel' % legacy)), [('c1', 'Convolution', [w, b])], legacy_shape=legacy)
        r = read_caffemodel(p)
        assert np.array_equal(r['c1'][0].reshape(w.shape), w) and np.array_equal(r['c1'][1].reshape(-1), b)
