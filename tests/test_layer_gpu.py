"""GPU: training-side layer through the C ABI -- label grids bit-exact, image within 1e-5."""
import os
import random

import numpy as np
import pytest

from oracle import layer_ref
from tests.golden import make_golden as G
from tests.util import GOLDEN, bits_equal

pytestmark = pytest.mark.gpu


def test_label_raster_sweep_bit_exact():
    from fcn_object_detector_b200.data_argumentation_layer import label_raster
    rs = np.random.RandomState(0)
    rects = []
    for i in range(96):
        w, h = int(rs.randint(1, 448)), int(rs.randint(1, 448))
        rects.append([int(rs.randint(-20, 448)), int(rs.randint(-20, 448)), w, h])
    rects += [[0, 0, 448, 448], [0, 0, 16, 16], [16, 16, 16, 16], [15, 15, 2, 2], [447, 447, 1, 1], [100, 100, 40, 40],
              [8, 8, 16, 16], [0, 0, 1, 1]]
    rects = np.asarray(rects, np.int32)[:, None, :]
    labels = np.zeros((rects.shape[0], 1), np.int32)
    got = label_raster(rects, labels)
    tot = 0
    for i in range(rects.shape[0]):
        exp = layer_ref.bounding_box_parameterized_labels(448, 448, 16, 1, [rects[i, 0]], [0])
        for g, e in zip(got, exp):
            assert bits_equal(g[i], e), (i, rects[i, 0])
        tot += int(exp[0].sum())
    assert tot > 1000


def test_label_raster_multi_rect_multi_class():
    from fcn_object_detector_b200.data_argumentation_layer import label_raster
    rs = np.random.RandomState(2)
    N, R, Cn = 6, 3, 3
    rects = np.zeros((N, R, 4), np.int32); labels = np.zeros((N, R), np.int32)
    for i in range(N):
        for r in range(R):
            w, h = int(rs.randint(20, 200)), int(rs.randint(20, 200))
            rects[i, r] = [int(rs.randint(0, 448 - w)), int(rs.randint(0, 448 - h)), w, h]
            labels[i, r] = rs.randint(0, Cn)
    nrect = np.asarray([3, 2, 1, 3, 0, 2], np.int32)
    got = label_raster(rects, labels, num_classes=Cn, nrect=nrect)
    for i in range(N):
        exp = layer_ref.bounding_box_parameterized_labels(448, 448, 16, Cn, rects[i, :nrect[i]], labels[i, :nrect[i]])
        for g, e in zip(got, exp):
            assert bits_equal(g[i], e)
    # other grid geometry
    got = label_raster(rects, labels, net_h=224, net_w=320, stride=8, num_classes=Cn)
    for i in range(N):
        exp = layer_ref.bounding_box_parameterized_labels(224, 320, 8, Cn, rects[i], labels[i])
        for g, e in zip(got, exp):
            assert bits_equal(g[i], e)


class _B(object):
    def __init__(self, arr):
        self.data = arr
    num = property(lambda s: s.data.shape[0])

    def reshape(self, *d):
        if tuple(d) != self.data.shape:
            self.data = np.zeros(d, np.float32)


def run_layer(data, lab, draws_dict):
    from fcn_object_detector_b200 import _lib
    from fcn_object_detector_b200.data_argumentation_layer import DataArgumentationLayer
    L = DataArgumentationLayer()
    L.param_str = '448, 448,16, 1'
    bottom = [_B(data), _B(lab)]
    top = [_B(np.zeros((1,), np.float32)) for _ in range(6)]
    L.setup(bottom, top)
    L.reshape(bottom, top)
    draws = []
    for d in draws_dict:
        a = _lib.AugDraw()
        for k, v in d.items():
            setattr(a, k, v)
        draws.append(a)
    L.forward(bottom, top, draws=draws)
    return [t.data for t in top], L.last_rects


def check_layer(data, lab, draws):
    exp, rects = layer_ref.layer_forward(data, lab, draws)
    got, grects = run_layer(data, lab, draws)
    assert np.array_equal(grects, np.asarray(rects, np.int32))
    for k in range(1, 6):
        assert bits_equal(got[k], exp[k]), k
    assert np.abs(got[0] - exp[0]).max() < 1e-5
    return exp


def test_layer_golden_fixture():
    g = np.load(os.path.join(GOLDEN, 'layer_small.npz'))
    data, lab, draws = G.layer_inputs()
    got, rects = run_layer(data, lab, draws)
    assert np.array_equal(rects, g['rects'])
    for k, name in zip(range(1, 6), ('fg', 'bbox', 'size', 'obj', 'cvg')):
        assert bits_equal(got[k], g[name]), name
    assert np.abs(got[0][:, :, ::37, ::41] - g['data_probe']).max() < 1e-5


def test_layer_config2_batch32():
    """BASELINE config 2: param_str '448, 448,16, 1', batch 32, 480x640 sources, one bbox per image."""
    rs = np.random.RandomState(1)
    N, Hs, Ws = 32, 480, 640
    data = rs.randint(0, 256, (N, 3, Hs, Ws)).astype(np.float32)
    lab = np.zeros((N, 1, 1, 5), np.float32)
    rng = random.Random(1234)
    draws = []
    for i in range(N):
        w, h = int(rs.randint(32, 257)), int(rs.randint(32, 257))
        rect = [int(rs.randint(0, Ws - w)), int(rs.randint(0, Hs - h)), w, h]
        lab[i, 0, 0] = rect + [0]
        draws.append(layer_ref.draw_params(rng, Ws, Hs, rect))
    exp = check_layer(data, lab, draws)
    assert exp[1].sum() > 100


def test_layer_all_flips_and_identity_size():
    rs = np.random.RandomState(4)
    N = 8
    data = rs.randint(0, 256, (N, 3, 448, 448)).astype(np.float32)
    lab = np.zeros((N, 1, 1, 5), np.float32)
    draws = []
    for i in range(N):
        rect = [100 + i, 120, 200, 180]
        lab[i, 0, 0] = rect + [0]
        draws.append(dict(flip=[-1, 0, 1, 2][i % 4], f1=1.0 + 0.1 * i, f2=1.0 + 0.07 * i, shift_x=i, shift_y=2 * i,
                          sign_x=i % 2, sign_y=(i // 2) % 2))
    check_layer(data, lab, draws)
