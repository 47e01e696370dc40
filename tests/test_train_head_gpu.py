"""GPU: training head of models/train_val.prototxt (mask / normalise chain + L1Loss x2.0 + EuclideanLoss) through the C ABI
against the CPU oracle, fed by the augmentation layer's own label grids."""
import random

import numpy as np
import pytest

from oracle import layer_ref, train_ref
from tests.util import bits_equal

pytestmark = pytest.mark.gpu


def _layer_tops(N, C=1, seed=1):
    rs = np.random.RandomState(seed)
    rects = np.zeros((N, 2, 4), np.int32); labels = np.zeros((N, 2), np.int32)
    for i in range(N):
        for r in range(2):
            w, h = int(rs.randint(32, 257)), int(rs.randint(32, 257))
            rects[i, r] = [int(rs.randint(0, 448 - w)), int(rs.randint(0, 448 - h)), w, h]
            labels[i, r] = rs.randint(0, C)
    from fcn_object_detector_b200.data_argumentation_layer import label_raster
    return label_raster(rects, labels, num_classes=C)


@pytest.mark.parametrize('N,C', [(32, 1), (5, 3), (1, 1)])
def test_train_head_losses_and_output_gradients(N, C):
    from fcn_object_detector_b200.train_head import train_head_forward
    tops = _layer_tops(N, C)
    rs = np.random.RandomState(N)
    cov = rs.uniform(0, 1, tops[0].shape).astype(np.float32)
    bb = rs.normal(0, 30, tops[1].shape).astype(np.float32)
    bb[:, :, ::7, ::5] = (tops[1] / np.maximum(tops[4], 1))[:, :, ::7, ::5]          # some exact hits: diff == 0 -> sign 0
    got = train_head_forward(cov, bb, tops)
    exp = train_ref.head_forward(cov, bb, *tops)
    assert abs(got['loss_bbox'] - exp['loss_bbox']) <= 1e-6 * max(1.0, abs(exp['loss_bbox']))
    assert abs(got['loss_coverage'] - exp['loss_coverage']) <= 1e-6 * max(1.0, abs(exp['loss_coverage']))
    assert abs(got['loss'] - exp['loss']) <= 2e-6 * max(1.0, abs(exp['loss']))
    assert bits_equal(got['d_coverage'], exp['d_coverage'])
    assert bits_equal(got['d_bboxes'], exp['d_bboxes'])
    assert exp['loss_bbox'] > 0 and (exp['d_bboxes'] != 0).any() and (exp['d_bboxes'] == 0).any()
    # deterministic: the float64 two-stage sum has a fixed order
    again = train_head_forward(cov, bb, tops)
    assert again['loss_bbox'] == got['loss_bbox'] and again['loss_coverage'] == got['loss_coverage']
    # no-gradient form
    ng = train_head_forward(cov, bb, tops, want_grads=False)
    assert ng['d_bboxes'] is None and ng['loss'] == got['loss']


def test_train_head_perfect_prediction_has_zero_bbox_loss():
    from fcn_object_detector_b200.train_head import train_head_forward
    tops = _layer_tops(4)
    got = train_head_forward(tops[0], tops[1], tops)          # coverage = its label, bboxes = bbox-label
    assert got['loss_coverage'] == 0.0 and got['loss_bbox'] == 0.0
    assert not got['d_bboxes'].any() and not got['d_coverage'].any()


def test_slice_points_match_the_prototxt():
    """slice_point 1 / 5 / 9 / 13 at num_classes = 1 (train_val.prototxt:46-51)."""
    lab = np.arange(2 * 17 * 4).reshape(2, 17, 2, 2)
    parts = train_ref.slice_label(lab, 1)
    assert [p.shape[1] for p in parts] == [1, 4, 4, 4, 4]
    assert parts[1][0, 0, 0, 0] == lab[0, 1, 0, 0] and parts[4][0, 0, 0, 0] == lab[0, 13, 0, 0]
