"""GPU: post-net stage (decode + groupRectangles + filters + rescale) through the C ABI,
bit-exact against the oracle."""
import os

import numpy as np
import pytest

from oracle import postnet_ref
from tests.golden import make_golden as G
from tests.util import GOLDEN

pytestmark = pytest.mark.gpu


def run_gpu(cov, bb, src_hw=None, **kw):
    from fcn_object_detector_b200.detector import postnet
    return postnet(cov, bb, src_hw=src_hw, **kw)


def oracle(cov, bb, shape, **kw):
    weights = []

    def grp(rects, thr, eps):
        r, w = postnet_ref.group_rectangles(rects, thr, eps)
        return r, w
    ob, ol = postnet_ref.postnet(cov, bb, shape, group_fn=grp, **kw)
    return ob, ol


def compare(cov, bb, shape=(448, 448, 3), **kw):
    B = cov.shape[0]
    okw = dict(prob_thresh=kw.get('prob_thresh', 0.5), min_boxes=kw.get('min_boxes', 3), eps=kw.get('eps', 0.2),
               stride=kw.get('stride', 16), skip_background=bool(kw.get('skip_background', 1)))
    hw = None if shape[:2] == (448, 448) else np.tile(np.asarray(shape[:2], np.int32), (B, 1))
    got = run_gpu(cov, bb, src_hw=hw, **kw)
    n = 0
    for b in range(B):
        ob, ol = oracle(cov[b], bb[b], shape, **okw)
        assert np.array_equal(got[b][0], ob), (b, got[b][0], ob)
        assert np.array_equal(got[b][1], ol)
        n += len(ol)
    return n


def test_golden_fixture():
    g = np.load(os.path.join(GOLDEN, 'postnet_planted.npz'))
    cov, bb = G.postnet_inputs()
    for tag, kw, hw in (('s16', dict(stride=16), None), ('s8', dict(stride=8), None), ('vga', dict(), (480, 640))):
        src = None if hw is None else np.tile(np.asarray(hw, np.int32), (cov.shape[0], 1))
        got = run_gpu(cov, bb, src_hw=src, **kw)
        for b in range(cov.shape[0]):
            k = int(g['cnt_' + tag][b])
            assert len(got[b][0]) == k
            assert np.array_equal(got[b][0], g['det_' + tag][b, :k]) and np.array_equal(got[b][1], g['lab_' + tag][b, :k])


def test_planted_batch_strides_and_params():
    rng = np.random.RandomState(3)
    maps = [postnet_ref.planted_maps(rng, n_obj=k % 5) for k in range(24)]
    cov = np.stack([m[0] for m in maps]); bb = np.stack([m[1] for m in maps])
    assert compare(cov, bb) > 10
    compare(cov, bb, stride=8)                                   # the literal upstream 16/2 (A2)
    compare(cov, bb, shape=(480, 640, 3))
    compare(cov, bb, shape=(1080, 1920, 3))
    compare(cov, bb, prob_thresh=0.3, min_boxes=1, eps=0.1)
    compare(cov, bb, prob_thresh=0.7, min_boxes=2, eps=0.3)
    compare(cov, bb, skip_background=0)


def test_random_maps_dense_candidates():
    """Unstructured maps: many candidates (up to all 784 cells), chains of similar rects."""
    rs = np.random.RandomState(9)
    cov = rs.uniform(0, 1, (6, 4, 28, 28)).astype(np.float32)
    bb = rs.normal(0, 6, (6, 16, 28, 28)).astype(np.float32)
    bb[:, 2::4] += 40; bb[:, 3::4] += 40                         # boxes ~40 px, neighbours overlap -> long chains
    compare(cov, bb, prob_thresh=0.0)
    compare(cov, bb, prob_thresh=0.5)
    compare(cov, bb, prob_thresh=0.9, min_boxes=1)


def test_empty_zero_and_threshold_edges():
    cov = np.zeros((3, 4, 28, 28), np.float32); bb = np.zeros((3, 16, 28, 28), np.float32)
    got = run_gpu(cov, bb)
    assert all(len(g[0]) == 0 for g in got)
    # all candidates but all-zero boxes: `not propose_boxes.any()` (:339) -> nothing
    cov[:] = 1.0
    bb[:, :, 1:, :] = 0; bb[:, :, :, 1:] = 0
    compare(cov, bb)
    # coverage exactly at threshold counts (>=)
    cov[:] = 0.5
    bb = np.random.RandomState(1).normal(0, 1, bb.shape).astype(np.float32)
    compare(cov, bb)


def test_large_batch_property():
    """At full batch (256 frames) every frame must equal the same frame run alone (frame independence)."""
    rng = np.random.RandomState(5)
    maps = [postnet_ref.planted_maps(rng, n_obj=k % 4) for k in range(8)]
    cov = np.stack([m[0] for m in maps] * 32); bb = np.stack([m[1] for m in maps] * 32)
    got = run_gpu(cov, bb)
    for b in range(8, 256):
        assert np.array_equal(got[b][0], got[b % 8][0]) and np.array_equal(got[b][1], got[b % 8][1])
    compare(cov[:8], bb[:8])


def test_overflow_is_reported_not_silent():
    """The reference keeps every cluster (fcn_object_detector.py:104-124); the C ABI stores FCNOD_MAX_DET per frame.
    More than that must come back as out_count > FCNOD_MAX_DET with the first FCNOD_MAX_DET rows intact."""
    from fcn_object_detector_b200 import _lib
    M = _lib.FCNOD_MAX_DET
    rs = np.random.RandomState(11)
    cov = np.zeros((3, 4, 28, 28), np.float32)
    bb = np.zeros((3, 16, 28, 28), np.float32)
    bb[:, 2::4] = 30.0; bb[:, 3::4] = 30.0                     # 30 x 30 boxes: pass the height filter
    bb += rs.uniform(-1, 1, bb.shape).astype(np.float32)
    cov[0, 1, ::2, ::2] = 0.9                                   # frame 0: 196 isolated candidates in class 0 alone
    cov[1, 1, :6, ::2] = 0.9; cov[1, 2, :4, ::2] = 0.9          # frame 1: 84 + 56 over two classes (class 0 alone overflows)
    cov[2, 3, 0, :10] = 0.9                                     # frame 2: 10, no overflow
    kw = dict(min_boxes=0, prob_thresh=0.5)
    overflow = []
    got = run_gpu(cov, bb, overflow=overflow, **kw)
    totals = {}
    for b in range(3):
        ob, ol = oracle(cov[b], bb[b], (448, 448, 3), prob_thresh=0.5, min_boxes=0, eps=0.2, stride=16, skip_background=True)
        totals[b] = len(ol)
        k = min(len(ol), M)
        assert len(got[b][0]) == k
        # (frame 1: the per-class lists are capped at M too, so the stored rows are class 0's first M)
        assert np.array_equal(got[b][0], ob[:k]) and np.array_equal(got[b][1], ol[:k])
    assert totals[0] == 196 and totals[1] == 140 and totals[2] == 10
    assert overflow == [(0, 196), (1, 140)], overflow


def test_gridbox_to_boxes_matches_reference_function():
    """Standalone decode kernel = gridbox_to_boxes (fcn_object_detector.py:357-394): boxes (float64), cvgs and mask
    bit-exact against the oracle, at stride 16, at the literal stride 8 (grid 56 clamped to the 28-row map) and on a
    map whose width is not a multiple of 4 (scalar loads)."""
    from fcn_object_detector_b200.detector import decode_batch, gridbox_to_boxes
    rs = np.random.RandomState(17)
    for (g, stride, thr) in ((28, 16, 0.5), (28, 8, 0.5), (28, 16, 0.0), (28, 16, 1.1), (27, 16, 0.4), (14, 32, 0.6)):
        cvg = rs.uniform(0, 1, (g, g)).astype(np.float32)
        box = rs.normal(0, 20, (4, g, g)).astype(np.float32)
        eb, ec, em = postnet_ref.gridbox_to_boxes(cvg, box, thr, 448, 448, stride)
        gb, gc, gm = gridbox_to_boxes(cvg, box, thr, 448, 448, stride)
        assert gb.dtype == np.float64 and gb.shape == eb.shape, (g, stride, thr, gb.shape, eb.shape)
        assert np.array_equal(gb.view(np.uint64), np.ascontiguousarray(eb, np.float64).view(np.uint64))
        assert np.array_equal(gc, np.asarray(ec, np.float64).reshape(-1, 3))
        assert np.array_equal(gm, em)
    # batch form, several classes, capped output: counts are untruncated, the first `cap` rows are stored
    cov = rs.uniform(0, 1, (5, 4, 28, 28)).astype(np.float32)
    bb = rs.normal(0, 20, (5, 16, 28, 28)).astype(np.float32)
    boxes, cell, score, cnt = decode_batch(cov, bb, cap=100, prob_thresh=0.5)
    for b in range(5):
        for c in range(3):
            eb, ec, em = postnet_ref.gridbox_to_boxes(cov[b, c + 1], bb[b, 4 * c:4 * c + 4], 0.5, 448, 448, 16)
            assert cnt[b, c] == len(eb) > 100
            assert np.array_equal(boxes[b, c], eb[:100])
            assert np.array_equal(score[b, c], np.asarray(ec)[:100, 2].astype(np.float32))
