"""GPU: copy-paste synthesis (ArgumentationEngineMapping.argument, argumentation_engine.py:651-746): the compositing kernel
against the reference's per-pixel loop, bit-exact, including overlaps (later proposals win), clipping at the right / bottom
edge and numpy's negative-index wrap."""
import random

import numpy as np
import pytest

from oracle import synth_ref

pytestmark = pytest.mark.gpu


def _objects(rs, n=5):
    objs = []
    for k in range(n):
        H, W = int(rs.randint(40, 90)), int(rs.randint(40, 90))
        img = rs.randint(0, 256, (H, W, 3)).astype(np.uint8)
        w, h = int(rs.randint(10, W - 8)), int(rs.randint(10, H - 8))
        x, y = int(rs.randint(0, W - w)), int(rs.randint(0, H - h))
        mask = np.zeros((H, W, 3), np.uint8)
        yy, xx = np.mgrid[0:H, 0:W]
        ell = ((xx - (x + w / 2.0)) / (w / 2.0)) ** 2 + ((yy - (y + h / 2.0)) / (h / 2.0)) ** 2 <= 1.0
        mask[ell] = int(rs.randint(1, 256))
        objs.append((img, mask, k % 3, [x, y, w, h]))
    return objs


def test_composite_matches_the_reference_loop():
    from fcn_object_detector_b200.synthesis import composite, plan_proposals
    rs = np.random.RandomState(2)
    objs = _objects(rs)
    N, H, W = 6, 120, 160
    bg = rs.randint(0, 256, (N, H, W, 3)).astype(np.uint8)
    msk = np.zeros((N, H, W), np.uint8)
    rng = random.Random(5)
    plans = [plan_proposals(objs, 4, W, H, rng, [[10, 10, 30, 30]] if n % 2 else None)[0] for n in range(N)]
    # hand-made hard cases on image 0: overlapping patches, a patch hanging over the right / bottom edge, and one wider than
    # the background placed at a negative offset (numpy wraps the negative index)
    roi = rs.randint(0, 256, (50, 70, 3)).astype(np.uint8); m = (rs.uniform(0, 1, (50, 70)) > 0.3).astype(np.uint8) * 255
    wide = rs.randint(0, 256, (20, 200, 3)).astype(np.uint8); mw = (rs.uniform(0, 1, (20, 200)) > 0.5).astype(np.uint8) * 7
    tall = rs.randint(0, 256, (150, 12, 3)).astype(np.uint8); mt = np.full((150, 12), 255, np.uint8)
    plans[0] = [(roi, m, 20, 30, 0), (roi[::-1].copy(), m[::-1].copy(), 45, 50, 1), (roi, m, 130, 100, 2), (wide, mw, -40, 60, 1),
                (tall, mt, 5, -30, 0)]
    assert sum(len(p) for p in plans) > 12
    got_i, got_m = composite(bg, msk, plans)
    exp_i, exp_m = synth_ref.composite(bg, msk, plans)
    assert np.array_equal(got_m, exp_m) and np.array_equal(got_i, exp_i)
    assert (exp_m > 0).sum() > 2000 and set(np.unique(exp_m)) >= {0, 1, 2, 3}


def test_argument_places_without_overlap_and_returns_the_reference_tuple():
    from fcn_object_detector_b200.synthesis import ArgumentationEngineMapping, IOU_THRESH, _iou
    rs = np.random.RandomState(3)
    eng = ArgumentationEngineMapping(_objects(rs), rng=random.Random(11))
    bg = rs.randint(0, 256, (240, 320, 3)).astype(np.uint8)
    img, mask, rects, labels = eng.argument(3, bg, None, [100, 80, 60, 60])
    assert img.shape == bg.shape and mask.shape == (240, 320, 1) and len(rects) == len(labels) + 1
    for a in range(len(rects)):
        for b in range(a):
            assert _iou(list(rects[b]), list(rects[a])) <= IOU_THRESH
    assert (mask > 0).any() and (img != bg).any()
    imgs, masks, rr, ll = eng.argument_batch(2, np.stack([bg, bg[::-1]]))
    assert imgs.shape == (2, 240, 320, 3) and masks.shape == (2, 240, 320, 1)
