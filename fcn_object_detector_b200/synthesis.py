"""Copy-paste data synthesis on the B200 path: `ArgumentationEngineMapping.argument`
(/root/reference/scripts/data_argumentation_layer/argumentation_engine.py:651-746).

Per background image the reference picks `num_proposals` object crops, flips / rescales them (cv2), searches a position
whose IoU with the rects already placed is <= 0.05 (:699-721), and then pastes the masked pixels with a per-pixel Python
loop (:723-728) -- the part that dominates its run time.  Here the draws, the cv2 flip / resize of the small object patch
and the position search stay on the host, in the reference's order; the compositing of a whole batch of backgrounds is
one kernel launch (fcnod_composite).
"""
import ctypes as C
import random

import numpy as np

from . import _lib

IOU_THRESH = 0.05        # ArgumentationEngineMapping.__iou_thresh (:618)
MAX_COUNTER = 100        # (:619)


def _iou(a, b):
    """JaccardCoeff.iou (:26-55) as used by the position search: fp32 areas and divides."""
    f = np.float32
    x = max(a[0], b[0]); y = max(a[1], b[1])
    w = min(a[0] + a[2], b[0] + b[2]) - x
    h = min(a[1] + a[3], b[1] + b[3]) - y
    if w < 0 or h < 0:
        return 0.0
    ux = min(a[0], b[0]); uy = min(a[1], b[1])
    uw = max(a[0] + a[2], b[0] + b[2]) - ux
    uh = max(a[1] + a[3], b[1] + b[3]) - uy
    with np.errstate(all='ignore'):
        score = (f(w * h) / f(uw * uh)) / (f(a[2] * a[3]) / f(b[2] * b[3]))
    return float(score)


def _flip_rect(rect, im_w, im_h, flag):
    p1 = (rect[0], rect[1]); p2 = (rect[0] + rect[2], rect[1] + rect[3])
    if flag == -1:
        p1 = (im_w - p1[0] - 1, im_h - p1[1] - 1); p2 = (im_w - p2[0] - 1, im_h - p2[1] - 1)
    elif flag == 0:
        p1 = (p1[0], im_h - p1[1] - 1); p2 = (p2[0], im_h - p2[1] - 1)
    elif flag == 1:
        p1 = (im_w - p1[0] - 1, p1[1]); p2 = (im_w - p2[0] - 1, p2[1])
    x = min(p1[0], p2[0]); y = min(p1[1], p2[1])
    return [max(x, 0), max(y, 0), abs(p2[0] - p1[0]), abs(p2[1] - p1[1])]


def plan_proposals(objects, num_proposals, im_x, im_y, rng, placed=None):
    """The host half of `argument` for ONE background: draws, patch preparation and position search, in the reference's
    order.  objects: list of (image uint8 HxWx3, mask uint8 HxWx3 or HxW, label, rect [x,y,w,h]).
    Returns (patches [(roi uint8 hxwx3, mask uint8 hxw, cx, cy, label)], flag_position list, labels list)."""
    import cv2
    flag_position = [list(r) for r in (placed or [])]
    patches, labels = [], []
    for _ in range(num_proposals):
        image, mask, label, rect = objects[rng.randint(0, len(objects) - 1)]          # :668-673
        mask = np.where(np.asarray(mask) > 0, 255, 0).astype(np.uint8)               # :677
        if mask.ndim == 2:
            mask = np.repeat(mask[:, :, None], 3, axis=2)
        x, y, w, h = [int(v) for v in rect]
        flip_flag = rng.randint(-1, 2)                                               # :679
        if -2 < flip_flag < 2:
            image = cv2.flip(image, flip_flag); mask = cv2.flip(mask, flip_flag)
            x, y, w, h = _flip_rect([x, y, w, h], image.shape[1], image.shape[0], flip_flag)
        im_roi = image[y:y + h, x:x + w].copy(); im_msk = mask[y:y + h, x:x + w].copy()
        if rng.randint(0, 1):                                                        # :689-696
            scale = rng.uniform(1.0, 2.2)
            w = int(w * scale); h = int(h * scale)
            im_roi = cv2.resize(im_roi, (int(w), int(h))); im_msk = cv2.resize(im_msk, (int(w), int(h)))

        def position():
            cx, cy = rng.randint(0, im_x - 1), rng.randint(0, im_y - 1)
            cx = cx - ((cx + w) - im_x) if cx + w > im_x - 1 else cx
            cy = cy - ((cy + h) - im_y) if cy + h > im_y - 1 else cy
            return cx, cy
        cx, cy = position()                                                          # :698-701
        nrect = [cx, cy, w, h]
        counter, found = 0, True
        for bbox in flag_position:                                                   # :705-721
            if _iou(bbox, nrect) > IOU_THRESH and found:
                is_ok = True
                while True:
                    cx, cy = position()
                    nrect = [cx, cy, w, h]
                    for bbox2 in flag_position:
                        if _iou(bbox2, nrect) > IOU_THRESH:
                            is_ok = False
                            break
                    if is_ok:
                        break
                    counter += 1
                    if counter > MAX_COUNTER:
                        found = False
                        break
        if found and im_roi.size:
            patches.append((np.ascontiguousarray(im_roi), np.ascontiguousarray(im_msk[:, :, 0]), int(cx), int(cy), int(label)))
            flag_position.append(nrect)
            labels.append(label)
    return patches, flag_position, labels


def composite(backgrounds, masks, patch_lists):
    """backgrounds (N,H,W,3) uint8 and masks (N,H,W) uint8 are composited IN PLACE copies (returned) with, per image, its list
    of patches (roi, mask, cx, cy, label) in application order.  One fcnod_composite call for the whole batch."""
    imgs = np.ascontiguousarray(backgrounds, dtype=np.uint8).copy()
    msk = np.ascontiguousarray(masks, dtype=np.uint8).copy()
    N, H, W, _ = imgs.shape
    table, begin, rois, rmasks, off = [], [0], [], [], 0
    for n in range(N):
        for roi, m, cx, cy, label in patch_lists[n]:
            h, w = m.shape
            if cx < -W or cy < -H:
                raise IndexError('proposal placed beyond the background (the reference raises IndexError here too)')
            table.append(_lib.Paste(off, w, h, cx, cy, label))
            rois.append(np.ascontiguousarray(roi, np.uint8).reshape(-1)); rmasks.append(np.ascontiguousarray(m, np.uint8).reshape(-1))
            off += h * w
        begin.append(len(table))
    if not table:
        return imgs, msk
    roi_buf = np.concatenate(rois); rm_buf = np.concatenate(rmasks)
    parr = (_lib.Paste * len(table))(*table)
    barr = np.asarray(begin, np.int32)
    _lib.check(_lib.lib().fcnod_composite(_lib.ptr(imgs), _lib.ptr(msk), N, H, W, _lib.ptr(roi_buf), _lib.ptr(rm_buf), off, parr,
                                          _lib.ptr(barr)))
    return imgs, msk


class ArgumentationEngineMapping(object):
    """The part of the reference class this path needs: `argument(num_proposals, im_bg, im_mk, mrect)` for one background,
    and `argument_batch` for many (one device launch)."""

    def __init__(self, objects, rng=None):
        self.objects = objects
        self.rng = rng or random.Random()

    def argument_batch(self, num_proposals, backgrounds, masks=None, mrects=None):
        backgrounds = np.asarray(backgrounds, np.uint8)
        N, H, W, _ = backgrounds.shape
        masks = np.zeros((N, H, W), np.uint8) if masks is None else np.asarray(masks, np.uint8).reshape(N, H, W)
        plans = [plan_proposals(self.objects, num_proposals, W, H, self.rng, None if mrects is None or mrects[n] is None else [mrects[n]])
                 for n in range(N)]
        imgs, msk = composite(backgrounds, masks, [p[0] for p in plans])
        return imgs, msk[..., None], [np.array(p[1]) for p in plans], [np.array(p[2]) for p in plans]

    def argument(self, num_proposals, im_bg, im_mk=None, mrect=None):
        imgs, msk, rects, labels = self.argument_batch(num_proposals, im_bg[None], None if im_mk is None else np.asarray(im_mk)[None],
                                                       [mrect])
        return imgs[0], msk[0], rects[0], labels[0]
