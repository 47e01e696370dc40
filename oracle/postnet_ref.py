"""CPU restatement of the detector's pre/post-net stage -- TEST INFRASTRUCTURE ONLY (oracle).

Follows /root/reference/scripts/fcn_object_detector.py with Python-2 semantics:
  demean_rgb_image   :407-413      gridbox_to_boxes  :357-394
  vote_boxes         :337-351      resize_detection  :396-405
  run_detector       :54-132 (per-class loop :104-118, blob reads :89-90)
cv.groupRectangles (OpenCV objdetect, cascadedetect.cpp; not vendored upstream,
version unpinned -- CMakeLists.txt:18) is restated from its published algorithm
in `group_rectangles` and pinned against the cv2 present in this container by
tests/test_oracle_postnet.py (integer rects; the float->int step of the 2.4/3.x
Python binding is round-half-even).

PARITY UNPINNED at the upstream boundary: the reference holds no golden vectors
(SURVEY.md 8c); the cv2 cross-check pins `group_rectangles` only.
"""
import math
import numpy as np


# ----------------------------------------------------------------------------- pre-net
def demean_rgb_image(im):
    """fcn_object_detector.py:407-413 (float64, BGR channel order, global min-max)."""
    im = im.astype(float)
    im[:, :, 0] -= float(104.0069879317889)
    im[:, :, 1] -= float(116.66876761696767)
    im[:, :, 2] -= float(122.6789143406786)
    return (im - im.min()) / (im.max() - im.min())


def resize_bilinear(img, out_w, out_h):
    """cv.resize(img, (w, h)) default INTER_LINEAR on a float image (:80):
    half-pixel centres, edge clamp.  numpy restatement (matches cv2 to ~1e-7)."""
    img = np.asarray(img)
    H, W = img.shape[:2]
    if (W, H) == (out_w, out_h):
        return img.copy()
    sx = W / float(out_w)
    sy = H / float(out_h)
    fx = (np.arange(out_w) + 0.5) * sx - 0.5
    fy = (np.arange(out_h) + 0.5) * sy - 0.5
    x0 = np.floor(fx).astype(int); ax = fx - x0
    y0 = np.floor(fy).astype(int); ay = fy - y0
    ax[x0 < 0] = 0.0; x0[x0 < 0] = 0
    ax[x0 >= W - 1] = 0.0; x0[x0 >= W - 1] = W - 1
    ay[y0 < 0] = 0.0; y0[y0 < 0] = 0
    ay[y0 >= H - 1] = 0.0; y0[y0 >= H - 1] = H - 1
    x1 = np.minimum(x0 + 1, W - 1); y1 = np.minimum(y0 + 1, H - 1)
    ax = ax.astype(img.dtype)[None, :, None]; ay = ay.astype(img.dtype)[:, None, None]
    top = img[y0][:, x0] * (1 - ax) + img[y0][:, x1] * ax
    bot = img[y1][:, x0] * (1 - ax) + img[y1][:, x1] * ax
    return top * (1 - ay) + bot * ay


def preprocess(bgr_u8, net_w, net_h):
    """:79-82 -> (3, net_h, net_w) float32 as written into blobs['data'].data[0]."""
    im = demean_rgb_image(bgr_u8)
    im = resize_bilinear(im, net_w, net_h)
    return im.transpose((2, 0, 1)).astype(np.float32)


# ----------------------------------------------------------------------------- decode
def gridbox_to_boxes(net_cvg, net_boxes, prob_thresh, im_w=448, im_h=448, stride=16):
    """:357-394.  `stride` is literally 16/2 == 8 upstream (:360, Python 2); the
    north-star net is stride 16 so it is a parameter here."""
    grid_sz_x = int(im_w // stride)
    grid_sz_y = int(im_h // stride)
    cell_width = im_w // grid_sz_x        # Py2 int / int
    cell_height = im_h // grid_sz_y
    cvg_val = net_cvg[0:grid_sz_y][0:grid_sz_x]      # double *row* slice, as written (:371)
    mask = (cvg_val >= prob_thresh)
    y, x = np.where(mask == 1)
    mx = x * cell_width
    my = y * cell_height
    x1 = np.asarray([net_boxes[0][y[i]][x[i]] for i in range(x.size)], dtype=np.float32) + mx
    y1 = np.asarray([net_boxes[1][y[i]][x[i]] for i in range(x.size)], dtype=np.float32) + my
    x2 = np.asarray([net_boxes[2][y[i]][x[i]] for i in range(x.size)], dtype=np.float32) + mx
    y2 = np.asarray([net_boxes[3][y[i]][x[i]] for i in range(x.size)], dtype=np.float32) + my
    # float32 array + int64 array -> float64 (same under legacy and NEP-50 promotion)
    assert x1.dtype == np.float64
    boxes = np.transpose(np.vstack((x1, y1, x2, y2)))
    cvgs = np.transpose(np.vstack((x, y, np.asarray([cvg_val[y[i]][x[i]] for i in range(x.size)]))))
    return boxes, cvgs, mask


# ----------------------------------------------------------------------------- groupRectangles
def _cv_round(v):
    """cv::saturate_cast<int>(float|double) == cvRound: round half to even."""
    return int(np.rint(v))


def _similar(r1, r2, eps):
    delta = eps * (min(r1[2], r2[2]) + min(r1[3], r2[3])) * 0.5
    return (abs(r1[0] - r2[0]) <= delta and abs(r1[1] - r2[1]) <= delta and
            abs(r1[0] + r1[2] - r2[0] - r2[2]) <= delta and
            abs(r1[1] + r1[3] - r2[1] - r2[3]) <= delta)


def group_rectangles(rects, group_threshold, eps=0.2):
    """Restatement of cv::groupRectangles(rectList, weights, groupThreshold, eps).
    rects: list of int (x, y, w, h).  Returns (rects, weights)."""
    rects = [tuple(int(v) for v in r) for r in rects]
    n = len(rects)
    if group_threshold <= 0 or n == 0:
        return [list(r) for r in rects], [1] * n
    # cv::partition: equivalence classes = connected components of the (symmetric)
    # predicate; class ids are handed out in order of first member index.
    parent = list(range(n))

    def find(i):
        while parent[i] != i:
            parent[i] = parent[parent[i]]
            i = parent[i]
        return i
    for i in range(n):
        for j in range(i + 1, n):
            if _similar(rects[i], rects[j], eps):
                a, b = find(i), find(j)
                if a != b:
                    parent[max(a, b)] = min(a, b)
    labels, ids = [], {}
    for i in range(n):
        r = find(i)
        if r not in ids:
            ids[r] = len(ids)
        labels.append(ids[r])
    nclasses = len(ids)
    sums = [[0, 0, 0, 0] for _ in range(nclasses)]
    cnt = [0] * nclasses
    for i in range(n):
        c = labels[i]
        for k in range(4):
            sums[c][k] += rects[i][k]
        cnt[c] += 1
    rr = []
    for c in range(nclasses):
        s = np.float32(1.0) / np.float32(cnt[c])
        rr.append([_cv_round(np.float32(sums[c][k]) * s) for k in range(4)])
    out_r, out_w = [], []
    for i in range(nclasses):
        r1, n1 = rr[i], cnt[i]
        if n1 <= group_threshold:
            continue
        keep = True
        for j in range(nclasses):
            n2 = cnt[j]
            if j == i or n2 <= group_threshold:
                continue
            r2 = rr[j]
            dx = _cv_round(r2[2] * eps)
            dy = _cv_round(r2[3] * eps)
            if (r1[0] >= r2[0] - dx and r1[1] >= r2[1] - dy and
                    r1[0] + r1[2] <= r2[0] + r2[2] + dx and
                    r1[1] + r1[3] <= r2[1] + r2[3] + dy and
                    (n2 > max(3, n1) or n1 < 3)):
                keep = False
                break
        if keep:
            out_r.append(r1)
            out_w.append(n1)
    return out_r, out_w


def boxes_to_int_rects(boxes):
    """The OpenCV 2.4/3.x Python binding converts each float64 with
    saturate_cast<int> (round half even) -- SURVEY.md A5."""
    return np.rint(np.asarray(boxes, dtype=np.float64)).astype(np.int64).clip(-2**31, 2**31 - 1).astype(np.int32)


def vote_boxes(propose_boxes, min_boxes=3, eps=0.2, group_fn=None):
    """:337-351.  Returns list of [x1, y1, x2, y2, confidence]."""
    detections = []
    propose_boxes = np.asarray(propose_boxes)
    if not propose_boxes.any():
        return detections
    rects = boxes_to_int_rects(propose_boxes).tolist()
    nboxes, weights = (group_fn or group_rectangles)(rects, min_boxes, eps)
    weights = np.ravel(np.asarray(weights))
    for rect, weight in zip(nboxes, weights):
        if (rect[3] - rect[1]) >= 20:
            confidence = math.log(weight)
            detections.append([rect[0], rect[1], rect[2], rect[3], confidence])
    return detections


def resize_detection(in_shape, bbox, net_w=448, net_h=448):
    """:396-405: float scale then assignment into an int array (truncation toward 0)."""
    diffx = float(in_shape[1]) / float(net_w)
    diffy = float(in_shape[0]) / float(net_h)
    out = bbox
    for index, box in enumerate(bbox):
        out[index, 0] = int(box[0] * diffx)
        out[index, 1] = int(box[1] * diffy)
        out[index, 2] = int(box[2] * diffx)
        out[index, 3] = int(box[3] * diffy)
    return out


def postnet(coverage, bboxes, in_shape, prob_thresh=0.5, min_boxes=3, eps=0.2,
            net_w=448, net_h=448, stride=16, skip_background=True, group_fn=None):
    """:89-132 for ONE frame.  coverage (C_cvg, G, G), bboxes (C_bbox, G, G) float32.
    Returns (object_boxes int (K,5), object_labels int (K,))."""
    probability_map = coverage[1:] if skip_background else coverage
    object_boxes, object_labels = [], []
    for index, p_map in enumerate(probability_map):
        idx = index * 4
        boxes, cvgs, mask = gridbox_to_boxes(p_map, bboxes[idx:idx + 4], prob_thresh,
                                             net_w, net_h, stride)
        for box in vote_boxes(boxes, min_boxes, eps, group_fn):
            object_boxes.append(box)
            object_labels.append(index)
    if not len(object_boxes):
        return np.zeros((0, 5), dtype=np.int64), np.zeros((0,), dtype=np.int64)
    ob = np.asarray([[int(v) for v in b] for b in object_boxes], dtype=np.int64)  # dtype=np.int truncates
    ol = np.asarray(object_labels, dtype=np.int64)
    ob = resize_detection(in_shape, ob, net_w, net_h)
    return ob, ol


def planted_maps(rng, n_obj=3, n_cvg=4, grid=28, cell=16, cls=None, jitter=1.5):
    """Synthetic coverage/bbox maps with planted objects (SURVEY.md 8d decode/NMS parity inputs)."""
    cov = rng.uniform(0.0, 0.3, size=(n_cvg, grid, grid)).astype(np.float32)
    bb = rng.normal(0.0, 1.0, size=(4 * n_cvg, grid, grid)).astype(np.float32)
    W = grid * cell
    for k in range(n_obj):
        c = (cls[k] if cls is not None else int(rng.randint(1, n_cvg)))
        w, h = int(rng.randint(48, 200)), int(rng.randint(48, 200))
        x, y = int(rng.randint(0, W - w)), int(rng.randint(0, W - h))
        gx0, gx1 = x // cell, (x + w - 1) // cell
        gy0, gy1 = y // cell, (y + h - 1) // cell
        for gy in range(gy0, gy1 + 1):
            for gx in range(gx0, gx1 + 1):
                cov[c, gy, gx] = rng.uniform(0.6, 1.0)
                j = rng.normal(0, jitter, size=4)
                ch = 4 * (c - 1)
                bb[ch + 0, gy, gx] = x - gx * cell + j[0]
                bb[ch + 1, gy, gx] = y - gy * cell + j[1]
                bb[ch + 2, gy, gx] = x + w - gx * cell + j[2]
                bb[ch + 3, gy, gx] = y + h - gy * cell + j[3]
    return cov, bb
