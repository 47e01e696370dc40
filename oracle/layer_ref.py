"""CPU restatement of the training-side augmentation layer -- TEST INFRASTRUCTURE ONLY (oracle).

Follows /root/reference/scripts/data_argumentation_layer/argumentation_engine.py
with the Python-2 / 2017-NumPy semantics it ran under:
  JaccardCoeff.iou                     :24-55
  bounding_box_parameterized_labels    :69-109
  resize_image_and_labels              :114-138
  random_argumentation                 :143-188   (colour aug :176 is photometric only; not restated)
  crop_image_dimension                 :193-236
  flip_image                           :241-267
  generate_box_labels / grid_region    :272-292
  demean_rgb_image (float32)           :297-303
and the layer contract of README.md:60-76 / data_argumentation_layer.py:16-72,101-111.
All random draws are INPUTS (`draws`), made once on the host and fed identically
to this oracle and to the CUDA path.

PARITY UNPINNED: the upstream repository holds no expected outputs for this path.
"""
import math
import numpy as np

from .postnet_ref import resize_bilinear

IOU_THRESH = 0.1      # ArgumentationEngine.FLT_EPSILON_  (:67)


def _f32(v):
    return np.float32(v)


def iou(a, b):
    """:26-35.  a = grid cell [x,y,s,s] (float64), b = rect (ints).  fp32 areas, fp32 divides."""
    x = max(a[0], b[0]); y = max(a[1], b[1])
    w = min(a[0] + a[2], b[0] + b[2]) - x
    h = min(a[1] + a[3], b[1] + b[3]) - y
    if w < 0 or h < 0:
        return 0
    ux = min(a[0], b[0]); uy = min(a[1], b[1])
    uw = max(a[0] + a[2], b[0] + b[2]) - ux
    uh = max(a[1] + a[3], b[1] + b[3]) - uy
    with np.errstate(all='ignore'):
        aub = _f32(uw * uh)
        anb = _f32(w * h)
        area_ratio = _f32(a[2] * a[3]) / _f32(b[2] * b[3])
        score = anb / aub
        score = score / area_ratio
    return score


def grid_region(im_h, im_w, stride):
    """:283-292 (Python-2 integer division)."""
    gh, gw = im_h // stride, im_w // stride
    boxes = np.zeros((gh, gw, 4))
    for j in range(gh):
        for i in range(gw):
            boxes[j][i] = (i * stride, j * stride, stride, stride)
    return boxes


def generate_box_labels(boxes, rect, iou_thresh=IOU_THRESH):
    """:272-278.  The compare mixes np.float32 with a Python float: evaluated in
    double on the reference's NumPy (A10) -> float(score) > 0.1."""
    lab = np.zeros((boxes.shape[0], boxes.shape[1]))
    for j in range(boxes.shape[0]):
        for i in range(boxes.shape[1]):
            if float(iou(boxes[j, i], rect)) > iou_thresh:
                lab[j, i] = 1.0
    return lab


def bounding_box_parameterized_labels(im_h, im_w, stride, num_classes, rects, labels):
    """:69-109 -> (foreground, boxes, size, obj, coverage) as float32 (the blob dtype)."""
    boxes = grid_region(im_h, im_w, stride)
    ch = 4 * num_classes
    gh, gw = boxes.shape[:2]
    fg = np.zeros((num_classes, gh, gw))
    bl = np.zeros((ch, gh, gw)); sl = np.zeros((ch, gh, gw))
    ol = np.zeros((ch, gh, gw)); cl = np.zeros((ch, gh, gw))
    for rect, label in zip(rects, labels):
        rect = [int(v) for v in rect]
        k = int(label * 4)
        region = generate_box_labels(boxes, rect)
        for j in range(gh):
            for i in range(gw):
                if region[j, i] == 1.0:
                    t = boxes[j, i]
                    bl[k + 0, j, i] = rect[0] - t[0]
                    bl[k + 1, j, i] = rect[1] - t[1]
                    bl[k + 2, j, i] = (rect[0] + rect[2]) - t[0]
                    bl[k + 3, j, i] = (rect[1] + rect[3]) - t[1]
                    sl[k + 0, j, i] = 1.0 / rect[2]
                    sl[k + 1, j, i] = 1.0 / rect[3]
                    sl[k + 2, j, i] = 1.0 / rect[2]
                    sl[k + 3, j, i] = 1.0 / rect[3]
                    ol[k:k + 4, j, i] = _f32(t[2] * t[3]) / _f32(rect[2] * rect[3])
                    cl[k:k + 4, j, i] = region[j, i]
                    fg[int(label), j, i] = 1.0
    return tuple(a.astype(np.float32) for a in (fg, bl, sl, ol, cl))


def flip_rect(rect, im_w, im_h, flag):
    """:241-267 rect part (off-by-one kept)."""
    p1 = (rect[0], rect[1]); p2 = (rect[0] + rect[2], rect[1] + rect[3])
    if flag == -1:
        p1 = (im_w - p1[0] - 1, im_h - p1[1] - 1); p2 = (im_w - p2[0] - 1, im_h - p2[1] - 1)
    elif flag == 0:
        p1 = (p1[0], im_h - p1[1] - 1); p2 = (p2[0], im_h - p2[1] - 1)
    elif flag == 1:
        p1 = (im_w - p1[0] - 1, p1[1]); p2 = (im_w - p2[0] - 1, p2[1])
    x = min(p1[0], p2[0]); y = min(p1[1], p2[1])
    w = abs(p2[0] - p1[0]); h = abs(p2[1] - p1[1])
    return [0 if x < 0 else x, 0 if y < 0 else y, w, h]


def flip_image(img, flag):
    """cv.flip semantics: 0 = around x axis (rows reversed), >0 = columns, <0 = both."""
    if flag == 0:
        return img[::-1]
    if flag > 0:
        return img[:, ::-1]
    return img[::-1, ::-1]


def crop_bounds(rect, f1, f2):
    """Upper bounds of the two randint(0, .) draws at :200 -- host-side helper."""
    w = rect[2] * f2 + int(rect[2] * f1)
    h = rect[3] * f2 + int(rect[3] * f1)
    return int(w / 2), int(h / 2)


def crop_window(im_w, im_h, rect, f1, f2, shift_x, shift_y, sign_x, sign_y):
    """:168-169 + :193-236 -> ((x, y, w, h) as the floats/ints the reference holds, new_rect)."""
    rx, ry, rw, rh = [int(v) for v in rect]
    widths = (int(rw * f1), rw * f2)
    heights = (int(rh * f1), rh * f2)
    x = (rx + rw // 2) - widths[0]
    y = (ry + rh // 2) - heights[0]
    w = widths[1] + widths[0]
    h = heights[1] + heights[0]
    cx, cy = rx + rw / 2.0, ry + rh / 2.0
    cx = (cx + shift_x) if sign_x else (cx - shift_x)
    cy = (cy + shift_y) if sign_y else (cy - shift_y)
    nx = int(cx - (w / 2)); ny = int(cy - (h / 2))
    nw = int(w); nh = int(h)
    if nx > x:
        nx = x
        nw -= abs(nx - x)
    if ny > y:
        ny = y
        nh -= abs(ny - y)
    if nx + nw < x + w:
        nx += ((x + w) - (nx + nw))
    if ny + nh < y + h:
        ny += ((y + h) - (ny + nh))
    x, y, w, h = nx, ny, nw, nh
    x = 0 if x < 0 else x
    y = 0 if y < 0 else y
    w = (w - (w + x) - im_w) if x > im_w else w
    h = (h - (h + y) - im_h) if y > im_h else h
    new_rect = [int(rx - x), int(ry - y), rw, rh]
    return (x, y, w, h), new_rect


def demean_f32(im):
    """:297-303 (float32)."""
    im = im.astype(np.float32)
    im[:, :, 0] -= _f32(104.0069879317889)
    im[:, :, 1] -= _f32(116.66876761696767)
    im[:, :, 2] -= _f32(122.6789143406786)
    return (im - im.min()) / (im.max() - im.min())


def resize_rect(rect, src_w, src_h, net_w, net_h):
    """:123-136 (fp32 arithmetic, int truncation)."""
    ratio_x = _f32(src_w) / _f32(net_w)
    ratio_y = _f32(src_h) / _f32(net_h)
    x, y, w, h = (_f32(v) for v in rect)
    xt = x / ratio_x; yt = y / ratio_y
    xb = (x + w) / ratio_x; yb = (y + h) / ratio_y
    return (int(xt), int(yt), int(xb - xt), int(yb - yt))


def augment_one(img_hwc, rect, draw, net_w, net_h, photo=None):
    """random_argumentation (:143-188, one rect) + resize_image_and_labels (:114-138).
    draw = dict(flip, f1, f2, shift_x, shift_y, sign_x, sign_y); photo = optional draws of the colour augmentation (:176,
    oracle.photo_ref), applied to the uint8 crop.  Returns (img float32 HWC, rect)."""
    H, W = img_hwc.shape[:2]
    flag = draw['flip']
    if -2 < flag < 2:
        im = flip_image(img_hwc, flag)
        r = flip_rect(rect, W, H, flag)
    else:
        im = img_hwc
        r = [int(v) for v in rect]
    (x, y, w, h), r = crop_window(W, H, r, draw['f1'], draw['f2'], draw['shift_x'],
                                  draw['shift_y'], draw['sign_x'], draw['sign_y'])
    roi = im[int(y):int(y + h), int(x):int(x + w)]
    if photo is not None and roi.size:
        from . import photo_ref
        roi = photo_ref.apply(np.clip(roi, 0, 255).astype(np.uint8), photo).astype(np.float32)     # :176
    roi = demean_f32(np.ascontiguousarray(roi))
    out = resize_bilinear(roi, net_w, net_h).astype(np.float32)
    r = resize_rect(r, roi.shape[1], roi.shape[0], net_w, net_h)
    return out, r, roi.shape[:2]


def layer_forward(data_in, label, draws, net_w=448, net_h=448, stride=16, num_classes=1, photo=None):
    """README.md:60-76 contract: data_in (N,3,Hs,Ws), label (N,1,1,5)=[x,y,w,h,label].
    Returns the 6 tops (float32): data, coverage-label, bbox-label, size-block, obj-block, coverage-block."""
    N = data_in.shape[0]
    gh, gw = net_h // stride, net_w // stride
    top = [np.zeros((N, 3, net_h, net_w), np.float32), np.zeros((N, num_classes, gh, gw), np.float32)]
    top += [np.zeros((N, 4 * num_classes, gh, gw), np.float32) for _ in range(4)]
    rects_out = []
    for i in range(N):
        img = np.ascontiguousarray(data_in[i].transpose(1, 2, 0))
        lab = np.asarray(label[i]).reshape(-1)
        rect = [int(v) for v in lab[:4]]
        cls = int(lab[4])
        im, r, _ = augment_one(img, rect, draws[i], net_w, net_h, None if photo is None else photo[i])
        fg, bl, sl, ol, cl = bounding_box_parameterized_labels(net_h, net_w, stride, num_classes, [r], [cls])
        top[0][i] = im.transpose(2, 0, 1)
        top[1][i] = fg; top[2][i] = bl; top[3][i] = sl; top[4][i] = ol; top[5][i] = cl
        rects_out.append(r)
    return top, rects_out


def draw_params(rng, im_w, im_h, rect):
    """The reference's draw order (:146,:165-166,:200-202) using a `random.Random`."""
    flip = rng.randint(-1, 2)
    sx = int(math.floor(_f32(im_w) / _f32(rect[2])))
    sy = int(math.floor(_f32(im_h) / _f32(rect[3])))
    f1 = rng.uniform(1.0, float(sx))
    f2 = rng.uniform(1.0, float(sy))
    r = flip_rect(rect, im_w, im_h, flip) if -2 < flip < 2 else list(rect)
    bx, by = crop_bounds(r, f1, f2)
    shift_x, shift_y = rng.randint(0, bx), rng.randint(0, by)
    sign_x, sign_y = rng.randint(0, 1), rng.randint(0, 1)
    return dict(flip=flip, f1=f1, f2=f2, shift_x=shift_x, shift_y=shift_y, sign_x=sign_x, sign_y=sign_y)
