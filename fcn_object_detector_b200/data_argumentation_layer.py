"""`DataArgumentationLayer` -- the Caffe Python layer of the training net, on the B200 path.

Drop-in for module `data_argumentation_layer`, layer `DataArgumentationLayer`
with the README contract (/root/reference/README.md:60-76):
    bottoms  data_in (N,3,Hs,Ws), label (N,1,1,5) = [x, y, w, h, class]
    tops     data, coverage-label, bbox-label, size-block, obj-block, coverage-block
    param_str '448, 448,16, 1'  = image_size_x, image_size_y, stride, num_classes
setup/reshape/forward/backward follow
/root/reference/scripts/data_argumentation_layer/data_argumentation_layer.py:16-55,57-72,75-127
(the checked-in HEAD variant reads private files and takes no bottoms; the label
maths is identical -- SURVEY.md F5/A4).  The random draws are made here, on the
host, in the reference's order (argumentation_engine.py:146,165-166,200-202) with
Python's `random`; everything else (flip, crop, demean, resize, label grids) runs
in libfcnod_b200 (fcnod_aug_forward).  Colour-space augmentation (the imgaug chain at
argumentation_engine.py:176,308-322) is photometric only -- it cannot change the label grids --
and is OFF by default so that the grids and the image top are reproducible from the geometric
draws alone; `layer.color_aug = True` (or passing `photo=` draws to forward) runs it on the device
where the reference has it, between the crop and the demeaning (photometric.py, fcnod_aug_forward_photo).
"""
import ctypes as C
import math
import random

import numpy as np

from . import _lib
from . import caffe_shim as caffe


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


def draw_augmentation(rng, im_w, im_h, rect):
    """The reference's random draws for one single-rect image, in its order."""
    d = _lib.AugDraw()
    d.flip = rng.randint(-1, 2)                                            # :146
    sx = int(math.floor(np.float32(im_w) / np.float32(rect[2])))           # :160-163
    sy = int(math.floor(np.float32(im_h) / np.float32(rect[3])))
    d.f1 = rng.uniform(1.0, float(sx))                                     # :165-166
    d.f2 = rng.uniform(1.0, float(sy))
    r = _flip_rect(rect, im_w, im_h, d.flip) if -2 < d.flip < 2 else list(rect)
    w = r[2] * d.f2 + int(r[2] * d.f1)                                     # :168-169,:196-197
    h = r[3] * d.f2 + int(r[3] * d.f1)
    d.shift_x = rng.randint(0, int(w / 2))                                 # :201
    d.shift_y = rng.randint(0, int(h / 2))
    d.sign_x = rng.randint(0, 1)                                           # :202-203
    d.sign_y = rng.randint(0, 1)
    return d


class DataArgumentationLayer(caffe.Layer):

    def setup(self, bottom, top):
        if len(bottom) != 2:
            raise Exception('This layer needs 2 bottoms: data_in and label')
        if len(top) < 6:
            raise Exception('Current Implementation needs 6 top blobs')
        try:
            plist = self.param_str.split(',')
            self.image_size_x = int(plist[0])
            self.image_size_y = int(plist[1])
            self.stride = int(plist[2])
            self.num_classes = int(plist[3])
        except (ValueError, IndexError):
            raise ValueError('Parameter string missing or data type is wrong!')
        self.rng = random.Random()
        self.rng.seed()
        self.color_aug = getattr(self, 'color_aug', False)
        self.np_rng = np.random.RandomState()          # imgaug draws from numpy's generator (imgaug/imgaug.py:30)
        self.last_draws = None
        self.last_photo = None
        self.last_rects = None

    def reshape(self, bottom, top):
        n_images = bottom[0].num
        out_size_x = int(self.image_size_x // self.stride)
        out_size_y = int(self.image_size_y // self.stride)
        channels = int(self.num_classes * 4)
        top[0].reshape(n_images, 3, self.image_size_y, self.image_size_x)
        top[1].reshape(n_images, self.num_classes, out_size_y, out_size_x)     # coverage-label
        top[2].reshape(n_images, channels, out_size_y, out_size_x)             # bbox-label
        top[3].reshape(n_images, channels, out_size_y, out_size_x)             # size-block
        top[4].reshape(n_images, channels, out_size_y, out_size_x)             # obj-block
        top[5].reshape(n_images, channels, out_size_y, out_size_x)             # coverage-block

    def forward(self, bottom, top, draws=None, photo=None):
        data_in = np.ascontiguousarray(bottom[0].data, dtype=np.float32)
        label = np.ascontiguousarray(np.asarray(bottom[1].data, dtype=np.float32).reshape(data_in.shape[0], -1)[:, :5])
        N, _, src_h, src_w = data_in.shape
        if draws is None:
            draws = [draw_augmentation(self.rng, src_w, src_h, [int(v) for v in label[i, :4]]) for i in range(N)]
        darr = (_lib.AugDraw * N)(*draws)
        if photo is None and self.color_aug:
            from .photometric import draw_photo
            photo = [draw_photo(self.np_rng) for _ in range(N)]
        parr = None
        if photo is not None:
            from .photometric import make_draw
            parr = (_lib.PhotoDraw * N)(*[make_draw(d) for d in photo])
        rects = np.zeros((N, 4), np.int32)
        outs = [top[i].data for i in range(6)]
        for o in outs:
            if o.dtype != np.float32 or not o.flags['C_CONTIGUOUS']:
                raise ValueError('top blobs must be contiguous float32')
        _lib.check(_lib.lib().fcnod_aug_forward_photo(_lib.ptr(data_in), _lib.ptr(label), darr, parr, N, src_h, src_w,
                                                      self.image_size_y, self.image_size_x, self.stride, self.num_classes,
                                                      _lib.ptr(outs[0]), _lib.ptr(outs[1]), _lib.ptr(outs[2]), _lib.ptr(outs[3]),
                                                      _lib.ptr(outs[4]), _lib.ptr(outs[5]), _lib.ptr(rects)))
        self.last_draws = draws
        self.last_photo = photo
        self.last_rects = rects

    def backward(self, top, propagate_down, bottom):
        pass


def label_raster(rects, labels, net_h=448, net_w=448, stride=16, num_classes=1, nrect=None):
    """Label rasteriser alone.  rects (N,R,4) int32 at net resolution, labels (N,R)."""
    rects = np.ascontiguousarray(rects, dtype=np.int32)
    labels = np.ascontiguousarray(labels, dtype=np.int32)
    N, R = labels.shape
    gh, gw = net_h // stride, net_w // stride
    fg = np.zeros((N, num_classes, gh, gw), np.float32)
    outs = [np.zeros((N, 4 * num_classes, gh, gw), np.float32) for _ in range(4)]
    nr = None if nrect is None else np.ascontiguousarray(nrect, dtype=np.int32)
    _lib.check(_lib.lib().fcnod_label_raster(_lib.ptr(rects), _lib.ptr(labels), _lib.ptr(nr), N, R, net_h, net_w, stride,
                                             num_classes, _lib.ptr(fg), _lib.ptr(outs[0]), _lib.ptr(outs[1]),
                                             _lib.ptr(outs[2]), _lib.ptr(outs[3])))
    return [fg] + outs
