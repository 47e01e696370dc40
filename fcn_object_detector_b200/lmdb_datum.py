"""Wire format of the two LMDBs that feed `DataArgumentationLayer`'s bottoms.

/root/reference/scripts/misc/create_training_lmdb.py:172-185 writes, per sample, an image `Datum` and a label
`Datum` with `caffe.io.array_to_datum` under the key '{:0>10d}'.format(counter):

  * image:  `im3` after demean_rgb_image, NOT transposed (:172-175) -> an H x W x 3 float array stored as
            channels = H, height = W, width = 3, float_data = H*W*3 values;
  * label:  `box_label` (1, 1, 5) = [x, y, w, h, class] -> channels 1, height 1, width 5, float_data.

Records are decoded by the library (fcnod_datum_decode: uint8 `data` or packed / unpacked `float_data`).  Reading the
LMDB container itself needs the `lmdb` package, which the training host has (the reference imports it) and this image
does not: `read_lmdb` imports it lazily.
"""
import ctypes as C

import numpy as np

from . import _lib


def decode_datum(buf):
    """bytes of one serialized Datum -> (float32 array shaped (channels, height, width) as stored, label or None)."""
    buf = bytes(buf)
    raw = (C.c_uint8 * len(buf)).from_buffer_copy(buf)
    dims = (C.c_int * 3)()
    label = C.c_int(-1)
    L = _lib.lib()
    _lib.check(L.fcnod_datum_decode(raw, len(buf), C.byref(dims), C.byref(label), None, 0))
    out = np.empty(tuple(dims), np.float32)
    _lib.check(L.fcnod_datum_decode(raw, len(buf), C.byref(dims), C.byref(label), _lib.ptr(out), out.size))
    return out, (None if label.value < 0 else int(label.value))


def image_to_chw(arr):
    """The image datum as the layer's `data_in` bottom wants it: (3, H, W).  create_training_lmdb.py stores H x W x 3
    un-transposed (channels = H, height = W, width = 3); a datum already stored C x H x W passes through."""
    if arr.ndim != 3:
        raise ValueError('image datum must be 3-d')
    if arr.shape[2] in (1, 3) and arr.shape[0] > 4:
        return np.ascontiguousarray(arr.transpose(2, 0, 1))
    return arr


def datums_to_bottoms(image_datums, label_datums):
    """Serialized image / label records of one minibatch -> (data_in (N,3,H,W) float32, label (N,1,1,5) float32)."""
    imgs = [image_to_chw(decode_datum(b)[0]) for b in image_datums]
    labs = [decode_datum(b)[0].reshape(-1)[:5] for b in label_datums]
    if len(imgs) != len(labs):
        raise ValueError('as many label records as image records are needed')
    return np.stack(imgs).astype(np.float32), np.stack(labs).astype(np.float32).reshape(len(labs), 1, 1, 5)


def read_lmdb(path, limit=None):
    """All (key, value) records of an LMDB directory in key order (needs the `lmdb` package)."""
    try:
        import lmdb
    except ImportError:
        raise _lib.FcnodError('reading an LMDB container needs the `lmdb` package (not installed here); '
                              'records already in memory decode with decode_datum')
    out = []
    with lmdb.open(path, readonly=True, lock=False) as env, env.begin() as txn:
        for k, v in txn.cursor():
            out.append((bytes(k), bytes(v)))
            if limit is not None and len(out) >= limit:
                break
    return out
