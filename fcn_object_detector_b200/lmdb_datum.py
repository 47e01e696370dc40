"""Wire format of the two LMDBs that feed `DataArgumentationLayer`'s bottoms.

/root/reference/scripts/misc/create_training_lmdb.py:172-185 writes, per sample, an image `Datum` and a label
`Datum` with `caffe.io.array_to_datum` under the key '{:0>10d}'.format(counter):

  * image:  `im3` after demean_rgb_image, NOT transposed (:172-175) -> an H x W x 3 float array stored as
            channels = H, height = W, width = 3, float_data = H*W*3 values;
  * label:  `box_label` (1, 1, 5) = [x, y, w, h, class] -> channels 1, height 1, width 5, float_data.

Records are decoded by the library (fcnod_datum_decode: uint8 `data` or packed / unpacked `float_data`).  The LMDB
container is read with the `lmdb` package when the host has it (the reference imports it; this image does not) and with
the package's own read-only reader of the published file layout (`lmdb_file.LMDBFile`) otherwise.  `minibatches` is the
pair of Data layers of /root/reference/models/train_val.prototxt:1-26: both databases walked in key order, `batch_size`
records at a time, starting over at the end.
"""
import ctypes as C
import itertools

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


def _records(path):
    """(key, value) records of one LMDB environment in key order: py-lmdb if importable, else lmdb_file.LMDBFile."""
    try:
        import lmdb
    except ImportError:
        lmdb = None
    if lmdb is not None:
        with lmdb.open(path, readonly=True, lock=False) as env, env.begin() as txn:
            for k, v in txn.cursor():
                yield bytes(k), bytes(v)
        return
    from .lmdb_file import LMDBFile
    with LMDBFile(path) as db:
        for k, v in db:
            yield bytes(k), bytes(v)


def read_lmdb(path, limit=None):
    """All (key, value) records of an LMDB directory in key order (the first `limit` of them if given)."""
    out = []
    for kv in _records(path):
        if limit is not None and len(out) >= limit:
            break
        out.append(kv)
    return out


def minibatches(features_path, labels_path, batch_size=1, epochs=None):
    """The `train_data` / `train_label` Data layers of models/train_val.prototxt:1-26 feeding the Python layer of
    README.md:60-76: yields (data_in (N,3,H,W), label (N,1,1,5)) float32 pairs, N = batch_size, both databases read in key
    order and rewound together at the end (Caffe's DataLayer cursor: Next, then SeekToFirst when exhausted).  The two
    databases were written with the same keys (create_training_lmdb.py:161,169,174,184); a mismatch raises.
    `epochs=None` cycles forever."""
    if batch_size < 1:
        raise ValueError('batch_size must be positive')
    imgs, labs, epoch = [], [], 0
    while epochs is None or epoch < epochs:
        n = 0
        fi, li = _records(features_path), _records(labels_path)
        for a, b in itertools.zip_longest(fi, li):
            if a is None or b is None:
                raise ValueError('the feature and label databases hold different numbers of records')
            (ki, vi), (kl, vl) = a, b
            if ki != kl:
                raise ValueError('feature record %r is paired with label record %r' % (ki, kl))
            imgs.append(vi)
            labs.append(vl)
            n += 1
            if len(imgs) == batch_size:
                yield datums_to_bottoms(imgs, labs)
                imgs, labs = [], []
        if n == 0:
            raise ValueError('empty LMDB')
        epoch += 1
