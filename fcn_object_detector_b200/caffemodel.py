"""Writer for Caffe `.caffemodel` files (binary protobuf `NetParameter`).

The upstream launch file points at `models/snapshot_iter_1000.caffemodel`
(/root/reference/launch/fcn_object_detector.launch:8) which is not shipped
(/root/reference/.MISSING_LARGE_BLOBS:1).  The product reads such files in C++
(csrc/netdef.cpp); this module only *writes* them, so that tests, smoke and
bench can synthesise weights with the fillers the prototxt declares (xavier
weights, constant bias -- /root/reference/models/deploy.prototxt:38-45).

Wire layout used (caffe.proto): NetParameter{name=1, layer=100};
LayerParameter{name=1, type=2, blobs=7}; BlobProto{data=5 packed float,
shape=7}; BlobShape{dim=1 packed int64}.
"""
import struct
import numpy as np


def _varint(n):
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _ld(field, payload):
    return _varint((field << 3) | 2) + _varint(len(payload)) + payload


RAW_TYPES = {'double': (0, '<f8'), 'float': (1, '<f4'), 'half': (2, '<f2')}     # NVCaffe enum Type [ext]


def _blob(arr, legacy=False, raw=None):
    arr = np.ascontiguousarray(arr, dtype='<f4')
    if raw is not None:       # NVCaffe: shape=7, raw_data_type=10 (varint), raw_data=12 (bytes)
        code, dt = RAW_TYPES[raw]
        shape = _ld(1, b''.join(_varint(d) for d in arr.shape))
        return _ld(7, shape) + _varint((10 << 3) | 0) + _varint(code) + _ld(12, arr.astype(dt).tobytes())
    if legacy:
        dims = list(arr.shape)
        dims = [1] * (4 - len(dims)) + dims
        head = b''.join(_varint((f << 3) | 0) + _varint(d) for f, d in zip((1, 2, 3, 4), dims))
        return head + _ld(5, arr.tobytes())
    shape = _ld(1, b''.join(_varint(d) for d in arr.shape))
    return _ld(7, shape) + _ld(5, arr.tobytes())


def write_caffemodel(path, layers, net_name='fcn_detector', legacy_shape=False, raw=None):
    """layers: iterable of (name, type, [ndarray blobs]).  raw = 'float' | 'half' | 'double' writes the blobs the way
    NVCaffe snapshots do (raw_data_type + raw_data instead of the packed float `data` field)."""
    body = _ld(1, net_name.encode())
    for name, typ, blobs in layers:
        lp = _ld(1, name.encode()) + _ld(2, typ.encode())
        for b in blobs:
            lp += _ld(7, _blob(b, legacy_shape, raw))
        body += _ld(100, lp)
    with open(path, 'wb') as f:
        f.write(body)
    return path


def xavier_weights(conv_shapes, seed=0):
    """Caffe 'xavier' filler: U(-s, s), s = sqrt(3 / fan_in); constant bias."""
    rng = np.random.RandomState(seed)
    out = []
    for name, cout, cin, k, bias in conv_shapes:
        s = np.sqrt(3.0 / (cin * k * k))
        w = rng.uniform(-s, s, size=(cout, cin, k, k)).astype(np.float32)
        b = np.full((cout,), bias, dtype=np.float32)
        out.append((name, 'Convolution', [w, b]))
    return out


def firing_heads(layers, box=(-40.0, -40.0, 56.0, 56.0), cvg_bias=-0.6):
    """Xavier heads give boxes a few pixels tall, which vote_boxes' `height >= 20` test (:346) always drops.  Give the
    regressor a bias that makes every cell propose a box of `box` (x1,y1,x2,y2 relative to its cell) plus the random
    part, and pull the coverage down so that only part of the grid fires: neighbouring proposals then cluster and
    survive, which is what a trained snapshot does.  For tests that need detections to compare."""
    out = []
    for name, typ, blobs in layers:
        w, b = blobs
        if name == 'bbox/regressor':
            b = np.tile(np.asarray(box, np.float32), w.shape[0] // 4)
        elif name == 'cvg/classifier':
            b = np.full_like(b, cvg_bias)
        out.append((name, typ, [w, b]))
    return out


def write_synthetic_caffemodel(path, seed=0, firing=False, **kw):
    from .netspec import conv_shapes
    layers = xavier_weights(conv_shapes(**kw), seed)
    return write_caffemodel(path, firing_heads(layers) if firing else layers)
