"""`.caffemodel` reader (protobuf wire format) -- TEST INFRASTRUCTURE ONLY (CPU oracle).

Independent of the product's C++ reader.  Accepts V2 `layer` (field 100) and V1
`layers` (field 2) messages, BlobProto.data packed or unpacked, `shape` or
legacy num/channels/height/width, `double_data` (field 8), and NVCaffe's
`raw_data_type` (field 10) / `raw_data` (field 12) [ext: NVCaffe caffe.proto].
Call site being restated: caffe.Net(proto, weights, caffe.TEST),
/root/reference/scripts/fcn_object_detector.py:317.
"""
import struct
import numpy as np


def _fields(buf):
    i, n = 0, len(buf)
    while i < n:
        key = 0
        shift = 0
        while True:
            b = buf[i]; i += 1
            key |= (b & 0x7F) << shift
            shift += 7
            if not b & 0x80:
                break
        field, wt = key >> 3, key & 7
        if wt == 0:
            v = 0; shift = 0
            while True:
                b = buf[i]; i += 1
                v |= (b & 0x7F) << shift
                shift += 7
                if not b & 0x80:
                    break
            yield field, wt, v
        elif wt == 1:
            yield field, wt, buf[i:i + 8]; i += 8
        elif wt == 2:
            ln = 0; shift = 0
            while True:
                b = buf[i]; i += 1
                ln |= (b & 0x7F) << shift
                shift += 7
                if not b & 0x80:
                    break
            yield field, wt, buf[i:i + ln]; i += ln
        elif wt == 5:
            yield field, wt, buf[i:i + 4]; i += 4
        else:
            raise ValueError('caffemodel: unsupported wire type %d' % wt)


def _varints(buf):
    out = []
    v = 0; shift = 0
    for b in buf:
        v |= (b & 0x7F) << shift
        shift += 7
        if not b & 0x80:
            out.append(v); v = 0; shift = 0
    return out


def _blob(buf):
    legacy = {}
    shape = None
    chunks = []
    dchunks = []
    raw, raw_type = None, None
    for f, wt, v in _fields(buf):
        if f == 10 and wt == 0:
            raw_type = v
            continue
        if f == 12 and wt == 2:
            raw = bytes(v)
            continue
        if f in (1, 2, 3, 4) and wt == 0:
            legacy[f] = v
        elif f == 5:
            chunks.append(bytes(v))
        elif f == 8:
            dchunks.append(bytes(v))
        elif f == 7:
            for f2, wt2, v2 in _fields(v):
                if f2 == 1:
                    shape = (shape or []) + (_varints(v2) if wt2 == 2 else [v2])
    if shape is None:
        shape = [legacy.get(k, 1) for k in (1, 2, 3, 4)]
    if chunks:
        data = np.frombuffer(b''.join(chunks), dtype='<f4').astype(np.float32)
    elif dchunks:
        data = np.frombuffer(b''.join(dchunks), dtype='<f8').astype(np.float32)
    else:       # NVCaffe raw storage: DOUBLE = 0, FLOAT = 1, FLOAT16 = 2
        dt = {0: '<f8', 1: '<f4', 2: '<f2'}[raw_type]
        data = np.frombuffer(raw, dtype=dt).astype(np.float32)
    return data.reshape(shape)


def read_caffemodel(path):
    """-> {layer name: [ndarray, ...]} for layers that carry blobs."""
    with open(path, 'rb') as f:
        buf = memoryview(f.read())
    out = {}
    for f, wt, v in _fields(buf):
        if f == 100 and wt == 2:      # LayerParameter: name=1, blobs=7
            name, blobs = None, []
            for f2, wt2, v2 in _fields(v):
                if f2 == 1:
                    name = bytes(v2).decode()
                elif f2 == 7:
                    blobs.append(_blob(v2))
            if blobs:
                out[name] = blobs
        elif f == 2 and wt == 2:      # V1LayerParameter: name=4, blobs=6
            name, blobs = None, []
            for f2, wt2, v2 in _fields(v):
                if f2 == 4:
                    name = bytes(v2).decode()
                elif f2 == 6:
                    blobs.append(_blob(v2))
            if blobs:
                out[name] = blobs
    return out
