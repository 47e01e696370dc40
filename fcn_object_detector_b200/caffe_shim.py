"""A `caffe`-shaped module backed by libfcnod_b200 -- just enough of pycaffe for
the detector node and the Python layer to run unchanged:

    import fcn_object_detector_b200.caffe_shim as caffe
    caffe.set_device(0); caffe.set_mode_gpu()
    net = caffe.Net(proto, weights, caffe.TEST)
    net.blobs['data'].reshape(1, 3, 448, 448)
    net.blobs['data'].data[0][...] = chw
    net.forward()
    cvg = net.blobs['coverage'].data[0]

Engine calls the reference makes (file:line in /root/reference):
  caffe.Net(proto, weights, caffe.TEST)      scripts/fcn_object_detector.py:317
  net.blobs['data'].data.shape / reshape     :324,:328,:199
  net.blobs['data'].data[0][...] = ...       :82
  net.forward()                              :87
  net.blobs[name].data[0]                    :89-90
  caffe.set_device / caffe.set_mode_gpu      :68-69
  caffe.Layer (setup/reshape/forward)        scripts/data_argumentation_layer/data_argumentation_layer.py:14
The script's blob names `pool_score` / `upscore_pool5_bbox` / `score` do not exist in
models/deploy.prototxt; they are accepted as aliases of `coverage` / `bboxes`
(SURVEY.md A1).
"""
import ctypes as C
import numpy as np

from . import _lib

TEST = 1
TRAIN = 0
_state = {'device': 0, 'precision': _lib.FCNOD_PRECISION_BF16}

BLOB_ALIASES = {'pool_score': 'coverage', 'score': 'coverage', 'upscore_pool5_bbox': 'bboxes'}


def set_device(device_id):
    _state['device'] = int(device_id)


def set_mode_gpu():
    pass


def set_mode_cpu():
    raise _lib.FcnodError('fcn_object_detector_b200 has no CPU mode')


def set_precision(name):
    """'bf16' (tcgen05 path, default) or 'fp32' (SIMT parity mode)."""
    _state['precision'] = {'bf16': _lib.FCNOD_PRECISION_BF16, 'fp32': _lib.FCNOD_PRECISION_FP32}[name]


class Blob(object):
    """Host mirror of a blob (NCHW float32).  For net blobs `.data` is refreshed lazily.
    pinned=True keeps `.data` in page-locked memory (fcnod_host_alloc), like Caffe's own SyncedMemory in GPU mode:
    the copies of the layer / detector entry points are then true DMA."""

    def __init__(self, shape=(0,), net=None, name=None, pinned=False):
        self._net = net
        self._name = name
        self._pinned = pinned
        self._data = self._alloc(shape)
        self._stale = False
        self.diff = None

    def _alloc(self, shape):
        if self._pinned:
            a = _lib.pinned_empty(tuple(shape), np.float32)
            a[...] = 0
            return a
        return np.zeros(shape, dtype=np.float32)

    @property
    def data(self):
        if self._stale and self._net is not None:
            self._net._read_blob(self._name, self._data)
            self._stale = False
        return self._data

    @property
    def shape(self):
        return self._data.shape

    @property
    def num(self):
        return self._data.shape[0]

    @property
    def channels(self):
        return self._data.shape[1]

    @property
    def height(self):
        return self._data.shape[2]

    @property
    def width(self):
        return self._data.shape[3]

    def reshape(self, *dims):
        dims = tuple(int(d) for d in dims)
        if self._net is not None:
            if self._name != self._net._input_name:
                raise _lib.FcnodError('only the input blob can be reshaped')
            cur = self._data.shape
            if dims[1:] != cur[1:]:
                raise _lib.FcnodError('only the batch dimension can change (fully convolutional '
                                      'resizing needs a new prototxt)')
            self._net._reshape(dims[0])
        elif dims != self._data.shape:
            self._data = self._alloc(dims)


class Net(object):
    def __init__(self, model_proto, weights, phase=TEST, device=None, precision=None):
        L = _lib.lib()
        self._lib = L
        self._h = C.c_void_p()
        dev = _state['device'] if device is None else device
        prec = _state['precision'] if precision is None else precision
        if isinstance(prec, str):
            prec = {'bf16': _lib.FCNOD_PRECISION_BF16, 'fp32': _lib.FCNOD_PRECISION_FP32}[prec]
        _lib.check(L.fcnod_net_create(str(model_proto).encode(), str(weights).encode(), int(dev), int(prec),
                                      C.byref(self._h)))
        self.device = dev
        self.blobs = {}
        self._names = [L.fcnod_net_blob_name(self._h, i).decode() for i in range(L.fcnod_net_num_blobs(self._h))]
        self._input_name = self._names[0]
        self._make_blobs()

    def _dims(self, name):
        d = (C.c_int * 4)()
        _lib.check(self._lib.fcnod_net_blob_dims(self._h, name.encode(), C.byref(d)))
        return tuple(d)

    def _make_blobs(self):
        class _Blobs(dict):
            def __missing__(s, key):
                if key in BLOB_ALIASES and BLOB_ALIASES[key] in s:
                    return s[BLOB_ALIASES[key]]
                raise KeyError(key)
        blobs = _Blobs()
        for n in self._names:
            blobs[n] = Blob(self._dims(n), self, n)
        self.blobs = blobs

    def _reshape(self, batch):
        _lib.check(self._lib.fcnod_net_reshape(self._h, int(batch)))
        for n, b in self.blobs.items():
            shp = self._dims(n)
            if shp != b._data.shape:
                b._data = np.zeros(shp, dtype=np.float32)
            b._stale = False

    def _sync_after_device_run(self, batch=None):
        """The library ran a forward on its own (fcnod_detect*): resize the host mirrors to the batch the net is at now
        (fcnod_detect runs large batches as two half-batches, so that may not be the caller's batch), mark them stale."""
        d = (C.c_int * 4)()
        _lib.check(self._lib.fcnod_net_input_dims(self._h, d))
        self._reshape(int(d[0]))
        for n, b in self.blobs.items():
            b._stale = True

    def _read_blob(self, name, out):
        _lib.check(self._lib.fcnod_net_blob_read(self._h, name.encode(), _lib.ptr(out), out.size))

    def forward(self):
        x = np.ascontiguousarray(self.blobs[self._input_name]._data, dtype=np.float32)
        _lib.check(self._lib.fcnod_net_forward(self._h, _lib.ptr(x)))
        for n, b in self.blobs.items():
            b._stale = n != self._input_name
        outs = {}
        return outs

    def describe(self):
        buf = C.create_string_buffer(1 << 17)
        _lib.check(self._lib.fcnod_net_describe(self._h, buf, len(buf)))
        return buf.value.decode()

    @property
    def handle(self):
        return self._h

    def close(self):
        if self._h:
            self._lib.fcnod_net_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Layer(object):
    """Base class with pycaffe's Python-layer protocol (setup / reshape / forward / backward)."""
    param_str = ''

    def setup(self, bottom, top):
        pass

    def reshape(self, bottom, top):
        pass

    def forward(self, bottom, top):
        pass

    def backward(self, top, propagate_down, bottom):
        pass
