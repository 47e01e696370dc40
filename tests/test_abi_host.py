"""CPU: the C-ABI library loads and exports every declared symbol; host-side logic
(prototxt/caffemodel readers, planning, Python shims) -- no compute calls."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from tests.util import ROOT


def declared_symbols():
    text = open(os.path.join(ROOT, 'include', 'fcnod_b200.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(fcnod_[a-z0-9_]+)\s*\(', text)))


def test_library_exports_every_declared_symbol(built_lib):
    from fcn_object_detector_b200 import _lib
    lib = C.CDLL(built_lib)
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(_lib.SIGNATURES) == names          # the ctypes table binds exactly the header
    assert _lib.lib().fcnod_version() == 100


def test_plan_only_handle(built_lib, model_files):
    from fcn_object_detector_b200 import _lib
    L = _lib.lib()
    proto, cm = model_files
    for prec in (_lib.FCNOD_PRECISION_FP32, _lib.FCNOD_PRECISION_BF16):
        h = C.c_void_p()
        _lib.check(L.fcnod_net_create(proto.encode(), cm.encode(), -1, prec, C.byref(h)))
        buf = C.create_string_buffer(1 << 16)
        _lib.check(L.fcnod_net_describe(h, buf, len(buf)))
        plan = buf.value.decode().splitlines()
        assert len(plan) == L.fcnod_net_launches_per_forward(h)
        if prec == _lib.FCNOD_PRECISION_FP32:
            # parity mode: one launch per layer (ReLU / Concat / Dropout folded), every blob materialised, no tensor cores
            assert len(plan) == 77 and not any('conv_tc' in l for l in plan)
            assert sum('conv_simt' in l for l in plan) == 59 and sum(l.split()[1] == 'pool' for l in plan) == 12
        else:
            # deploy mode: every convolution on tcgen05 (conv1 through the space-to-depth fold), fused pool/LRN pairs
            assert not any('conv_simt' in l for l in plan) and len(plan) < 77
            assert any('+s2d+power' in l for l in plan) and any(l.split()[1] == 's2d' for l in plan)
            assert any(l.split()[1] == 'pool+lrn' for l in plan) and any('+lrn(conv2/norm2)' in l for l in plan)
        assert not any('relu' == l.split()[1] for l in plan)          # all 57 ReLUs fused into conv epilogues
        assert not any('concat-copy' in l for l in plan)              # all 9 Concats eliminated
        d = (C.c_int * 4)()
        _lib.check(L.fcnod_net_blob_dims(h, b'coverage', C.byref(d))); assert tuple(d) == (1, 4, 28, 28)
        _lib.check(L.fcnod_net_blob_dims(h, b'bboxes', C.byref(d))); assert tuple(d) == (1, 16, 28, 28)
        _lib.check(L.fcnod_net_blob_dims(h, b'inception_5b/output', C.byref(d))); assert tuple(d) == (1, 1024, 28, 28)
        # compute on a plan-only handle must fail loudly -- there is no CPU fallback
        x = np.zeros((1, 3, 448, 448), np.float32)
        assert L.fcnod_net_forward(h, _lib.ptr(x)) != 0
        assert b'no CPU fallback' in L.fcnod_last_error()
        _lib.check(L.fcnod_net_destroy(h))


def test_create_errors(built_lib, model_files, tmp_path):
    from fcn_object_detector_b200 import _lib
    L = _lib.lib()
    proto, cm = model_files
    h = C.c_void_p()
    assert L.fcnod_net_create(b'/nonexistent.prototxt', cm.encode(), -1, 1, C.byref(h)) != 0
    assert b'cannot open prototxt' in L.fcnod_last_error()
    assert L.fcnod_net_create(proto.encode(), b'/nonexistent.caffemodel', -1, 1, C.byref(h)) != 0
    bad = tmp_path / 'bad.prototxt'
    bad.write_text('input: "data"\ninput_shape { dim: 1 dim: 3 dim: 8 dim: 8 }\nlayer { name: "x" type: "Softmax" bottom: "data" top: "y" }\n')
    assert L.fcnod_net_create(str(bad).encode(), cm.encode(), -1, 1, C.byref(h)) != 0
    assert b'unsupported type' in L.fcnod_last_error()
    trunc = tmp_path / 'trunc.caffemodel'
    trunc.write_bytes(open(cm, 'rb').read()[:1000])
    assert L.fcnod_net_create(proto.encode(), str(trunc).encode(), -1, 1, C.byref(h)) != 0
    # weights with a wrong shape
    from fcn_object_detector_b200.caffemodel import write_caffemodel
    wrong = write_caffemodel(str(tmp_path / 'wrong.caffemodel'), [('conv1/7x7_s2', 'Convolution', [np.zeros((64, 3, 5, 5), np.float32), np.zeros(64, np.float32)])])
    assert L.fcnod_net_create(proto.encode(), wrong.encode(), -1, 1, C.byref(h)) != 0
    assert b'wrong size' in L.fcnod_last_error()
    assert L.fcnod_net_create(proto.encode(), cm.encode(), -1, 7, C.byref(h)) != 0


def test_cpp_caffemodel_reader_accepts_legacy_shape(built_lib, model_files, tmp_path):
    """V2 layer + legacy num/channels/height/width blob header (SURVEY.md hard part 7)."""
    from fcn_object_detector_b200 import _lib
    from fcn_object_detector_b200.caffemodel import write_caffemodel, xavier_weights
    from fcn_object_detector_b200.netspec import conv_shapes
    proto, _ = model_files
    p = write_caffemodel(str(tmp_path / 'legacy.caffemodel'), xavier_weights(conv_shapes(), 0), legacy_shape=True)
    h = C.c_void_p()
    _lib.check(_lib.lib().fcnod_net_create(proto.encode(), p.encode(), -1, 1, C.byref(h)))
    _lib.lib().fcnod_net_destroy(h)


def test_missing_library_fails_loudly(monkeypatch):
    from fcn_object_detector_b200 import _lib
    monkeypatch.setattr(_lib, '_lib', None)
    monkeypatch.setattr(_lib, 'LIB_PATH', '/nonexistent/libfcnod_b200.so')
    with pytest.raises(_lib.FcnodError):
        _lib.lib()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, 'fcn_object_detector_b200')
    for dp, dn, fn in os.walk(pkg):
        for f in fn:
            if f.endswith(('.py', '.cu', '.cuh', '.cpp', '.h')):
                text = open(os.path.join(dp, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', text, flags=re.M), f
                assert 'oracle/' not in text or f == 'build.py', f


def test_shim_net_blobs_and_aliases(built_lib, model_files):
    import fcn_object_detector_b200.caffe_shim as caffe
    proto, cm = model_files
    net = caffe.Net(proto, cm, caffe.TEST, device=-1)
    assert net.blobs['data'].data.shape == (1, 3, 448, 448)
    assert net.blobs['pool_score'] is net.blobs['coverage']
    assert net.blobs['upscore_pool5_bbox'] is net.blobs['bboxes'] and net.blobs['score'].data.shape[1] == 4
    with pytest.raises(KeyError):
        net.blobs['nope']
    with pytest.raises(Exception):
        net.blobs['data'].reshape(1, 3, 224, 224)
    with pytest.raises(Exception):
        caffe.set_mode_cpu()


class _B(object):
    def __init__(self, arr):
        self.data = arr
    num = property(lambda s: s.data.shape[0])

    def reshape(self, *d):
        self.data = np.zeros(d, np.float32)


def test_layer_setup_reshape_contract():
    from fcn_object_detector_b200.data_argumentation_layer import DataArgumentationLayer
    L = DataArgumentationLayer()
    L.param_str = '448, 448,16, 1'
    bottom = [_B(np.zeros((32, 3, 480, 640), np.float32)), _B(np.zeros((32, 1, 1, 5), np.float32))]
    top = [_B(np.zeros((1,), np.float32)) for _ in range(6)]
    L.setup(bottom, top)
    assert (L.image_size_x, L.image_size_y, L.stride, L.num_classes) == (448, 448, 16, 1)
    L.reshape(bottom, top)
    assert top[0].data.shape == (32, 3, 448, 448) and top[1].data.shape == (32, 1, 28, 28)
    assert all(t.data.shape == (32, 4, 28, 28) for t in top[2:])
    with pytest.raises(Exception, match='6 top blobs'):
        L.setup(bottom, top[:5])
    L.param_str = '448, x'
    with pytest.raises(ValueError, match='Parameter string missing or data type is wrong!'):
        L.setup(bottom, top)
    L.backward(top, None, bottom)


def test_host_draws_match_oracle_draws():
    import random
    from fcn_object_detector_b200.data_argumentation_layer import draw_augmentation
    from oracle.layer_ref import draw_params
    a, b = random.Random(7), random.Random(7)
    rs = np.random.RandomState(0)
    for i in range(200):
        w, h = int(rs.randint(32, 257)), int(rs.randint(32, 257))
        rect = [int(rs.randint(0, 640 - w)), int(rs.randint(0, 480 - h)), w, h]
        d, o = draw_augmentation(a, 640, 480, rect), draw_params(b, 640, 480, rect)
        assert (d.flip, d.f1, d.f2, d.shift_x, d.shift_y, d.sign_x, d.sign_y) == \
            (o['flip'], o['f1'], o['f2'], o['shift_x'], o['shift_y'], o['sign_x'], o['sign_y'])


def test_detector_without_files_is_inert():
    from fcn_object_detector_b200.detector import FCNObjectDetector
    d = FCNObjectDetector(pretrained_weights=None, deployment_prototxt=None)
    assert d.net is None and d.labels == []


# ----------------------------------------------------------------------------- on-disk formats
def _cpp_blob(path, layer, index):
    from fcn_object_detector_b200 import _lib
    L = _lib.lib()
    dims = (C.c_int * 4)(); nd = C.c_int()
    _lib.check(L.fcnod_caffemodel_blob(path.encode(), layer.encode(), index, C.byref(dims), C.byref(nd), None, 0))
    out = np.empty(tuple(dims)[:nd.value], np.float32)
    _lib.check(L.fcnod_caffemodel_blob(path.encode(), layer.encode(), index, C.byref(dims), C.byref(nd), _lib.ptr(out), out.size))
    return out


def test_caffemodel_nvcaffe_raw_data(built_lib, model_files, tmp_path):
    """NVCaffe snapshots (the reference trained with NVCaffe: train/train.sh:19) store blobs as raw_data_type + raw_data.
    float32 / IEEE half / float64 payloads: the C++ loader, the oracle's reader and the written arrays agree."""
    from fcn_object_detector_b200 import _lib
    from fcn_object_detector_b200.caffemodel import write_caffemodel, xavier_weights
    from fcn_object_detector_b200.netspec import conv_shapes
    from oracle.caffemodel_io import read_caffemodel
    proto, _ = model_files
    layers = xavier_weights(conv_shapes(), 3)
    # include values that exercise half subnormals / exact halves
    layers[0][2][1][:6] = [0.0, -0.0, 6.1035e-05, 5.96e-08, 65504.0, -1.5]
    for raw, tol in (('float', 0.0), ('double', 0.0), ('half', None)):
        p = write_caffemodel(str(tmp_path / ('raw_%s.caffemodel' % raw)), layers, raw=raw)
        ref = read_caffemodel(p)
        for name, _t, blobs in (layers[0], layers[7], layers[-1]):
            for i, arr in enumerate(blobs):
                got = _cpp_blob(p, name, i)
                want = arr.astype('<f2').astype(np.float32) if raw == 'half' else arr
                assert got.shape == arr.shape
                assert np.array_equal(got.view(np.uint32), np.ascontiguousarray(want).view(np.uint32)), (raw, name, i)
                assert np.array_equal(ref[name][i], want)
        h = C.c_void_p()
        _lib.check(_lib.lib().fcnod_net_create(proto.encode(), p.encode(), -1, 1, C.byref(h)))     # and the planner takes it
        _lib.lib().fcnod_net_destroy(h)
    # a declared type that contradicts the payload width is refused, not mis-read
    good = open(str(tmp_path / 'raw_float.caffemodel'), 'rb').read()
    bad = good.replace(bytes([10 << 3, 1]), bytes([10 << 3, 2]), 1)
    (tmp_path / 'bad_type.caffemodel').write_bytes(bad)
    h = C.c_void_p()
    assert _lib.lib().fcnod_net_create(proto.encode(), str(tmp_path / 'bad_type.caffemodel').encode(), -1, 1, C.byref(h)) != 0
    assert b'raw_data_type' in _lib.lib().fcnod_last_error()


def test_prototxt_unsupported_params_are_refused(built_lib, model_files, tmp_path):
    """dilation / group / per-axis kernels would silently compute something else: the loader must say so."""
    from fcn_object_detector_b200 import _lib
    L = _lib.lib()
    _, cm = model_files
    head = 'input: "data"\ninput_shape { dim: 1 dim: 3 dim: 8 dim: 8 }\n'
    conv = 'layer { name: "c" type: "Convolution" bottom: "data" top: "c" convolution_param { num_output: 8 kernel_size: 3 %s } }\n'
    pool = 'layer { name: "p" type: "Pooling" bottom: "data" top: "p" pooling_param { pool: MAX kernel_size: 3 %s } }\n'
    for body, word in ((conv % 'dilation: 2', b'dilation'), (conv % 'group: 2', b'group'), (conv % 'kernel_h: 3', b'kernel_h'),
                       (conv % 'pad_w: 1', b'pad_w'), (pool % 'stride_h: 2', b'stride_h'), (pool % 'global_pooling: true', b'global_pooling')):
        f = tmp_path / 'bad.prototxt'
        f.write_text(head + body)
        h = C.c_void_p()
        assert L.fcnod_net_create(str(f).encode(), cm.encode(), -1, 1, C.byref(h)) != 0
        assert word in L.fcnod_last_error(), (word, L.fcnod_last_error())


def _datum_bytes(arr, label=None, packed=False):
    """caffe.io.array_to_datum(arr, label).SerializeToString(), restated: uint8 -> data (4), else float_data (6)."""
    from fcn_object_detector_b200.caffemodel import _ld, _varint
    import struct
    c, h, w = arr.shape
    out = b''.join(_varint((f << 3) | 0) + _varint(v) for f, v in ((1, c), (2, h), (3, w)))
    if arr.dtype == np.uint8:
        out += _ld(4, arr.tobytes())
    elif packed:
        out += _ld(6, arr.astype('<f4').tobytes())
    else:       # proto2 `repeated float` without [packed = true]: one tag + 4 bytes per element
        out += b''.join(bytes([(6 << 3) | 5]) + struct.pack('<f', float(v)) for v in arr.astype(np.float32).flat)
    if label is not None:
        out += _varint((5 << 3) | 0) + _varint(label)
    return out


def test_lmdb_datum_records_decode_into_the_layer_bottoms(built_lib):
    """create_training_lmdb.py:172-185: image datum = float H x W x 3 stored un-transposed, label datum = (1,1,5)."""
    from fcn_object_detector_b200.lmdb_datum import datums_to_bottoms, decode_datum, image_to_chw
    rs = np.random.RandomState(0)
    imgs = [rs.uniform(0, 1, (12, 16, 3)).astype(np.float32) for _ in range(3)]
    labs = [np.asarray([[[4 + i, 5, 6, 7, 0]]], np.float64) for i in range(3)]
    for packed in (False, True):
        data_in, label = datums_to_bottoms([_datum_bytes(a, packed=packed) for a in imgs], [_datum_bytes(l, packed=packed) for l in labs])
        assert data_in.shape == (3, 3, 12, 16) and label.shape == (3, 1, 1, 5)
        for i in range(3):
            assert np.array_equal(data_in[i], imgs[i].transpose(2, 0, 1))
            assert np.array_equal(label[i, 0, 0], labs[i].reshape(-1).astype(np.float32))
    # uint8 data + label field; an already C x H x W datum passes through
    u8 = rs.randint(0, 256, (3, 10, 14)).astype(np.uint8)
    arr, lab = decode_datum(_datum_bytes(u8, label=2))
    assert lab == 2 and np.array_equal(arr, u8.astype(np.float32)) and image_to_chw(arr) is arr
    # malformed records fail loudly
    from fcn_object_detector_b200._lib import FcnodError
    with pytest.raises(FcnodError):
        decode_datum(_datum_bytes(u8)[:-5])
    with pytest.raises(FcnodError):
        decode_datum(_datum_bytes(u8) + bytes([(7 << 3) | 0, 1]))        # encoded = true


def test_lane_plan_orders_every_dependency(built_lib, model_files):
    """The forward runs independent launches on three lanes (streams / graph branches) joined by event waits.  Structural
    check on the plan-only handle: for every launch, each launch it depends on through a buffer region is ordered before it
    -- by lane order, by one of its waits, or transitively -- and the inception branches really are spread over the lanes."""
    from fcn_object_detector_b200 import _lib
    L = _lib.lib()
    proto, cm = model_files
    h = C.c_void_p()
    _lib.check(L.fcnod_net_create(proto.encode(), cm.encode(), -1, 1, C.byref(h)))
    n = L.fcnod_net_launches_per_forward(h)
    lanes, deps, waits = [], [], []
    for i in range(n):
        lane, nd, nw = C.c_int(), C.c_int(), C.c_int()
        d = (C.c_int * 64)(); w = (C.c_int * 8)()
        _lib.check(L.fcnod_net_step_lane(h, i, C.byref(lane), d, 64, C.byref(nd), w, 8, C.byref(nw)))
        assert nd.value <= 64 and nw.value <= 8
        lanes.append(lane.value); deps.append(list(d[:nd.value])); waits.append(list(w[:nw.value]))
    done_before = []            # launches guaranteed complete when launch i starts
    last_on = {}
    for i in range(n):
        s = set()
        if lanes[i] in last_on:
            p = last_on[lanes[i]]
            s |= done_before[p] | {p}
        for w in waits[i]:
            assert w < i and lanes[w] != lanes[i]
            s |= done_before[w] | {w}
        for d in deps[i]:
            assert d in s, 'launch %d depends on %d, which nothing orders before it' % (i, d)
        done_before.append(s)
        last_on[lanes[i]] = i
    assert set(lanes) == {0, 1, 2}
    desc = (C.c_char * (1 << 17))()
    _lib.check(L.fcnod_net_describe(h, desc, len(desc)))
    plan = desc.value.decode().splitlines()
    # the critical chain (fused 1x1 -> 3x3) of every module stays on lane 0; 5x5 and pool_proj hang off on lanes 1 and 2
    for i, line in enumerate(plan):
        if '/3x3 3x3/1' in line or '/1x1 + ' in line:
            assert lanes[i] == 0, line
        if '/5x5 5x5/1' in line:
            assert lanes[i] == 1, line
        if '+maxpool3x3' in line:
            assert lanes[i] == 2, line
    L.fcnod_net_destroy(h)
