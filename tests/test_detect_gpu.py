"""GPU: the detector's predict-on-image entry point (prelude + net + post-net) through the C ABI."""
import numpy as np
import pytest

from oracle import net_ref, postnet_ref
from tests.util import rel_err

pytestmark = pytest.mark.gpu


def test_preprocess_matches_oracle():
    from fcn_object_detector_b200.detector import preprocess
    rs = np.random.RandomState(0)
    for shape in ((448, 448), (480, 640), (300, 200), (301, 203)):     # the last: frames off the 16-byte grid
        imgs = rs.randint(0, 256, (3,) + shape + (3,)).astype(np.uint8)
        got = preprocess(imgs)
        for b in range(3):
            exp = postnet_ref.preprocess(imgs[b], 448, 448)
            assert np.abs(got[b] - exp).max() < 1e-6, shape
    # identity size: exact
    imgs = rs.randint(0, 256, (1, 448, 448, 3)).astype(np.uint8)
    assert np.array_equal(preprocess(imgs)[0], postnet_ref.preprocess(imgs[0], 448, 448))


def _planted_head_weights(cm_path, tmp_path):
    """Synthetic caffemodel whose heads fire: random-init heads give coverage ~0.1-0.7 everywhere with
    incoherent boxes; scale the coverage head so some cells clear 0.5 and clusters form."""
    return cm_path


def test_run_detector_config1_fp32(model_files):
    """BASELINE config 1: batch 1, 448x448 synthetic BGR frame, synthetic weights seed 0, thresholds 0.5/3/0.2.
    fp32 mode; the GPU's detections must equal the oracle post-net applied to the GPU's own coverage/bboxes
    (bit-exact), and those maps must match the oracle net to 1e-5."""
    from fcn_object_detector_b200.detector import FCNObjectDetector
    proto, cm = model_files
    img = np.random.RandomState(0).randint(0, 256, (448, 448, 3)).astype(np.uint8)
    det = FCNObjectDetector(pretrained_weights=cm, deployment_prototxt=proto, precision='fp32')
    boxes, labels = det.predict(img)
    net = det.net
    cov = net.blobs['pool_score'].data[0].copy(); bb = net.blobs['upscore_pool5_bbox'].data[0].copy()
    ob, ol = postnet_ref.postnet(cov, bb, img.shape)
    assert np.array_equal(boxes, ob) and np.array_equal(labels, ol)
    from oracle import prototxt
    from oracle.caffemodel_io import read_caffemodel
    x = postnet_ref.preprocess(img, 448, 448)[None]
    exp = net_ref.forward(prototxt.layers(prototxt.load(proto)), read_caffemodel(cm), x, want=['coverage', 'bboxes'])
    assert rel_err(cov, exp['coverage'][0]) < 1e-5 and rel_err(bb, exp['bboxes'][0]) < 1e-5
    # run_detector returns None when nothing is detected, like the reference (:128-130)
    r = det.run_detector(img)
    assert (r is None) == (len(ob) == 0)


def test_detect_batch_bf16_consistency(model_files):
    from fcn_object_detector_b200.detector import FCNObjectDetector
    proto, cm = model_files
    rs = np.random.RandomState(3)
    imgs = rs.randint(0, 256, (4, 480, 640, 3)).astype(np.uint8)
    det = FCNObjectDetector(pretrained_weights=cm, deployment_prototxt=proto, precision='bf16', detection_threshold=0.45,
                            min_boxes=1)
    res = det.predict_batch(imgs)
    net = det.net
    cov = net.blobs['coverage'].data.copy(); bb = net.blobs['bboxes'].data.copy()
    for b in range(4):
        ob, ol = postnet_ref.postnet(cov[b], bb[b], imgs[b].shape, prob_thresh=0.45, min_boxes=1)
        assert np.array_equal(res[b][0], ob) and np.array_equal(res[b][1], ol)


def test_predict_stream_equals_predict_batch(firing_model_files):
    """Streaming submit/collect (two batches in flight, upload overlapped with compute) returns, batch for batch,
    exactly what the one-call path returns -- including across a batch-size change and a source-size change."""
    from fcn_object_detector_b200.detector import FCNObjectDetector
    proto, cm = firing_model_files
    rs = np.random.RandomState(11)
    batches = [rs.randint(0, 256, (4, 480, 640, 3)).astype(np.uint8) for _ in range(4)]
    batches += [rs.randint(0, 256, (2, 448, 448, 3)).astype(np.uint8) for _ in range(3)]
    batches += [rs.randint(0, 256, (2, 300, 200, 3)).astype(np.uint8)]
    det = FCNObjectDetector(pretrained_weights=cm, deployment_prototxt=proto, precision='bf16', detection_threshold=0.45,
                            min_boxes=1)
    want = [det.predict_batch(b) for b in batches]
    got = list(det.predict_stream(iter(batches)))
    assert len(got) == len(want)
    ndet = 0
    for g, w in zip(got, want):
        assert len(g) == len(w)
        for (gb, gl), (wb, wl) in zip(g, w):
            assert np.array_equal(gb, wb) and np.array_equal(gl, wl)
            ndet += len(wb)
    assert ndet >= len(batches)          # the biased heads fire on every frame
    # the one-call path still works after streaming, and an abandoned stream leaves nothing in flight
    it = det.predict_stream(iter(batches[:3]))
    next(it)
    it.close()
    again = det.predict_batch(batches[0])
    assert all(np.array_equal(a[0], b[0]) for a, b in zip(again, want[0]))


def test_detect_submit_collect_errors(model_files):
    import ctypes as C
    from fcn_object_detector_b200 import _lib
    from fcn_object_detector_b200.detector import FCNObjectDetector
    proto, cm = model_files
    det = FCNObjectDetector(pretrained_weights=cm, deployment_prototxt=proto, precision='bf16')
    L = _lib.lib()
    h = det.net.handle
    p = _lib.default_postnet_params()
    imgs = np.zeros((2, 448, 448, 3), np.uint8)
    M = _lib.FCNOD_MAX_DET
    out = [np.zeros((2, M, 5), np.int32), np.zeros((2, M), np.int32), np.zeros((2, M), np.int32), np.zeros((2,), np.int32)]
    ptrs = [_lib.ptr(o) for o in out]
    t = [C.c_int32(-1) for _ in range(3)]
    assert L.fcnod_detect_collect(h, 0, *ptrs) != 0                       # nothing in flight
    assert L.fcnod_detect_submit(h, _lib.ptr(imgs), 2, 448, 448, C.byref(p), C.byref(t[0])) == 0
    assert L.fcnod_detect_submit(h, _lib.ptr(imgs), 2, 448, 448, C.byref(p), C.byref(t[1])) == 0
    assert L.fcnod_detect_submit(h, _lib.ptr(imgs), 2, 448, 448, C.byref(p), C.byref(t[2])) != 0   # a third one
    assert b'in flight' in L.fcnod_last_error()
    assert L.fcnod_detect_collect(h, t[1].value, *ptrs) != 0              # out of order
    assert L.fcnod_detect_collect(h, t[0].value, *ptrs) == 0
    imgs4 = np.zeros((4, 448, 448, 3), np.uint8)
    assert L.fcnod_detect_submit(h, _lib.ptr(imgs4), 4, 448, 448, C.byref(p), C.byref(t[2])) != 0  # batch change in flight
    assert L.fcnod_detect_collect(h, t[1].value, *ptrs) == 0


def test_run_detector_firing_heads_matches_oracle(firing_model_files):
    """fp32, heads that fire: boxes and labels equal the oracle post-net on the GPU's own maps, non-empty, at a
    source size that exercises resize_detection; the maps match the oracle net to 1e-5."""
    from fcn_object_detector_b200.detector import FCNObjectDetector
    from oracle import prototxt
    from oracle.caffemodel_io import read_caffemodel
    proto, cm = firing_model_files
    img = np.random.RandomState(5).randint(0, 256, (480, 640, 3)).astype(np.uint8)
    det = FCNObjectDetector(pretrained_weights=cm, deployment_prototxt=proto, precision='fp32', detection_threshold=0.45)
    boxes, labels = det.predict(img)
    cov = det.net.blobs['coverage'].data[0].copy(); bb = det.net.blobs['bboxes'].data[0].copy()
    ob, ol = postnet_ref.postnet(cov, bb, img.shape, prob_thresh=0.45)
    assert len(ob) > 0
    assert np.array_equal(boxes, ob) and np.array_equal(labels, ol)
    x = postnet_ref.preprocess(img, 448, 448)[None]
    exp = net_ref.forward(prototxt.layers(prototxt.load(proto)), read_caffemodel(cm), x, want=['coverage', 'bboxes'])
    assert rel_err(cov, exp['coverage'][0]) < 1e-5 and rel_err(bb, exp['bboxes'][0]) < 1e-5
    eb, el = postnet_ref.postnet(exp['coverage'][0], exp['bboxes'][0], img.shape, prob_thresh=0.45)
    assert len(eb) == len(ob)


def test_predict_stream_first_call_in_fresh_process(firing_model_files):
    """fcnod_detect_submit captures forward + post-net into a graph on first use: it must work when it is the very
    first post-net call of the process (per-device post-net set-up happens before the capture starts)."""
    import os
    import subprocess
    import sys
    from tests.util import ROOT
    proto, cm = firing_model_files
    code = (
        "import sys, numpy as np\n"
        "sys.path.insert(0, %r)\n"
        "from fcn_object_detector_b200.detector import FCNObjectDetector\n"
        "det = FCNObjectDetector(pretrained_weights=%r, deployment_prototxt=%r, precision='bf16', detection_threshold=0.45, min_boxes=1)\n"
        "rs = np.random.RandomState(11)\n"
        "batches = [rs.randint(0, 256, (2, 448, 448, 3)).astype(np.uint8) for _ in range(3)]\n"
        "got = list(det.predict_stream(iter(batches)))\n"
        "want = [det.predict_batch(b) for b in batches]\n"
        "assert all(np.array_equal(g[0], w[0]) for gg, ww in zip(got, want) for g, w in zip(gg, ww))\n"
        "print('stream-first ok', sum(len(g[0]) for gg in got for g in gg))\n" % (ROOT, cm, proto))
    r = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, FCNOD_AUTOTUNE='0'))
    assert r.returncode == 0 and 'stream-first ok' in r.stdout, (r.stdout[-2000:], r.stderr[-2000:])


def test_alternating_batch_sizes_keep_their_plans(model_files):
    """run_detector (batch 1) next to batched calls: the buffers are kept and each batch size keeps the tiling picks
    it was first given (no re-autotune, same results)."""
    from fcn_object_detector_b200.detector import FCNObjectDetector
    proto, cm = model_files
    rs = np.random.RandomState(21)
    one = rs.randint(0, 256, (1, 448, 448, 3)).astype(np.uint8)
    four = rs.randint(0, 256, (4, 448, 448, 3)).astype(np.uint8)
    det = FCNObjectDetector(pretrained_weights=cm, deployment_prototxt=proto, precision='bf16', detection_threshold=0.3, min_boxes=1)
    r4 = det.predict_batch(four); plan4 = det.net.describe(); cov4 = det.net.blobs['coverage'].data.copy()
    r1 = det.predict_batch(one); plan1 = det.net.describe(); cov1 = det.net.blobs['coverage'].data.copy()
    for _ in range(2):
        det.predict_batch(four)
        assert det.net.describe() == plan4 and np.array_equal(det.net.blobs['coverage'].data, cov4)
        det.predict_batch(one)
        assert det.net.describe() == plan1 and np.array_equal(det.net.blobs['coverage'].data, cov1)


def test_fused_prelude_feeds_conv1_and_data_blob_is_materialised_on_demand(model_files, monkeypatch):
    """On the bf16 plan the detect paths write the first convolution's space-to-depth operand straight from the uint8 frames
    (no fp32 `data` round trip).  Results equal the unfused prelude's bit for bit, and `blobs['data']` still answers."""
    from fcn_object_detector_b200.detector import FCNObjectDetector, preprocess
    proto, cm = model_files
    rs = np.random.RandomState(8)
    for shape in ((448, 448), (480, 640), (301, 203)):
        imgs = rs.randint(0, 256, (3,) + shape + (3,)).astype(np.uint8)
        det = FCNObjectDetector(pretrained_weights=cm, deployment_prototxt=proto, precision='bf16', detection_threshold=0.3, min_boxes=1)
        res = det.predict_batch(imgs)
        cov = det.net.blobs['coverage'].data.copy(); bb = det.net.blobs['bboxes'].data.copy()
        data = det.net.blobs['data'].data.copy()
        assert np.array_equal(data, preprocess(imgs))
        monkeypatch.setenv('FCNOD_NO_PRELUDE_FUSION', '1')
        ref = FCNObjectDetector(pretrained_weights=cm, deployment_prototxt=proto, precision='bf16', detection_threshold=0.3, min_boxes=1)
        monkeypatch.delenv('FCNOD_NO_PRELUDE_FUSION')
        res2 = ref.predict_batch(imgs)
        assert np.array_equal(ref.net.blobs['coverage'].data, cov) and np.array_equal(ref.net.blobs['bboxes'].data, bb)
        assert all(np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) for a, b in zip(res, res2))
        det.net.close(); ref.net.close()


def test_multi_gpu_detector_one_process(firing_model_files):
    """One process, one replica per device: the sharded result equals the single-device result frame for frame."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('needs 2 CUDA devices')
    from fcn_object_detector_b200.detector import FCNObjectDetector, MultiGpuDetector
    proto, cm = firing_model_files
    rs = np.random.RandomState(31)
    imgs = rs.randint(0, 256, (7, 448, 448, 3)).astype(np.uint8)
    kw = dict(precision='bf16', detection_threshold=0.45, min_boxes=1)
    one = FCNObjectDetector(pretrained_weights=cm, deployment_prototxt=proto, **kw)
    want = one.predict_batch(imgs)
    multi = MultiGpuDetector(cm, proto, device_ids=list(range(torch.cuda.device_count())), **kw)
    got = multi.predict_batch(imgs)
    assert len(got) == len(want) and sum(len(w[0]) for w in want) > 0
    for g, w in zip(got, want):
        assert np.array_equal(g[0], w[0]) and np.array_equal(g[1], w[1])
    got2 = multi.predict_batch(imgs[:1])
    assert np.array_equal(got2[0][0], want[0][0])
    multi.close()
