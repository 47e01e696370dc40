"""GPU: the detector's predict-on-image entry point (prelude + net + post-net) through the C ABI."""
import numpy as np
import pytest

from oracle import net_ref, postnet_ref
from tests.util import rel_err

pytestmark = pytest.mark.gpu


def test_preprocess_matches_oracle():
    from fcn_object_detector_b200.detector import preprocess
    rs = np.random.RandomState(0)
    for shape in ((448, 448), (480, 640), (300, 200)):
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
