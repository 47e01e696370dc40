"""Short program for ncu: the non-network kernels of the hot path at BASELINE sizes --
the training-side layer forward (config 2: batch 32, 480x640 sources), the detector prelude + whole chain from uint8 frames
(batch 64), and the post-net on planted maps (batch 64).  No oracle involved: this is a profiling driver."""
import os
import random
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fcn_object_detector_b200 import _lib  # noqa: E402
from fcn_object_detector_b200.caffemodel import write_synthetic_caffemodel  # noqa: E402
from fcn_object_detector_b200.data_argumentation_layer import DataArgumentationLayer, draw_augmentation  # noqa: E402
from fcn_object_detector_b200.detector import FCNObjectDetector, postnet  # noqa: E402
from fcn_object_detector_b200.netspec import write_deploy_prototxt  # noqa: E402


class Blob(object):
    def __init__(self, arr):
        self.data = arr
    num = property(lambda s: s.data.shape[0])

    def reshape(self, *d):
        if tuple(d) != self.data.shape:
            self.data = np.zeros(d, np.float32)


rs = np.random.RandomState(1)
# ---- layer forward, config 2
N, Hs, Ws = 32, 480, 640
data = rs.randint(0, 256, (N, 3, Hs, Ws)).astype(np.float32)
lab = np.zeros((N, 1, 1, 5), np.float32)
for i in range(N):
    w, h = int(rs.randint(32, 257)), int(rs.randint(32, 257))
    lab[i, 0, 0] = [int(rs.randint(0, Ws - w)), int(rs.randint(0, Hs - h)), w, h, 0]
L = DataArgumentationLayer()
L.param_str = '448, 448,16, 1'
bottom = [Blob(data), Blob(lab)]
top = [Blob(np.zeros((1,), np.float32)) for _ in range(6)]
L.setup(bottom, top)
L.reshape(bottom, top)
random.seed(1234)
for _ in range(2):
    L.forward(bottom, top)
print('layer ok', [t.data.shape for t in top])
# ---- detector from uint8 frames (prelude kernels + net + post-net), batch 64
d = '/tmp/fcnod_prof'
os.makedirs(d, exist_ok=True)
proto = write_deploy_prototxt(os.path.join(d, 'deploy.prototxt'))
cm = write_synthetic_caffemodel(os.path.join(d, 'w.caffemodel'), seed=0)
det = FCNObjectDetector(pretrained_weights=cm, deployment_prototxt=proto, precision='bf16')
imgs = rs.randint(0, 256, (64, 480, 640, 3)).astype(np.uint8)
for _ in range(2):
    res = det.predict_batch(imgs)
print('detect ok', sum(len(r[0]) for r in res))
# ---- post-net on planted maps, batch 64
cov = rs.uniform(0, 0.3, (64, 4, 28, 28)).astype(np.float32)
bb = np.zeros((64, 16, 28, 28), np.float32)
for b in range(64):
    for c in range(3):
        for k in range(3):
            y0, x0 = rs.randint(2, 20, 2)
            cov[b, c + 1, y0:y0 + 5, x0:x0 + 5] = rs.uniform(0.6, 1.0, (5, 5))
            for yy in range(y0, y0 + 5):
                for xx in range(x0, x0 + 5):
                    bb[b, 4 * c:4 * c + 4, yy, xx] = np.array([x0 * 16 - xx * 16, y0 * 16 - yy * 16, (x0 + 5) * 16 - xx * 16, (y0 + 5) * 16 - yy * 16]) + rs.normal(0, 1.5, 4)
for _ in range(2):
    out = postnet(cov, bb)
print('postnet ok', sum(len(o[0]) for o in out))
# ---- bandwidth kernels at the roofline batch (4096 frames / images per launch): standalone decode + label raster
import torch  # noqa: E402
import bench  # noqa: E402
bw = bench.bandwidth_kernels(torch, _lib.lib(), _lib, torch.cuda.current_stream().cuda_stream, bench.peaks(), reps=1)
print('bandwidth ok', {k: round(v['frac_of_hbm_peak'], 3) for k, v in bw.items() if isinstance(v, dict)})
