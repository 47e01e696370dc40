"""Timeline of CTA 0 of tensor-core convolution steps (clock64 stamps): python tools/tc_timeline.py B step [step ...]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fcn_object_detector_b200 import _lib  # noqa: E402
from fcn_object_detector_b200.caffemodel import write_synthetic_caffemodel  # noqa: E402
from fcn_object_detector_b200.detector import FCNObjectDetector  # noqa: E402
from fcn_object_detector_b200.netspec import write_deploy_prototxt  # noqa: E402

B = int(sys.argv[1])
steps = [int(a) for a in sys.argv[2:]]
d = '/tmp/fcnod_prof'
os.makedirs(d, exist_ok=True)
proto = write_deploy_prototxt(os.path.join(d, 'deploy.prototxt'))
cm = write_synthetic_caffemodel(os.path.join(d, 'w.caffemodel'), seed=0)
det = FCNObjectDetector(pretrained_weights=cm, deployment_prototxt=proto, precision='bf16')
L = _lib.lib()
h = det.net.handle
data = torch.rand((B, 3, 448, 448), device='cuda')
p = _lib.default_postnet_params()
_lib.check(L.fcnod_detect_device(h, data.data_ptr(), B, C.byref(p), None, None, None, None, torch.cuda.current_stream().cuda_stream))
torch.cuda.synchronize()
plan = det.net.describe().splitlines()
for s in steps:
    out = np.zeros((4, 8, 64), np.uint64)
    _lib.check(L.fcnod_net_profile_conv_step(h, s, _lib.ptr(out), out.size))
    print(plan[s][:200])
    t0 = int(out[out > 0].min())
    rel = lambda v: int(v) - t0 if v else -1
    for tile in range(8):
        n = 14
        print(' tile %d  B issue   %s' % (tile, [rel(v) for v in out[0, tile, :60] if v][:n]))
        print('         B landed  %s' % ([rel(v) for v in out[1, tile, :60] if v][:n]))
        print('         tap sent  %s' % ([rel(v) for v in out[3, tile, :60] if v][:n]))
        print('         mma start %d tempty-ok %d commit %d' % (rel(out[1, tile, 61]), rel(out[1, tile, 62]), rel(out[1, tile, 63])))
        print('         epi  wait %d tfull-ok %d tmem-released %d done %d' % tuple(rel(out[2, tile, i]) for i in range(4)))
        if out[2, tile, 8:].any():      # POOL mode: raw-patch TMA issue per chunk; pooling warp 0: start / raw landed / A stage free / pooled
            print('         raw issue %s' % [rel(v) for v in out[0, tile, 16:32] if v])
            print('         pool      %s' % [tuple(rel(out[2, tile, 8 + 4 * c + k]) for k in range(4)) for c in range(14) if out[2, tile, 8 + 4 * c]])
