"""Short program for ncu: builds the bf16 net at batch B and runs the whole hot path `iters` times (no graph).
Only the LAST pass lies between cudaProfilerStart/Stop: run ncu with `--profile-from-start off` so that the launches
of the tiling autotuner (fcnod_net_reshape) and the warm-up passes are not captured."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fcn_object_detector_b200 import _lib  # noqa: E402
from fcn_object_detector_b200.caffemodel import write_synthetic_caffemodel  # noqa: E402
from fcn_object_detector_b200.detector import FCNObjectDetector  # noqa: E402
from fcn_object_detector_b200.netspec import write_deploy_prototxt  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 2
d = '/tmp/fcnod_prof'
os.makedirs(d, exist_ok=True)
proto = write_deploy_prototxt(os.path.join(d, 'deploy.prototxt'))
cm = write_synthetic_caffemodel(os.path.join(d, 'w.caffemodel'), seed=0)
det = FCNObjectDetector(pretrained_weights=cm, deployment_prototxt=proto, precision='bf16')
L = _lib.lib()
data = torch.rand((B, 3, 448, 448), device='cuda')
p = _lib.default_postnet_params()
s = torch.cuda.current_stream().cuda_stream
for i in range(iters):
    if i == iters - 1:
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
    _lib.check(L.fcnod_detect_device(det.net.handle, data.data_ptr(), B, C.byref(p), None, None, None, None, s))
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print('ok', L.fcnod_net_launches_per_forward(det.net.handle) + 2, 'launches per pass')
