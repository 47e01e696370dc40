"""Time the post-net kernels on several kinds of maps (B frames): python tools/postnet_probe.py [B]"""
import ctypes as C, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fcn_object_detector_b200 import _lib
from oracle import postnet_ref
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
L = _lib.lib()
p = _lib.default_postnet_params()
M = _lib.FCNOD_MAX_DET
scr = torch.empty(int(L.fcnod_postnet_scratch_bytes(B, 4)), dtype=torch.uint8, device='cuda')
o = [torch.empty((B, M, 5), dtype=torch.int32, device='cuda'), torch.empty((B, M), dtype=torch.int32, device='cuda'),
     torch.empty((B, M), dtype=torch.int32, device='cuda'), torch.empty((B,), dtype=torch.int32, device='cuda')]
s = torch.cuda.current_stream().cuda_stream
def run(cov, bb, name):
    cov = cov.contiguous().cuda(); bb = bb.contiguous().cuda()
    f = lambda: _lib.check(L.fcnod_postnet_device(cov.data_ptr(), bb.data_ptr(), B, 4, 16, 28, 28, None, C.byref(p), scr.data_ptr(),
                                                  o[0].data_ptr(), o[1].data_ptr(), o[2].data_ptr(), o[3].data_ptr(), s))
    f(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10): f()
    b.record(); torch.cuda.synchronize()
    n = (cov[:, 1:] >= 0.5).sum(dim=(2, 3)).float()
    print('%-28s %8.1f us   cand/class mean %.0f max %.0f  dets %d' % (name, a.elapsed_time(b) * 100, n.mean().item(), n.max().item(), int(o[3].sum().item())))
g = torch.Generator().manual_seed(0)
run(torch.zeros(B, 4, 28, 28), torch.zeros(B, 16, 28, 28), 'empty')
run(torch.rand(B, 4, 28, 28, generator=g), torch.randn(B, 16, 28, 28, generator=g) * 0.3, 'uniform cov, bbox~0 (bench-like)')
run(torch.rand(B, 4, 28, 28, generator=g), torch.randn(B, 16, 28, 28, generator=g) * 200, 'uniform cov, bbox wild')
run(torch.ones(B, 4, 28, 28), torch.zeros(B, 16, 28, 28), 'all cells, bbox 0')
pl = [postnet_ref.planted_maps(np.random.RandomState(3 + i)) for i in range(B)]
run(torch.tensor(np.stack([c for c, b in pl])), torch.tensor(np.stack([b for c, b in pl])), 'planted (3 objects)')
