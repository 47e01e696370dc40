"""GPU: the allocation-free layer forward (`fcnod_aug_forward_device_noalloc`) against the existing entry point.

The variant keeps the per-image geometry records in a fixed `static __device__` array inside the library instead of
caller-provided scratch (include/fcnod_b200.h; no counterpart in the reference, added for callers that may not call
cudaMalloc).  Called through the C ABI twice in a row -- the second call with a different, smaller batch, so records the
first call left in the array lie behind it -- and compared with `fcnod_aug_forward_device` on the same inputs: label
grids and rects bit for bit, the image top within the layer tests' 1e-5.  Collected last on purpose (file name).
"""
import ctypes as C
import random

import numpy as np
import pytest

from oracle import layer_ref

pytestmark = pytest.mark.gpu

NET_H, NET_W, STRIDE, NCLS = 448, 448, 16, 1


def _inputs(seed, N, Hs, Ws):
    rs = np.random.RandomState(seed)
    rng = random.Random(seed)
    data = rs.randint(0, 256, (N, 3, Hs, Ws)).astype(np.float32)
    lab = np.zeros((N, 5), np.float32)
    draws = []
    for i in range(N):
        w, h = int(rs.randint(24, Ws // 2)), int(rs.randint(24, Hs // 2))
        rect = [int(rs.randint(0, Ws - w)), int(rs.randint(0, Hs - h)), w, h]
        lab[i] = rect + [0]
        draws.append(layer_ref.draw_params(rng, Ws, Hs, rect))
    return data, lab, draws


def _device_args(torch, _lib, data, lab, draws):
    N = data.shape[0]
    arr = (_lib.AugDraw * N)()
    for a, d in zip(arr, draws):
        for k, v in d.items():
            setattr(a, k, v)
    d_in = torch.from_numpy(data).cuda()
    d_lab = torch.from_numpy(lab).cuda()
    d_draw = torch.from_numpy(np.frombuffer(bytes(arr), dtype=np.uint8).copy()).cuda()
    return d_in, d_lab, d_draw


def _tops(torch, N):
    gh, gw = NET_H // STRIDE, NET_W // STRIDE
    shapes = [(N, 3, NET_H, NET_W), (N, NCLS, gh, gw)] + [(N, 4 * NCLS, gh, gw)] * 4
    tops = [torch.full(s, float('nan'), device='cuda') for s in shapes]            # poisoned: every element must be written
    rects = torch.full((N, 4), -12345, dtype=torch.int32, device='cuda')
    return tops, rects


def _run_both(torch, _lib, seed, N, Hs, Ws):
    L = _lib.lib()
    data, lab, draws = _inputs(seed, N, Hs, Ws)
    d_in, d_lab, d_draw = _device_args(torch, _lib, data, lab, draws)
    stream = torch.cuda.current_stream().cuda_stream
    # existing entry point, caller scratch
    ref_tops, ref_rects = _tops(torch, N)
    scr = torch.empty(int(L.fcnod_aug_scratch_bytes(N)), dtype=torch.uint8, device='cuda')
    _lib.check(L.fcnod_aug_forward_device(d_in.data_ptr(), d_lab.data_ptr(), d_draw.data_ptr(), N, Hs, Ws, NET_H, NET_W, STRIDE, NCLS,
                                          *[t.data_ptr() for t in ref_tops], ref_rects.data_ptr(), scr.data_ptr(), stream))
    # the variant: outputs allocated BEFORE the call, nothing is allocated inside it
    tops, rects = _tops(torch, N)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    _lib.check(L.fcnod_aug_forward_device_noalloc(d_in.data_ptr(), d_lab.data_ptr(), d_draw.data_ptr(), N, Hs, Ws, NET_H, NET_W, STRIDE,
                                                  NCLS, *[t.data_ptr() for t in tops], rects.data_ptr(), stream))
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    assert np.array_equal(rects.cpu().numpy(), ref_rects.cpu().numpy())
    for k in range(1, 6):
        a, b = tops[k].cpu().numpy(), ref_tops[k].cpu().numpy()
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), k
    a, b = tops[0].cpu().numpy(), ref_tops[0].cpu().numpy()
    assert np.isfinite(a).all() and np.abs(a - b).max() < 1e-5
    assert float(ref_tops[1].sum()) > 0            # the rects label cells: not a comparison of empty grids
    return tops


def test_noalloc_layer_forward_matches_existing_entry_point_twice():
    import torch
    from fcn_object_detector_b200 import _lib
    torch.cuda.set_device(0)
    _run_both(torch, _lib, seed=11, N=12, Hs=240, Ws=320)
    # second call in a row: smaller batch, other source size; records 5..11 of the first call are still in the array
    _run_both(torch, _lib, seed=12, N=5, Hs=300, Ws=260)
    # and the first workload again gives the first answer (the array carries nothing from call to call)
    _run_both(torch, _lib, seed=11, N=12, Hs=240, Ws=320)


def test_noalloc_layer_forward_rejects_batches_beyond_the_limit():
    from fcn_object_detector_b200 import _lib
    L = _lib.lib()
    N = _lib.FCNOD_AUG_NOALLOC_MAX_N + 1
    rc = L.fcnod_aug_forward_device_noalloc(None, None, None, N, 240, 320, NET_H, NET_W, STRIDE, NCLS,
                                            None, None, None, None, None, None, None, None)
    assert rc != 0
    msg = L.fcnod_last_error().decode()
    assert 'FCNOD_AUG_NOALLOC_MAX_N' in msg and str(_lib.FCNOD_AUG_NOALLOC_MAX_N) in msg, msg
