"""Host-side multi-GPU logic on CPU: contiguous frame sharding + gather over a world-size-2 gloo group.
The detector is replaced by a stub that tags frames (the CUDA path itself is covered by the -m gpu tests);
what is checked here is the partitioning, ordering and the absence of any data-path collective."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_range_partitions():
    from fcn_object_detector_b200.sharding import shard_range
    for n in (0, 1, 7, 64, 255, 256):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    assert [shard_range(256, r, 8) for r in range(8)] == [(32 * r, 32 * r + 32) for r in range(8)]   # BASELINE config 4
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


class _StubDetector(object):
    """predict_batch stand-in: one (boxes, labels) per frame derived from the frame's content."""
    def __init__(self, rank):
        self.rank = rank
        self.calls = []

    def predict_batch(self, images):
        self.calls.append(len(images))
        return [(np.array([[int(im[0, 0, 0]), self.rank, 0, 0, 0]], np.int64), np.array([int(im[0, 0, 1])], np.int64))
                for im in images]


def _worker(rank, world, port, n_frames, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from fcn_object_detector_b200.sharding import predict_sharded, shard_range
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        frames = np.zeros((n_frames, 4, 4, 3), np.uint8)
        frames[:, 0, 0, 0] = np.arange(n_frames)
        frames[:, 0, 0, 1] = np.arange(n_frames) % 3
        det = _StubDetector(rank)
        res = predict_sharded(det, frames, rank, world)
        lo, hi = shard_range(n_frames, rank, world)
        ok = det.calls == ([hi - lo] if hi > lo else [])
        ok = ok and len(res) == n_frames
        for i, (boxes, labels) in enumerate(res):
            owner = [r for r in range(world) if shard_range(n_frames, r, world)[0] <= i < shard_range(n_frames, r, world)[1]][0]
            ok = ok and int(boxes[0, 0]) == i and int(boxes[0, 1]) == owner and int(labels[0]) == i % 3
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('n_frames', [8, 5, 1])
def test_predict_sharded_world2_gloo(n_frames):
    s = socket.socket(); s.bind(('127.0.0.1', 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_frames, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got == [(0, True), (1, True)]


def test_multi_gpu_detector_splits_and_orders(monkeypatch):
    """MultiGpuDetector (one process, one replica per device) on CPU with the C ABI stubbed out: the split is the
    contiguous shard_range one, every shard is SUBMITTED before any is collected, results come back in frame order."""
    import ctypes as C
    from fcn_object_detector_b200 import _lib, detector as D
    log = []

    class FakeNet(object):
        def __init__(self, dev):
            self.handle = dev
        def _sync_after_device_run(self, batch=None):
            pass
        def close(self):
            pass

    class FakeDet(object):
        def __init__(self, pretrained_weights=None, deployment_prototxt=None, device_id=0, **kw):
            self.net = FakeNet(device_id)
        def _params(self):
            return _lib.PostnetParams()

    class FakeLib(object):
        def __init__(self):
            self.shards = {}
        def fcnod_detect_submit(self, h, images, B, H, W, p, ticket):
            arr = np.ctypeslib.as_array(C.cast(images, C.POINTER(C.c_uint8)), shape=(B, H, W, 3))
            self.shards[h] = arr[:, 0, 0, 0].copy()
            log.append(('submit', h, B))
            ticket._obj.value = 7
            return 0
        def fcnod_detect_collect(self, h, ticket, det, lab, wgt, cnt):
            tags = self.shards[h]
            n = len(tags)
            d = np.ctypeslib.as_array(C.cast(det, C.POINTER(C.c_int32)), shape=(n, _lib.FCNOD_MAX_DET, 5))
            c = np.ctypeslib.as_array(C.cast(cnt, C.POINTER(C.c_int32)), shape=(n,))
            la = np.ctypeslib.as_array(C.cast(lab, C.POINTER(C.c_int32)), shape=(n, _lib.FCNOD_MAX_DET))
            c[:] = 1; d[:, 0, 0] = tags; d[:, 0, 1] = h; la[:, 0] = 0
            log.append(('collect', h, n))
            return 0
        def fcnod_last_error(self):
            return b''
    fake = FakeLib()
    monkeypatch.setattr(D, 'FCNObjectDetector', FakeDet)
    monkeypatch.setattr(_lib, 'lib', lambda: fake)
    m = D.MultiGpuDetector('w', 'p', device_ids=[0, 1, 2])
    imgs = np.zeros((8, 4, 4, 3), np.uint8)
    imgs[:, 0, 0, 0] = np.arange(8) + 10
    res = m.predict_batch(imgs)
    assert [int(r[0][0, 0]) for r in res] == list(range(10, 18))                  # frame order kept
    assert [int(r[0][0, 1]) for r in res] == [0, 0, 0, 1, 1, 1, 2, 2]             # 3 + 3 + 2, contiguous
    assert [e[0] for e in log] == ['submit'] * 3 + ['collect'] * 3               # all devices busy before any wait
    res = m.predict_batch(imgs[:2])                                               # fewer frames than devices
    assert len(res) == 2 and [e for e in log[6:]] == [('submit', 0, 1), ('submit', 1, 1), ('collect', 0, 1), ('collect', 1, 1)]
