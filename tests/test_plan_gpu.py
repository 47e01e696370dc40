"""GPU: the plan that bench.py times is the plan that is parity-tested.

BASELINE config 3 (`deploy.prototxt inference batch 64 @448x448 bf16 on 1 B200`): the tensor-core
tiling of every convolution launch is chosen per batch size by timing the best-ranked candidates on the device, so
the batch-64 plan differs from the plans the small-batch tests exercise.  Here:

  * the autotuned batch-64 plan runs config 3's input (64x3x448x448 uniform[0,1), seed 2) and its `coverage` /
    `bboxes` are compared with the fp32 oracle (1e-2, metric max|a-b|/max|b| per blob) on a sample of the frames;
  * EVERY candidate the autotuner may choose at batch 64 (the first FCNOD_AUTOTUNE of each launch's list) is forced
    in turn and its output blobs must equal the validated plan's bit for bit (every tiling accumulates in the same
    K order), compared through a device-side checksum;
  * the picks are cached per batch size: going to another batch and back gives the same plan.

The measured errors go to gpurun_out/batch64_errors.json (copied into profiles/ per round).
"""
import ctypes as C
import json
import os

import numpy as np
import pytest

from tests.util import ROOT, oracle_forward, rel_err

pytestmark = pytest.mark.gpu

B64 = 64
SAMPLE = [0, 9, 18, 27, 36, 45, 54, 63]
TOL_BF16 = 1e-2


@pytest.fixture(scope='module')
def net64(model_files):
    """One bf16 net at batch 64 with config 3's input already run forward (shared by the tests of this file)."""
    import fcn_object_detector_b200.caffe_shim as caffe
    proto, cm = model_files
    x = np.random.RandomState(2).uniform(0, 1, (B64, 3, 448, 448)).astype(np.float32)
    net = caffe.Net(proto, cm, caffe.TEST, device=0, precision='bf16')
    net.blobs['data'].reshape(*x.shape)
    net.blobs['data'].data[...] = x
    net.forward()
    yield net, x
    net.close()


def _checksum(net, blob):
    from fcn_object_detector_b200 import _lib
    v = C.c_ulonglong()
    _lib.check(_lib.lib().fcnod_net_blob_checksum(net.handle, blob.encode(), C.byref(v)))
    return int(v.value)


def _step_outputs(net, step):
    from fcn_object_detector_b200 import _lib
    buf = C.create_string_buffer(1024)
    _lib.check(_lib.lib().fcnod_net_step_outputs(net.handle, step, buf, len(buf)))
    return [n for n in buf.value.decode().split(',') if n]


def test_bf16_full_net_batch64_config3(net64, model_files):
    proto, cm = model_files
    net, x = net64
    plan = net.describe()
    assert 'conv_simt' not in plan and '+s2d+power' in plan
    want = ['inception_4e/output', 'inception_5b/output', 'coverage', 'bboxes']
    got = {k: net.blobs[k].data[SAMPLE].copy() for k in want}
    exp = oracle_forward(proto, cm, x[SAMPLE], want=want)
    errs = {k: rel_err(got[k], exp[k]) for k in want}
    os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
    with open(os.path.join(ROOT, 'gpurun_out', 'batch64_errors.json'), 'w') as f:
        json.dump({'config': 'BASELINE config 3: batch 64, 448x448, bf16 plan after the batch-64 autotune; uniform[0,1) seed 2; '
                             'oracle = fp32 CPU restatement on frames %s' % SAMPLE,
                   'metric': 'max|a-b| / max|b| per blob', 'tolerance': TOL_BF16, 'errors': errs,
                   'plan': plan.splitlines()}, f, indent=1)
    assert errs['coverage'] < TOL_BF16 and errs['bboxes'] < TOL_BF16, errs
    assert errs['inception_4e/output'] < 3e-2 and errs['inception_5b/output'] < 3e-2, errs
    # frames of a batch are independent: the 64-frame batch agrees with the same frames run as a batch of 8
    import fcn_object_detector_b200.caffe_shim as caffe
    small = caffe.Net(proto, cm, caffe.TEST, device=0, precision='bf16')
    try:
        small.blobs['data'].reshape(len(SAMPLE), 3, 448, 448)
        small.blobs['data'].data[...] = x[SAMPLE]
        small.forward()
        for k in ('coverage', 'bboxes'):
            assert np.array_equal(small.blobs[k].data, got[k]), k
    finally:
        small.close()


def test_every_autotune_candidate_gives_the_same_bits(net64):
    """Force each candidate the batch-64 autotuner may pick, launch by launch; the launch's output blobs must keep the
    checksum of the plan validated above."""
    from fcn_object_detector_b200 import _lib
    L = _lib.lib()
    net, _ = net64
    h = net.handle
    ntry = max(1, int(L.fcnod_net_autotune_candidates()))
    nsteps = L.fcnod_net_launches_per_forward(h)
    swept, families = 0, set()
    plan = net.describe().splitlines()
    for step in range(nsteps):
        ncand = L.fcnod_net_tc_candidates(h, step)
        if ncand == 0:
            continue
        outs = _step_outputs(net, step)
        assert outs, plan[step]
        base_pick = L.fcnod_net_tc_pick(h, step)
        _lib.check(L.fcnod_net_run_step(h, step, 1, None))
        base = [_checksum(net, o) for o in outs]
        for cand in range(min(ncand, ntry)):
            _lib.check(L.fcnod_net_set_tc_pick(h, step, cand))
            _lib.check(L.fcnod_net_run_step(h, step, 1, None))
            got = [_checksum(net, o) for o in outs]
            assert got == base, ('step %d candidate %d differs from pick %d' % (step, cand, base_pick), plan[step],
                                 net.describe().splitlines()[step])
            swept += 1
        _lib.check(L.fcnod_net_set_tc_pick(h, step, base_pick))
        _lib.check(L.fcnod_net_run_step(h, step, 1, None))
        assert [_checksum(net, o) for o in outs] == base
        d = plan[step]
        families.add(('s2d' if '+s2d' in d else 'halo' if '[halo' in d else 'heads' if '+nchw' in d else '1x1',
                      d.split('CK=')[1].split(' ')[0]))
    assert swept >= 40
    # conv1 (space-to-depth), 3x3 HALO CK=64, 5x5 HALO CK=16/32, fused 1x1, heads
    assert {('s2d', '16'), ('halo', '64'), ('halo', '32'), ('halo', '16'), ('1x1', '64'), ('heads', '64')} <= families, families
    assert net.describe().splitlines() == plan


def test_picks_are_cached_per_batch(net64):
    from fcn_object_detector_b200 import _lib
    L = _lib.lib()
    net, x = net64
    plan64 = net.describe()
    net.blobs['data'].reshape(1, 3, 448, 448)
    plan1 = net.describe()
    assert plan1 != plan64
    net.blobs['data'].reshape(B64, 3, 448, 448)
    assert net.describe() == plan64
    net.blobs['data'].reshape(1, 3, 448, 448)
    assert net.describe() == plan1
    # and the buffers survived the round trip: batch 64 still computes
    net.blobs['data'].reshape(B64, 3, 448, 448)
    net.blobs['data'].data[...] = x
    net.forward()
    assert np.isfinite(net.blobs['coverage'].data).all()
