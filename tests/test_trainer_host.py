"""CPU: host side of the training step -- the train_val / solver definitions, the oracle's ADAM against hand-derived
answers, the data-parallel averaging rule under a world-size-2 gloo group, and the no-CPU-fallback rule of the trainer."""
import math
import os

import numpy as np
import pytest

from oracle import prototxt, train_net_ref

REF = '/root/reference'


def _records(text):
    keep = ('name', 'type', 'bottom', 'top', 'num_output', 'kernel', 'stride', 'pad', 'dropout_ratio', 'shift')
    return [{k: L.get(k) for k in keep} for L in prototxt.layers(prototxt.parse(text))]


def test_generated_train_val_matches_the_reference_layer_for_layer():
    from fcn_object_detector_b200 import netspec
    path = os.path.join(REF, 'models', 'train_val.prototxt')
    if not os.path.exists(path):
        pytest.skip('upstream tree not reachable')
    ours, theirs = _records(netspec.train_val_prototxt()), _records(open(path).read())
    assert len(ours) == len(theirs) == 152
    for a, b in zip(ours, theirs):
        assert a == b, (a, b)
    # and the pieces the record view does not show: slice points, Eltwise operation, loss weight
    up = prototxt.parse(open(path).read())['layer']
    mine = prototxt.parse(netspec.train_val_prototxt())['layer']
    for a, b in zip(mine, up):
        assert a.get('slice_param') == b.get('slice_param') or [dict(x, slice_dim=x.get('slice_dim')) for x in a.get('slice_param')] == b.get('slice_param')
        assert a.get('eltwise_param') == b.get('eltwise_param')
        assert a.get('loss_weight') == b.get('loss_weight')
        assert [p for p in a.get('param', [])] == [p for p in b.get('param', [])], a.get('name')


def test_generated_three_class_train_val_matches_train_val2():
    from fcn_object_detector_b200 import netspec
    path = os.path.join(REF, 'models', 'train_val2.prototxt')
    if not os.path.exists(path):
        pytest.skip('upstream tree not reachable')
    ours, theirs = _records(netspec.train_val_prototxt(num_classes=3)), _records(open(path).read())
    assert len(ours) == len(theirs)
    for a, b in zip(ours, theirs):
        assert a == b, (a, b)
    sl = [x for x in prototxt.parse(netspec.train_val_prototxt(num_classes=3))['layer'] if x.get('type') == ['Slice']][0]
    up = [x for x in prototxt.parse(open(path).read())['layer'] if x.get('type') == ['Slice']][0]
    assert sl['slice_param'][0]['slice_point'] == up['slice_param'][0]['slice_point'] == [3, 15, 27, 39]


def test_solver_prototxt_values(tmp_path):
    from fcn_object_detector_b200 import netspec
    from fcn_object_detector_b200.trainer import read_solver_prototxt
    p = tmp_path / 'solver.prototxt'
    p.write_text(netspec.solver_prototxt())
    v = read_solver_prototxt(str(p))
    assert (v['base_lr'], v['gamma'], v['stepsize'], v['momentum'], v['weight_decay']) == (1e-4, 0.1, 10000, 0.9, 1e-7)
    assert (v['momentum2'], v['delta']) == (0.999, 1e-8)          # caffe.proto defaults
    up = os.path.join(REF, 'train', 'bounding_box', 'solver.prototxt')
    if os.path.exists(up):
        assert read_solver_prototxt(up) == v
    p.write_text(netspec.solver_prototxt().replace('solver_type: ADAM', 'solver_type: SGD'))
    with pytest.raises(ValueError):
        read_solver_prototxt(str(p))
    p.write_text(netspec.solver_prototxt().replace('lr_policy: "step"', 'lr_policy: "poly"'))
    with pytest.raises(ValueError):
        read_solver_prototxt(str(p))


def test_oracle_adam_known_answers():
    """Hand-derived from adam_solver.cpp's formulas with b1 = 0.9, b2 = 0.999, delta = 1e-8, rate 1e-4, no decay (bias blob):
      t = 1: m = 0.1 g, v = 0.001 g^2, correction = sqrt(1 - 0.999) / (1 - 0.9) = 0.316228
             step = 2 * 1e-4 * 0.316228 * 0.1 g / (0.0316228 |g| + 1e-8) = 2e-4 * sign(g) * 1 / (1 + 3.16e-7 / |g|)
      t = 2 with the same g: m = 0.19 g, v = 0.001999 g^2, correction = sqrt(1 - 0.998001) / (1 - 0.81) = 0.235320
             step = 2e-4 * 0.235320 * 0.19 / sqrt(0.001999) = 2e-4 * 1.000000 (for |g| >> delta)."""
    w = {'c': [np.zeros((1, 1, 1, 1), np.float32), np.asarray([1.0, -2.0, 0.5], np.float32)]}
    g = {'c': [np.zeros((1, 1, 1, 1), np.float32), np.asarray([0.3, -4.0, 1e-3], np.float32)]}
    st = {}
    w1 = train_net_ref.adam_update(w, g, st, 0)
    assert np.allclose(w['c'][1] - w1['c'][1], 2e-4 * np.sign(g['c'][1]), rtol=1e-3)
    w2 = train_net_ref.adam_update(w1, g, st, 1)
    assert np.allclose(w1['c'][1] - w2['c'][1], 2e-4 * np.sign(g['c'][1]), rtol=1e-3)
    # weights: lr_mult 1 and L2 decay folded into the gradient: g = 0 -> the decay alone drives the step
    w = {'c': [np.full((1, 1, 1, 1), 0.5, np.float32), np.zeros(1, np.float32)]}
    g = {'c': [np.zeros((1, 1, 1, 1), np.float32), np.zeros(1, np.float32)]}
    w1 = train_net_ref.adam_update(w, g, {}, 0)
    geff = 1e-7 * 0.5
    want = 1e-4 * geff / (geff + 1e-8 / math.sqrt(0.001))
    assert abs(float(w['c'][0].ravel()[0] - w1['c'][0].ravel()[0]) - want) < 7e-8             # one ulp of 0.5 in float32 is 6e-8
    assert float(w1['c'][1][0]) == 0.0                      # bias: decay_mult 0, zero gradient -> no movement
    # lr_policy "step": rate drops by gamma at iteration 10000
    assert train_net_ref.learning_rate(9999) == 1e-4 and abs(train_net_ref.learning_rate(10000) - 1e-5) < 1e-12


def _rank_main(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from fcn_object_detector_b200.trainer import allreduce_scale
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    g = torch.from_numpy(np.random.RandomState(100 + rank).randn(1000).astype(np.float32))
    dist.all_reduce(g, op=dist.ReduceOp.SUM)
    g *= allreduce_scale(world)
    q.put((rank, g.numpy().copy()))
    dist.destroy_process_group()


def test_gradient_averaging_rule_world_size_2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    per_rank = [{'c': [np.random.RandomState(100 + r).randn(1000).astype(np.float32)]} for r in range(2)]
    want = train_net_ref.average_gradients(per_rank)['c'][0]
    assert np.allclose(got[0], want, atol=1e-6) and np.array_equal(got[0], got[1])


def test_trainer_has_no_cpu_fallback(built_lib, tmp_path):
    """Without a CUDA device the trainer must refuse loudly (this container has none; on a GPU box the test is moot)."""
    import ctypes as C
    import torch
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    from fcn_object_detector_b200 import _lib, netspec
    from fcn_object_detector_b200.caffemodel import write_synthetic_caffemodel
    proto = netspec.write_train_val_prototxt(str(tmp_path / 'tv.prototxt'))
    cm = write_synthetic_caffemodel(str(tmp_path / 'w.caffemodel'), num_cvg=1, num_bbox=4)
    h = C.c_void_p()
    rc = _lib.lib().fcnod_trainer_create(proto.encode(), cm.encode(), 0, 2, 128, 128, None, C.byref(h))
    assert rc != 0 and b'no CUDA device' in _lib.lib().fcnod_last_error()
