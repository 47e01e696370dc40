"""GPU: one training iteration of models/train_val.prototxt (csrc/train_net.cu through the C ABI) against the CPU oracle
(oracle/train_net_ref.py: torch autograd over the restated forward graph + Caffe's ADAM ApplyUpdate).

Sizes the oracle finishes in seconds: the full network body on 2 frames of 128x160 (label grids 8x10).  Gates (fp32 engine):
activations and their gradients 1e-4, weight / bias gradients 2e-4, losses 1e-5 (all max|a-b| / max|b| per blob), updated
weights 1e-6 absolute after the first ApplyUpdate.  The dropout mask is drawn on the device and read back, so both sides
use the same one.
"""
import os

import numpy as np
import pytest

from oracle import prototxt, train_net_ref
from tests.util import rel_err

pytestmark = pytest.mark.gpu

N, H, W = 2, 128, 160
GH, GW = H // 16, W // 16


@pytest.fixture(scope='module')
def train_files(tmp_path_factory):
    from fcn_object_detector_b200 import netspec
    from fcn_object_detector_b200.caffemodel import write_synthetic_caffemodel
    d = tmp_path_factory.mktemp('train')
    proto = netspec.write_train_val_prototxt(str(d / 'train_val.prototxt'))
    cm = write_synthetic_caffemodel(str(d / 'init.caffemodel'), seed=3, num_cvg=1, num_bbox=4)
    solver = str(d / 'solver.prototxt')
    with open(solver, 'w') as f:
        f.write(netspec.solver_prototxt())
    return proto, cm, solver


def _inputs(seed=0):
    """Synthetic tops of the augmentation layer: an image and label grids with the layer's structure (a rectangle of
    foreground cells; bbox-label in pixels; size-block = 1 / rect size; obj-block = 1 / cells; coverage-block = mask)."""
    rs = np.random.RandomState(seed)
    data = rs.uniform(0, 255, (N, 3, H, W)).astype(np.float32)
    cvg = np.zeros((N, 1, GH, GW), np.float32)
    bbox = np.zeros((N, 4, GH, GW), np.float32)
    size = np.zeros((N, 4, GH, GW), np.float32)
    obj = np.zeros((N, 4, GH, GW), np.float32)
    cblk = np.zeros((N, 4, GH, GW), np.float32)
    for n in range(N):
        y0, x0 = rs.randint(0, GH - 3), rs.randint(0, GW - 4)
        hh, ww = rs.randint(2, 4), rs.randint(2, 5)
        cvg[n, 0, y0:y0 + hh, x0:x0 + ww] = 1
        cblk[n, :, y0:y0 + hh, x0:x0 + ww] = 1
        bbox[n, :, y0:y0 + hh, x0:x0 + ww] = rs.uniform(-40, 40, (4, hh, ww))
        size[n, 0::2, y0:y0 + hh, x0:x0 + ww] = 1.0 / (16 * ww)
        size[n, 1::2, y0:y0 + hh, x0:x0 + ww] = 1.0 / (16 * hh)
        obj[n, :, y0:y0 + hh, x0:x0 + ww] = 1.0 / (hh * ww)
    return data, (cvg, bbox, size, obj, cblk)


def _oracle_weights(cm):
    from oracle.caffemodel_io import read_caffemodel
    return read_caffemodel(cm)


@pytest.fixture(scope='module')
def solver(train_files):
    from fcn_object_detector_b200.trainer import Solver
    proto, cm, sp = train_files
    s = Solver(proto, cm, batch=N, height=H, width=W, solver_prototxt=sp, seed=7)
    yield s
    s.close()


def test_solver_reads_the_reference_solver_values(solver):
    p = solver.params
    f = np.float32
    assert (f(p.base_lr), f(p.gamma), p.stepsize) == (f(1e-4), f(0.1), 10000)
    assert (f(p.momentum), f(p.momentum2), f(p.delta), f(p.weight_decay)) == (f(0.9), f(0.999), f(1e-8), f(1e-7))
    assert (solver.num_classes, solver.grid_h, solver.grid_w) == (1, GH, GW)
    assert len(solver.layers) == 59 and solver.layers[0] == 'conv1/7x7_s2' and solver.layers[-1] == 'bbox/regressor'


def test_forward_backward_and_update_match_the_oracle(solver, train_files):
    import torch
    proto, cm, _ = train_files
    layers = prototxt.layers(prototxt.parse(open(proto).read()))
    weights = {k: [np.asarray(a, np.float32) for a in v] for k, v in _oracle_weights(cm).items()}
    data, grids = _inputs(0)
    dev = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (data,) + grids]
    stream = torch.cuda.current_stream().cuda_stream
    solver.forward_backward_device([t.data_ptr() for t in dev], stream=stream)
    torch.cuda.synchronize()
    got_loss = solver.losses(stream)
    mask = solver.dropout_mask()
    assert mask.shape == (N, 1024, GH, GW) and 0.5 < mask.mean() < 0.7        # keep probability 0.6

    want = ['conv1/7x7_s2', 'pool1/norm1', 'conv2/norm2', 'pool2/3x3_s2', 'inception_3a/output', 'inception_3b/pool', 'pool3/3x3_s2',
            'inception_4e/output', 'inception_5b/output', 'pool5/drop_s1', 'coverage', 'bboxes']
    # Max-pool routing.  (1) The device's selection obeys Caffe's rule on the device's own inputs, bit for bit: first maximum
    # of the clipped window in row-major order.  (2) The oracle then backpropagates through the SAME selection: where two
    # inputs of a window differ by less than the two forwards do (1e-6), the other forward may select the other element, and
    # gradients routed to another pixel are not an error of either side.
    pools = {L['name']: L for L in layers if L['type'] == 'Pooling'}
    assert len(pools) == 12
    argmax = {name: solver.pool_argmax(name) for name in pools}
    for name in ('pool1/3x3_s2', 'pool2/3x3_s2', 'inception_3b/pool', 'pool3/3x3_s2', 'inception_5a/pool'):
        L = pools[name]
        y, idx = train_net_ref.caffe_pool_argmax(solver.blob(L['bottom'][0]), L['kernel'], L['stride'], L['pad'])
        assert np.array_equal(y.view(np.uint32), solver.blob(name).view(np.uint32)), name
        assert np.array_equal(idx, argmax[name]), name
    ref = train_net_ref.forward_backward(layers, weights, data, grids, dropout_mask=mask, want=want, pool_argmax=argmax)
    free = train_net_ref.forward_backward(layers, weights, data, grids, dropout_mask=mask, want=('coverage',))
    assert abs(free['loss_coverage'] - ref['loss_coverage']) < 1e-6 * abs(ref['loss_coverage'])      # same forward values

    # ---- forward
    for k in want:
        e = rel_err(solver.blob(k), ref['blobs'][k])
        assert e < 1e-4, ('data', k, e)
    assert abs(got_loss['loss_bbox'] - ref['loss_bbox']) <= 1e-5 * max(1.0, abs(ref['loss_bbox']))
    assert abs(got_loss['loss_coverage'] - ref['loss_coverage']) <= 1e-5 * max(1.0, abs(ref['loss_coverage']))
    # ---- gradients of the activations (ReLU is in place: Caffe's diff of a convolution top is the gradient AFTER the ReLU
    # mask, autograd's retained gradient of the ReLU output is the one before it; compare blobs that are not ReLU tops)
    for k in ('pool1/norm1', 'conv2/norm2', 'pool2/3x3_s2', 'inception_3b/pool', 'pool3/3x3_s2', 'pool5/drop_s1', 'coverage', 'bboxes'):
        e = rel_err(solver.blob(k, 'diff'), ref['blob_grads'][k])
        assert e < 1e-4, ('diff', k, e)
    # ---- gradients of all 59 x 2 parameter blobs
    worst = 0.0
    for name in solver.layers:
        gw, gb = ref['grads'][name]
        ew = rel_err(solver.param(name, 0, 'grad').reshape(gw.shape), gw)
        eb = rel_err(solver.param(name, 1, 'grad'), gb)
        worst = max(worst, ew, eb)
        assert ew < 2e-4 and eb < 2e-4, (name, ew, eb)
    # ---- ApplyUpdate, twice (the second one exercises the moments and the bias correction at t = 2)
    state = {}
    grads = {k: [solver.param(k, 0, 'grad').reshape(weights[k][0].shape), solver.param(k, 1, 'grad')] for k in solver.layers}
    solver.apply_update(1.0, stream)
    mults = train_net_ref.multipliers(layers)
    assert mults['inception_5b/3x3_reduce'][0] == [1.0, 1.0] and mults['conv1/7x7_s2'][0] == [1.0, 2.0]
    new_w = train_net_ref.adam_update(weights, grads, state, 0, mults=mults)
    for name in solver.layers:
        assert np.abs(solver.param(name, 0).reshape(new_w[name][0].shape) - new_w[name][0]).max() < 1e-6, name
        assert np.abs(solver.param(name, 1) - new_w[name][1]).max() < 1e-6, name
    # first ADAM step: every weight with a gradient well above delta moves by rate * lr_mult (known answer)
    # (m = 0.1 g, v = 0.001 g^2, correction = sqrt(0.001) / 0.1: step = 1e-4 * g / (|g| + 1e-8 / sqrt(0.001)) -> 1e-4 * sign(g))
    g = grads['conv2/3x3'][0].astype(np.float64)
    moved = weights['conv2/3x3'][0].astype(np.float64) - solver.param('conv2/3x3', 0).reshape(g.shape)
    big = np.abs(g) > 1e-5            # well above weight_decay * w (4e-9)
    assert big.sum() > 100
    assert np.allclose(moved[big], 1e-4 * g[big] / (np.abs(g[big]) + 1e-8 / np.sqrt(0.001)), rtol=2e-3, atol=1e-8)
    assert solver.iter == 1
    solver.forward_backward_device([t.data_ptr() for t in dev], stream=stream)
    grads2 = {k: [solver.param(k, 0, 'grad').reshape(weights[k][0].shape), solver.param(k, 1, 'grad')] for k in solver.layers}
    solver.apply_update(1.0, stream)
    new_w2 = train_net_ref.adam_update(new_w, grads2, state, 1, mults=mults)
    for name in ('conv1/7x7_s2', 'inception_4c/5x5', 'inception_5b/3x3_reduce', 'cvg/classifier', 'bbox/regressor'):
        assert np.abs(solver.param(name, 0).reshape(new_w2[name][0].shape) - new_w2[name][0]).max() < 2e-6, name
        assert np.abs(solver.param(name, 1) - new_w2[name][1]).max() < 2e-6, name
    os.makedirs('gpurun_out', exist_ok=True)
    with open('gpurun_out/train_parity.txt', 'w') as f:
        f.write('worst relative error over the 118 parameter gradients (fp32 engine vs torch autograd, %dx%dx%d): %.3g\n' % (N, H, W, worst))


def test_dropout_masks_are_reproducible_and_change_per_iteration(train_files):
    import torch
    from fcn_object_detector_b200.trainer import Solver
    proto, cm, sp = train_files
    data, grids = _inputs(1)
    dev = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (data,) + grids]
    masks = []
    for seed in (7, 7, 8):
        s = Solver(proto, cm, batch=N, height=H, width=W, solver_prototxt=sp, seed=seed)
        try:
            s.forward_backward_device([t.data_ptr() for t in dev])
            m0 = s.dropout_mask()
            s.apply_update()
            s.forward_backward_device([t.data_ptr() for t in dev])
            masks.append((m0, s.dropout_mask()))
        finally:
            s.close()
    assert np.array_equal(masks[0][0], masks[1][0]) and np.array_equal(masks[0][1], masks[1][1])
    assert not np.array_equal(masks[0][0], masks[0][1]) and not np.array_equal(masks[0][0], masks[2][0])


def test_training_reduces_the_loss_and_the_snapshot_loads_in_the_detector(train_files, tmp_path):
    """Host-buffer iterations on a fixed minibatch: the loss goes down; the snapshot written afterwards is a .caffemodel the
    inference engine loads, and its fp32 forward reproduces the trainer's own TEST-phase coverage (dropout off)."""
    from fcn_object_detector_b200 import netspec
    from fcn_object_detector_b200.trainer import Solver
    import fcn_object_detector_b200.caffe_shim as caffe
    proto, cm, sp = train_files
    data, grids = _inputs(2)
    s = Solver(proto, cm, batch=N, height=H, width=W, solver_prototxt=sp, seed=3, dropout_ratio_override=0.0, base_lr=1e-3)
    try:
        losses = [s.step(data, grids)['loss'] for _ in range(12)]
        assert losses[-1] < 0.7 * losses[0], losses
        assert all(np.isfinite(losses))
        snap = s.snapshot(str(tmp_path / 'snapshot_iter_12.caffemodel'), netspec.conv_shape_dict(num_cvg=1, num_bbox=4))
        # one more forward (no update needed: compare activations of the current weights)
        import torch
        dev = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (data,) + grids]
        s.forward_backward_device([t.data_ptr() for t in dev])
        cov_train, box_train = s.blob('coverage'), s.blob('bboxes')
    finally:
        s.close()
    deploy = netspec.write_deploy_prototxt(str(tmp_path / 'deploy.prototxt'), batch=N, height=H, width=W, num_cvg=1, num_bbox=4)
    net = caffe.Net(deploy, snap, caffe.TEST, device=0, precision='fp32')
    try:
        net.blobs['data'].reshape(N, 3, H, W)
        net.blobs['data'].data[...] = data
        net.forward()
        cov_deploy, box_deploy = net.blobs['coverage'].data.copy(), net.blobs['bboxes'].data.copy()
    finally:
        net.close()
    # two fp32 engines (different tilings and sigmoid formulations): the linear head at 1e-5, the sigmoid output at 1e-4
    assert rel_err(box_deploy, box_train) < 1e-5
    assert rel_err(cov_deploy, cov_train) < 1e-4


def test_bf16_tensor_core_engine_against_the_oracle_and_the_fp32_engine(train_files):
    """The bf16 engine (activations / gradients in bf16, convolutions as tcgen05 GEMMs, fp32 master weights): forward within
    1e-2 of the oracle on the outputs, parameter gradients within the accumulated bf16 rounding of the backward chain (see
    below; the pooling selection of the device is given to the oracle), and training makes the same kind of progress as the
    fp32 engine."""
    import torch
    from fcn_object_detector_b200.trainer import Solver
    proto, cm, sp = train_files
    layers = prototxt.layers(prototxt.parse(open(proto).read()))
    weights = {k: [np.asarray(a, np.float32) for a in v] for k, v in _oracle_weights(cm).items()}
    data, grids = _inputs(0)
    dev = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (data,) + grids]
    s = Solver(proto, cm, batch=N, height=H, width=W, solver_prototxt=sp, seed=7, precision='bf16')
    try:
        s.forward_backward_device([t.data_ptr() for t in dev])
        torch.cuda.synchronize()
        mask = s.dropout_mask()
        pools = [L['name'] for L in layers if L['type'] == 'Pooling']
        argmax = {name: s.pool_argmax(name) for name in pools}
        want = ['conv1/7x7_s2', 'conv2/norm2', 'inception_3a/output', 'inception_4e/output', 'inception_5b/output', 'coverage', 'bboxes']
        ref = train_net_ref.forward_backward(layers, weights, data, grids, dropout_mask=mask, want=want, pool_argmax=argmax)
        errs = {k: rel_err(s.blob(k), ref['blobs'][k]) for k in want}
        assert errs['coverage'] < 1e-2 and errs['bboxes'] < 1e-2, errs
        assert all(e < 3e-2 for e in errs.values()), errs
        got = s.losses()
        assert abs(got['loss_coverage'] - ref['loss_coverage']) < 1e-2 * ref['loss_coverage']
        # Parameter gradients.  bf16 rounds every activation and every gradient blob on the way down (8 mantissa bits, ~0.4 %
        # per element and stage) and the error accumulates over the 22 convolution stages between the loss and conv1: measured
        # as the relative L2 error of each gradient blob it grows from 1e-2 at inception_5b to 6e-2 at inception_3a and 0.15
        # at conv1 (profiles/r02w_train_parity_bf16.txt).  Gate: 0.1 per blob (0.2 for conv1) and a cosine of at least 0.98
        # with the oracle's gradient -- the fp32 engine is the parity engine (1e-4, test above).
        def l2(a, b):
            return float(np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-30))

        def cos(a, b):
            return float(np.dot(a.ravel(), b.ravel()) / max(np.linalg.norm(a.ravel()) * np.linalg.norm(b.ravel()), 1e-30))
        worst, worst_name, report = 0.0, None, []
        for name in s.layers:
            gw, gb = ref['grads'][name]
            a = s.param(name, 0, 'grad').reshape(gw.shape)
            e = l2(a, gw)
            report.append('%-28s dW rel-L2 %.3e cos %.5f  db rel-L2 %.3e' % (name, e, cos(a, gw), l2(s.param(name, 1, 'grad'), gb)))
            assert e < (0.2 if name == 'conv1/7x7_s2' else 0.1) and cos(a, gw) > 0.98, report[-1]
            if e > worst:
                worst, worst_name = e, name
        with open('gpurun_out/train_parity_bf16.txt', 'w') as f:
            f.write('bf16 engine vs fp32 oracle (%dx%dx%d): forward errors %s; worst parameter-gradient relative L2 error %.3g (%s)\n' % (N, H, W, errs, worst, worst_name))
            f.write('\n'.join(report) + '\n')
    finally:
        s.close()
    # a few iterations on a fixed minibatch: the loss goes down like the fp32 engine's
    out = {}
    for prec in ('fp32', 'bf16'):
        s = Solver(proto, cm, batch=N, height=H, width=W, solver_prototxt=sp, seed=3, dropout_ratio_override=0.0, base_lr=1e-3, precision=prec)
        try:
            out[prec] = [s.step(data, grids)['loss'] for _ in range(8)]
        finally:
            s.close()
    assert out['bf16'][-1] < 0.8 * out['bf16'][0], out
    assert abs(out['bf16'][0] - out['fp32'][0]) < 2e-2 * out['fp32'][0], out


def test_three_class_network_of_train_val2(tmp_path):
    """models/train_val2.prototxt: the same network with 3 classes (heads of 3 / 12 channels, slice points 3 / 15 / 27 / 39).
    fp32 engine against the oracle on the losses, the head gradients and the first convolution's gradient; the bf16 engine
    runs the same step (its 12-channel regressor stays on the SIMT kernels) within 2e-2 of the fp32 losses."""
    import torch
    from fcn_object_detector_b200 import netspec
    from fcn_object_detector_b200.caffemodel import write_synthetic_caffemodel
    from fcn_object_detector_b200.trainer import Solver
    C3 = 3
    proto = netspec.write_train_val_prototxt(str(tmp_path / 'train_val2.prototxt'), num_classes=C3)
    cm = write_synthetic_caffemodel(str(tmp_path / 'init3.caffemodel'), seed=5, num_cvg=C3, num_bbox=4 * C3)
    layers = prototxt.layers(prototxt.parse(open(proto).read()))
    weights = {k: [np.asarray(a, np.float32) for a in v] for k, v in _oracle_weights(cm).items()}
    rs = np.random.RandomState(9)
    data = rs.uniform(0, 255, (N, 3, H, W)).astype(np.float32)
    cvg = (rs.uniform(0, 1, (N, C3, GH, GW)) > 0.8).astype(np.float32)
    blk = np.repeat(cvg, 4, axis=1)
    grids = (cvg, (rs.uniform(-30, 30, (N, 4 * C3, GH, GW)) * blk).astype(np.float32), (blk / 64.0).astype(np.float32),
             (blk / 6.0).astype(np.float32), blk.astype(np.float32))
    dev = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (data,) + grids]
    losses = {}
    for prec in ('fp32', 'bf16'):
        s = Solver(proto, cm, batch=N, height=H, width=W, seed=11, precision=prec)
        try:
            assert s.num_classes == C3
            s.forward_backward_device([t.data_ptr() for t in dev])
            torch.cuda.synchronize()
            losses[prec] = s.losses()
            if prec == 'fp32':
                argmax = {L['name']: s.pool_argmax(L['name']) for L in layers if L['type'] == 'Pooling'}
                ref = train_net_ref.forward_backward(layers, weights, data, grids, dropout_mask=s.dropout_mask(), pool_argmax=argmax)
                assert abs(losses[prec]['loss_bbox'] - ref['loss_bbox']) <= 1e-5 * max(1.0, abs(ref['loss_bbox']))
                assert abs(losses[prec]['loss_coverage'] - ref['loss_coverage']) <= 1e-5 * max(1.0, abs(ref['loss_coverage']))
                for name in ('bbox/regressor', 'cvg/classifier', 'inception_4c/3x3', 'conv1/7x7_s2'):
                    gw, gb = ref['grads'][name]
                    assert rel_err(s.param(name, 0, 'grad').reshape(gw.shape), gw) < 2e-4, name
                    assert rel_err(s.param(name, 1, 'grad'), gb) < 2e-4, name
        finally:
            s.close()
    for k in ('loss_bbox', 'loss_coverage'):
        assert abs(losses['bf16'][k] - losses['fp32'][k]) < 2e-2 * max(1.0, abs(losses['fp32'][k])), (k, losses)


_IM2COL_SCRIPT = r"""
import sys
import numpy as np
import torch
sys.path.insert(0, %(root)r)
from fcn_object_detector_b200.trainer import Solver
from tests.test_trainer_gpu import _inputs, N, H, W
proto, cm, sp, out = sys.argv[1:5]
data, grids = _inputs(0)
dev = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (data,) + grids]
s = Solver(proto, cm, batch=N, height=H, width=W, solver_prototxt=sp, seed=7, precision='bf16')
s.forward_backward_device([t.data_ptr() for t in dev])
torch.cuda.synchronize()
np.savez(out, conv1=s.blob('conv1/7x7_s2'), cov=s.blob('coverage'), gw=s.param('conv1/7x7_s2', 0, 'grad'), gb=s.param('conv1/7x7_s2', 1, 'grad'),
         allb=np.concatenate([s.param(n, 1, 'grad') for n in s.layers]), w3a=s.param('inception_3a/3x3', 0, 'grad'), w5b=s.param('inception_5b/pool_proj', 0, 'grad'))
s.close()
"""


def test_staged_im2col_of_the_image_writes_what_the_generic_gather_writes(train_files, tmp_path):
    """conv1's column matrix: im2col_c3_kernel (input rows staged in shared memory) against im2col_kernel<1, 7>
    (FCNOD_TRAIN_IM2COL_C3=0), one process each since the knob is read once.  The forward is deterministic, so conv1's output
    and the network's coverage map must be bit-identical; conv1's weight gradient (the GEMM over the kept column matrix,
    fp32 reduction order not fixed) to 1e-5."""
    import subprocess
    import sys
    from tests.util import ROOT
    proto, cm, sp = train_files
    script = tmp_path / 'im2col_variant.py'
    script.write_text(_IM2COL_SCRIPT % {'root': ROOT})
    got = {}
    for knob in ('0', '1'):
        out = str(tmp_path / ('im2col_%s.npz' % knob))
        env = dict(os.environ, FCNOD_TRAIN_IM2COL_C3=knob)
        r = subprocess.run([sys.executable, str(script), proto, cm, sp, out], cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        got[knob] = np.load(out)
    assert np.abs(got['1']['conv1']).max() > 0
    assert np.array_equal(got['0']['conv1'], got['1']['conv1'])
    assert np.array_equal(got['0']['cov'], got['1']['cov'])
    assert rel_err(got['1']['gw'], got['0']['gw']) < 1e-5 and rel_err(got['1']['gb'], got['0']['gb']) < 1e-5


def test_relu_and_bias_gradient_pass_per_concat_blob_equals_the_pass_per_convolution(train_files, tmp_path):
    """FCNOD_TRAIN_RB_FUSE=1 (one ReLU-mask / bias-gradient launch over a Concat blob, bias sums routed to the writing
    convolutions by a segment table) against one launch per convolution: every bias gradient of the network and the weight
    gradients of three layers (which see every masked dY below them) to 1e-5 (fp32 atomics, order not fixed)."""
    import subprocess
    import sys
    from tests.util import ROOT
    proto, cm, sp = train_files
    script = tmp_path / 'rb_variant.py'
    script.write_text(_IM2COL_SCRIPT % {'root': ROOT})
    got = {}
    for knob in ('0', '1'):
        out = str(tmp_path / ('rb_%s.npz' % knob))
        env = dict(os.environ, FCNOD_TRAIN_RB_FUSE=knob)
        r = subprocess.run([sys.executable, str(script), proto, cm, sp, out], cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        got[knob] = np.load(out)
    assert np.abs(got['0']['allb']).max() > 0
    for k in ('allb', 'gw', 'w3a', 'w5b'):
        assert rel_err(got['1'][k], got['0'][k]) < 1e-5, (k, rel_err(got['1'][k], got['0'][k]))
    assert np.array_equal(got['0']['cov'], got['1']['cov'])


def test_peer_memory_update_matches_the_nccl_path_on_two_gpus():
    """Data-parallel update: reduce-scatter + ADAM + all-gather in one kernel over NVLink peer memory (CUDA IPC) against
    NCCL all-reduce + update, two ranks under torchrun (tools/dp_check.py): replicas bit-identical, reduced gradients of the
    rank's shard within 1e-4, parameters equal to 1e-6.  Needs two GPUs."""
    import subprocess
    import sys
    import torch
    from tests.util import ROOT
    if torch.cuda.device_count() < 2:
        pytest.skip('needs 2 GPUs')
    r = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr', '127.0.0.1',
                        '--master-port', '29517', os.path.join(ROOT, 'tools', 'dp_check.py'), 'bf16'], cwd=ROOT, capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert 'replicas identical nccl=True peer=True' in r.stdout


def test_full_size_448_step_against_the_oracle(train_files):
    """BASELINE config 5's geometry (448x448, 28x28 label grids), one frame: fp32 engine losses at 1e-5 and the gradients of the
    first, a middle and the last convolutions at 2e-4; bf16 engine losses at 1e-2."""
    import torch
    from fcn_object_detector_b200.trainer import Solver
    proto, cm, sp = train_files
    layers = prototxt.layers(prototxt.parse(open(proto).read()))
    weights = {k: [np.asarray(a, np.float32) for a in v] for k, v in _oracle_weights(cm).items()}
    rs = np.random.RandomState(21)
    data = rs.uniform(0, 255, (1, 3, 448, 448)).astype(np.float32)
    cvg = np.zeros((1, 1, 28, 28), np.float32)
    cvg[0, 0, 9:15, 11:19] = 1
    blk = np.repeat(cvg, 4, axis=1)
    grids = (cvg, (rs.uniform(-60, 60, (1, 4, 28, 28)) * blk).astype(np.float32), (blk / 128.0).astype(np.float32), (blk / 48.0).astype(np.float32), blk)
    dev = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (data,) + grids]
    ref = None
    for prec in ('fp32', 'bf16'):
        s = Solver(proto, cm, batch=1, height=448, width=448, solver_prototxt=sp, seed=13, precision=prec)
        try:
            assert (s.grid_h, s.grid_w) == (28, 28)
            s.forward_backward_device([t.data_ptr() for t in dev])
            torch.cuda.synchronize()
            got = s.losses()
            if prec == 'fp32':
                argmax = {L['name']: s.pool_argmax(L['name']) for L in layers if L['type'] == 'Pooling'}
                ref = train_net_ref.forward_backward(layers, weights, data, grids, dropout_mask=s.dropout_mask(), pool_argmax=argmax)
                for k in ('loss_bbox', 'loss_coverage'):
                    assert abs(got[k] - ref[k]) <= 1e-5 * max(1.0, abs(ref[k])), (k, got[k], ref[k])
                for name in ('conv1/7x7_s2', 'inception_4a/5x5', 'bbox/regressor'):
                    gw, gb = ref['grads'][name]
                    assert rel_err(s.param(name, 0, 'grad').reshape(gw.shape), gw) < 2e-4, name
                    assert rel_err(s.param(name, 1, 'grad'), gb) < 2e-4, name
            else:
                # (its dropout mask is the same function of (seed, iteration, element), so the oracle run above applies)
                for k in ('loss_bbox', 'loss_coverage'):
                    assert abs(got[k] - ref[k]) <= 1e-2 * max(1.0, abs(ref[k])), (k, got[k], ref[k])
        finally:
            s.close()


def test_trainer_refuses_a_head_or_a_network_it_does_not_implement(train_files, tmp_path):
    """The fused head is the one of train_val.prototxt: other slice points, an Eltwise SUM, another bbox loss weight than the
    solver's, or an in-place-less ReLU are refused with an error that names the layer (never trained differently)."""
    import ctypes as C
    from fcn_object_detector_b200 import _lib
    proto, cm, _ = train_files
    text = open(proto).read()
    cases = {
        'slice': (text.replace('slice_point: 5', 'slice_point: 6'), 'slice-label'),
        'eltwise': (text.replace('operation: PROD', 'operation: SUM', 1), 'bb-label-norm'),
        'loss_weight': (text.replace('loss_weight: 2.0', 'loss_weight: 3.0'), 'bbox_loss'),
        'layer_type': (text.replace('type: "LRN"', 'type: "BatchNorm"', 1), 'BatchNorm'),
    }
    L = _lib.lib()
    for tag, (bad, needle) in cases.items():
        path = str(tmp_path / ('bad_%s.prototxt' % tag))
        with open(path, 'w') as f:
            f.write(bad)
        h = C.c_void_p()
        rc = L.fcnod_trainer_create(path.encode(), cm.encode(), 0, 1, 64, 64, None, C.byref(h))
        msg = L.fcnod_last_error().decode()
        assert rc != 0 and needle in msg, (tag, rc, msg)
    # wrong label-grid shapes are refused by the host mirror before anything is launched
    from fcn_object_detector_b200.trainer import Solver
    s = Solver(proto, cm, batch=1, height=64, width=64)
    try:
        with pytest.raises(ValueError):
            s.step(np.zeros((1, 3, 64, 64), np.float32), [np.zeros((1, 1, 4, 4), np.float32)] * 5)
    finally:
        s.close()
