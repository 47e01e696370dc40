"""GPU: network forward through the C ABI against the oracle.
fp32 mode (SIMT): 1e-5 relative.  bf16 mode (tcgen05): 1e-2 relative on coverage / bboxes
(metric max|a-b| / max|b| per blob)."""
import os

import numpy as np
import pytest

from tests.golden import make_golden as G
from tests.util import GOLDEN, bf16_round, gpu_forward, oracle_forward, rel_err, write_mini_net, write_stem_net

pytestmark = pytest.mark.gpu

TOL_FP32 = 1e-5
TOL_BF16 = 1e-2


def test_fp32_full_net_matches_oracle(model_files):
    proto, cm = model_files
    x = np.random.RandomState(2).uniform(0, 1, (2, 3, 448, 448)).astype(np.float32)
    want = ['conv1/7x7_s2', 'pool1/3x3_s2', 'pool1/norm1', 'conv2/3x3', 'conv2/norm2', 'pool2/3x3_s2',
            'inception_3a/output', 'inception_3b/output', 'pool3/3x3_s2', 'inception_4a/output',
            'inception_4e/output', 'inception_5b/output', 'cvg/classifier', 'coverage', 'bboxes']
    exp = oracle_forward(proto, cm, x, want=want)
    got, plan = gpu_forward(proto, cm, x, 'fp32', want)
    assert 'conv_tc' not in plan
    for k in want:
        assert got[k].shape == exp[k].shape, k
        assert rel_err(got[k], exp[k]) < TOL_FP32, (k, rel_err(got[k], exp[k]))


def test_fp32_golden_fixture(model_files):
    proto, cm = model_files
    g = np.load(os.path.join(GOLDEN, 'net_seed2.npz'))
    got, _ = gpu_forward(proto, cm, G.net_inputs(), 'fp32', ['coverage', 'bboxes', 'pool1/norm1', 'inception_3a/output'])
    assert rel_err(got['coverage'], g['coverage']) < TOL_FP32 and rel_err(got['bboxes'], g['bboxes']) < TOL_FP32
    assert rel_err(got['pool1/norm1'][:, ::8, ::16, ::16], g['norm1_probe']) < TOL_FP32
    assert rel_err(got['inception_3a/output'][:, ::16, ::8, ::8], g['i3a_probe']) < TOL_FP32


TC_CASES = [
    # tag, H, W, cin, branches [(cout, k)], batch
    ('k1_c64', 28, 28, 64, [(64, 1)], 2),
    ('k3_c64', 28, 28, 64, [(128, 3)], 2),
    ('k5_c16', 28, 28, 16, [(32, 5)], 2),
    ('k3_c96_sw64', 56, 56, 96, [(128, 3)], 1),
    ('k5_c24', 28, 28, 24, [(64, 5)], 1),
    ('k1_c480', 28, 28, 480, [(192, 1), (96, 1)], 1),
    ('k3_c160_n320', 28, 28, 160, [(320, 3), (24, 1)], 1),
    ('k3_w112', 112, 112, 64, [(192, 3)], 1),
    ('mix_3a', 56, 56, 192, [(64, 1), (128, 3), (32, 5), (32, 1)], 2),
    ('odd_hw', 20, 36, 32, [(48, 3), (16, 5)], 3),
    ('cin48_k5', 28, 28, 48, [(128, 5)], 1),
    ('big_w', 12, 160, 64, [(64, 3)], 1),
    # max-pool 3x3/1 fused into the A-operand path of its 1x1 consumer (POOL mode)
    ('poolproj_28', 28, 28, 192, [(64, 1), (32, 'p')], 2),
    ('poolproj_56', 56, 56, 192, [(32, 'p'), (64, 3)], 1),
    ('poolproj_odd', 20, 36, 72, [(48, 'p')], 3),              # channel tail (72 = 64 + 8), ragged tiles
    ('poolproj_n160', 28, 28, 128, [(160, 'p')], 1),           # two N tiles
    ('poolproj_k832', 28, 28, 832, [(128, 'p')], 4),           # 13 channel chunks, streamed weights
]


@pytest.mark.parametrize('case', TC_CASES, ids=[c[0] for c in TC_CASES])
def test_conv_tc_mini_nets(case, tmp_path):
    """Every tcgen05 tiling the deploy net uses (CK 16/32/64, BN 16..128, 1x1/3x3/5x5, concat slices),
    against the oracle with the same bf16 operand rounding: only fp32 accumulation order differs."""
    tag, H, W, cin, branches, batch = case
    proto, cm = write_mini_net(str(tmp_path), tag, H, W, cin, branches, batch=batch, seed=len(tag))
    x = np.random.RandomState(1).uniform(0, 1, (batch, 3, H, W)).astype(np.float32)
    names = ['br%d' % i for i in range(len(branches))]
    tc_layers = set(names + ['final'])
    want = names + ['out']
    exp = oracle_forward(proto, cm, x, bf16=True, want=want, tc_layers=tc_layers)
    got, plan = gpu_forward(proto, cm, x, 'bf16', want)
    assert 'conv_simt br' not in plan and 'conv_simt final' not in plan, plan     # every branch + the head on tcgen05
    if any(k == 'p' for _c, k in branches):
        assert '+maxpool3x3' in plan and '\npool ' not in plan and ': pool ' not in plan, plan
    assert all(('br%d ' % i) in plan or ('br%d\n' % i) in plan for i in range(len(branches))), plan
    for k in names:
        # stored bf16: allow one bf16 ulp of the blob's magnitude
        assert rel_err(got[k], exp[k]) < 1.0 / 128, (k, rel_err(got[k], exp[k]))
    assert rel_err(got['out'], exp['out']) < 5e-3, rel_err(got['out'], exp['out'])
    # and against the true fp32 oracle at the bf16 tolerance
    ref = oracle_forward(proto, cm, x, want=['out'])
    assert rel_err(got['out'], ref['out']) < TOL_BF16


def test_pool_of_signed_input_is_not_fused(tmp_path):
    """POOL mode substitutes zero fill for Caffe's clipped pooling windows, which is only valid on non-negative input:
    a pool behind a convolution WITHOUT ReLU must stay a separate launch (and still match the oracle)."""
    from fcn_object_detector_b200.caffemodel import write_caffemodel
    from tests.util import _conv_txt
    rs = np.random.RandomState(5)
    txt = 'input: "data"\ninput_shape { dim: 2 dim: 3 dim: 28 dim: 28 }\n'
    txt += 'layer { name: "xf" type: "Power" bottom: "data" top: "x" power_param { shift: -0.5 } }\n'
    txt += _conv_txt('pre', 'x', 'pre', 64, 1, relu=False)
    txt += 'layer { name: "pl" type: "Pooling" bottom: "pre" top: "pl" pooling_param { pool: MAX kernel_size: 3 stride: 1 pad: 1 } }\n'
    txt += _conv_txt('proj', 'pl', 'out', 32, 1, relu=False)
    layers = [('pre', 'Convolution', [rs.uniform(-1, 1, (64, 3, 1, 1)).astype(np.float32), rs.uniform(-0.5, 0.1, 64).astype(np.float32)]),
              ('proj', 'Convolution', [rs.uniform(-0.2, 0.2, (32, 64, 1, 1)).astype(np.float32), rs.uniform(-0.2, 0.2, 32).astype(np.float32)])]
    proto = str(tmp_path / 'signed.prototxt')
    open(proto, 'w').write(txt)
    cm = write_caffemodel(str(tmp_path / 'signed.caffemodel'), layers)
    x = rs.uniform(0, 1, (2, 3, 28, 28)).astype(np.float32)
    got, plan = gpu_forward(proto, cm, x, 'bf16', ['pl', 'out'])
    assert '+maxpool3x3' not in plan and 'pool pl' in plan, plan
    ref = oracle_forward(proto, cm, x, want=['pl', 'out'])
    assert (ref['pl'] < 0).any()
    assert rel_err(got['pl'], ref['pl']) < TOL_BF16 and rel_err(got['out'], ref['out']) < TOL_BF16


STEM_CASES = [
    # tag, H, W, batch, k, pad, cout, shift, scale, c2
    ('deploy_like', 64, 96, 2, 7, 3, 64, -127.0, 1.0, 192),
    ('two_boxes', 40, 300, 1, 7, 3, 64, -127.0, 1.0, 64),      # W/2 = 150 -> two 75-pixel boxes per row
    ('k3_p1', 36, 52, 2, 3, 1, 32, -0.5, 2.0, 64),               # ka = 2 taps
    ('k5_p2', 48, 48, 1, 5, 2, 48, 0.25, 0.5, 64),               # ka = 3 -> padded to 4
    ('full_rows', 448, 64, 1, 7, 3, 64, -127.0, 1.0, 64),        # all 224 output rows, every border class
]


@pytest.mark.parametrize('case', STEM_CASES, ids=[c[0] for c in STEM_CASES])
def test_s2d_conv_and_fused_pool_lrn(case, tmp_path, monkeypatch):
    """bf16 plan of the stem: Power folded into the stride-2 convolution (space-to-depth, tcgen05), fused
    Pool->LRN and LRN->Pool, against the fp32 oracle and against the unfused bf16 plan."""
    tag, H, W, batch, k, pad, cout, shift, scale, c2 = case
    proto, cm = write_stem_net(str(tmp_path), tag, H, W, batch=batch, seed=len(tag), k=k, pad=pad, cout=cout, shift=shift,
                               scale=scale, c2=c2)
    x = np.random.RandomState(4).uniform(0, 1, (batch, 3, H, W)).astype(np.float32)
    want = ['c1', 'n1', 'n2', 'p2', 'out']
    ref = oracle_forward(proto, cm, x, want=want)
    got, plan = gpu_forward(proto, cm, x, 'bf16', want)
    # Pool -> LRN is one launch; the second LRN runs in the epilogue of the 3x3 convolution before it (its pool is then a
    # plain launch), or, where that does not apply, LRN -> Pool is one launch
    assert '+s2d+power' in plan and 'pool+lrn' in plan and ('+lrn(n2)' in plan or 'lrn+pool' in plan) and 'conv_simt' not in plan, plan
    errs = {n: rel_err(got[n], ref[n]) for n in want}
    # c1 is stored in bf16 (1 ulp = 2^-8 of the value); later blobs accumulate a few such roundings
    assert errs['c1'] < 1.0 / 128, errs
    assert all(e < 2e-2 for e in errs.values()), errs
    # the fused-away blobs answer with a clear error, the unfused plan materialises them and agrees
    import fcn_object_detector_b200.caffe_shim as caffe
    from fcn_object_detector_b200._lib import FcnodError
    net = caffe.Net(proto, cm, caffe.TEST, device=0, precision='bf16')
    net.blobs['data'].reshape(*x.shape); net.blobs['data'].data[...] = x; net.forward()
    with pytest.raises(FcnodError):
        net.blobs['p1'].data
    if '+lrn(n2)' in plan:
        with pytest.raises(FcnodError):
            net.blobs['c2'].data            # the un-normalised convolution output is never materialised
    net.close()
    monkeypatch.setenv('FCNOD_NO_FUSION', '1')
    unf, plan2 = gpu_forward(proto, cm, x, 'bf16', want + ['p1', 'c2'])
    assert 'pool+lrn' not in plan2 and 'conv_simt' in plan2
    for n in ('n1', 'n2', 'p2', 'out'):
        assert rel_err(got[n], unf[n]) < 2e-2, (n, rel_err(got[n], unf[n]))
    # and with the epilogue LRN switched off the LRN -> Pool launch takes over, with the same result
    monkeypatch.delenv('FCNOD_NO_FUSION')
    monkeypatch.setenv('FCNOD_NO_CONV_LRN', '1')
    alt, plan3 = gpu_forward(proto, cm, x, 'bf16', ['p2', 'out'])
    assert 'lrn+pool' in plan3 and '+lrn(' not in plan3
    for n in ('p2', 'out'):
        assert rel_err(got[n], alt[n]) < 2e-2, (n, rel_err(got[n], alt[n]))


def test_bf16_full_net_within_tolerance(model_files):
    proto, cm = model_files
    x = np.random.RandomState(2).uniform(0, 1, (2, 3, 448, 448)).astype(np.float32)
    want = ['conv1/7x7_s2', 'pool1/norm1', 'conv2/norm2', 'pool2/3x3_s2', 'inception_3a/output', 'inception_4e/output',
            'inception_5b/output', 'coverage', 'bboxes']
    exp = oracle_forward(proto, cm, x, want=want)
    got, plan = gpu_forward(proto, cm, x, 'bf16', want)
    assert 'conv_simt' not in plan and '+s2d+power' in plan
    errs = {k: rel_err(got[k], exp[k]) for k in want}
    print(errs)
    assert errs['coverage'] < TOL_BF16 and errs['bboxes'] < TOL_BF16, errs
    for k in want[:-2]:
        assert errs[k] < 3e-2, errs


def test_batch_independence_bf16(model_files):
    """Size-independent property at a larger batch: frame i of a batch equals frame i run alone."""
    proto, cm = model_files
    x = np.random.RandomState(7).uniform(0, 1, (8, 3, 448, 448)).astype(np.float32)
    full, _ = gpu_forward(proto, cm, x, 'bf16', ['coverage', 'bboxes'])
    one, _ = gpu_forward(proto, cm, x[5:6], 'bf16', ['coverage', 'bboxes'])
    assert np.array_equal(full['coverage'][5], one['coverage'][0]) and np.array_equal(full['bboxes'][5], one['bboxes'][0])
