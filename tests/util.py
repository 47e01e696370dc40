import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, 'tests', 'golden')
REFERENCE = '/root/reference'


def rel_err(a, b):
    """max(abs(a-b)) / max(abs(b)) per blob -- the parity metric of BASELINE.md section 4."""
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def bits_equal(a, b):
    a = np.ascontiguousarray(a, np.float32); b = np.ascontiguousarray(b, np.float32)
    return a.shape == b.shape and np.array_equal(a.view(np.uint32), b.view(np.uint32))


# ----------------------------------------------------------------------------- mini networks for kernel tests
def _conv_txt(name, bottom, top, cout, k, pad=0, stride=1, relu=True):
    s = ('layer { name: "%s" type: "Convolution" bottom: "%s" top: "%s" convolution_param { num_output: %d '
         'kernel_size: %d pad: %d stride: %d } }\n' % (name, bottom, top, cout, k, pad, stride))
    if relu:
        s += 'layer { name: "%s/relu" type: "ReLU" bottom: "%s" top: "%s" }\n' % (name, top, top)
    return s


def write_mini_net(dirpath, tag, H, W, cin, branches, batch=1, seed=0, final=8):
    """data(3) -> Power -> pre 1x1 (3->cin, relu) -> [conv k (cin->cout) relu]* -> Concat -> final 1x1 (no relu).
    branches: list of (cout, k).  Returns (prototxt path, caffemodel path)."""
    from fcn_object_detector_b200.caffemodel import write_caffemodel
    rs = np.random.RandomState(seed)
    txt = 'input: "data"\ninput_shape { dim: %d dim: 3 dim: %d dim: %d }\n' % (batch, H, W)
    txt += 'layer { name: "xf" type: "Power" bottom: "data" top: "x" power_param { shift: -0.25 } }\n'
    txt += _conv_txt('pre', 'x', 'pre', cin, 1)
    layers = [('pre', 'Convolution', [rs.uniform(-1, 1, (cin, 3, 1, 1)).astype(np.float32),
                                      rs.uniform(-0.2, 0.5, cin).astype(np.float32)])]
    tops = []
    for i, (cout, k) in enumerate(branches):
        n = 'br%d' % i
        if k == 'p':      # max-pool 3x3 / stride 1 / pad 1 -> 1x1 (the inception `pool -> pool_proj` branch)
            txt += ('layer { name: "%sp" type: "Pooling" bottom: "pre" top: "%sp" pooling_param { pool: MAX kernel_size: 3 '
                    'stride: 1 pad: 1 } }\n' % (n, n))
            txt += _conv_txt(n, n + 'p', n, cout, 1)
            k = 1
        else:
            txt += _conv_txt(n, 'pre', n, cout, k, pad=(k - 1) // 2)
        s = np.sqrt(3.0 / (cin * k * k))
        layers.append((n, 'Convolution', [rs.uniform(-s, s, (cout, cin, k, k)).astype(np.float32),
                                          rs.uniform(-0.2, 0.2, cout).astype(np.float32)]))
        tops.append(n)
    ctot = sum(c for c, _k in branches)
    if len(tops) > 1:
        txt += 'layer { name: "cat" type: "Concat" %s top: "cat" }\n' % ' '.join('bottom: "%s"' % t for t in tops)
        last = 'cat'
    else:
        last = tops[0]
    txt += _conv_txt('final', last, 'out', final, 1, relu=False)
    s = np.sqrt(3.0 / ctot)
    layers.append(('final', 'Convolution', [rs.uniform(-s, s, (final, ctot, 1, 1)).astype(np.float32),
                                            rs.uniform(-0.2, 0.2, final).astype(np.float32)]))
    proto = os.path.join(dirpath, tag + '.prototxt')
    with open(proto, 'w') as f:
        f.write(txt)
    cm = write_caffemodel(os.path.join(dirpath, tag + '.caffemodel'), layers)
    return proto, cm


def write_stem_net(dirpath, tag, H, W, batch=1, seed=0, k=7, pad=3, cout=64, shift=-127.0, scale=1.0, mid=64, c2=64):
    """The deploy net's stem in miniature (deploy.prototxt:8-162): data -> Power(shift) -> conv kxk/2 relu -> pool 3/2
    -> LRN -> 1x1 (mid) relu -> 3x3 (c2) relu -> LRN -> pool 3/2 -> 1x1 out.  Exercises the bf16 plan's
    space-to-depth convolution (Power folded, border-aware bias) and the fused Pool->LRN / LRN->Pool kernels."""
    from fcn_object_detector_b200.caffemodel import write_caffemodel
    rs = np.random.RandomState(seed)
    txt = 'input: "data"\ninput_shape { dim: %d dim: 3 dim: %d dim: %d }\n' % (batch, H, W)
    txt += 'layer { name: "xf" type: "Power" bottom: "data" top: "x" power_param { shift: %r scale: %r } }\n' % (shift, scale)
    txt += _conv_txt('c1', 'x', 'c1', cout, k, pad=pad, stride=2)
    txt += 'layer { name: "p1" type: "Pooling" bottom: "c1" top: "p1" pooling_param { pool: MAX kernel_size: 3 stride: 2 } }\n'
    txt += 'layer { name: "n1" type: "LRN" bottom: "p1" top: "n1" lrn_param { local_size: 5 alpha: 0.0001 beta: 0.75 } }\n'
    txt += _conv_txt('r2', 'n1', 'r2', mid, 1)
    txt += _conv_txt('c2', 'r2', 'c2', c2, 3, pad=1)
    txt += 'layer { name: "n2" type: "LRN" bottom: "c2" top: "n2" lrn_param { local_size: 5 alpha: 0.0001 beta: 0.75 } }\n'
    txt += 'layer { name: "p2" type: "Pooling" bottom: "n2" top: "p2" pooling_param { pool: MAX kernel_size: 3 stride: 2 } }\n'
    txt += _conv_txt('final', 'p2', 'out', 8, 1, relu=False)

    def wb(co, ci, kk, bias_lo=-0.2, bias_hi=0.2):
        s_ = np.sqrt(3.0 / (ci * kk * kk))
        return [rs.uniform(-s_, s_, (co, ci, kk, kk)).astype(np.float32), rs.uniform(bias_lo, bias_hi, co).astype(np.float32)]
    layers = [('c1', 'Convolution', wb(cout, 3, k)), ('r2', 'Convolution', wb(mid, cout, 1)),
              ('c2', 'Convolution', wb(c2, mid, 3)), ('final', 'Convolution', wb(8, c2, 1))]
    proto = os.path.join(dirpath, tag + '.prototxt')
    with open(proto, 'w') as f:
        f.write(txt)
    cm = write_caffemodel(os.path.join(dirpath, tag + '.caffemodel'), layers)
    return proto, cm


def bf16_round(a):
    import torch
    return torch.as_tensor(np.asarray(a, np.float32)).to(torch.bfloat16).to(torch.float32).numpy()


def oracle_forward(proto, cm, x, bf16=False, want=None, tc_layers=None):
    """Oracle forward; bf16=True predicts the bf16 path: weights of tensor-core convs and every stored
    activation rounded to bf16 (fp32 accumulate)."""
    import torch
    from oracle import net_ref, prototxt
    from oracle.caffemodel_io import read_caffemodel
    Ls = prototxt.layers(prototxt.load(proto))
    W = read_caffemodel(cm)
    if not bf16:
        return net_ref.forward(Ls, W, x, want=want)
    Wb = {}
    for k, v in W.items():
        Wb[k] = [bf16_round(v[0]) if (tc_layers is None or k in tc_layers) else v[0]] + list(v[1:])

    def rnd(t):
        return t.to(torch.bfloat16).to(torch.float32)
    return net_ref.forward(Ls, Wb, x, want=want, act_round=rnd)


def gpu_forward(proto, cm, x, precision, want):
    import fcn_object_detector_b200.caffe_shim as caffe
    net = caffe.Net(proto, cm, caffe.TEST, device=0, precision=precision)
    try:
        net.blobs['data'].reshape(*x.shape)
        net.blobs['data'].data[...] = x
        net.forward()
        return {k: net.blobs[k].data.copy() for k in want}, net.describe()
    finally:
        net.close()
