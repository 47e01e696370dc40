"""Generator for the detector's network definition (Caffe prototxt text).

The upstream repository ships the network as `models/deploy.prototxt`
(/root/reference/models/deploy.prototxt:1-2178): a GoogLeNet-v1 body without
pool4/pool5 (stride 16) ending in a 1x1 coverage head (+sigmoid) and a 1x1
bbox-regression head.  That file is not copied here; this module *emits* an
equivalent definition from a compact table so tests, the smoke test and the
benchmark can run on a box where the upstream tree is absent.
tests/test_netspec.py checks layer-by-layer equivalence against the upstream
file whenever it is reachable.
"""

# (name, 1x1, 3x3_reduce, 3x3, 5x5_reduce, 5x5, pool_proj)
_INCEPTION = [
    ('3a', 64, 96, 128, 16, 32, 32),
    ('3b', 128, 128, 192, 32, 96, 64),
    'pool3',
    ('4a', 192, 96, 208, 16, 48, 64),
    ('4b', 160, 112, 224, 24, 64, 64),
    ('4c', 128, 128, 256, 24, 64, 64),
    ('4d', 112, 144, 288, 32, 64, 64),
    ('4e', 256, 160, 320, 32, 128, 128),
    ('5a', 256, 160, 320, 32, 128, 128),
    ('5b', 384, 192, 384, 48, 128, 128),
]


def _conv(name, bottom, top, num_output, kernel, pad=0, stride=1, bias=0.2):
    s = ['layer {', '  name: "%s"' % name, '  type: "Convolution"',
         '  bottom: "%s"' % bottom, '  top: "%s"' % top,
         '  param { lr_mult: 1.0 decay_mult: 1.0 }',
         '  param { lr_mult: 2.0 decay_mult: 0.0 }',
         '  convolution_param {', '    num_output: %d' % num_output]
    if pad:
        s.append('    pad: %d' % pad)
    s.append('    kernel_size: %d' % kernel)
    if stride != 1:
        s.append('    stride: %d' % stride)
    s += ['    weight_filler { type: "xavier" }',
          '    bias_filler { type: "constant" value: %s }' % repr(float(bias)),
          '  }', '}']
    return s


def _simple(name, typ, bottoms, top, body=()):
    s = ['layer {', '  name: "%s"' % name, '  type: "%s"' % typ]
    s += ['  bottom: "%s"' % b for b in bottoms]
    s.append('  top: "%s"' % top)
    s += ['  ' + b for b in body]
    s.append('}')
    return s


def _relu(name, blob):
    return _simple(name, 'ReLU', [blob], blob)


def _pool(name, bottom, kernel, stride, pad=0):
    body = ['pooling_param {', '  pool: MAX', '  kernel_size: %d' % kernel, '  stride: %d' % stride]
    if pad:
        body.append('  pad: %d' % pad)
    body.append('}')
    return _simple(name, 'Pooling', [bottom], name, body)


def _lrn(name, bottom):
    return _simple(name, 'LRN', [bottom], name,
                   ['lrn_param {', '  local_size: 5', '  alpha: 0.0001', '  beta: 0.75', '}'])


def deploy_prototxt(batch=1, height=448, width=448, num_cvg=4, num_bbox=16):
    """Text of the stride-16 detection network (deploy form)."""
    o = ['input: "data"', 'input_shape {', '  dim: %d' % batch, '  dim: 3',
         '  dim: %d' % height, '  dim: %d' % width, '}']
    o += _simple('deploy_transform', 'Power', ['data'], 'transformed_data',
                 ['power_param {', '  shift: -127.0', '}'])
    o += _conv('conv1/7x7_s2', 'transformed_data', 'conv1/7x7_s2', 64, 7, pad=3, stride=2)
    o += _relu('conv1/relu_7x7', 'conv1/7x7_s2')
    o += _pool('pool1/3x3_s2', 'conv1/7x7_s2', 3, 2)
    o += _lrn('pool1/norm1', 'pool1/3x3_s2')
    o += _conv('conv2/3x3_reduce', 'pool1/norm1', 'conv2/3x3_reduce', 64, 1)
    o += _relu('conv2/relu_3x3_reduce', 'conv2/3x3_reduce')
    o += _conv('conv2/3x3', 'conv2/3x3_reduce', 'conv2/3x3', 192, 3, pad=1)
    o += _relu('conv2/relu_3x3', 'conv2/3x3')
    o += _lrn('conv2/norm2', 'conv2/3x3')
    o += _pool('pool2/3x3_s2', 'conv2/norm2', 3, 2)
    prev = 'pool2/3x3_s2'
    for spec in _INCEPTION:
        if spec == 'pool3':
            o += _pool('pool3/3x3_s2', prev, 3, 2)
            prev = 'pool3/3x3_s2'
            continue
        tag, c1, c3r, c3, c5r, c5, cp = spec
        p = 'inception_%s/' % tag
        o += _conv(p + '1x1', prev, p + '1x1', c1, 1)
        o += _relu(p + 'relu_1x1', p + '1x1')
        o += _conv(p + '3x3_reduce', prev, p + '3x3_reduce', c3r, 1)
        o += _relu(p + 'relu_3x3_reduce', p + '3x3_reduce')
        o += _conv(p + '3x3', p + '3x3_reduce', p + '3x3', c3, 3, pad=1)
        o += _relu(p + 'relu_3x3', p + '3x3')
        o += _conv(p + '5x5_reduce', prev, p + '5x5_reduce', c5r, 1)
        o += _relu(p + 'relu_5x5_reduce', p + '5x5_reduce')
        o += _conv(p + '5x5', p + '5x5_reduce', p + '5x5', c5, 5, pad=2)
        o += _relu(p + 'relu_5x5', p + '5x5')
        o += _pool(p + 'pool', prev, 3, 1, pad=1)
        o += _conv(p + 'pool_proj', p + 'pool', p + 'pool_proj', cp, 1)
        o += _relu(p + 'relu_pool_proj', p + 'pool_proj')
        o += _simple(p + 'output', 'Concat',
                     [p + '1x1', p + '3x3', p + '5x5', p + 'pool_proj'], p + 'output')
        prev = p + 'output'
    o += _simple('pool5/drop_s1', 'Dropout', [prev], 'pool5/drop_s1',
                 ['dropout_param {', '  dropout_ratio: 0.4', '}'])
    o += _conv('cvg/classifier', 'pool5/drop_s1', 'cvg/classifier', num_cvg, 1, bias=0.0)
    o += _simple('coverage/sig', 'Sigmoid', ['cvg/classifier'], 'coverage')
    o += _conv('bbox/regressor', 'pool5/drop_s1', 'bboxes', num_bbox, 1, bias=0.0)
    return '\n'.join(o) + '\n'


def write_deploy_prototxt(path, **kw):
    with open(path, 'w') as f:
        f.write(deploy_prototxt(**kw))
    return path


def conv_shapes(num_cvg=4, num_bbox=16):
    """[(layer name, Cout, Cin, k, bias filler value)] in definition order."""
    out = [('conv1/7x7_s2', 64, 3, 7, 0.2), ('conv2/3x3_reduce', 64, 64, 1, 0.2),
           ('conv2/3x3', 192, 64, 3, 0.2)]
    cin = 192
    for spec in _INCEPTION:
        if spec == 'pool3':
            continue
        tag, c1, c3r, c3, c5r, c5, cp = spec
        p = 'inception_%s/' % tag
        out += [(p + '1x1', c1, cin, 1, 0.2), (p + '3x3_reduce', c3r, cin, 1, 0.2),
                (p + '3x3', c3, c3r, 3, 0.2), (p + '5x5_reduce', c5r, cin, 1, 0.2),
                (p + '5x5', c5, c5r, 5, 0.2), (p + 'pool_proj', cp, cin, 1, 0.2)]
        cin = c1 + c3 + c5 + cp
    out += [('cvg/classifier', num_cvg, cin, 1, 0.0), ('bbox/regressor', num_bbox, cin, 1, 0.0)]
    return out
