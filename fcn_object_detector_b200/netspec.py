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


# the one convolution whose bias learns at lr_mult 1.0 upstream (models/train_val.prototxt, models/deploy.prototxt): kept
_BIAS_LR_MULT = {'inception_5b/3x3_reduce': 1.0}


def _conv(name, bottom, top, num_output, kernel, pad=0, stride=1, bias=0.2):
    s = ['layer {', '  name: "%s"' % name, '  type: "Convolution"',
         '  bottom: "%s"' % bottom, '  top: "%s"' % top,
         '  param { lr_mult: 1.0 decay_mult: 1.0 }',
         '  param { lr_mult: %s decay_mult: 0.0 }' % repr(_BIAS_LR_MULT.get(name, 2.0)),
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


def _body(num_cvg, num_bbox):
    """The layers between `deploy_transform` and the two heads (shared by the deploy and the train_val forms)."""
    text = deploy_prototxt(num_cvg=num_cvg, num_bbox=num_bbox)
    return text[text.index('layer {'):]


def train_val_prototxt(num_classes=1, bbox_loss_weight=2.0, lmdb_features='lmdb/features', lmdb_labels='lmdb/labels'):
    """Text of the training network in the form of /root/reference/models/train_val.prototxt:1-2281: two LMDB Data layers
    (:1-26), the input Power (:27-35), the label Slice (:36-52), the two label normalisers (:53-72), the network body with
    num_classes / 4 num_classes head channels, the three Eltwise masks (:2237-2266) and the two losses (:2267-2281)."""
    c = num_classes
    o = []
    for name, top, src in (('train_data', 'data', lmdb_features), ('train_label', 'label', lmdb_labels)):
        o += ['layer {', '  name: "%s"' % name, '  type: "Data"', '  top: "%s"' % top, '  include { phase: TRAIN }',
              '  data_param {', '    source: "%s"' % src, '    batch_size: 1', '    backend: LMDB', '  }', '}']
    body = _body(c, 4 * c).rstrip('\n').split('\n')
    # the Power layer is the first layer record of the body: the Slice and the normalisers follow it as in the upstream file
    end_power = body.index('}') + 1
    o += body[:end_power]
    o += ['layer {', '  name: "slice-label"', '  type: "Slice"', '  bottom: "label"', '  top: "coverage-label"',
          '  top: "bbox-label"', '  top: "size-block"', '  top: "obj-block"', '  top: "coverage-block"', '  slice_param {',
          '    slice_dim: 1', '    slice_point: %d' % c, '    slice_point: %d' % (5 * c), '    slice_point: %d' % (9 * c),
          '    slice_point: %d' % (13 * c), '  }', '}']
    prod = ['eltwise_param {', '  operation: PROD', '}']
    o += _simple('bb-label-norm', 'Eltwise', ['bbox-label', 'size-block'], 'bbox-label-norm', prod)
    o += _simple('bb-obj-norm', 'Eltwise', ['bbox-label-norm', 'obj-block'], 'bbox-obj-label-norm', prod)
    o += body[end_power:]
    o += _simple('bbox_mask', 'Eltwise', ['bboxes', 'coverage-block'], 'bboxes-masked', prod)
    o += _simple('bbox-norm', 'Eltwise', ['bboxes-masked', 'size-block'], 'bboxes-masked-norm', prod)
    o += _simple('bbox-obj-norm', 'Eltwise', ['bboxes-masked-norm', 'obj-block'], 'bboxes-obj-masked-norm', prod)
    o += _simple('bbox_loss', 'L1Loss', ['bboxes-obj-masked-norm', 'bbox-obj-label-norm'], 'loss_bbox',
                 ['loss_weight: %s' % repr(float(bbox_loss_weight))])
    o += _simple('coverage_loss', 'EuclideanLoss', ['coverage', 'coverage-label'], 'loss_coverage')
    return '\n'.join(o) + '\n'


def write_train_val_prototxt(path, **kw):
    with open(path, 'w') as f:
        f.write(train_val_prototxt(**kw))
    return path


def solver_prototxt():
    """Text with the values of /root/reference/train/bounding_box/solver.prototxt:3-15."""
    return ('net: "bounding_box/train_val.prototxt"\nsnapshot_prefix: "snapshots/snapshot"\nbase_lr: 1e-04\nmax_iter: 1\n'
            'iter_size: 1\nlr_policy: "step"\ngamma: 0.1\nmomentum: 0.9\nweight_decay: 1e-07\nstepsize: 10000\nsnapshot: 5000\n'
            'solver_type: ADAM\nsolver_mode: GPU\ndisplay: 20\ndebug_info: false\n')


def conv_shape_dict(num_cvg=4, num_bbox=16):
    return {name: (cout, cin, k, k) for name, cout, cin, k, _ in conv_shapes(num_cvg, num_bbox)}


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
