"""Minimal Caffe prototxt (protobuf text format) reader -- TEST INFRASTRUCTURE ONLY.

Part of the CPU oracle (see oracle/README.md).  The product has its own C++
parser in fcn_object_detector_b200/csrc/prototxt.cpp; the two are deliberately
independent so that one checks the other.

Returns nested dicts: every field maps to a *list* of values (text-format
fields may repeat).  Follows the layout of /root/reference/models/deploy.prototxt:1-2178.
"""
import re

_TOKEN = re.compile(r'''
    \s*(?:
      (?P<comment>\#[^\n]*) |
      (?P<brace>[{}]) |
      (?P<colon>:) |
      (?P<string>"(?:[^"\\]|\\.)*"|'(?:[^'\\]|\\.)*') |
      (?P<atom>[^\s{}:#"']+)
    )''', re.X)


def _tokens(text):
    pos = 0
    n = len(text)
    while pos < n:
        m = _TOKEN.match(text, pos)
        if not m:
            if text[pos:].strip() == '':
                return
            raise ValueError('prototxt: bad token at offset %d' % pos)
        pos = m.end()
        if m.group('comment') is not None:
            continue
        for kind in ('brace', 'colon', 'string', 'atom'):
            v = m.group(kind)
            if v is not None:
                yield kind, v
                break


def _scalar(kind, v):
    if kind == 'string':
        return v[1:-1]
    try:
        return int(v)
    except ValueError:
        pass
    try:
        return float(v)
    except ValueError:
        pass
    if v == 'true':
        return True
    if v == 'false':
        return False
    return v  # enum identifier


def parse(text):
    toks = list(_tokens(text))
    i = 0

    def message(depth):
        nonlocal i
        out = {}
        while i < len(toks):
            kind, v = toks[i]
            if kind == 'brace' and v == '}':
                if depth == 0:
                    raise ValueError('prototxt: unbalanced }')
                i += 1
                return out
            if kind != 'atom':
                raise ValueError('prototxt: expected field name, got %r' % (v,))
            name = v
            i += 1
            kind, v = toks[i]
            if kind == 'colon':
                i += 1
                kind, v = toks[i]
            if kind == 'brace' and v == '{':
                i += 1
                out.setdefault(name, []).append(message(depth + 1))
            else:
                out.setdefault(name, []).append(_scalar(kind, v))
                i += 1
        if depth != 0:
            raise ValueError('prototxt: missing }')
        return out

    return message(0)


def _one(d, key, default=None):
    v = d.get(key)
    return v[0] if v else default


def layers(net):
    """Flatten a parsed NetParameter to a list of simple layer records."""
    out = []
    for L in net.get('layer', []):
        rec = {'name': _one(L, 'name'), 'type': _one(L, 'type'),
               'bottom': list(L.get('bottom', [])), 'top': list(L.get('top', []))}
        t = rec['type']
        if t == 'Convolution':
            p = _one(L, 'convolution_param', {})
            rec.update(num_output=_one(p, 'num_output'), kernel=_one(p, 'kernel_size'),
                       stride=_one(p, 'stride', 1), pad=_one(p, 'pad', 0),
                       bias_term=_one(p, 'bias_term', True),
                       bias_value=_one(_one(p, 'bias_filler', {}), 'value', 0.0),
                       weight_filler=_one(_one(p, 'weight_filler', {}), 'type', 'constant'))
            pr = L.get('param', [])
            rec.update(lr_mult=[_one(x, 'lr_mult', 1.0) for x in pr] + [1.0] * (2 - len(pr)),
                       decay_mult=[_one(x, 'decay_mult', 1.0) for x in pr] + [1.0] * (2 - len(pr)))
        elif t == 'Pooling':
            p = _one(L, 'pooling_param', {})
            rec.update(pool=_one(p, 'pool', 'MAX'), kernel=_one(p, 'kernel_size'),
                       stride=_one(p, 'stride', 1), pad=_one(p, 'pad', 0))
        elif t == 'LRN':
            p = _one(L, 'lrn_param', {})
            rec.update(local_size=_one(p, 'local_size', 5), alpha=_one(p, 'alpha', 1.0),
                       beta=_one(p, 'beta', 0.75), k=_one(p, 'k', 1.0))
        elif t == 'Power':
            p = _one(L, 'power_param', {})
            rec.update(power=_one(p, 'power', 1.0), scale=_one(p, 'scale', 1.0),
                       shift=_one(p, 'shift', 0.0))
        elif t == 'Dropout':
            p = _one(L, 'dropout_param', {})
            rec.update(dropout_ratio=_one(p, 'dropout_ratio', 0.5))
        out.append(rec)
    return out


def input_spec(net):
    name = _one(net, 'input')
    shp = _one(net, 'input_shape')
    if shp is not None:
        dims = list(shp.get('dim', []))
    else:
        dims = list(net.get('input_dim', []))
    return name, dims


def load(path):
    with open(path) as f:
        return parse(f.read())
