"""Generates the golden fixtures in this directory from the CPU oracle (oracle/).

Run from the repo root:  python tests/golden/make_golden.py
The upstream repository holds no golden vectors for this path, so
these are produced by the oracle restatement; `group_rectangles` is additionally
cross-checked here against the cv2 of this container before anything is written.
Inputs are regenerated from seeds by the tests, only expected outputs are stored.
"""
import os
import random
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))

from oracle import layer_ref, net_ref, postnet_ref, prototxt  # noqa: E402
from oracle.caffemodel_io import read_caffemodel  # noqa: E402
from fcn_object_detector_b200.caffemodel import write_synthetic_caffemodel  # noqa: E402
from fcn_object_detector_b200.netspec import deploy_prototxt  # noqa: E402


def postnet_inputs(seed=3, B=4):
    rng = np.random.RandomState(seed)
    cov, bb = [], []
    for b in range(B):
        c, x = postnet_ref.planted_maps(rng, n_obj=1 + b % 4)
        cov.append(c); bb.append(x)
    return np.stack(cov), np.stack(bb)


def pack_dets(dets, M=64):
    B = len(dets)
    out = np.zeros((B, M, 5), np.int32); lab = np.zeros((B, M), np.int32); cnt = np.zeros((B,), np.int32)
    for b, (ob, ol) in enumerate(dets):
        cnt[b] = len(ob)
        out[b, :len(ob)] = ob; lab[b, :len(ob)] = ol
    return out, lab, cnt


def layer_inputs(seed=1, N=8, Hs=120, Ws=160):
    rs = np.random.RandomState(seed)
    data = rs.randint(0, 256, (N, 3, Hs, Ws)).astype(np.float32)
    lab = np.zeros((N, 1, 1, 5), np.float32)
    rng = random.Random(1234)
    draws = []
    for i in range(N):
        w, h = int(rs.randint(8, 64)), int(rs.randint(8, 64))
        x, y = int(rs.randint(0, Ws - w)), int(rs.randint(0, Hs - h))
        lab[i, 0, 0] = [x, y, w, h, 0]
        draws.append(layer_ref.draw_params(rng, Ws, Hs, [x, y, w, h]))
    return data, lab, draws


def net_inputs(seed=2, B=1):
    return np.random.RandomState(seed).uniform(0, 1, (B, 3, 448, 448)).astype(np.float32)


def main():
    import cv2
    # --- pin group_rectangles against cv2 before trusting it
    rng = np.random.RandomState(0)
    for it in range(300):
        rects = []
        for c in range(rng.randint(0, 4)):
            x, y, w, h = rng.randint(0, 300), rng.randint(0, 300), rng.randint(20, 150), rng.randint(20, 150)
            for m in range(rng.randint(1, 10)):
                j = rng.randint(-8, 9, size=4)
                rects.append([int(x + j[0]), int(y + j[1]), int(max(1, w + j[2])), int(max(1, h + j[3]))])
        a = postnet_ref.group_rectangles(rects, 3, 0.2)
        b = cv2.groupRectangles([list(r) for r in rects], 3, 0.2)
        assert a[0] == [list(map(int, r)) for r in b[0]] and a[1] == [int(v) for v in np.ravel(b[1])]

    cov, bb = postnet_inputs()
    out = {}
    for stride in (16, 8):
        dets = [postnet_ref.postnet(cov[b], bb[b], (448, 448, 3), stride=stride) for b in range(cov.shape[0])]
        d, l, c = pack_dets(dets)
        out['det_s%d' % stride] = d; out['lab_s%d' % stride] = l; out['cnt_s%d' % stride] = c
    dets = [postnet_ref.postnet(cov[b], bb[b], (480, 640, 3)) for b in range(cov.shape[0])]
    out['det_vga'], out['lab_vga'], out['cnt_vga'] = pack_dets(dets)
    np.savez_compressed(os.path.join(HERE, 'postnet_planted.npz'), **out)

    data, lab, draws = layer_inputs()
    top, rects = layer_ref.layer_forward(data, lab, draws)
    np.savez_compressed(os.path.join(HERE, 'layer_small.npz'), fg=top[1], bbox=top[2], size=top[3], obj=top[4],
                        cvg=top[5], rects=np.asarray(rects, np.int32),
                        data_sum=np.asarray([top[0].astype(np.float64).sum()]),
                        data_probe=top[0][:, :, ::37, ::41].copy())

    cm = '/tmp/_golden_syn.caffemodel'
    write_synthetic_caffemodel(cm, seed=0)
    W = read_caffemodel(cm)
    Ls = prototxt.layers(prototxt.parse(deploy_prototxt()))
    o = net_ref.forward(Ls, W, net_inputs(), want=['coverage', 'bboxes', 'pool1/norm1', 'inception_3a/output'])
    np.savez_compressed(os.path.join(HERE, 'net_seed2.npz'), coverage=o['coverage'], bboxes=o['bboxes'],
                        norm1_probe=o['pool1/norm1'][:, ::8, ::16, ::16].copy(),
                        i3a_probe=o['inception_3a/output'][:, ::16, ::8, ::8].copy())
    print('golden fixtures written to', HERE)


if __name__ == '__main__':
    main()
