"""Photometric augmentation of the training-side layer on the B200 path: `color_space_argumentation`
(/root/reference/scripts/data_argumentation_layer/argumentation_engine.py:308-322), the imgaug chain
OneOf(GaussianBlur | AverageBlur | MedianBlur) -> Sharpen -> Add -> Multiply -> Grayscale on uint8 images.

The random draws are made here, on the host, from a `numpy.random.RandomState` with the distributions imgaug's
parameters have (Uniform / DiscreteUniform / Binomial; augmenters.py:2779-3876); the pixels are processed in
libfcnod_b200 (fcnod_photo_augment: three kernels).  imgaug's own random stream is not reproduced (draws are
not a parity target) -- given the same draws the device result is bit-identical to what imgaug's scipy / cv2 calls give.
"""
import ctypes as C

import numpy as np

from . import _lib


def draw_photo(rs):
    """One image's draws as a plain dict (what oracle.photo_ref.apply and make_draw consume)."""
    d = dict(blur_kind=0, blur_k=0, sigma=0.0)
    which = int(rs.randint(0, 3))                                   # OneOf: one child, uniformly
    if which == 0:
        sigma = float(rs.uniform(0.0, 3.0))                         # GaussianBlur((0, 3.0)); eps = 0.001
        if sigma > 0.001:
            d.update(blur_kind=1, sigma=sigma)
    elif which == 1:
        k = int(rs.randint(2, 8))                                   # AverageBlur(k=(2, 7)): DiscreteUniform, same k both axes
        d.update(blur_kind=2, blur_k=k)
    else:
        k = int(rs.randint(3, 8))                                   # MedianBlur(k=(3, 7)): even draws are raised by one
        d.update(blur_kind=3, blur_k=k + 1 if k % 2 == 0 else k)
    d['sharpen_alpha'] = float(rs.uniform(0.0, 1.0))                # Sharpen(alpha=(0, 1.0), lightness=(0.75, 1.5))
    d['sharpen_lightness'] = float(rs.uniform(0.75, 1.5))
    if rs.binomial(1, 0.5):                                         # Add((-2, 21), per_channel=0.5)
        d['add'] = [int(v) for v in rs.randint(-2, 22, 3)]
    else:
        d['add'] = [int(rs.randint(-2, 22))] * 3
    if rs.binomial(1, 0.5):                                         # Multiply((0.75, 1.25), per_channel=0.5)
        d['mul'] = [float(v) for v in rs.uniform(0.75, 1.25, 3)]
    else:
        d['mul'] = [float(rs.uniform(0.75, 1.25))] * 3
    d['gray_alpha'] = float(rs.uniform(0.0, 0.5))                   # Grayscale(alpha=(0.0, 0.50))
    return d


IDENTITY = dict(blur_kind=0, blur_k=0, sigma=0.0, sharpen_alpha=0.0, sharpen_lightness=1.0, add=[0, 0, 0], mul=[1.0, 1.0, 1.0],
                gray_alpha=0.0)


def gaussian_weights(sigma, truncate=4.0):
    """scipy.ndimage's 1-d kernel: radius int(truncate * sigma + 0.5), exp(-x^2 / (2 sigma^2)) normalised by its sum.
    Returns (radius, weights[0..radius]) with weights[0] the centre tap."""
    radius = int(truncate * float(sigma) + 0.5)
    x = np.arange(-radius, radius + 1)
    phi = np.exp(-0.5 / (sigma * sigma) * x ** 2)
    phi = phi / phi.sum()
    return radius, phi[radius:]


def sharpen_matrix(alpha, lightness):
    """Sharpen's filter2D matrix (augmenters.py:3196-3215), float32 as under the NumPy the reference ran on."""
    f = np.float32
    nochange = np.array([[0, 0, 0], [0, 1, 0], [0, 0, 0]], f)
    effect = np.array([[-1, -1, -1], [-1, 8 + lightness, -1], [-1, -1, -1]], f)
    return (f(1 - alpha) * nochange + f(alpha) * effect).astype(f)


def make_draw(d):
    """dict -> the C struct."""
    s = _lib.PhotoDraw()
    s.blur_kind = int(d['blur_kind']); s.blur_k = int(d.get('blur_k', 0))
    if s.blur_kind == 1:
        r, w = gaussian_weights(d['sigma'])
        if r > 12:
            raise ValueError('gaussian sigma too large (radius %d > 12)' % r)
        s.gauss_radius = r
        for i, v in enumerate(w):
            s.gauss_w[i] = float(v)
    if s.blur_kind == 2 and not 2 <= s.blur_k <= 16:
        raise ValueError('AverageBlur k out of range')
    if s.blur_kind == 3 and s.blur_k not in (3, 5, 7):
        raise ValueError('MedianBlur k must be 3, 5 or 7')
    m = sharpen_matrix(d['sharpen_alpha'], d['sharpen_lightness'])
    for i, v in enumerate(m.reshape(-1)):
        s.sharpen[i] = float(v)
    for c in range(3):
        s.add[c] = int(d['add'][c])
        s.mul[c] = float(np.float32(d['mul'][c]))
    s.gray_alpha = float(d['gray_alpha'])
    return s


def photo_augment(images, draws, sizes=None):
    """images (N,H,W,3) uint8, draws: list of dicts (draw_photo), sizes: optional (N,2) valid (h, w) per image (rows packed
    at the image's own width inside the H x W pitch).  Returns the augmented batch, same layout."""
    images = np.ascontiguousarray(images, dtype=np.uint8)
    if images.ndim != 4 or images.shape[3] != 3:
        raise ValueError('images must be (N,H,W,3) uint8')
    N, H, W, _ = images.shape
    darr = (_lib.PhotoDraw * N)(*[make_draw(d) for d in draws])
    sz = None if sizes is None else np.ascontiguousarray(sizes, dtype=np.int32).reshape(N, 2)
    out = np.zeros_like(images)
    _lib.check(_lib.lib().fcnod_photo_augment(_lib.ptr(images), _lib.ptr(sz), darr, N, H, W, _lib.ptr(out)))
    return out
