"""CPU restatement of the layer's photometric augmentation -- TEST INFRASTRUCTURE ONLY (oracle).

/root/reference/scripts/data_argumentation_layer/argumentation_engine.py:308-322 builds an imgaug Sequential; imgaug is
vendored upstream (scripts/data_argumentation_layer/imgaug/imgaug/augmenters.py) but not importable here (skimage is
absent), so each augmenter's `_augment_images` body is restated with the SAME third-party calls it makes:
  GaussianBlur   :2829-2842  scipy.ndimage.gaussian_filter per channel (uint8 in / out)
  AverageBlur    :2932-2946  cv2.blur(img, (k, k))
  MedianBlur     :3013-3023  cv2.medianBlur(img, k)
  Sharpen        :3196-3217 + Convolve :3100-3126  cv2.filter2D(channel, -1, matrix), np.clip
  Add            :3563-3587  int32 add, np.clip, astype(uint8)
  Multiply       :3853-3876  float32 multiply, np.clip, astype(uint8)
  Grayscale      :2680-2726  cv2.cvtColor(image, COLOR_RGB2GRAY), (alpha*gray + (1-alpha)*image).astype(uint8)
scipy and cv2 are the versions of this container (the reference pins neither): PARITY UNPINNED at the upstream boundary.
"""
import cv2
import numpy as np
from scipy import ndimage


def sharpen_matrix(alpha, lightness):
    f = np.float32
    nochange = np.array([[0, 0, 0], [0, 1, 0], [0, 0, 0]], f)
    effect = np.array([[-1, -1, -1], [-1, 8 + lightness, -1], [-1, -1, -1]], f)
    return (f(1 - alpha) * nochange + f(alpha) * effect).astype(f)       # float32 under the reference's NumPy


def apply(image, d):
    """image: (h, w, 3) uint8; d: the draw dict of fcn_object_detector_b200.photometric.draw_photo."""
    img = np.ascontiguousarray(image).copy()
    if d['blur_kind'] == 1:
        for c in range(3):
            img[:, :, c] = ndimage.gaussian_filter(img[:, :, c], d['sigma'])
    elif d['blur_kind'] == 2:
        img = cv2.blur(img, (d['blur_k'], d['blur_k']))
    elif d['blur_kind'] == 3:
        img = cv2.medianBlur(img, d['blur_k'])
    m = sharpen_matrix(d['sharpen_alpha'], d['sharpen_lightness'])
    img = np.ascontiguousarray(img)
    for c in range(3):
        img[..., c] = cv2.filter2D(np.ascontiguousarray(img[..., c]), -1, m)
    img = np.clip(img, 0, 255).astype(np.uint8)
    a = img.astype(np.int32)
    for c in range(3):
        a[..., c] += d['add'][c]
    img = np.clip(a, 0, 255).astype(np.uint8)
    f = img.astype(np.float32)
    for c in range(3):
        f[..., c] *= np.float32(d['mul'][c])
    img = np.clip(f, 0, 255).astype(np.uint8)
    alpha = d['gray_alpha']
    if alpha > 0.001:
        g = cv2.cvtColor(img, cv2.COLOR_RGB2GRAY)
        g = np.tile(g[:, :, None], (1, 1, 3))
        img = (alpha * g + (1 - alpha) * img).astype(np.uint8)
    return img
