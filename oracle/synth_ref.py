"""CPU restatement of the copy-paste compositing loop -- TEST INFRASTRUCTURE ONLY (oracle).
/root/reference/scripts/data_argumentation_layer/argumentation_engine.py:723-728, statement for statement (the per-pixel
Python loop), including numpy's negative-index wrap when a proposal is wider than the background.  PARITY UNPINNED."""
import numpy as np


def paste(img_output, mask_output, im_roi, im_msk, cx, cy, label):
    """In place, like the reference.  im_msk: (h, w) = mask[..., 0]."""
    im_y, im_x = img_output.shape[:2]
    h, w = im_msk.shape
    for j in range(0, h, 1):
        for i in range(0, w, 1):
            nx, ny = i + cx, j + cy
            if im_msk[j, i] > 0 and nx < im_x and ny < im_y:
                img_output[ny, nx] = im_roi[j, i]
                mask_output[ny, nx] = label + 1


def composite(backgrounds, masks, patch_lists):
    imgs = np.array(backgrounds, np.uint8, copy=True); msk = np.array(masks, np.uint8, copy=True)
    for n, patches in enumerate(patch_lists):
        for roi, m, cx, cy, label in patches:
            paste(imgs[n], msk[n], roi, m, cx, cy, label)
    return imgs, msk
