"""CPU restatement of the training head of /root/reference/models/train_val.prototxt -- TEST INFRASTRUCTURE ONLY (oracle).

PARITY UNPINNED (no upstream vectors; Caffe / NVCaffe cannot run here).  Layers restated, with their prototxt lines:
  slice-label   Slice, slice_dim 1, points 1/5/9/13 (num_classes = 1)      :36-52
  bb-label-norm, bb-obj-norm          Eltwise PROD                          :53-72
  bbox_mask, bbox-norm, bbox-obj-norm Eltwise PROD                          :2237-2266
  bbox_loss     L1Loss, loss_weight 2.0                                     :2267-2274
  coverage_loss EuclideanLoss                                               :2275-2281
Loss semantics [ext: Caffe / NVCaffe layer sources, not vendored upstream]:
  EuclideanLoss  = sum((a - b)^2) / (2 * num)            backward: (a - b) / num
  L1Loss (NVCaffe, DIGITS DetectNet) = sum(|a - b|) / num   backward: sign(a - b) / num
Eltwise PROD backward: bottom_diff_i = top_diff * prod of the other bottoms.
"""
import numpy as np


def slice_label(label, num_classes=1):
    """:36-52 -- label (N, 17C, G, G) -> the five grids."""
    c = num_classes
    pts = [0, c, 5 * c, 9 * c, 13 * c, 17 * c]
    return [label[:, pts[i]:pts[i + 1]] for i in range(5)]


def head_forward(coverage, bboxes, cvg_label, bbox_label, size_block, obj_block, cvg_block, bbox_loss_weight=2.0):
    f = np.float32
    coverage, bboxes = coverage.astype(f), bboxes.astype(f)
    n = coverage.shape[0]
    target = (bbox_label.astype(f) * size_block.astype(f)) * obj_block.astype(f)                  # :53-72
    pred = ((bboxes * cvg_block.astype(f)) * size_block.astype(f)) * obj_block.astype(f)          # :2237-2266
    diff = pred - target
    loss_bbox = np.abs(diff.astype(np.float64)).sum() / n                                         # :2267-2274
    cdiff = coverage - cvg_label.astype(f)
    loss_cvg = ((cdiff * cdiff).astype(np.float64)).sum() / (2.0 * n)                              # :2275-2281
    inv_n = f(1.0) / f(n)
    g = (f(bbox_loss_weight) * inv_n) * np.sign(diff).astype(f)
    d_bboxes = ((g * obj_block.astype(f)) * size_block.astype(f)) * cvg_block.astype(f)
    d_coverage = cdiff * inv_n
    return dict(loss_bbox=loss_bbox, loss_coverage=loss_cvg, loss=loss_cvg + bbox_loss_weight * loss_bbox,
                d_coverage=d_coverage, d_bboxes=d_bboxes)
