"""Predict-on-image entry point of the detector node, on the B200 path.

Mirrors `FCNObjectDetector` of /root/reference/scripts/fcn_object_detector.py:25-461
without ROS: the rosparams of :33-38 are constructor arguments with the same
names and defaults, `run_detector(cv_img)` is the body of :54-132 (everything
after :132 is drawing / HighGUI and is dropped), `callback` forwards to it as
in :307-308.  All arithmetic runs in libfcnod_b200 (prelude, network, decode,
groupRectangles clustering, rescale); this class only marshals buffers.
"""
import ctypes as C
import logging
import os

import numpy as np

from . import _lib
from . import caffe_shim as caffe

log = logging.getLogger('fcn_object_detector')


def _unpack(det, lab, wgt, cnt, weights=False, overflow=None):
    """Per-frame arrays out of the fixed-size C-ABI buffers.  `cnt[b]` is the number of detections the reference would
    return; the library stores the first FCNOD_MAX_DET of them.  A frame with more is reported, never silently cut:
    a warning is logged and (frame index, total) is appended to `overflow`."""
    M = _lib.FCNOD_MAX_DET
    out = []
    for b in range(len(cnt)):
        total = int(cnt[b])
        n = min(total, M)
        if total > M:
            log.warning('frame %d: %d detections, only the first %d are returned (FCNOD_MAX_DET)', b, total, M)
            if overflow is not None:
                overflow.append((b, total))
        row = (det[b, :n].astype(np.int64), lab[b, :n].astype(np.int64))
        out.append(row + (wgt[b, :n].astype(np.int64),) if weights else row)
    return out


class FCNObjectDetector(object):
    def __init__(self, pretrained_weights=None, deployment_prototxt=None, detection_threshold=0.5, min_boxes=3,
                 nms_eps=0.2, device_id=0, manifest=None, stride=16, precision='bf16'):
        self.__net = None
        self.__im_width = None
        self.__im_height = None
        self.__prob_thresh = detection_threshold
        self.__min_bbox_thresh = min_boxes
        self.__group_eps_thresh = nms_eps
        self.__weights = pretrained_weights
        self.__model_proto = deployment_prototxt
        self.__device_id = device_id
        self.__stride = stride            # upstream literal is 16/2 (:360); the deploy net is stride 16 (A2)
        self.__precision = precision
        self.overflow = []                # (frame index, untruncated count) of the last call's frames that exceeded FCNOD_MAX_DET
        if self.is_file_valid():
            self.load_caffe_model()
            log.info('DETECTOR SETUP SUCCESSFUL')
        self.__label_manifest = manifest
        self.load_label_manifest()

    # ------------------------------------------------------------------ :315-328
    def load_caffe_model(self):
        log.info('LOADING CAFFE MODEL..')
        caffe.set_device(self.__device_id)
        caffe.set_mode_gpu()
        self.__net = caffe.Net(self.__model_proto, self.__weights, caffe.TEST, precision=self.__precision)
        shape = self.__net.blobs['data'].data.shape
        self.__im_height = shape[2]
        self.__im_width = shape[3]
        self.__net.blobs['data'].reshape(1, 3, self.__im_height, self.__im_width)

    @property
    def net(self):
        return self.__net

    def _params(self):
        return _lib.default_postnet_params(prob_thresh=float(self.__prob_thresh), min_boxes=int(self.__min_bbox_thresh),
                                           eps=float(self.__group_eps_thresh), net_w=int(self.__im_width),
                                           net_h=int(self.__im_height), stride=int(self.__stride))

    # ------------------------------------------------------------------ :54-132
    def run_detector(self, cv_img):
        """cv_img: HxWx3 uint8 BGR (what CvBridge.imgmsg_to_cv2(msg, 'bgr8') returns, :57).
        Returns (object_boxes int (K,5) [x1,y1,x2,y2,int(conf)] in source pixels, object_labels int (K,))
        or None when nothing is detected (the reference logs and returns, :128-130)."""
        if cv_img is None:
            return None
        boxes, labels = self.predict_batch(np.asarray(cv_img)[None])[0]
        if not len(boxes):
            log.warning('not detection')
            return None
        return boxes, labels

    def predict(self, cv_img):
        """Like run_detector but always returns (boxes, labels) (possibly empty)."""
        return self.predict_batch(np.asarray(cv_img)[None])[0]

    def predict_batch(self, images):
        """images: (B,H,W,3) uint8 BGR.  One fcnod_detect call: H2D, prelude, net, post-net, D2H."""
        images = np.ascontiguousarray(images, dtype=np.uint8)
        if images.ndim != 4 or images.shape[3] != 3:
            raise ValueError('images must be (B,H,W,3) uint8 BGR')
        B, H, W, _ = images.shape
        M = _lib.FCNOD_MAX_DET
        det = np.zeros((B, M, 5), np.int32)
        lab = np.zeros((B, M), np.int32)
        wgt = np.zeros((B, M), np.int32)
        cnt = np.zeros((B,), np.int32)
        p = self._params()
        _lib.check(_lib.lib().fcnod_detect(self.__net.handle, _lib.ptr(images), B, H, W, C.byref(p), _lib.ptr(det),
                                           _lib.ptr(lab), _lib.ptr(wgt), _lib.ptr(cnt)))
        self.__net._sync_after_device_run(B)
        self.overflow = []
        return _unpack(det, lab, wgt, cnt, overflow=self.overflow)

    def predict_stream(self, batches):
        """Generator over an iterable of (B,H,W,3) uint8 BGR batches (the node's image stream, :134-147, batched).
        Yields predict_batch's result for every batch, in order, keeping two batches in flight
        (fcnod_detect_submit / fcnod_detect_collect): the upload of batch i+1 overlaps the compute of batch i.
        A batch's array must not be overwritten before its result is yielded; page-locked arrays overlap."""
        L = _lib.lib()
        M = _lib.FCNOD_MAX_DET
        p = self._params()
        pending = []

        def collect():
            ticket, images, B = pending.pop(0)
            det = np.empty((B, M, 5), np.int32)
            lab = np.empty((B, M), np.int32)
            wgt = np.empty((B, M), np.int32)
            cnt = np.empty((B,), np.int32)
            _lib.check(L.fcnod_detect_collect(self.__net.handle, ticket, _lib.ptr(det), _lib.ptr(lab), _lib.ptr(wgt),
                                              _lib.ptr(cnt)))
            self.overflow = []
            return _unpack(det, lab, wgt, cnt, overflow=self.overflow)

        try:
            for images in batches:
                images = np.ascontiguousarray(images, dtype=np.uint8)
                if images.ndim != 4 or images.shape[3] != 3:
                    raise ValueError('images must be (B,H,W,3) uint8 BGR')
                B, H, W, _ = images.shape
                if pending and pending[0][2] != B:           # the batch size changes only with nothing in flight
                    while pending:
                        yield collect()
                ticket = C.c_int32(-1)
                _lib.check(L.fcnod_detect_submit(self.__net.handle, _lib.ptr(images), B, H, W, C.byref(p),
                                                 C.byref(ticket)))
                pending.append((ticket.value, images, B))
                if len(pending) == 2:
                    yield collect()
            while pending:
                yield collect()
        finally:
            # an abandoned generator must not leave tickets in flight
            while pending:
                try:
                    collect()
                except Exception:
                    pending.clear()
            self.__net._sync_after_device_run()

    def predict_data(self, data):
        """data: (B,3,H,W) float32 blob contents (post-prelude).  Returns per-frame (boxes, labels, weights)."""
        data = np.ascontiguousarray(data, dtype=np.float32)
        B = data.shape[0]
        M = _lib.FCNOD_MAX_DET
        det = np.zeros((B, M, 5), np.int32)
        lab = np.zeros((B, M), np.int32)
        wgt = np.zeros((B, M), np.int32)
        cnt = np.zeros((B,), np.int32)
        p = self._params()
        _lib.check(_lib.lib().fcnod_detect_from_data(self.__net.handle, _lib.ptr(data), B, C.byref(p), _lib.ptr(det),
                                                     _lib.ptr(lab), _lib.ptr(wgt), _lib.ptr(cnt)))
        self.__net._sync_after_device_run(B)
        self.overflow = []
        return _unpack(det, lab, wgt, cnt, weights=True, overflow=self.overflow)

    def callback(self, cv_img):
        return self.run_detector(cv_img)

    # ------------------------------------------------------------------ :428-461
    def is_file_valid(self):
        if self.__model_proto is None or self.__weights is None:
            log.critical('PROVIDE PRETRAINED MODEL! KILLING NODE...')
            return False
        if not os.path.isfile(str(self.__model_proto)) or not os.path.isfile(str(self.__weights)):
            log.critical('NOT SUCH FILES')
            return False
        return True

    def load_label_manifest(self):
        self.labels = []
        if self.__label_manifest is None or not os.path.isfile(self.__label_manifest):
            log.warning('Object Label Names not Provided!')
            if self.__net is not None:
                out_size = self.__net.blobs['score'].data.shape[1]
                self.labels = ['object_' + str(i - 1) for i in range(out_size)]
            return
        for line in open(self.__label_manifest):
            idx, _, oname = line.rstrip('\n').split(' ')
            self.labels.append(str(oname))


def postnet(coverage, bboxes, src_hw=None, overflow=None, **params):
    """Post-net stage alone on host arrays: coverage (B,Cc,G,G), bboxes (B,Cb,G,G) float32.
    Returns per-frame (boxes (K,5) int, labels (K,), weights (K,)).  Frames with more than FCNOD_MAX_DET detections
    are cut to the first FCNOD_MAX_DET, logged, and appended to `overflow` (a list) as (frame, total)."""
    coverage = np.ascontiguousarray(coverage, dtype=np.float32)
    bboxes = np.ascontiguousarray(bboxes, dtype=np.float32)
    B, Cc, gh, gw = coverage.shape
    Cb = bboxes.shape[1]
    M = _lib.FCNOD_MAX_DET
    det = np.zeros((B, M, 5), np.int32); lab = np.zeros((B, M), np.int32)
    wgt = np.zeros((B, M), np.int32); cnt = np.zeros((B,), np.int32)
    hw = None if src_hw is None else np.ascontiguousarray(src_hw, dtype=np.int32).reshape(B, 2)
    p = _lib.default_postnet_params(**params)
    _lib.check(_lib.lib().fcnod_postnet(_lib.ptr(coverage), _lib.ptr(bboxes), B, Cc, Cb, gh, gw, _lib.ptr(hw), C.byref(p),
                                        _lib.ptr(det), _lib.ptr(lab), _lib.ptr(wgt), _lib.ptr(cnt)))
    return _unpack(det, lab, wgt, cnt, weights=True, overflow=overflow)


class MultiGpuDetector(object):
    """One process, several GPUs (the reference node is ONE process: scripts/fcn_object_detector.py:307-308).
    A replica of the detector (own handle, own streams, own CUDA graph, replicated weights) lives on every device of
    `device_ids`; `predict_batch` splits the frames contiguously over the replicas (sharding.shard_range), submits every
    shard with the asynchronous fcnod_detect_submit -- uploads and compute of all devices run at once -- and then
    collects them in order.  Frames are independent: no data-path collective, results are concatenated on the host."""

    def __init__(self, pretrained_weights, deployment_prototxt, device_ids, **kw):
        if not device_ids:
            raise ValueError('device_ids must name at least one CUDA device')
        self.device_ids = list(device_ids)
        self.replicas = [FCNObjectDetector(pretrained_weights=pretrained_weights, deployment_prototxt=deployment_prototxt,
                                           device_id=d, **kw) for d in self.device_ids]
        self.overflow = []

    def predict_batch(self, images):
        from .sharding import shard_range
        images = np.ascontiguousarray(images, dtype=np.uint8)
        if images.ndim != 4 or images.shape[3] != 3:
            raise ValueError('images must be (B,H,W,3) uint8 BGR')
        B, H, W, _ = images.shape
        L = _lib.lib()
        M = _lib.FCNOD_MAX_DET
        world = len(self.replicas)
        pending = []
        for r, det in enumerate(self.replicas):
            lo, hi = shard_range(B, r, world)
            if hi == lo:
                continue
            p = det._params()
            ticket = C.c_int32(-1)
            shard = images[lo:hi]
            _lib.check(L.fcnod_detect_submit(det.net.handle, _lib.ptr(shard), hi - lo, H, W, C.byref(p), C.byref(ticket)))
            pending.append((det, ticket.value, lo, hi, shard))
        out, self.overflow = [], []
        for det, ticket, lo, hi, _shard in pending:
            n = hi - lo
            d = np.empty((n, M, 5), np.int32); lab = np.empty((n, M), np.int32)
            wgt = np.empty((n, M), np.int32); cnt = np.empty((n,), np.int32)
            _lib.check(L.fcnod_detect_collect(det.net.handle, ticket, _lib.ptr(d), _lib.ptr(lab), _lib.ptr(wgt), _lib.ptr(cnt)))
            ov = []
            out.extend(_unpack(d, lab, wgt, cnt, overflow=ov))
            self.overflow.extend((lo + b, t) for b, t in ov)
            det.net._sync_after_device_run()
        return out

    def run_detector(self, cv_img):
        return self.replicas[0].run_detector(cv_img)

    def close(self):
        for det in self.replicas:
            if det.net is not None:
                det.net.close()


def gridbox_to_boxes(net_cvg, net_boxes, prob_thresh=0.5, im_w=448, im_h=448, stride=16):
    """`gridbox_to_boxes` of scripts/fcn_object_detector.py:357-394 on the device (fcnod_decode), one class map:
    net_cvg (G,G) float32, net_boxes (4,G,G) float32.  Returns (boxes (n,4) float64 [x1,y1,x2,y2],
    cvgs (n,3) [x, y, score], mask (rows_used, G) bool) exactly as the reference does."""
    net_cvg = np.ascontiguousarray(net_cvg, dtype=np.float32)
    net_boxes = np.ascontiguousarray(net_boxes, dtype=np.float32)
    gh, gw = net_cvg.shape
    cap = gh * gw
    boxes = np.zeros((cap, 4), np.float64); cell = np.zeros((cap,), np.int32)
    score = np.zeros((cap,), np.float32); cnt = np.zeros((1,), np.int32)
    p = _lib.default_postnet_params(prob_thresh=float(prob_thresh), net_w=int(im_w), net_h=int(im_h), stride=int(stride),
                                    skip_background=0)
    _lib.check(_lib.lib().fcnod_decode(_lib.ptr(net_cvg), _lib.ptr(net_boxes), 1, 1, 4, gh, gw, C.byref(p), cap,
                                       _lib.ptr(boxes), _lib.ptr(cell), _lib.ptr(score), _lib.ptr(cnt)))
    n = int(cnt[0])
    rows_used = min(gh, int(im_h // stride), int(im_w // stride))
    mask = np.zeros((rows_used, gw), bool)
    y, x = np.divmod(cell[:n], gw)
    mask[y, x] = True
    cvgs = np.transpose(np.vstack((x, y, score[:n].astype(np.float64)))) if n else np.zeros((0, 3))
    return boxes[:n].copy(), cvgs, mask


def decode_batch(coverage, bboxes, cap=None, **params):
    """fcnod_decode on a batch: coverage (B,Cc,G,G), bboxes (B,Cb,G,G).  Returns (boxes (B,ncls,cap,4) float64,
    cell (B,ncls,cap) int32, score (B,ncls,cap) float32, count (B,ncls) int32)."""
    coverage = np.ascontiguousarray(coverage, dtype=np.float32)
    bboxes = np.ascontiguousarray(bboxes, dtype=np.float32)
    B, Cc, gh, gw = coverage.shape
    p = _lib.default_postnet_params(**params)
    ncls = Cc - (1 if p.skip_background else 0)
    cap = gh * gw if cap is None else int(cap)
    boxes = np.zeros((B, ncls, cap, 4), np.float64); cell = np.zeros((B, ncls, cap), np.int32)
    score = np.zeros((B, ncls, cap), np.float32); cnt = np.zeros((B, ncls), np.int32)
    _lib.check(_lib.lib().fcnod_decode(_lib.ptr(coverage), _lib.ptr(bboxes), B, Cc, bboxes.shape[1], gh, gw, C.byref(p), cap,
                                       _lib.ptr(boxes), _lib.ptr(cell), _lib.ptr(score), _lib.ptr(cnt)))
    return boxes, cell, score, cnt


def preprocess(images, net_h=448, net_w=448):
    """Prelude alone: (B,H,W,3) uint8 BGR -> (B,3,net_h,net_w) float32."""
    images = np.ascontiguousarray(images, dtype=np.uint8)
    B, H, W, _ = images.shape
    out = np.zeros((B, 3, net_h, net_w), np.float32)
    _lib.check(_lib.lib().fcnod_preprocess(_lib.ptr(images), B, H, W, net_h, net_w, _lib.ptr(out)))
    return out
