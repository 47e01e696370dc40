"""ctypes binding of libfcnod_b200.so (the C ABI declared in include/fcnod_b200.h).

There is no CPU fallback anywhere in this package: if the shared library is
missing, or a compute entry point fails (e.g. no CUDA device), an exception is
raised.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libfcnod_b200.so')

FCNOD_PRECISION_FP32 = 0
FCNOD_PRECISION_BF16 = 1
FCNOD_MAX_DET = 64
FCNOD_AUG_NOALLOC_MAX_N = 4096


class FcnodError(RuntimeError):
    pass


class PostnetParams(C.Structure):
    _fields_ = [('prob_thresh', C.c_float), ('min_boxes', C.c_int), ('eps', C.c_double),
                ('net_w', C.c_int), ('net_h', C.c_int), ('stride', C.c_int),
                ('skip_background', C.c_int), ('min_height', C.c_int)]


class Paste(C.Structure):
    _fields_ = [('roi_offset', C.c_int32), ('w', C.c_int32), ('h', C.c_int32), ('cx', C.c_int32), ('cy', C.c_int32),
                ('label', C.c_int32)]


class PhotoDraw(C.Structure):
    _fields_ = [('blur_kind', C.c_int32), ('blur_k', C.c_int32), ('gauss_radius', C.c_int32), ('_pad0', C.c_int32),
                ('gauss_w', C.c_double * 13), ('sharpen', C.c_float * 9), ('add', C.c_int32 * 3), ('mul', C.c_float * 3),
                ('_pad1', C.c_int32), ('gray_alpha', C.c_double)]


class SolverParams(C.Structure):
    _fields_ = [('base_lr', C.c_float), ('gamma', C.c_float), ('stepsize', C.c_int32), ('momentum', C.c_float),
                ('momentum2', C.c_float), ('delta', C.c_float), ('weight_decay', C.c_float), ('bbox_loss_weight', C.c_float),
                ('dropout_ratio_override', C.c_float), ('seed', C.c_uint64), ('precision', C.c_int32), ('reserved', C.c_int32)]


class AugDraw(C.Structure):
    _fields_ = [('flip', C.c_int32), ('shift_x', C.c_int32), ('shift_y', C.c_int32),
                ('sign_x', C.c_int32), ('sign_y', C.c_int32), ('_pad', C.c_int32),
                ('f1', C.c_double), ('f2', C.c_double)]


_P = C.c_void_p
_I = C.c_int
_SZ = C.c_size_t
_STR = C.c_char_p

# name -> (restype, argtypes); every symbol include/fcnod_b200.h declares
SIGNATURES = {
    'fcnod_version': (_I, []),
    'fcnod_last_error': (_STR, []),
    'fcnod_device_count': (_I, [C.POINTER(_I)]),
    'fcnod_host_alloc': (_I, [_SZ, C.POINTER(_P)]),
    'fcnod_host_free': (_I, [_P]),
    'fcnod_net_create': (_I, [_STR, _STR, _I, _I, C.POINTER(_P)]),
    'fcnod_net_destroy': (_I, [_P]),
    'fcnod_net_reshape': (_I, [_P, _I]),
    'fcnod_net_input_dims': (_I, [_P, C.POINTER(_I * 4)]),
    'fcnod_net_blob_dims': (_I, [_P, _STR, C.POINTER(_I * 4)]),
    'fcnod_net_num_blobs': (_I, [_P]),
    'fcnod_net_blob_name': (_STR, [_P, _I]),
    'fcnod_net_describe': (_I, [_P, _P, _SZ]),
    'fcnod_net_launches_per_forward': (_I, [_P]),
    'fcnod_net_forward': (_I, [_P, _P]),
    'fcnod_net_forward_device': (_I, [_P, _P, _P]),
    'fcnod_net_blob_read': (_I, [_P, _STR, _P, _SZ]),
    'fcnod_net_output_device': (_I, [_P, _STR, C.POINTER(_P)]),
    'fcnod_net_run_step': (_I, [_P, _I, _I, _P]),
    'fcnod_net_profile_conv_step': (_I, [_P, _I, _P, _SZ]),
    'fcnod_net_tc_candidates': (_I, [_P, _I]),
    'fcnod_net_tc_pick': (_I, [_P, _I]),
    'fcnod_net_set_tc_pick': (_I, [_P, _I, _I]),
    'fcnod_net_step_outputs': (_I, [_P, _I, _P, _SZ]),
    'fcnod_net_step_lane': (_I, [_P, _I, C.POINTER(_I), _P, _I, C.POINTER(_I), _P, _I, C.POINTER(_I)]),
    'fcnod_net_autotune_candidates': (_I, []),
    'fcnod_net_blob_checksum': (_I, [_P, _STR, C.POINTER(C.c_ulonglong)]),
    'fcnod_postnet_default_params': (None, [C.POINTER(PostnetParams)]),
    'fcnod_postnet_scratch_bytes': (_SZ, [_I, _I]),
    'fcnod_postnet_prepare': (_I, [C.POINTER(PostnetParams)]),
    'fcnod_postnet_device': (_I, [_P, _P, _I, _I, _I, _I, _I, _P, C.POINTER(PostnetParams), _P, _P, _P, _P, _P, _P]),
    'fcnod_decode_device': (_I, [_P, _P, _I, _I, _I, _I, _I, C.POINTER(PostnetParams), _I, _P, _P, _P, _P, _P]),
    'fcnod_decode': (_I, [_P, _P, _I, _I, _I, _I, _I, C.POINTER(PostnetParams), _I, _P, _P, _P, _P]),
    'fcnod_postnet': (_I, [_P, _P, _I, _I, _I, _I, _I, _P, C.POINTER(PostnetParams), _P, _P, _P, _P]),
    'fcnod_preprocess': (_I, [_P, _I, _I, _I, _I, _I, _P]),
    'fcnod_preprocess_device': (_I, [_P, _I, _I, _I, _I, _I, _P, _P, _P]),
    'fcnod_detect': (_I, [_P, _P, _I, _I, _I, C.POINTER(PostnetParams), _P, _P, _P, _P]),
    'fcnod_detect_submit': (_I, [_P, _P, _I, _I, _I, C.POINTER(PostnetParams), C.POINTER(C.c_int32)]),
    'fcnod_detect_collect': (_I, [_P, C.c_int32, _P, _P, _P, _P]),
    'fcnod_detect_from_data': (_I, [_P, _P, _I, C.POINTER(PostnetParams), _P, _P, _P, _P]),
    'fcnod_detect_device': (_I, [_P, _P, _I, C.POINTER(PostnetParams), C.POINTER(_P), C.POINTER(_P),
                                 C.POINTER(_P), C.POINTER(_P), _P]),
    'fcnod_detect_graph_build': (_I, [_P, _P, _I, C.POINTER(PostnetParams), _P]),
    'fcnod_detect_graph_launch': (_I, [_P, _P]),
    'fcnod_caffemodel_blob': (_I, [_STR, _STR, _I, C.POINTER(_I * 4), C.POINTER(_I), _P, _SZ]),
    'fcnod_datum_decode': (_I, [_P, _SZ, C.POINTER(_I * 3), C.POINTER(_I), _P, _SZ]),
    'fcnod_label_raster_device': (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P]),
    'fcnod_label_raster': (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P]),
    'fcnod_aug_scratch_bytes': (_SZ, [_I]),
    'fcnod_aug_forward_device': (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    'fcnod_aug_forward': (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P]),
    'fcnod_aug_forward_device_noalloc': (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P]),
    'fcnod_aug_photo_scratch_bytes': (_SZ, [_I, _I, _I]),
    'fcnod_aug_forward_photo_device': (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    'fcnod_aug_forward_photo': (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P]),
    'fcnod_composite_device': (_I, [_P, _P, _I, _I, _I, _P, _P, _P, _P, _P]),
    'fcnod_composite': (_I, [_P, _P, _I, _I, _I, _P, _P, _SZ, _P, _P]),
    'fcnod_photo_augment_device': (_I, [_P, _P, _P, _I, _I, _I, _P, _P, _P]),
    'fcnod_photo_augment': (_I, [_P, _P, _P, _I, _I, _I, _P]),
    'fcnod_train_head_scratch_bytes': (_SZ, []),
    'fcnod_train_head_forward_device': (_I, [_P] * 7 + [_I, _I, _I, _I, C.c_float, _P, _P, _P, _P, _P]),
    'fcnod_train_head_forward': (_I, [_P] * 7 + [_I, _I, _I, _I, C.c_float, _P, _P, _P]),
    'fcnod_solver_defaults': (_I, [_P]),
    'fcnod_trainer_create': (_I, [_STR, _STR, _I, _I, _I, _I, _P, C.POINTER(_P)]),
    'fcnod_trainer_destroy': (_I, [_P]),
    'fcnod_trainer_info': (_I, [_P, C.POINTER(_I), C.POINTER(_SZ)]),
    'fcnod_trainer_buffers': (_I, [_P, C.POINTER(_P), C.POINTER(_P), C.POINTER(_P), C.POINTER(_P)]),
    'fcnod_trainer_forward_backward_device': (_I, [_P] * 7 + [_P]),
    'fcnod_trainer_apply_update': (_I, [_P, C.c_float, _P]),
    'fcnod_trainer_losses': (_I, [_P, _P, _P]),
    'fcnod_trainer_ipc_handles': (_I, [_P, _P, _P]),
    'fcnod_trainer_open_peers': (_I, [_P, _I, _I, _P, _P]),
    'fcnod_trainer_apply_update_peer': (_I, [_P, _P]),
    'fcnod_trainer_iteration': (_I, [_P]),
    'fcnod_trainer_step': (_I, [_P] * 8),
    'fcnod_trainer_param_read': (_I, [_P, _STR, _I, _I, _P, C.POINTER(_SZ)]),
    'fcnod_trainer_param_layer': (_STR, [_P, _I]),
    'fcnod_trainer_blob_read': (_I, [_P, _STR, _I, C.POINTER(_I), _P, _SZ]),
    'fcnod_trainer_dropout_mask': (_I, [_P, _P, _SZ, C.POINTER(_I)]),
    'fcnod_trainer_pool_argmax': (_I, [_P, _STR, _P, _SZ, C.POINTER(_I)]),
    'fcnod_trainer_set_timing': (_I, [_P, _I]),
    'fcnod_trainer_gemm_times': (_I, [_P, _P, _P, _P]),
    'fcnod_gemm_bf16_device': (_I, [_I, _P, _I, _P, _I, _I, _I, _I, _P, _I, _I, _P, _I, _I, _P]),
}

_lib = None


def lib():
    """The loaded library; raises if it has not been built (python -m fcn_object_detector_b200.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FcnodError('%s is missing: build it with `python -m fcn_object_detector_b200.build` '
                             '(there is no CPU fallback)' % LIB_PATH)
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc):
    if rc != 0:
        msg = lib().fcnod_last_error()
        raise FcnodError(msg.decode() if msg else 'fcnod call failed')


def ptr(a):
    """Data pointer of a C-contiguous numpy array (or None)."""
    if a is None:
        return None
    if not a.flags['C_CONTIGUOUS']:
        raise ValueError('array must be C-contiguous')
    return a.ctypes.data_as(C.c_void_p)


def default_postnet_params(**kw):
    p = PostnetParams()
    lib().fcnod_postnet_default_params(C.byref(p))
    for k, v in kw.items():
        if not hasattr(p, k):
            raise AttributeError(k)
        setattr(p, k, v)
    return p


def pinned_empty(shape, dtype):
    """numpy array in page-locked host memory (fcnod_host_alloc): copies to / from the device are then true DMA.
    The memory is released when the array (and every view of it) is garbage collected."""
    import weakref

    import numpy as np
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) if len(shape) else 1
    nbytes = max(1, n * dtype.itemsize)
    p = C.c_void_p()
    check(lib().fcnod_host_alloc(nbytes, C.byref(p)))
    buf = (C.c_char * nbytes).from_address(p.value)
    weakref.finalize(buf, lib().fcnod_host_free, C.c_void_p(p.value))
    return np.frombuffer(buf, dtype=dtype, count=n).reshape(shape)
