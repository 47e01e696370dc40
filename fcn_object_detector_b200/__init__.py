"""B200-native (sm_100a) hot path of iKrishneel/fcn_object_detector.

Host classes mirror the reference's two interfaces:
  detector.FCNObjectDetector            predict-on-image (scripts/fcn_object_detector.py)
  data_argumentation_layer.DataArgumentationLayer   Caffe Python layer (README.md:60-76)
  caffe_shim                            the slice of pycaffe both of them use
All compute is in libfcnod_b200.so (C ABI: include/fcnod_b200.h); there is no CPU fallback.
"""
__version__ = '0.1.0'
