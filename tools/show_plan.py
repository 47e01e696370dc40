"""Print the bf16 deploy plan at batch B without a GPU (plan-only handle): python tools/show_plan.py [B]"""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fcn_object_detector_b200 import _lib
from fcn_object_detector_b200.caffemodel import write_synthetic_caffemodel
from fcn_object_detector_b200.netspec import write_deploy_prototxt
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
d = '/tmp/fcnod_plan'; os.makedirs(d, exist_ok=True)
proto = write_deploy_prototxt(os.path.join(d, 'deploy.prototxt'))
cm = os.path.join(d, 'w.caffemodel')
if not os.path.exists(cm): write_synthetic_caffemodel(cm, seed=0)
L = _lib.lib(); h = C.c_void_p()
_lib.check(L.fcnod_net_create(proto.encode(), cm.encode(), -1, _lib.FCNOD_PRECISION_BF16, C.byref(h)))
_lib.check(L.fcnod_net_reshape(h, B))
buf = C.create_string_buffer(1 << 16)
_lib.check(L.fcnod_net_describe(h, buf, len(buf)))
for l in buf.value.decode().splitlines():
    if 'conv_tc' in l:
        print(l.split(' -> ')[0][:190])
