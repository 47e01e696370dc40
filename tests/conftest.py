import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (B200, sm_100a); run with -m gpu on the GPU box')


@pytest.fixture(scope='session')
def built_lib():
    """Path of libfcnod_b200.so (built in-tree if stale).  Never falls back to anything else."""
    from fcn_object_detector_b200 import build
    return build.build()


@pytest.fixture(scope='session')
def model_files(tmp_path_factory):
    """(prototxt, caffemodel): generated deploy net + synthetic xavier weights, seed 0."""
    from fcn_object_detector_b200.caffemodel import write_synthetic_caffemodel
    from fcn_object_detector_b200.netspec import write_deploy_prototxt
    d = tmp_path_factory.mktemp('model')
    proto = write_deploy_prototxt(str(d / 'deploy.prototxt'))
    cm = write_synthetic_caffemodel(str(d / 'snapshot_iter_1000.caffemodel'), seed=0)
    return proto, cm


@pytest.fixture(scope='session')
def firing_model_files(tmp_path_factory):
    """Same net, heads biased so that random frames yield detections (caffemodel.firing_heads)."""
    from fcn_object_detector_b200.caffemodel import write_synthetic_caffemodel
    from fcn_object_detector_b200.netspec import write_deploy_prototxt
    d = tmp_path_factory.mktemp('model_firing')
    proto = write_deploy_prototxt(str(d / 'deploy.prototxt'))
    cm = write_synthetic_caffemodel(str(d / 'snapshot_iter_1000.caffemodel'), seed=0, firing=True)
    return proto, cm


@pytest.fixture(scope='session')
def oracle_net(model_files):
    from oracle import prototxt
    from oracle.caffemodel_io import read_caffemodel
    proto, cm = model_files
    return prototxt.layers(prototxt.load(proto)), read_caffemodel(cm)


