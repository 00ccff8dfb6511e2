import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _ensure_oracle():
    port = os.path.join(ROOT, "oracle", "_build", "libfbo_port.so")
    if not os.path.exists(port):
        from binless_b200 import build as b
        b.build_oracle()


@pytest.fixture(scope="session")
def oracle_port():
    _ensure_oracle()
    from oracle.loader import oracle_lib
    return oracle_lib("port")


@pytest.fixture(scope="session")
def oracle_ref():
    """The restated driver linked against the reference's own gfl_graph_fl.c / gfl_tf.c (built from /root/reference)."""
    _ensure_oracle()
    path = os.path.join(ROOT, "oracle", "_ref", "libfbo_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    from oracle.loader import oracle_lib
    return oracle_lib("reference")


@pytest.fixture(scope="session")
def oracle(oracle_port):
    from oracle.loader import oracle_lib
    return oracle_lib("auto")


@pytest.fixture(scope="session")
def gpu_lib():
    """The product library on a GPU box. No fallback: a missing library or device is a failure, not a skip."""
    import ctypes
    from binless_b200 import cuda_lib
    lib = cuda_lib()
    cudart_ok = ctypes.c_int(0)
    try:
        import torch
        assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    except ImportError:
        pass
    return lib
