import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle_py
    oracle_py.lib()
    return oracle_py


class _ModuleProxy:
    """The pybind11 module under the same tests as the ctypes mirror: a name the C++ module does not bind (rows outside the hot path's
    host classes, e.g. TSDF integration or the shaders) skips the test instead of failing it."""

    def __init__(self, module):
        object.__setattr__(self, "_module", module)

    def __getattr__(self, name):
        import types
        module = object.__getattribute__(self, "_module")
        if not hasattr(module, name):
            pytest.skip(f"{module.__name__}.{name} is not bound in the pybind11 module")
        value = getattr(module, name)
        return _ModuleProxy(value) if isinstance(value, types.ModuleType) else value


@pytest.fixture(scope="session")
def nnrt_mirror():
    """The ctypes mirror alone, for tests that reach into its internals (the fitter's C handle)."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from dynamicfuion_python_b200 import nnrt as module
    return module


@pytest.fixture(scope="session", params=["ctypes_mirror", "pybind11_module"])
def nnrt(request):
    """Both front ends of libnnrt_b200.so: the Python mirror over ctypes (dynamicfuion_python_b200.nnrt) and the pybind11 module `nnrt`
    with the C++ host classes (dynamicfuion_python_b200/pybind). Every test that takes this fixture runs against both."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    if request.param == "ctypes_mirror":
        from dynamicfuion_python_b200 import nnrt as module
        return module
    from dynamicfuion_python_b200 import build
    return _ModuleProxy(build.import_pybind_module())
