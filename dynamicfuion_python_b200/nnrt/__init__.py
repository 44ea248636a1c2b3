"""`nnrt` — B200-native drop-in for the dense-depth path of the reference's pybind11 module of the same name
(cpp/pybind/nnrt_pybind.cpp:23-40): same submodule tree and function names for that path, tensors in/out as torch CUDA tensors."""
from . import alignment, core, geometry, io, rendering  # noqa: F401
