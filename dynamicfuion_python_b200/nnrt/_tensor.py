"""Tensor plumbing between Python callers and the C ABI: torch owns device memory and streams, nothing else.

Accepts torch tensors, numpy arrays and anything exposing __dlpack__ (e.g. open3d.core.Tensor when Open3D is
present); returns torch CUDA tensors. A mesh argument is duck-typed like the reference's bindings accept an
open3d.t.geometry.TriangleMesh: an object with .vertex["positions"|"normals"] and .triangle["indices"], or a
(positions, normals, triangles) tuple, or a dict with those keys.
"""
import numpy as np
import torch

from .. import _lib


_cuda_checked = False


def device():
    global _cuda_checked
    if not _cuda_checked:  # the availability probe costs ~5 us per call (environment lookups); once per process is enough
        if not torch.cuda.is_available():
            raise RuntimeError("nnrt (B200 build) needs a CUDA device; there is no CPU path")
        _cuda_checked = True
    return torch.device("cuda", torch.cuda.current_device())


def to_torch(x):
    if isinstance(x, torch.Tensor):
        return x
    if isinstance(x, np.ndarray):
        return torch.from_numpy(np.ascontiguousarray(x))
    if hasattr(x, "__dlpack__"):
        return torch.from_dlpack(x)
    return torch.as_tensor(x)


def cuda(x, dtype):
    t = to_torch(x)
    if t.is_cuda and t.dtype == dtype and t.is_contiguous():  # the common case on the hot path: already where and what it must be
        return t
    if t.dtype == torch.bool and dtype == torch.uint8:
        t = t.to(torch.uint8)
    # a pinned host tensor is uploaded asynchronously on the current stream (stream order protects the consumer kernels; the caller
    # must not rewrite the pinned buffer before the stream has passed the copy, as with any cudaMemcpyAsync)
    return t.to(device=device(), dtype=dtype, non_blocking=(not t.is_cuda) and t.is_pinned()).contiguous()


def host_f64(x, shape):
    """Host-side float64 matrix (intrinsics / extrinsics live on CPU:0 in the reference API)."""
    t = to_torch(x).detach().to("cpu", torch.float64).contiguous()
    if tuple(t.shape) != tuple(shape):
        raise RuntimeError(f"expected a tensor of shape {tuple(shape)}, got {tuple(t.shape)}")
    return t


_EMPTY = {}


def ptr(t):
    """Device address for the C ABI. An empty tensor has no storage (data_ptr() == 0), which the library would read as a missing
    argument: it gets the address of a small per-device placeholder instead (the entry points return early on a zero count)."""
    if t is None:
        return None
    address = t.data_ptr()
    if address == 0 and t.numel() == 0 and t.is_cuda:
        index = t.device.index
        if index not in _EMPTY:
            _EMPTY[index] = torch.zeros(64, dtype=torch.float32, device=t.device)
        return _EMPTY[index].data_ptr()
    return address


def stream():
    """Raw handle of torch's current stream on the current device (the C ABI takes a cudaStream_t as void*)."""
    return torch._C._cuda_getCurrentRawStream(torch.cuda.current_device())


def empty(shape, dtype):
    return torch.empty(shape, dtype=dtype, device=device())


def call(name, *args):
    _lib.check(getattr(_lib.lib(), name)(*args))


class MeshView:
    """positions [V,3] f32, normals [V,3] f32 or None, triangles [F,3] int64 on the device."""

    def __init__(self, mesh):
        pos = nrm = tri = None
        self.triangle_normals = None
        if isinstance(mesh, MeshView):
            pos, nrm, tri = mesh.positions, mesh.normals, mesh.triangles
            self.triangle_normals = mesh.triangle_normals
        elif isinstance(mesh, (tuple, list)):
            pos, nrm, tri = (list(mesh) + [None, None])[:3]
        elif isinstance(mesh, dict):
            pos, nrm, tri = mesh.get("positions"), mesh.get("normals"), mesh.get("triangles", mesh.get("indices"))
            if mesh.get("triangle_normals") is not None:
                self.triangle_normals = cuda(mesh["triangle_normals"], torch.float32)
        elif hasattr(mesh, "vertex") and hasattr(mesh, "triangle"):
            pos = mesh.vertex["positions"]
            nrm = mesh.vertex["normals"] if "normals" in mesh.vertex else None
            tri = mesh.triangle["indices"] if "indices" in mesh.triangle else None
            if "normals" in mesh.triangle:
                self.triangle_normals = cuda(mesh.triangle["normals"], torch.float32)
        else:
            raise RuntimeError("mesh must be an Open3D t.geometry.TriangleMesh-like object, a dict or a (positions, normals, triangles) tuple")
        if pos is None:
            raise RuntimeError("Argument mesh needs to have vertex positions defined.")
        self.positions = cuda(pos, torch.float32)
        self.normals = None if nrm is None else cuda(nrm, torch.float32)
        self.triangles = None if tri is None else cuda(tri, torch.int64)

    def with_vertices(self, positions, normals):
        out = MeshView.__new__(MeshView)
        out.positions, out.normals, out.triangles, out.triangle_normals = positions, normals, self.triangles, None
        return out

    # minimal Open3D-like accessors so results can be consumed the same way
    @property
    def vertex(self):
        d = {"positions": self.positions}
        if self.normals is not None:
            d["normals"] = self.normals
        return d

    @property
    def triangle(self):
        d = {} if self.triangles is None else {"indices": self.triangles}
        if self.triangle_normals is not None:
            d["normals"] = self.triangle_normals
        return d


def store_mesh_attribute(mesh, group, name, value):
    """Writes a computed attribute back into the caller's mesh the way the reference's in-place operations do: a MeshView attribute,
    a dict key, or `mesh.vertex[name]` / `mesh.triangle[name]` of an Open3D-like object (converted through DLPack when Open3D is
    importable). Tuples are immutable: the caller uses the returned tensor."""
    if isinstance(mesh, MeshView):
        setattr(mesh, "normals" if group == "vertex" else "triangle_normals", value)
    elif isinstance(mesh, dict):
        mesh["normals" if group == "vertex" else "triangle_normals"] = value
    elif hasattr(mesh, group) and not isinstance(mesh, (tuple, list)):
        try:
            import open3d
            value = open3d.core.Tensor.from_dlpack(torch.utils.dlpack.to_dlpack(value))
        except ImportError:
            pass
        getattr(mesh, group)[name] = value
