"""nnrt.io — read_tensor / write_tensor in the reference's on-disk format (cpp/pybind/io.cpp:36-52, cpp/io/TensorIO.cpp:32-60).

Layout (little endian): shape (u32 count, i64 each), strides in elements (same encoding), dtype (i32 Open3D DtypeCode, i64 byte size, u8 name
length, name), blob (i32 device type = CPU 0, i32 device id, i64 byte count, the contiguous bytes); `compressed` wraps the stream in gzip
(3rd-party zstr over zlib in the reference). The reader applies the stored strides to the contiguous blob, like the reference. Host-side only."""
import gzip
import struct

import numpy as np
import torch

# torch dtype -> (Open3D v0.17 DtypeCode: Bool 1, Int 2, UInt 3, Float 4; byte size; Dtype::ToString())
_DTYPES = {torch.float32: (4, 4, "Float32"), torch.float64: (4, 8, "Float64"), torch.int8: (2, 1, "Int8"), torch.int16: (2, 2, "Int16"),
           torch.int32: (2, 4, "Int32"), torch.int64: (2, 8, "Int64"), torch.uint8: (3, 1, "UInt8"), torch.uint16: (3, 2, "UInt16"),
           torch.uint32: (3, 4, "UInt32"), torch.uint64: (3, 8, "UInt64"), torch.bool: (1, 1, "Bool")}
_BY_RECORD = {v: k for k, v in _DTYPES.items()}


def _to_torch(x):
    if isinstance(x, torch.Tensor):
        return x
    if isinstance(x, np.ndarray):
        return torch.from_numpy(x)
    if hasattr(x, "__dlpack__"):
        return torch.from_dlpack(x)
    return torch.as_tensor(x)


def _size_vector(values):
    return struct.pack("<I", len(values)) + b"".join(struct.pack("<q", int(v)) for v in values)


def _tensor_record(tensor):
    """The bytes TensorIO.cpp's operator<< writes for one tensor."""
    t = _to_torch(tensor)
    if t.dtype not in _DTYPES:
        raise RuntimeError(f"write_tensor: unsupported dtype {t.dtype}")
    code, size, name = _DTYPES[t.dtype]
    host = t.detach().contiguous().cpu()
    payload = host.reshape(-1).view(torch.uint8).numpy().tobytes() if host.numel() else b""
    return (_size_vector(t.shape) + _size_vector(t.stride()) + struct.pack("<iqB", code, size, len(name)) + name.encode() +
            struct.pack("<iiq", 0, 0, len(payload)) + payload)


def _write_stream(path, data, compressed):
    try:
        with (gzip.open(path, "wb") if compressed else open(path, "wb")) as fh:
            fh.write(data)
    except OSError as error:
        raise RuntimeError(f'Failed to open file at "{path}" for writing.') from error


def _read_stream(path, compressed):
    try:
        with open(path, "rb") as fh:
            raw = fh.read()
    except OSError as error:
        raise RuntimeError(f'Failed to open file at "{path}" for reading.') from error
    if compressed and raw[:2] == b"\x1f\x8b":  # zstr::ifstream auto-detects; a plain file passes through
        raw = gzip.decompress(raw)
    return raw


class _Reader:
    def __init__(self, raw):
        self.raw, self.offset = raw, 0

    def take(self, fmt):
        size = struct.calcsize(fmt)
        if self.offset + size > len(self.raw):
            raise RuntimeError("Failure reading from istream.")
        values = struct.unpack_from(fmt, self.raw, self.offset)
        self.offset += size
        return values

    def bytes(self, count):
        if count < 0 or self.offset + count > len(self.raw):
            raise RuntimeError("Failure reading from istream.")
        out = self.raw[self.offset:self.offset + count]
        self.offset += count
        return out

    def size_vector(self):
        (count,) = self.take("<I")
        if count > 64:
            raise RuntimeError("Failure reading from istream.")
        return list(self.take(f"<{count}q")) if count else []

    def tensor(self):
        """One tensor as TensorIO.cpp's operator>> reads it: the stored strides are applied to the contiguous blob."""
        shape, strides = self.size_vector(), self.size_vector()
        code, size, name_size = self.take("<iqB")
        name = self.bytes(name_size).decode("ascii", "replace")
        if (code, size, name) not in _BY_RECORD:
            raise RuntimeError(f'read_tensor: unsupported dtype record (code {code}, {size} bytes, "{name}")')
        device_type, device_id, byte_size = self.take("<iiq")
        if (device_type, device_id) != (0, 0):
            raise RuntimeError(f"Blob istream input for blobs with device type {device_type}:{device_id} not supported (only CPU:0 device is supported).")
        if byte_size < 0 or byte_size % size or len(shape) != len(strides):
            raise RuntimeError("Failure reading from istream.")
        dtype = _BY_RECORD[(code, size, name)]
        blob = self.bytes(byte_size)
        flat = torch.frombuffer(bytearray(blob), dtype=dtype) if byte_size else torch.empty((0,), dtype=dtype)
        count, last = 1, 0
        for extent, stride in zip(shape, strides):
            if extent < 0 or stride < 0:
                raise RuntimeError("Failure reading from istream.")
            count *= extent
            last += (extent - 1) * stride if extent > 0 else 0
        if count > 0 and last >= flat.numel():
            raise RuntimeError("read_tensor: the stored strides address beyond the stored data")
        return torch.as_strided(flat, shape, strides)


def write_tensor(path, tensor, compressed=True):
    """Write tensor in binary form."""
    _write_stream(path, _tensor_record(tensor), compressed)


def read_tensor(path, compressed=True):
    """Read tensor from binary form (returned on the CPU, like the reference's ReadTensor)."""
    return _Reader(_read_stream(path, compressed)).tensor()


# ---- voxel block grids (cpp/pybind/io.cpp:54-79, cpp/io/VoxelBlockGridIO.cpp:26-140) ---------------------------------------------
# Layout: f32 voxel size, i64 block resolution, i32 attribute count, per attribute (u32 name length, name, i32 attribute index) -- the
# reference walks a std::unordered_map, so the order of these records is unspecified and a reader must go by the stored index; this
# writer emits them in index order --, the hash map's device (i32 type, i32 id; 1 = CUDA), i64 hash map capacity, then as tensors
# (layout above) the ACTIVE block keys [B,3] Int32 and one value tensor per attribute [B,R,R,R,channels] in attribute-index order.
def _grid_record(grid):
    from .geometry.voxel_block_grid import VoxelBlockGrid
    if not isinstance(grid, VoxelBlockGrid):
        raise RuntimeError("write_voxel_block_grid: expected a nnrt.geometry.VoxelBlockGrid")
    names = grid.attribute_names()
    data = struct.pack("<fqi", float(grid.voxel_size), int(grid.block_resolution), len(names))
    for index, name in enumerate(names):
        data += struct.pack("<I", len(name)) + name.encode() + struct.pack("<i", index)
    device = grid.get_device()
    data += struct.pack("<iiq", 1 if device.type == "cuda" else 0, device.index or 0, int(grid.hashmap().capacity()))
    size = grid.hashmap().size()  # blocks are appended: the active buffer indices are 0 .. size-1 (geometry/voxel_block_grid.py)
    data += _tensor_record(grid.hashmap().key_tensor()[:size])
    for index in range(len(names)):
        data += _tensor_record(grid.hashmap().value_tensor(index)[:size])
    return data


def _read_grid(path, compressed, cls):
    reader = _Reader(_read_stream(path, compressed))
    voxel_size, block_resolution, attribute_count = reader.take("<fqi")
    if attribute_count < 0 or attribute_count > 64:
        raise RuntimeError("Failure reading from istream.")
    names = {}
    for _ in range(attribute_count):
        (length,) = reader.take("<I")
        name = reader.bytes(length).decode("utf-8", "replace")
        (index,) = reader.take("<i")
        names[index] = name
    if sorted(names) != list(range(attribute_count)):
        raise RuntimeError("read_voxel_block_grid: attribute indices are not 0..count-1")
    device_type, device_id, capacity = reader.take("<iiq")
    keys = reader.tensor()
    values = [reader.tensor() for _ in range(attribute_count)]
    block_count = int(keys.shape[0])
    for value in values:
        if value.dim() < 4 or value.shape[0] != block_count or any(int(e) != block_resolution for e in value.shape[1:4]):
            raise RuntimeError("read_voxel_block_grid: a value tensor does not have shape [blocks, R, R, R, ...]")
    grid = cls([names[i] for i in range(attribute_count)], [v.dtype for v in values], [tuple(v.shape[4:]) or (1,) for v in values], voxel_size,
               block_resolution, max(int(capacity), block_count, 1))
    grid.hashmap().insert(keys, values)
    return grid


def write_voxel_block_grid(path, voxel_block_grid, compressed=True):
    """Write voxel block hash in binary form."""
    _write_stream(path, _grid_record(voxel_block_grid), compressed)


def read_voxel_block_grid(path, compressed=True):
    """Read a voxel block grid written by write_voxel_block_grid (blocks land on the current CUDA device)."""
    from .geometry.voxel_block_grid import VoxelBlockGrid
    return _read_grid(path, compressed, VoxelBlockGrid)


def write_non_rigid_surface_voxel_block_grid(path, non_rigid_surface_voxel_block_grid, compressed=True):
    """Write voxel block hash in binary form."""
    _write_stream(path, _grid_record(non_rigid_surface_voxel_block_grid), compressed)


def read_non_rigid_surface_voxel_block_grid(path, compressed=True):
    from .geometry.voxel_block_grid import NonRigidSurfaceVoxelBlockGrid
    return _read_grid(path, compressed, NonRigidSurfaceVoxelBlockGrid)
