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


def write_tensor(path, tensor, compressed=True):
    """Write tensor in binary form."""
    t = _to_torch(tensor)
    if t.dtype not in _DTYPES:
        raise RuntimeError(f"write_tensor: unsupported dtype {t.dtype}")
    code, size, name = _DTYPES[t.dtype]
    host = t.detach().contiguous().cpu()
    payload = host.reshape(-1).view(torch.uint8).numpy().tobytes() if host.numel() else b""
    data = (_size_vector(t.shape) + _size_vector(t.stride()) + struct.pack("<iqB", code, size, len(name)) + name.encode() +
            struct.pack("<iiq", 0, 0, len(payload)) + payload)
    try:
        with (gzip.open(path, "wb") if compressed else open(path, "wb")) as fh:
            fh.write(data)
    except OSError as error:
        raise RuntimeError(f'Failed to open file at "{path}" for writing.') from error


def read_tensor(path, compressed=True):
    """Read tensor from binary form (returned on the CPU, like the reference's ReadTensor)."""
    try:
        with open(path, "rb") as fh:
            raw = fh.read()
    except OSError as error:
        raise RuntimeError(f'Failed to open file at "{path}" for reading.') from error
    if compressed and raw[:2] == b"\x1f\x8b":  # zstr::ifstream auto-detects; a plain file passes through
        raw = gzip.decompress(raw)
    offset = 0

    def take(fmt):
        nonlocal offset
        size = struct.calcsize(fmt)
        if offset + size > len(raw):
            raise RuntimeError("Failure reading from istream.")
        values = struct.unpack_from(fmt, raw, offset)
        offset += size
        return values

    def size_vector():
        (count,) = take("<I")
        if count > 64:
            raise RuntimeError("Failure reading from istream.")
        return list(take(f"<{count}q")) if count else []

    shape, strides = size_vector(), size_vector()
    code, size, name_size = take("<iqB")
    name = raw[offset:offset + name_size].decode("ascii", "replace")
    offset += name_size
    if (code, size, name) not in _BY_RECORD:
        raise RuntimeError(f'read_tensor: unsupported dtype record (code {code}, {size} bytes, "{name}")')
    device_type, device_id, byte_size = take("<iiq")
    if (device_type, device_id) != (0, 0):
        raise RuntimeError(f"Blob istream input for blobs with device type {device_type}:{device_id} not supported (only CPU:0 device is supported).")
    if byte_size < 0 or byte_size % size or len(shape) != len(strides) or offset + byte_size > len(raw):
        raise RuntimeError("Failure reading from istream.")
    dtype = _BY_RECORD[(code, size, name)]
    flat = torch.frombuffer(bytearray(raw[offset:offset + byte_size]), dtype=dtype) if byte_size else torch.empty((0,), dtype=dtype)
    count, last = 1, 0
    for extent, stride in zip(shape, strides):
        if extent < 0 or stride < 0:
            raise RuntimeError("Failure reading from istream.")
        count *= extent
        last += (extent - 1) * stride if extent > 0 else 0
    if count > 0 and last >= flat.numel():
        raise RuntimeError("read_tensor: the stored strides address beyond the stored data")
    return torch.as_strided(flat, shape, strides)
