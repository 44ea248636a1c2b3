"""nnrt.io.read_tensor / write_tensor (cpp/pybind/io.cpp:36-52, cpp/io/TensorIO.cpp): the bytes follow the reference's stream operators
(SizeVectorIO / DtypeIO / DeviceIO / BlobIO .cpp), checked against a hand-assembled expected byte string, round trips through both
front ends (pybind11 C++ and the Python mirror, each reading the other's files), gzip framing, the reference's stride quirk and errors.
Host-side: no GPU needed."""
import gzip
import struct

import numpy as np
import pytest
import torch

from dynamicfuion_python_b200 import build
from dynamicfuion_python_b200.nnrt import io as mirror_io


@pytest.fixture(scope="module")
def cpp_io():
    build.build_pybind()
    return build.import_pybind_module().io


def _expected_bytes(array, code, name):
    shape, itemsize = array.shape, array.dtype.itemsize
    strides = [s // itemsize for s in array.strides]
    out = struct.pack("<I", len(shape)) + b"".join(struct.pack("<q", d) for d in shape)
    out += struct.pack("<I", len(strides)) + b"".join(struct.pack("<q", s) for s in strides)
    out += struct.pack("<i", code) + struct.pack("<q", itemsize) + struct.pack("<B", len(name)) + name.encode()
    out += struct.pack("<i", 0) + struct.pack("<i", 0)  # Device: CPU, id 0
    payload = np.ascontiguousarray(array).tobytes()
    return out + struct.pack("<q", len(payload)) + payload


CASES = [(np.arange(24, dtype=np.float32).reshape(2, 3, 4), 4, "Float32"), (np.linspace(-1, 1, 7), 4, "Float64"),
         (np.arange(-5, 5, dtype=np.int64).reshape(5, 2), 2, "Int64"), (np.arange(6, dtype=np.int32), 2, "Int32"),
         (np.array([[1, 2], [3, 250]], np.uint8), 3, "UInt8"), (np.array([True, False, True]), 1, "Bool"),
         (np.array([1000, 65535], np.uint16), 3, "UInt16"), (np.zeros((0, 3), np.float32), 4, "Float32"), (np.array(2.5, np.float32), 4, "Float32")]


@pytest.mark.parametrize("writer", ["cpp", "mirror"])
@pytest.mark.parametrize("array,code,name", CASES, ids=[c[2] + str(c[0].shape) for c in CASES])
def test_written_bytes_follow_the_reference_layout(tmp_path, cpp_io, writer, array, code, name):
    io = cpp_io if writer == "cpp" else mirror_io
    path = str(tmp_path / "t.dat")
    io.write_tensor(path, torch.from_numpy(array) if array.dtype != np.uint16 else torch.from_numpy(array.view(np.int16)).view(torch.uint16), False)
    assert open(path, "rb").read() == _expected_bytes(array, code, name)
    io.write_tensor(path, array if array.dtype != np.uint16 else torch.from_numpy(array.view(np.int16)).view(torch.uint16))  # compressed by default
    raw = open(path, "rb").read()
    assert raw[:2] == b"\x1f\x8b" and gzip.decompress(raw) == _expected_bytes(array, code, name)  # gzip framing, as zstr::ofstream writes


@pytest.mark.parametrize("compressed", [True, False])
def test_round_trip_across_the_two_front_ends(tmp_path, cpp_io, compressed):
    rng = np.random.default_rng(4)
    tensors = [torch.from_numpy(rng.normal(size=(37, 5, 3)).astype(np.float32)), torch.arange(100, dtype=torch.int64).reshape(10, 10),
               torch.from_numpy(rng.integers(0, 2, (8, 8)).astype(bool)), torch.from_numpy(rng.normal(size=(4,)))]
    for k, t in enumerate(tensors):
        for writer, reader in ((cpp_io, mirror_io), (mirror_io, cpp_io), (cpp_io, cpp_io)):
            path = str(tmp_path / f"t{k}.dat")
            writer.write_tensor(path, t, compressed)
            back = reader.read_tensor(path, compressed)
            assert back.dtype == t.dtype and back.device.type == "cpu" and torch.equal(back, t)
    # a plain file read with compressed=True passes through (zstr::ifstream auto-detects)
    path = str(tmp_path / "plain.dat")
    cpp_io.write_tensor(path, tensors[0], False)
    assert torch.equal(cpp_io.read_tensor(path, True), tensors[0]) and torch.equal(mirror_io.read_tensor(path, True), tensors[0])


def test_non_contiguous_tensor_reproduces_the_reference_quirk(tmp_path, cpp_io):
    """The writer stores the ORIGINAL strides next to the CONTIGUOUS data and the reader applies them (TensorIO.cpp:33-42,55): a
    transposed tensor does not survive the reference's round trip; both front ends return what the reference returns."""
    t = torch.arange(12, dtype=torch.float32).reshape(3, 4).t()  # shape (4, 3), strides (1, 4)
    want = torch.as_strided(t.contiguous().reshape(-1), (4, 3), (1, 4))
    for io in (cpp_io, mirror_io):
        path = str(tmp_path / "nc.dat")
        io.write_tensor(path, t)
        assert torch.equal(io.read_tensor(path), want) and not torch.equal(want, t)


def test_errors(tmp_path, cpp_io):
    for io in (cpp_io, mirror_io):
        with pytest.raises(RuntimeError, match="Failed to open file"):
            io.read_tensor(str(tmp_path / "missing.dat"))
        with pytest.raises(RuntimeError, match="Failed to open file"):
            io.write_tensor(str(tmp_path / "no_such_dir" / "t.dat"), torch.zeros(3))
        path = str(tmp_path / "cuda_blob.dat")
        good = _expected_bytes(np.arange(4, dtype=np.float32), 4, "Float32")
        cut = good.index(b"Float32") + 7
        open(path, "wb").write(good[:cut] + struct.pack("<ii", 1, 0) + good[cut + 8:])  # DeviceType::CUDA
        with pytest.raises(RuntimeError, match="only CPU:0 device is supported"):
            io.read_tensor(path, False)
        open(path, "wb").write(good[:-5])  # truncated payload
        with pytest.raises(RuntimeError, match="Failure reading"):
            io.read_tensor(path, False)
