"""nnrt_b200_compute_vertex_normals_static_scratch (include/nnrt_b200.h): the allocation-free variant of compute_vertex_normals, whose
CSR scratch is a file-scope __device__ array. Called through the C ABI and compared with the pool-allocating entry point; both sum the
triangle normals of a vertex in ascending face order, so the tolerance is the one tests/test_normals.py uses: bit-exact.
Collected last on purpose."""
import numpy as np
import pytest
import torch

from dynamicfuion_python_b200 import _lib

MAX_VERTICES, MAX_FACES = 1 << 17, 1 << 18  # NNRT_B200_STATIC_SCRATCH_MAX_{VERTICES,FACES}


def _mesh(seed, vertex_count, face_count, referenced):
    rng = np.random.default_rng(seed)
    triangles = rng.integers(0, referenced, (face_count, 3)).astype(np.int64)
    triangles[: face_count // 8, 0] = 3  # one high-degree vertex
    normals = rng.normal(0, 1, (face_count, 3)).astype(np.float32)
    return torch.from_numpy(triangles).cuda(), torch.from_numpy(normals).cuda(), vertex_count


def _run(name, triangles, triangle_normals, vertex_count, normalized, out=None):
    if out is None:
        out = torch.full((vertex_count, 3), float("nan"), dtype=torch.float32, device="cuda")
    stream = torch._C._cuda_getCurrentRawStream(torch.cuda.current_device())
    _lib.check(getattr(_lib.lib(), name)(triangles.data_ptr(), triangles.shape[0], triangle_normals.data_ptr(), vertex_count, int(normalized),
                                         out.data_ptr(), stream))
    return out


@pytest.mark.gpu
def test_static_scratch_vertex_normals_match_the_allocating_entry_point_twice_in_a_row():
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    # second call is smaller and shaped differently: stale counters / corner lists of the first call must not leak into it
    for seed, (vertex_count, face_count, referenced), normalized in ((1, (60000, 150000, 55000), True), (2, (900, 2500, 700), False)):
        triangles, triangle_normals, _ = _mesh(seed, vertex_count, face_count, referenced)
        want = _run("nnrt_b200_compute_vertex_normals", triangles, triangle_normals, vertex_count, normalized)
        got = _run("nnrt_b200_compute_vertex_normals_static_scratch", triangles, triangle_normals, vertex_count, normalized)
        torch.cuda.synchronize()
        assert torch.equal(got.view(torch.int32), want.view(torch.int32))
        assert not got[referenced:].any()  # unreferenced vertices: zero sum (and (0,0,0) stays as it is under normalisation)


@pytest.mark.gpu
def test_static_scratch_rejects_meshes_beyond_its_limit_and_runs_under_stream_capture():
    torch.cuda.set_device(0)
    triangles, triangle_normals, vertex_count = _mesh(3, 5000, 12000, 5000)
    out = torch.zeros((vertex_count, 3), dtype=torch.float32, device="cuda")
    fn = _lib.lib().nnrt_b200_compute_vertex_normals_static_scratch
    assert fn(triangles.data_ptr(), triangles.shape[0], triangle_normals.data_ptr(), MAX_VERTICES + 1, 1, out.data_ptr(), None) == 1
    assert b"static scratch limit" in _lib.lib().nnrt_b200_last_error()
    assert fn(triangles.data_ptr(), MAX_FACES + 1, triangle_normals.data_ptr(), vertex_count, 1, out.data_ptr(), None) == 1
    # an allocation inside a global-mode capture invalidates it; this entry point records and replays
    want = _run("nnrt_b200_compute_vertex_normals", triangles, triangle_normals, vertex_count, True)
    _run("nnrt_b200_compute_vertex_normals_static_scratch", triangles, triangle_normals, vertex_count, True, out)  # address lookup outside capture
    out.zero_()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, capture_error_mode="global"):
        _run("nnrt_b200_compute_vertex_normals_static_scratch", triangles, triangle_normals, vertex_count, True, out)
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(out.view(torch.int32), want.view(torch.int32))
