"""Device time of the TSDF mesh extraction against the oracle on one host thread. usage (GPU box):
python tests/measure/mesh_extraction_times.py >> profiles/r2_mesh_extraction.jsonl"""
import itertools, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from dynamicfuion_python_b200 import nnrt
from oracle import oracle_py as oracle

R = 8
def blocks(keys, center, radius, voxel_size):
    keys = np.asarray(keys, np.int32)
    z, y, x = np.meshgrid(np.arange(R), np.arange(R), np.arange(R), indexing="ij")
    tsdf = np.empty((len(keys), R, R, R), np.float32)
    for b, k in enumerate(keys):
        p = np.stack([(k[0] * R + x) * voxel_size, (k[1] * R + y) * voxel_size, (k[2] * R + z) * voxel_size], -1)
        tsdf[b] = np.clip((np.linalg.norm(p - center, axis=-1) - radius) / (4 * voxel_size), -1, 1)
    return keys, tsdf, np.full(tsdf.shape, 10.0, np.float32)

for side, radius in ((8, 0.25), (16, 0.5)):
    voxel_size = 0.01
    # the blocks a fusion pipeline would hold: those within the truncation band of the surface
    allkeys = np.array(list(itertools.product(range(-side // 2, side // 2), repeat=3)), np.int32)
    centres = (allkeys + 0.5) * R * voxel_size
    keep = np.abs(np.linalg.norm(centres, axis=1) - radius) < 0.1
    keys, tsdf, weight = blocks(allkeys[keep], np.zeros(3), radius, voxel_size)
    dk, dt, dw = (torch.as_tensor(a).cuda() for a in (keys, tsdf, weight))
    f = nnrt.geometry.functional
    for _ in range(3):
        mesh = f.extract_triangle_mesh(dk, {"tsdf": dt, "weight": dw}, voxel_size)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        mesh = f.extract_triangle_mesh(dk, {"tsdf": dt, "weight": dw}, voxel_size)
    e1.record(); torch.cuda.synchronize()
    t0 = time.perf_counter(); ref = oracle.extract_triangle_mesh(keys, tsdf, weight, None, voxel_size, 3.0); cpu_ms = (time.perf_counter() - t0) * 1e3
    voxels = int(tsdf.size)
    print(json.dumps({"op": "extract_triangle_mesh", "blocks": len(keys), "voxels": voxels, "vertices": int(mesh.positions.shape[0]), "triangles": int(mesh.triangles.shape[0]),
                      "ms": e0.elapsed_time(e1) / 20, "algorithmic_MB": voxels * 8 / 1e6, "oracle_cpu_ms_1_thread": cpu_ms,
                      "equal_to_oracle": bool(np.array_equal(mesh.positions.cpu().numpy(), ref[0]) and np.array_equal(mesh.triangles.cpu().numpy(), ref[3]))}))
