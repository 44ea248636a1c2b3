"""CPU tests of the host logic: C ABI surface, Python mirror argument checks that need no device, frame-batch sharding over gloo (world 2)."""
import os
import re
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_c_abi_exports_every_declared_symbol():
    """libnnrt_b200.so loads (no GPU needed for dlopen) and exports every function include/nnrt_b200.h declares; the ctypes table in
    _lib.py binds exactly that set."""
    from dynamicfuion_python_b200 import _lib, build
    build.build()
    header = open(os.path.join(ROOT, "include", "nnrt_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(nnrt_b200_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 25
    lib = _lib.lib()
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/nnrt_b200.h but not exported"
    assert declared == set(_lib.exported_symbols())
    assert lib.nnrt_b200_version() >= 100


def test_no_cpu_fallback_in_product_path():
    """Neither the product package nor tools/ ever imports the oracle (measurement scripts that time it live under tests/measure),
    and asking for a device without CUDA raises instead of falling back."""
    for top in ("dynamicfuion_python_b200", "tools", "include"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, top)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".sh")):
                    text = open(os.path.join(dirpath, f)).read()
                    assert "oracle_py" not in text and "liboracle" not in text, f"{top}/{f} references the oracle"
    if not torch.cuda.is_available():
        from dynamicfuion_python_b200.nnrt import _tensor
        with pytest.raises(RuntimeError):
            _tensor.device()


def test_argument_errors_without_device():
    """LogError semantics of the reference constructors are host-side checks (DeformableMeshToImageFitter.cpp:79-82)."""
    from dynamicfuion_python_b200.nnrt.alignment import DeformableMeshToImageFitter, IterationMode
    with pytest.raises(RuntimeError):
        DeformableMeshToImageFitter(preconditioning_dampening_factor=-0.1)
    with pytest.raises(RuntimeError):
        DeformableMeshToImageFitter(preconditioning_dampening_factor=1.01)
    f = DeformableMeshToImageFitter()
    assert f.max_iteration_count == 100 and f.iteration_mode_sequence == [IterationMode.ALL]
    assert abs(f.options.arap_term_weight - 200.0) < 1e-6 and abs(f.options.max_depth - 10.0) < 1e-6


def test_shard_sequences_cover_and_balance():
    from dynamicfuion_python_b200.frame_batch import shard_sequences
    for world in (1, 2, 4, 8, 3):
        shards = [shard_sequences(64, r, world) for r in range(world)]
        assert sorted(sum(shards, [])) == list(range(64))
        assert max(map(len, shards)) - min(map(len, shards)) <= 1
    with pytest.raises(ValueError):
        shard_sequences(64, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from dynamicfuion_python_b200 import frame_batch
    seqs = frame_batch.shard_sequences(6, rank, world)
    log = []
    n = frame_batch.run_sequences(seqs, 3, lambda s, f: log.append((s, f)))
    fps, total, slowest = frame_batch.frames_per_second(n, 100.0 * (rank + 1))
    out[rank] = dict(seqs=seqs, log=log, fps=fps, total=total, slowest=slowest)
    dist.barrier()
    dist.destroy_process_group()


def test_frame_batch_two_ranks_gloo():
    """world_size 2 over gloo: disjoint sequence shards, serial frames inside a sequence, throughput = all frames / slowest rank."""
    world, port = 2, _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    assert sorted(out[0]["seqs"] + out[1]["seqs"]) == list(range(6))
    for r in range(world):
        assert out[r]["log"] == [(s, f) for s in out[r]["seqs"] for f in range(3)]
        assert out[r]["total"] == 18 and out[r]["slowest"] == 200.0
        assert abs(out[r]["fps"] - 18 / 0.2) < 1e-9


def test_pixel_rows_partition_the_image():
    """Bands of the pixel-sharded mode: disjoint, cover every row, multiples of the 8-row assembly tile except the last."""
    from dynamicfuion_python_b200.pixel_shard import pixel_rows
    for height in (120, 480, 1080, 2160):
        for world in (1, 2, 4, 8):
            rows = [pixel_rows(r, world, height) for r in range(world)]
            assert rows[0][0] == 0 and rows[-1][1] == height
            for (b0, e0), (b1, e1) in zip(rows, rows[1:]):
                assert e0 == b1 and b0 < e0 and b0 % 8 == 0
    with pytest.raises(ValueError):
        pixel_rows(2, 2, 480)
    with pytest.raises(ValueError):
        pixel_rows(0, 64, 120)


def test_stripes_partition_the_image():
    """Round-robin stripes of the pixel-sharded mode: every row belongs to exactly one rank, stripes start on the 8-row assembly tile
    grid, and the shares differ by at most one stripe."""
    from dynamicfuion_python_b200.pixel_shard import STRIPE_ROWS, stripe_rows_of
    assert STRIPE_ROWS % 8 == 0
    for height in (120, 480, 1080, 2160):
        for world in (1, 2, 3, 4, 8):
            for stripe in (8, 16, 32, 40):
                owner = -np.ones(height, int)
                counts = []
                for r in range(world):
                    ranges = stripe_rows_of(r, world, height, stripe)
                    counts.append(len(ranges))
                    for b, e in ranges:
                        assert b % 8 == 0 and b < e <= height and (owner[b:e] == -1).all()
                        owner[b:e] = r
                assert (owner >= 0).all() and max(counts) - min(counts) <= 1
                assert all(owner[v] == (v // stripe) % world for v in range(height))  # csrc/fitter_kernels.cuh::row_owned
    with pytest.raises(ValueError):
        stripe_rows_of(0, 2, 480, 12)
    with pytest.raises(ValueError):
        stripe_rows_of(2, 2, 480)


def test_shared_divisor_correction_yields_the_ieee_quotient():
    """The arithmetic of csrc/common.cuh::SharedDivisor restated in numpy (float32 values, each FMA evaluated exactly in float64 and
    rounded once -- a float32 product is exact in float64): r = RN(1 / d), q = RN(a r), e = RN(a - q d), a / d = RN(q + e r) must be the
    correctly rounded quotient, whereas the plain reciprocal multiply q is not (about a quarter of the pairs differ by an ulp: the
    reason the Jacobian chain cannot use it). The device-side check over 2^24 pairs is tests/test_gpu_stages.py."""
    rng = np.random.default_rng(5)
    n = 2_000_000
    a = (rng.standard_normal(n) * np.exp(rng.uniform(-10, 10, n))).astype(np.float32)
    d = (np.abs(rng.standard_normal(n)) * np.exp(rng.uniform(-10, 10, n)) + 1e-20).astype(np.float32)
    d[::2] *= -1
    ieee = a / d
    f64 = lambda x: x.astype(np.float64)
    r = (1.0 / f64(d)).astype(np.float32)
    q = (f64(a) * f64(r)).astype(np.float32)
    e = (f64(a) - f64(q) * f64(d)).astype(np.float32)
    corrected = (f64(q) + f64(e) * f64(r)).astype(np.float32)
    assert np.array_equal(corrected.view(np.int32), ieee.view(np.int32))
    assert (q != ieee).mean() > 0.1
