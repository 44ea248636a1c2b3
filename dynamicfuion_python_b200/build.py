"""In-tree build of libnnrt_b200.so (hand-written sm_100a CUDA behind the C ABI in include/nnrt_b200.h).

nvcc cross-compiles without a GPU. -fmad=false is deliberate: anchor selection and raster coverage are decided by
float comparisons that must round like the reference's non-FMA x86-64 build (DESIGN.md "bit-exact indices").
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libnnrt_b200.so")
SOURCES = ["c_api.cu", "geometry.cu", "warp_field_build.cu", "mesh_normals.cu", "tsdf.cu", "raster.cu", "fitter_kernels.cu", "assemble_kernel.cu", "solver_kernels.cu", "corner_cluster.cu", "linalg_api.cu", "linalg_blocks.cu", "assembly_stages.cu", "node_store.cu", "mesh_extraction.cu", "fitter.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-fmad=false", "-std=c++17", "--extended-lambda",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "-DNNRT_B200_BUILD"]
# Per-file overrides of -fmad=false. Empty on purpose: contraction + fast division in assemble_kernel.cu (-fmad=true -prec-div=false)
# removed a third of its instructions (3144 -> 1928 SASS) but gained only 2 us per iteration once the residual chain was kept
# exact, and pushed J^T J past the 1e-4 parity bar on entries with cancellation (measured, round 1).
RELAXED_ARITHMETIC = {}


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.exists(cand) or cand == "nvcc"):
            return cand
    raise RuntimeError("nvcc not found")


def _stale():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "nnrt_b200.h"), __file__]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not _stale():
        return LIB_PATH
    os.makedirs(LIB_DIR, exist_ok=True)
    obj_dir = os.path.join(HERE, "build")
    os.makedirs(obj_dir, exist_ok=True)
    nvcc = _nvcc()
    procs = []
    for src in SOURCES:
        obj = os.path.join(obj_dir, src.replace(".cu", ".o"))
        flags = [f for f in NVCC_FLAGS if f != "-fmad=false"] + RELAXED_ARITHMETIC[src] if src in RELAXED_ARITHMETIC else NVCC_FLAGS
        cmd = [nvcc] + flags + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((src, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    objs = []
    for src, obj, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{out}")
        if verbose:
            print(out)
        objs.append(obj)
    link = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB_PATH] + objs + ["-lcudart"]
    subprocess.check_call(link)
    return LIB_PATH


# ---- pybind11 module `nnrt` (C++ host classes over the C ABI): dynamicfuion_python_b200/pybind/nnrt.<abi>.so ---------------------------
PYBIND_DIR = os.path.join(HERE, "pybind")
PYBIND_SOURCES = ["bridge.cpp", "warp_field.cpp", "geometry_bind.cpp", "geometry_functional_bind.cpp", "tsdf_bind.cpp", "alignment_bind.cpp", "alignment_functional_bind.cpp", "core_rendering_bind.cpp", "linalg_blocks_bind.cpp", "io_bind.cpp", "nnrt_module.cpp"]


def pybind_module_path():
    import sysconfig
    return os.path.join(PYBIND_DIR, "nnrt" + sysconfig.get_config_var("EXT_SUFFIX"))


def build_pybind(force=False):
    """g++ + pybind11 headers; links libnnrt_b200.so by relative rpath. No CUDA toolkit needed at this step."""
    import sysconfig
    import pybind11
    out = pybind_module_path()
    deps = [os.path.join(PYBIND_DIR, f) for f in os.listdir(PYBIND_DIR) if f.endswith((".cpp", ".h"))] + [os.path.join(HERE, "..", "include", "nnrt_b200.h")]
    if not force and os.path.exists(out) and all(os.path.getmtime(d) <= os.path.getmtime(out) for d in deps):
        return out
    obj_dir = os.path.join(HERE, "build", "pybind")
    os.makedirs(obj_dir, exist_ok=True)
    flags = ["-O2", "-std=c++17", "-fPIC", "-fvisibility=hidden", "-I", pybind11.get_include(), "-I", sysconfig.get_paths()["include"]]
    procs = []
    for src in PYBIND_SOURCES:
        obj = os.path.join(obj_dir, src.replace(".cpp", ".o"))
        procs.append((src, obj, subprocess.Popen(["g++"] + flags + ["-c", os.path.join(PYBIND_DIR, src), "-o", obj], stdout=subprocess.PIPE,
                                                 stderr=subprocess.STDOUT, text=True)))
    objs = []
    for src, obj, p in procs:
        text, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f"g++ failed for {src}:\n{text}")
        objs.append(obj)
    subprocess.check_call(["g++", "-shared", "-o", out] + objs + ["-L", LIB_DIR, "-lnnrt_b200", "-lz", "-Wl,-rpath,$ORIGIN/../lib"])
    return out


def import_pybind_module():
    """`import nnrt` with the in-tree module directory on sys.path (what a user of the reference does after installing its wheel)."""
    if PYBIND_DIR not in sys.path:
        sys.path.insert(0, PYBIND_DIR)
    import nnrt
    return nnrt


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_pybind(force="--force" in sys.argv))
