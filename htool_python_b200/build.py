"""In-tree build of the native pieces (no JIT cache: the built .so files travel to the GPU box with the repo).

  libhmb200.so        CUDA kernels + planner/packer + C ABI (include/hmb200.h), nvcc, sm_100a only
  _htool_bridge*.so   pybind11 flattening shim compiled against the reference's header-only core
                      (needs /root/reference; on the GPU box the prebuilt file is used)
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / "csrc"
REF = Path(os.environ.get("HTOOL_REFERENCE_ROOT", "/root/reference"))
GXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"

LIB_HMB200 = PKG / "libhmb200.so"
EXT_SUFFIX = sysconfig.get_config_var("EXT_SUFFIX")
BRIDGE = PKG / f"_htool_bridge{EXT_SUFFIX}"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xptxas=-v",
    "-Xcompiler", "-fPIC,-fopenmp,-O3", "-shared", "-ccbin", GXX,
]


def _newer(target: Path, sources) -> bool:
    if not target.exists():
        return False
    t = target.stat().st_mtime
    return all(Path(s).stat().st_mtime <= t for s in sources)


def openblas_dir() -> Path:
    import importlib.util

    spec = importlib.util.find_spec("cv2")
    if spec is None or spec.origin is None:
        raise RuntimeError("opencv_python_headless (bundled OpenBLAS) not found")
    return Path(spec.origin).resolve().parent.parent / "opencv_python_headless.libs"


def openblas_lib() -> Path:
    d = openblas_dir()
    libs = sorted(d.glob("libopenblasp-*.so"))
    if not libs:
        raise RuntimeError(f"no libopenblasp-*.so under {d}")
    return libs[0]


def build_cuda(force: bool = False, verbose: bool = False) -> Path:
    srcs = sorted((CSRC / "capi").glob("*.cu")) + sorted((CSRC / "capi").glob("*.cpp"))
    deps = srcs + sorted(CSRC.rglob("*.cuh")) + sorted(CSRC.rglob("*.h")) + sorted((ROOT / "include").glob("*.h"))
    if not force and _newer(LIB_HMB200, deps):
        return LIB_HMB200
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc, *NVCC_FLAGS, "-I", str(ROOT / "include"), "-I", str(CSRC), *map(str, srcs), "-o", str(LIB_HMB200), "-lcudart"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed building libhmb200.so")
    (PKG / "libhmb200.ptxas.log").write_text(res.stdout + res.stderr)
    return LIB_HMB200


def build_bridge(force: bool = False, verbose: bool = False) -> Path:
    src = CSRC / "bridge" / "htool_bridge.cpp"
    if not (REF / "lib" / "htool" / "include").exists():
        if BRIDGE.exists():
            return BRIDGE  # GPU box: the prebuilt shim travels with the repo
        raise RuntimeError(f"{BRIDGE.name} is not built and the reference headers are not available at {REF}")
    if not force and _newer(BRIDGE, [src]):
        return BRIDGE
    ob = openblas_lib()
    cmd = [
        GXX, "-O3", "-march=x86-64-v3", "-std=c++17", "-fPIC", "-shared", "-fopenmp", "-fvisibility=hidden",
        "-Wno-deprecated-declarations",
        "-I", str(REF / "lib/htool/include"), "-I", str(REF / "lib/pybind11/include"),
        "-I", sysconfig.get_paths()["include"], str(src), "-o", str(BRIDGE), str(ob), f"-Wl,-rpath,{ob.parent}",
    ]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("g++ failed building the htool bridge")
    return BRIDGE


def build_oracle(verbose: bool = False) -> None:
    """Builds the CHECKER (oracle/); building it is not using it."""
    odir = ROOT / "oracle"
    target = "all" if (REF / "lib" / "htool" / "include").exists() else "_ref/libflat_oracle.so"
    res = subprocess.run(["make", "-C", str(odir), target, f"REF={REF}"], capture_output=True, text=True)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("make failed in oracle/")


if __name__ == "__main__":
    force = "--force" in sys.argv
    what = [a for a in sys.argv[1:] if not a.startswith("-")] or ["cuda", "bridge", "oracle"]
    if "cuda" in what:
        print(build_cuda(force, verbose=True))
    if "bridge" in what:
        print(build_bridge(force, verbose=True))
    if "oracle" in what:
        build_oracle(verbose=True)
