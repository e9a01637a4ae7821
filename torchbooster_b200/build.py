"""In-tree build of the C-ABI library (``torchbooster_b200/lib/libtorchbooster_b200.so``).

``nvcc`` cross-compiles for sm_100a without a GPU.  The library links the static CUDA runtime and
has no dependency on torch: PyTorch only owns the buffers whose pointers are passed in.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent
CSRC = ROOT / "csrc"
LIB_DIR = ROOT / "lib"
LIB_PATH = LIB_DIR / "libtorchbooster_b200.so"

SOURCES = ["tb_abi.cu", "tb_lut.cu", "tb_transform.cu", "tb_resize.cu", "tb_augment.cu", "tb_image.cu", "tb_metrics.cu", "tb_lmdb.cpp", "tb_augment_host.cpp"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC,-fvisibility=hidden,-Wall",
    "-cudart", "static",
]
# translation units whose float arithmetic must not be contracted into FMAs
NO_FMAD = {"tb_lut.cu", "tb_image.cu"}
# host translation units that restate CPython float expressions: no contraction on the host either
NO_CONTRACT = {"tb_augment_host.cpp"}


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; torchbooster_b200 has no CPU fallback and cannot be built without it")


def needs_build() -> bool:
    if not LIB_PATH.exists():
        return True
    t = LIB_PATH.stat().st_mtime
    deps = list(CSRC.glob("*")) + [ROOT.parent / "include" / "torchbooster_b200.h"]
    return any(d.stat().st_mtime > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> Path:
    if not force and not needs_build():
        return LIB_PATH
    nvcc = _nvcc()
    LIB_DIR.mkdir(exist_ok=True)
    objs = []
    obj_dir = ROOT / "build"
    obj_dir.mkdir(exist_ok=True)
    for src in SOURCES:
        path = CSRC / src
        if not path.exists():
            continue
        obj = obj_dir / (path.stem + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-c", str(path), "-o", str(obj)]
        if src in NO_FMAD:
            cmd.insert(1, "-fmad=false")
        if src in NO_CONTRACT:
            cmd[1:1] = ["-Xcompiler", "-ffp-contract=off"]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
            print(" ".join(cmd), file=sys.stderr)
        subprocess.run(cmd, check=True)
        objs.append(str(obj))
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "static", "-o", str(LIB_PATH), *objs, "-Xlinker", "--no-undefined",
           "-lpthread", "-ldl", "-lrt"]
    subprocess.run(cmd, check=True)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
