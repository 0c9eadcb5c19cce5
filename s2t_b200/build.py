"""Build libs2t_b200.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

Used by ``__graft_entry__.build()`` and importable on its own: ``python -m s2t_b200.build``.
The library is compiled only for ``sm_100a`` — there is no other-arch or CPU fallback.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
CSRC = PKG_DIR / "csrc"
BUILD_DIR = PKG_DIR / "build"
LIB_PATH = PKG_DIR / "libs2t_b200.so"
STAMP_PATH = PKG_DIR / "libs2t_b200.stamp"     # digest of the sources the .so was built from (travels with it)

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; the s2t_b200 CUDA library cannot be built")


def _digest(paths) -> str:
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    for p in sorted(paths):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def _inputs():
    sources = sorted(CSRC.glob("*.cu"))
    headers = sorted(CSRC.glob("*.cuh")) + [PKG_DIR.parent / "include" / "s2t_b200.h"]
    return sources, headers


def source_digest() -> str:
    sources, headers = _inputs()
    return _digest(headers) + _digest(sources)


def is_stale() -> bool:
    """True when libs2t_b200.so is missing or was built from other sources than the ones in the tree."""
    return not (LIB_PATH.exists() and STAMP_PATH.exists() and STAMP_PATH.read_text() == source_digest())


def build(force: bool = False, verbose: bool = False) -> Path:
    sources, headers = _inputs()
    BUILD_DIR.mkdir(exist_ok=True)
    stamp = BUILD_DIR / "stamp.txt"
    hdr_digest = _digest(headers)
    nvcc = _nvcc()

    def obj_for(src: Path) -> Path:
        return BUILD_DIR / (src.stem + ".o")

    def needs(src: Path) -> bool:
        tag = BUILD_DIR / (src.stem + ".tag")
        want = hdr_digest + _digest([src])
        return force or not obj_for(src).exists() or not tag.exists() or tag.read_text() != want

    def compile_one(src: Path) -> None:
        cmd = [nvcc, *NVCC_FLAGS, "-c", str(src), "-o", str(obj_for(src))]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{res.stdout}\n{res.stderr}")
        if verbose:
            sys.stderr.write(res.stderr)
        (BUILD_DIR / (src.stem + ".tag")).write_text(hdr_digest + _digest([src]))

    todo = [s for s in sources if needs(s)]
    if todo:
        with ThreadPoolExecutor(max_workers=min(8, len(todo))) as ex:
            list(ex.map(compile_one, todo))
    if todo or not LIB_PATH.exists():
        objs = [str(obj_for(s)) for s in sources]
        cmd = [nvcc, "-shared", "-o", str(LIB_PATH), *objs, "-gencode", "arch=compute_100a,code=sm_100a",
               "-lcudart"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
        stamp.write_text(hdr_digest)
    STAMP_PATH.write_text(source_digest())
    return LIB_PATH


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(p)
