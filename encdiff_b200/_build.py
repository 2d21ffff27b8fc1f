"""In-tree nvcc build of libencdiff_b200.so (sm_100a only).

The shared object is a plain C-ABI library (include/encdiff_b200.h); it is loaded with ctypes, not as a torch
extension, so the boundary carries no torch types.  Built files live next to the sources so they travel to the
GPU box with the repo snapshot.
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
BUILD_DIR = PKG_DIR / "_build"
LIB_PATH = PKG_DIR / "libencdiff_b200.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: encdiff_b200 has no CPU fallback and needs its CUDA library built")


def sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def _digest(paths) -> str:
    h = hashlib.sha256()
    for p in sorted(paths):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def _headers() -> list[Path]:
    """Everything a .cu may #include: *.cuh, *.inc (unet_bwd.inc is the whole backward launch sequence) and the ABI."""
    return list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.inc")) + [PKG_DIR.parent / "include" / "encdiff_b200.h"]


def _includes(src: Path, seen=None) -> list[Path]:
    """Files under csrc/ or include/ that ``src`` #includes with quotes, transitively (so a change to one .inc or .cuh
    recompiles only the translation units that see it)."""
    import re
    seen = {} if seen is None else seen
    for name in re.findall(r'^\s*#\s*include\s+"([^"]+)"', src.read_text(), flags=re.M):
        for base in (CSRC, PKG_DIR.parent / "include"):
            cand = base / name
            if cand.exists() and cand not in seen:
                seen[cand] = True
                _includes(cand, seen)
    return list(seen)


def is_fresh() -> bool:
    stamp = BUILD_DIR / "stamp"
    deps = list(CSRC.glob("*.cu")) + _headers()
    return LIB_PATH.exists() and stamp.exists() and stamp.read_text() == _digest(deps)


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every .cu under csrc/ for sm_100a and link libencdiff_b200.so.  No-op when up to date."""
    if not force and is_fresh():
        return LIB_PATH
    nvcc = _nvcc()
    BUILD_DIR.mkdir(exist_ok=True)
    srcs = sources()
    headers = _headers()

    def compile_one(src: Path):
        obj = BUILD_DIR / (src.stem + ".o")
        tag = BUILD_DIR / (src.stem + ".tag")
        dig = hashlib.sha256((_digest(_includes(src)) + _digest([src])).encode()).hexdigest()
        if not force and obj.exists() and tag.exists() and tag.read_text() == dig:
            return obj, ""
        cmd = [nvcc, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
        tag.write_text(dig)
        return obj, r.stderr

    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        results = list(ex.map(compile_one, srcs))
    log = "\n".join(s for _, s in results if s)
    (BUILD_DIR / "ptxas.log").write_text(log)
    if verbose:
        print(log, file=sys.stderr)
    objs = [str(o) for o, _ in results]
    cmd = [nvcc, "-shared", "-o", str(LIB_PATH), *objs, "-gencode", "arch=compute_100a,code=sm_100a",
           "-Xcompiler", "-fPIC"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    deps = list(CSRC.glob("*.cu")) + headers
    (BUILD_DIR / "stamp").write_text(_digest(deps))
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
