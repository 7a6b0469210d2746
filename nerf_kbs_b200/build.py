"""In-tree build of libnkbs.so (all sm_100a kernels + the C ABI) with nvcc.

Incremental: every object carries a digest of its own source, of the headers that source includes (followed
transitively through ``#include "..."``) and of the flags; only stale objects are recompiled."""
from __future__ import annotations

import hashlib
import os
import re
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
INC = HERE.parent / "include"
LIB = HERE / "libnkbs.so"
OBJ = HERE / "build"
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "--extended-lambda"]
_INCLUDE = re.compile(r'^\s*#\s*include\s*"([^"]+)"', re.M)


def _nvcc() -> str:
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.sep not in c or os.path.exists(c)):
            return c
    raise RuntimeError("nvcc not found")


def _deps(src: Path, seen=None):
    """``src`` and every local header it includes, transitively."""
    seen = seen if seen is not None else {}
    if src in seen:
        return seen
    text = src.read_text()
    seen[src] = text
    for name in _INCLUDE.findall(text):
        for base in (src.parent, CSRC, INC):
            p = base / name
            if p.exists():
                _deps(p, seen)
                break
    return seen


def _digest(src: Path) -> str:
    h = hashlib.sha256()
    for p, text in sorted(_deps(src).items()):
        h.update(p.name.encode())
        h.update(text.encode())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> Path:
    srcs = sorted(CSRC.glob("*.cu"))
    OBJ.mkdir(exist_ok=True)
    nvcc = _nvcc()
    digests = {s: _digest(s) for s in srcs}

    def stale(src: Path) -> bool:
        obj, stamp = OBJ / (src.stem + ".o"), OBJ / (src.stem + ".stamp")
        return force or not obj.exists() or not stamp.exists() or stamp.read_text() != digests[src]

    todo = [s for s in srcs if stale(s)]
    # objects of sources that no longer exist must not be linked
    keep = {s.stem for s in srcs}
    for o in OBJ.glob("*.o"):
        if o.stem not in keep:
            o.unlink()

    def one(src: Path) -> None:
        obj = OBJ / (src.stem + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-I", str(CSRC), "-I", str(INC), "-c", str(src), "-o", str(obj)]
        if verbose:
            print(" ".join(cmd), file=sys.stderr)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src.name}:\n{r.stdout}\n{r.stderr}")
        if verbose and r.stderr.strip():
            print(r.stderr, file=sys.stderr)
        (OBJ / (src.stem + ".stamp")).write_text(digests[src])

    if todo:
        with ThreadPoolExecutor(max_workers=min(8, len(todo))) as ex:
            list(ex.map(one, todo))
    objs = [OBJ / (s.stem + ".o") for s in srcs]
    if todo or not LIB.exists() or any(o.stat().st_mtime > LIB.stat().st_mtime for o in objs):
        cmd = [nvcc, "-shared", "-o", str(LIB), *map(str, objs)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
