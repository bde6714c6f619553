"""In-tree build of ``libscaletorch_b200.so`` (sm_100a only).

The native library is plain CUDA C++ behind a C ABI (``include/scaletorch_b200.h``):
no torch headers, so every translation unit cross-compiles in seconds on a box
without a GPU and the resulting ``.so`` travels with the source tree.

Usage: ``python -m scaletorch_b200._build [--force] [--verbose]``
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
BUILD = ROOT / "build" / "sm100a"
LIB_DIR = Path(__file__).resolve().parent / "lib"
LIB_PATH = LIB_DIR / "libscaletorch_b200.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "--use_fast_math",
    "-Xcompiler", "-fPIC",
    "-Xcompiler", "-fvisibility=default",
    "--expt-relaxed-constexpr",
    f"-I{CSRC}", f"-I{ROOT / 'include'}",
]


def _nvcc() -> str:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not Path(cand).exists():
        raise RuntimeError("nvcc not found; set NVCC=/path/to/nvcc")
    return cand


def sources() -> list[Path]:
    """The ``.cu`` / ``.cpp`` files that make up ``libscaletorch_b200.so``."""
    return sorted(p for p in CSRC.rglob("*.cu")) + sorted(p for p in CSRC.rglob("*.cpp"))


def _digest(src: Path) -> str:
    h = hashlib.sha1()
    h.update(src.read_bytes())
    for hdr in sorted(list(CSRC.rglob("*.cuh")) + list(CSRC.rglob("*.h")) + list((ROOT / "include").glob("*.h"))):
        h.update(hdr.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def _compile_one(src: Path, verbose: bool) -> tuple[Path, str]:
    rel = src.relative_to(CSRC)
    obj = BUILD / (str(rel).replace("/", "__") + ".o")
    stamp = obj.with_suffix(".sha1")
    dig = _digest(src)
    if obj.exists() and stamp.exists() and stamp.read_text() == dig:
        return obj, ""
    cmd = [_nvcc(), *NVCC_FLAGS]
    if src.suffix == ".cpp":
        cmd += ["-x", "cu"]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += ["-c", str(src), "-o", str(obj)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    stamp.write_text(dig)
    return obj, r.stderr


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile (``nvcc … -gencode arch=compute_100a,code=sm_100a -lineinfo``) and link the library in-tree; objects are cached by content hash unless ``force``."""
    BUILD.mkdir(parents=True, exist_ok=True)
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    if force:
        for f in BUILD.glob("*"):
            f.unlink()
    srcs = sources()
    objs: list[Path] = []
    logs: list[str] = []
    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs) or 1)) as ex:
        for obj, log in ex.map(lambda s: _compile_one(s, verbose), srcs):
            objs.append(obj)
            if log:
                logs.append(log)
    newest = max((o.stat().st_mtime for o in objs), default=0)
    if force or not LIB_PATH.exists() or LIB_PATH.stat().st_mtime < newest:
        # -arch at link time too: without it nvcc adds an (empty) device-link stub for its default arch (sm_52)
        cmd = [_nvcc(), "-shared", "-arch=sm_100a", "-o", str(LIB_PATH), *map(str, objs)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    if verbose:
        print("\n".join(logs))
    return LIB_PATH


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(p)
