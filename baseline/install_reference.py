"""Puts the UNMODIFIED reference package under ``baseline/_ref`` so ``bench.py --impl reference`` can run it.

``baseline/_ref`` is git-ignored (the reference is not product source) but travels to the GPU box with the tree
snapshot, exactly like the built ``.so``.  It is therefore (re)created wherever ``/root/reference`` is visible:
by ``__graft_entry__.build()`` on the build box, and lazily by ``bench.py`` itself.

Order of attempts (the first that leaves ``baseline/_ref/scaletorch/__init__.py`` wins):
  1. ``pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target baseline/_ref <src>``
  2. the same with ``--no-deps`` (dependency resolution fails here only on ``colorama``, which the reference uses
     to colour log lines; ``baseline/colorama.py`` stands in for it on the reference arm's ``sys.path``)
  3. a plain copy of the pure-Python package directory.
The source tree may be read-only, so pip builds from a scratch copy.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import tempfile
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
REF_DIR = ROOT / "baseline" / "_ref"
REF_SRC = Path(os.environ.get("SCALETORCH_REFERENCE_SRC", "/root/reference"))
WHEELHOUSE = "/opt/wheelhouse"


def installed() -> bool:
    return (REF_DIR / "scaletorch" / "__init__.py").is_file()


def _pip(src: Path, no_deps: bool) -> tuple[bool, str]:
    cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links", WHEELHOUSE,
           "--target", str(REF_DIR), "--upgrade", "--quiet"]
    if no_deps:
        cmd.append("--no-deps")
    cmd.append(str(src))
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    except Exception as e:  # pip missing, timeout ...
        return False, f"{type(e).__name__}: {e}"
    lines = (r.stderr or r.stdout).strip().splitlines()
    return r.returncode == 0 and installed(), (lines[-1] if lines else "")


def ensure(verbose: bool = False) -> tuple[bool, str]:
    """Returns (available, how).  Never raises: a missing reference only disables the reference arm."""
    if installed():
        return True, "present"
    if not (REF_SRC / "scaletorch" / "__init__.py").is_file():
        return False, f"{REF_SRC} not visible and baseline/_ref/scaletorch absent"
    REF_DIR.mkdir(parents=True, exist_ok=True)
    how = ""
    try:
        with tempfile.TemporaryDirectory(prefix="stref_") as tmp:
            work = Path(tmp) / "reference"
            shutil.copytree(REF_SRC, work, ignore=shutil.ignore_patterns(".git", "__pycache__", "*.pyc", "docs", "docker"))
            for no_deps in (False, True):
                ok, msg = _pip(work, no_deps)
                if verbose:
                    print(f"[reference] pip install {'--no-deps ' if no_deps else ''}-> {'ok' if ok else 'failed: ' + msg}")
                if ok:
                    how = "pip --no-deps" if no_deps else "pip"
                    break
    except Exception as e:
        if verbose:
            print(f"[reference] pip path failed: {type(e).__name__}: {e}")
    if not installed():
        try:
            shutil.copytree(REF_SRC / "scaletorch", REF_DIR / "scaletorch", dirs_exist_ok=True,
                            ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
            how = "copytree"
        except Exception as e:
            return False, f"copy failed: {type(e).__name__}: {e}"
    return installed(), how


if __name__ == "__main__":
    ok, how = ensure(verbose=True)
    print(f"reference {'available' if ok else 'UNAVAILABLE'} at {REF_DIR} ({how})")
    sys.exit(0 if ok else 1)
