"""The public C header must declare every ``extern "C"`` entry point of the native library with the same parameter
types, in the same order -- and the ctypes signatures in ``scaletorch_b200/_native.py`` must have the same arity.

Pure text checks: they run on the CPU-only build box and catch the classic drift (a parameter added to a kernel
launcher, the header or the ctypes table forgotten), which otherwise shows up as stack garbage on the GPU.
"""
from __future__ import annotations

import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "scaletorch_b200.h"
SOURCES = (sorted((ROOT / "csrc").glob("*.cu")) + sorted((ROOT / "csrc" / "common").glob("*.cu"))
           + sorted((ROOT / "csrc" / "runtime").glob("*.cpp")))

_DEF = re.compile(r'extern\s+"C"\s+([A-Za-z_][\w\s\*]*?)\b(stb_\w+)\s*\(([^)]*)\)\s*\{', re.S)
_BLOCK = re.compile(r'extern\s+"C"\s*\{(.*?)\n\}\s*//\s*extern "C"', re.S)     # `extern "C" { ... }  // extern "C"`
_BLOCK_DEF = re.compile(r'^([A-Za-z_][\w \*]*?)\b(stb_\w+)\s*\(([^)]*)\)\s*\{', re.S | re.M)
_DECL = re.compile(r'^\s*([A-Za-z_][\w\s\*]*?)\b(stb_\w+)\s*\(([^)]*)\)\s*;', re.S | re.M)


def _strip_comments(text: str) -> str:
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    return re.sub(r"//[^\n]*", " ", text)


def _types(params: str) -> list[str]:
    """Parameter list -> normalised type strings (names dropped)."""
    params = " ".join(params.split())
    if params in ("", "void"):
        return []
    out = []
    for p in params.split(","):
        p = p.strip()
        m = re.match(r"^(.*?)(\b[A-Za-z_]\w*)$", p)          # trailing identifier = the parameter name
        ty = m.group(1) if m and m.group(1).strip() else p
        ty = ty.replace(" *", "*").replace("* ", "*").strip()
        ty = re.sub(r"\s+", " ", ty)
        for a, b in (("uint64_t", "unsigned long long"), ("int64_t", "long long"), ("uint32_t", "unsigned"),
                     ("int32_t", "int"), ("unsigned int", "unsigned")):
            ty = re.sub(rf"\b{a}\b", b, ty)
        out.append(ty)
    return out


def _definitions() -> dict[str, tuple[str, list[str]]]:
    found = {}
    for src in SOURCES:
        raw = src.read_text()
        for m in _DEF.finditer(_strip_comments(raw)):
            found[m.group(2)] = (src.name, _types(m.group(3)))
        for blk in _BLOCK.finditer(raw):                       # matched before comment stripping: the closing marker
            for m in _BLOCK_DEF.finditer(_strip_comments(blk.group(1))):
                found[m.group(2)] = (src.name, _types(m.group(3)))
    return found


def _declarations() -> dict[str, list[str]]:
    return {m.group(2): _types(m.group(3)) for m in _DECL.finditer(_strip_comments(HEADER.read_text()))}


def test_sources_found():
    defs = _definitions()
    assert len(defs) >= 50, f"only {len(defs)} extern \"C\" definitions parsed -- the regex lost track of the sources"
    assert "stb_gemm_bf16" in defs and "stb_loader_create" in defs


def test_every_entry_point_is_declared():
    defs, decls = _definitions(), _declarations()
    missing = sorted(set(defs) - set(decls))
    assert not missing, f"not declared in include/scaletorch_b200.h: {missing}"


def test_no_stale_declarations():
    defs, decls = _definitions(), _declarations()
    stale = sorted(set(decls) - set(defs))
    assert not stale, f"declared in the header but defined nowhere under csrc/: {stale}"


@pytest.mark.parametrize("name", sorted(_definitions()))
def test_parameter_types_match(name):
    src, want = _definitions()[name]
    got = _declarations().get(name)
    if got is None:
        pytest.skip("covered by test_every_entry_point_is_declared")
    assert got == want, f"{name}: header {got} != {src} {want}"


def test_ctypes_table_arity():
    """``_native.py`` lists argtypes per entry point; a wrong count corrupts the call silently."""
    text = (ROOT / "scaletorch_b200" / "_native.py").read_text()
    defs = _definitions()
    checked = 0
    for m in re.finditer(r'"(stb_\w+)"\s*:\s*\(?\s*\[(.*?)\]', text, re.S):
        name, args = m.group(1), m.group(2)
        if name not in defs:
            continue
        n = len([a for a in re.split(r",(?![^()]*\))", args) if a.strip()])
        assert n == len(defs[name][1]), f"{name}: _native.py lists {n} argtypes, {defs[name][0]} takes {len(defs[name][1])}"
        checked += 1
    if checked == 0:
        pytest.skip("_native.py does not use a literal name -> [argtypes] table")
