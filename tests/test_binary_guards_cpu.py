"""Properties of the BUILT library that a source review cannot show and no GPU is needed to check (``cuobjdump``):

* every embedded cubin is sm_100a -- no other architecture, no PTX-only fallback that would JIT something else;
* the GEMM-shaped kernels contain tensor-core instructions of the tcgen05 generation (``UTC*MMA``), TMEM loads (``LDTM``)
  and TMA copies (``UTMALDG`` / ``UTMASTG`` / ``UTMAREDG``); no ``HMMA`` (mma.sync) tensor instruction is left anywhere;
* register / stack budgets: the tensor-core kernels stay within the budgets their role layouts were designed for and
  nothing spills more than the two attention kernels' known few bytes;
* the NVLS kernels contain ``multimem`` reductions (``LDGMC``).

``profiles/sass/INDEX.md`` is the full census (``tools/dump_sass.py``); this test keeps the headline facts true for
whatever library is about to travel to the GPU box.
"""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "scaletorch_b200", "lib", "libscaletorch_b200.so")

pytestmark = pytest.mark.skipif(shutil.which("cuobjdump") is None or not os.path.isfile(LIB),
                                reason="needs cuobjdump and the built library")


def _run(*args):
    return subprocess.run(["cuobjdump", *args, LIB], capture_output=True, text=True, check=True).stdout


@pytest.fixture(scope="module")
def res_usage():
    """{demangled-ish function name: (registers, stack bytes)} from ``cuobjdump -res-usage``."""
    out = _run("-res-usage")
    table = {}
    for m in re.finditer(r"Function (\S+):\s*\n\s*REG:(\d+) STACK:(\d+)", out):
        table[m.group(1)] = (int(m.group(2)), int(m.group(3)))
    assert len(table) >= 60, len(table)
    return table


@pytest.fixture(scope="module")
def sass_by_function():
    """{mangled function name: set of opcode mnemonics (first token of each instruction)}."""
    out = _run("-sass")
    funcs, cur = {}, None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = funcs.setdefault(m.group(1), set())
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
        if m and cur is not None:
            cur.add(m.group(1))
    assert len(funcs) >= 60, len(funcs)
    return funcs


def test_only_sm_100a_code_is_embedded():
    elfs = re.findall(r"ELF file\s+\d+: (\S+)", _run("-lelf"))
    assert len(elfs) >= 10 and all(e.endswith(".sm_100a.cubin") for e in elfs), elfs
    ptx = subprocess.run(["cuobjdump", "-lptx", LIB], capture_output=True, text=True).stdout
    assert "PTX file" not in ptx, ptx[:400]


def _match(funcs, *needles):
    hit = {k: v for k, v in funcs.items() if all(n in k for n in needles)}
    assert hit, needles
    return hit


def test_tensor_core_kernels_are_tcgen05_with_tmem_and_tma(sass_by_function):
    families = ["gemm_pair_kernel", "gemm_bf16_kernel", "grouped_gemm_kernel", "flash_attn_fwd_kernel", "attn_bwd_fused_kernel",
                "attn_bwd_dkdv_kernel", "attn_bwd_dq_kernel"]
    for fam in families:
        for name, ops in _match(sass_by_function, fam).items():
            heads = {o.split(".")[0] for o in ops}
            assert any(h.startswith("UTC") and h.endswith("MMA") for h in heads), (name, sorted(heads))
            assert "LDTM" in heads, name
            assert "UTMALDG" in heads, name
    # the epilogues that store through TMA
    for fam in ("gemm_pair_kernel", "grouped_gemm_kernel"):
        for name, ops in _match(sass_by_function, fam).items():
            assert {"UTMASTG", "UTMAREDG"} & {o.split(".")[0] for o in ops}, name
    # nothing anywhere uses the warp-level mma.sync tensor path
    for name, ops in sass_by_function.items():
        assert not any(o.split(".")[0] in ("HMMA", "IMMA", "QMMA") for o in ops), name


def test_nvls_kernels_contain_multimem_reductions(sass_by_function):
    for fam in ("allreduce_f32_kernel", "allreduce_bf16_kernel", "reduce_scatter_f32_kernel"):
        per_variant = [any(o.startswith("LDGMC") for o in ops) for ops in _match(sass_by_function, fam).values()]
        assert any(per_variant) and not all(per_variant), (fam, per_variant)     # the <true> variant has it, <false> has peer loads


def test_register_and_stack_budgets(res_usage):
    def fam(needle):
        hit = {k: v for k, v in res_usage.items() if needle in k}
        assert hit, needle
        return hit

    for name, (regs, stack) in fam("gemm_pair_kernel").items():
        assert regs <= 128 and stack == 0, (name, regs, stack)             # 256 threads x 128 regs: one CTA of the pair per SM
    for name, (regs, stack) in fam("gemm_bf16_kernel").items():
        assert regs <= 128 and stack == 0, (name, regs, stack)
    for name, (regs, stack) in fam("grouped_gemm_kernel").items():
        assert regs <= 128 and stack == 0, (name, regs, stack)
    for name, (regs, stack) in fam("flash_attn_fwd_kernel").items():
        assert regs <= 168 and stack <= 16, (name, regs, stack)            # 384 threads x 168 = 64512 of 65536 registers
    for name, (regs, stack) in fam("attn_bwd_fused_kernel").items():
        assert regs <= 128 and stack <= 64, (name, regs, stack)            # 512 threads x 128 = the whole register file
    # nothing else spills at all
    known = ("flash_attn_fwd_kernel", "attn_bwd_fused_kernel")
    spilling = {k: v for k, v in res_usage.items() if v[1] > 0 and not any(n in k for n in known)}
    assert not spilling, spilling


def test_committed_sass_index_describes_this_library(res_usage):
    """``profiles/sass/INDEX.md`` (the evidence a reviewer reads) must have been generated from the library that is in the
    tree: per kernel family the same number of instantiations with the same register and stack figures."""
    index = open(os.path.join(ROOT, "profiles", "sass", "INDEX.md")).read()
    by_family = {}
    for m in re.finditer(r"^\| ([\w:]+) \| `[^`]*` \| \d+ \| (\d+) \| (\d+) \|", index, re.M):
        by_family.setdefault(m.group(1), []).append((int(m.group(2)), int(m.group(3))))
    assert len(by_family) >= 25, sorted(by_family)
    names = list(res_usage)
    pretty = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    assert len(pretty) == len(names)
    for family, rows in by_family.items():                      # tools/dump_sass.py: a family is a substring of the demangled name
        built = sorted(res_usage[n] for n, p in zip(names, pretty) if family in p)
        assert built == sorted(rows), (family, built, sorted(rows))
