"""The public header is usable by its consumers: it compiles as C11 and as C++17, and a program that calls an entry point
links against the in-tree library (no GPU is needed to compile or link; the program is not run)."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CUDA_INC = next((p for p in ("/usr/local/cuda/include", os.path.join(os.environ.get("CUDA_HOME", "/nonexistent"), "include"))
                 if os.path.isfile(os.path.join(p, "cuda_runtime_api.h"))), None)
LIB = os.path.join(ROOT, "scaletorch_b200", "lib", "libscaletorch_b200.so")

SRC = r"""
#include "scaletorch_b200.h"
#include <stdio.h>
int main(void) {
    /* take the address of a few entry points from different families: the declarations must be complete prototypes */
    int (*gemm)(const void*, const void*, void*, int, int, int, int, int, int, int, int, int, int, int, cudaStream_t) = stb_gemm_bf16;
    printf("abi %d %d %s\n", stb_abi_version(), gemm != 0, stb_last_error());
    return 0;
}
"""

needs_toolchain = pytest.mark.skipif(CUDA_INC is None or shutil.which("gcc") is None or shutil.which("g++") is None,
                                     reason="needs gcc / g++ and the CUDA runtime headers")


@needs_toolchain
@pytest.mark.parametrize("compiler,std,ext", [("gcc", "-std=c11", "c"), ("g++", "-std=c++17", "cpp")])
def test_header_compiles_standalone(tmp_path, compiler, std, ext):
    src = tmp_path / f"use_header.{ext}"
    src.write_text(SRC)
    cmd = [compiler, std, "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), "-I", CUDA_INC,
           "-c", str(src), "-o", str(tmp_path / "use_header.o")]
    p = subprocess.run(cmd, capture_output=True, text=True)
    assert p.returncode == 0, p.stderr[-3000:]


@needs_toolchain
def test_program_links_against_the_library(tmp_path):
    if not os.path.isfile(LIB):
        pytest.skip("library not built (python -m scaletorch_b200._build)")
    src = tmp_path / "use_header.c"
    src.write_text(SRC)
    exe = tmp_path / "use_header"
    cmd = ["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), "-I", CUDA_INC, str(src), "-o", str(exe),
           "-L", os.path.dirname(LIB), "-lscaletorch_b200", "-Wl,-rpath," + os.path.dirname(LIB), "-Wl,--allow-shlib-undefined"]
    p = subprocess.run(cmd, capture_output=True, text=True)
    assert p.returncode == 0, p.stderr[-3000:]
    # every function the header declares is exported by the library
    import re

    header = open(os.path.join(ROOT, "include", "scaletorch_b200.h")).read()
    declared = set(re.findall(r"\b(stb_[a-z0-9_]+)\s*\(", header))
    nm = subprocess.run(["nm", "-D", "--defined-only", LIB], capture_output=True, text=True).stdout
    exported = set(re.findall(r"\b(stb_[a-z0-9_]+)\b", nm))
    assert len(declared) >= 50 and declared <= exported, sorted(declared - exported)


@needs_toolchain
def test_c_example_compiles_and_links(tmp_path):
    """``examples/c_abi/gemm_rmsnorm.c`` builds with strict warnings against the header, the library and the CUDA runtime."""
    cudart = next((d for d in ("/usr/local/cuda/lib64", os.path.join(os.environ.get("CUDA_HOME", "/nonexistent"), "lib64"))
                   if os.path.isfile(os.path.join(d, "libcudart.so"))), None)
    if cudart is None or not os.path.isfile(LIB):
        pytest.skip("needs libcudart and the built library")
    exe = tmp_path / "gemm_rmsnorm"
    cmd = ["gcc", "-std=c11", "-O1", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"), "-I", CUDA_INC,
           os.path.join(ROOT, "examples", "c_abi", "gemm_rmsnorm.c"), "-L", os.path.dirname(LIB), "-lscaletorch_b200",
           "-Wl,-rpath," + os.path.dirname(LIB), "-L", cudart, "-lcudart", "-lm", "-o", str(exe)]
    p = subprocess.run(cmd, capture_output=True, text=True)
    assert p.returncode == 0, p.stderr[-3000:]
    assert exe.is_file()
