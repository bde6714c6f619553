"""``pip install .`` builds the sm_100a library in-tree first (``scaletorch_b200/lib/libscaletorch_b200.so``) so the
wheel carries it; ``python -m scaletorch_b200._build`` does the same without installing.  No torch headers are needed:
the library is plain CUDA C++ behind a C ABI (``include/scaletorch_b200.h``)."""
import os
import sys

from setuptools import setup
from setuptools.command.build_py import build_py


class BuildWithNative(build_py):
    def run(self):
        if os.environ.get("SCALETORCH_B200_SKIP_NATIVE", "0") != "1":
            sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
            try:
                from scaletorch_b200 import _build

                print("building", _build.build())
            except Exception as e:  # no nvcc on this machine: install the Python layer only
                print(f"warning: native library not built ({e}); CUDA ops will raise until it is")
        super().run()
        # the vendored model configs live at the repository root; an installed package finds them under scaletorch_b200/configs
        here = os.path.dirname(os.path.abspath(__file__))
        src = os.path.join(here, "configs")
        if os.path.isdir(src):
            import shutil

            shutil.copytree(src, os.path.join(self.build_lib, "scaletorch_b200", "configs"), dirs_exist_ok=True)


setup(cmdclass={"build_py": BuildWithNative})
