#!/usr/bin/env python
"""Root-level training entry point: ``torchrun … train.py <flags>`` is the same as ``torchrun … tools/train.py <flags>``.

The reference keeps a copy of its trainer at the repository root (``train.py``) next to ``tools/train.py``; here the root
script only forwards to the one implementation in ``tools/train.py`` (see its docstring for the flags, or
``docs/training_guide.md``).
"""
import os
import runpy
import sys

if __name__ == "__main__":
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, here)
    runpy.run_path(os.path.join(here, "tools", "train.py"), run_name="__main__")
