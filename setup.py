"""Builds the native sm_100a library with nvcc before packaging (`pip install .` or
`python setup.py build_py`).  The reference is pure Python under Poetry (reference
pyproject.toml:1-30) and has no build step."""
import os
import sys

from setuptools import setup
from setuptools.command.build_py import build_py

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


class BuildWithNative(build_py):
    def run(self):
        from ddl_b200.build import build_native

        build_native(verbose=True)
        super().run()


setup(cmdclass={"build_py": BuildWithNative})
