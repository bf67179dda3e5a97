"""Packaging (parity: reference setup.py:20-50 -- a plain setuptools package; metadata lives in pyproject.toml).

The CUDA library is not a Python extension module: it is a C-ABI shared object built in-tree by
`python -m llmss_b200.ops.build` (nvcc, sm_100a) and shipped as package data.  `python setup.py build_kernels`
(or BUILD_KERNELS=1 pip install .) runs that build first.
"""
import os
import subprocess
import sys

from setuptools import Command, setup
from setuptools.command.build_py import build_py


class BuildKernels(Command):
    description = "compile csrc/*.cu into llmss_b200/_C/libllmss_b200.so (nvcc, sm_100a)"
    user_options = []

    def initialize_options(self):
        pass

    def finalize_options(self):
        pass

    def run(self):
        subprocess.check_call([sys.executable, "-m", "llmss_b200.ops.build"], cwd=os.path.dirname(os.path.abspath(__file__)))


class BuildPy(build_py):
    def run(self):
        if os.environ.get("BUILD_KERNELS") == "1":
            self.run_command("build_kernels")
        super().run()


setup(cmdclass={"build_kernels": BuildKernels, "build_py": BuildPy})
