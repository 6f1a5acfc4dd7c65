"""Packaging for grace_b200 (+ the grace_dl import shim).  Unlike the reference's setup.py, which
lists only the top-level package, every sub-package is discovered, and the CUDA
library is built in-tree for sm_100a by ``python setup.py build_native`` (or ``__graft_entry__.build()``)."""
import os
import subprocess

from setuptools import Command, find_packages, setup

HERE = os.path.dirname(os.path.abspath(__file__))


class BuildNative(Command):
    description = "compile grace_b200/lib/libgrace_b200.so with nvcc for sm_100a"
    user_options = []

    def initialize_options(self):
        pass

    def finalize_options(self):
        pass

    def run(self):
        subprocess.check_call(["bash", os.path.join(HERE, "grace_b200", "csrc", "build.sh")])


setup(
    name="grace_b200",
    version="0.1.0",
    description="Blackwell-native gradient compression for data-parallel PyTorch (GRACE capabilities)",
    packages=find_packages(include=["grace_b200", "grace_b200.*", "grace_dl", "grace_dl.*"]),
    package_data={"grace_b200": ["lib/*.so", "csrc/*"]},
    python_requires=">=3.9",
    install_requires=["torch"],
    cmdclass={"build_native": BuildNative},
)
