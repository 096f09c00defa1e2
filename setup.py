"""Packaging for idc_models_b200.

The native code is built IN-TREE by ``idc_models_b200/build.py`` (nvcc for ``sm_100a`` + g++; objects cached under
``build/``; the two shared libraries land in ``idc_models_b200/_native/`` and are loaded with ctypes). This file only
gives that build the usual front doors:

    python setup.py build_ext --inplace        # = python -m idc_models_b200.build
    pip install --no-build-isolation -e .      # editable install; the libraries stay in the tree

A regular (non-editable) install copies the built libraries as package data.
"""
import os
import sys

from setuptools import Command, find_packages, setup
from setuptools.command.build_py import build_py

HERE = os.path.dirname(os.path.abspath(__file__))


def _build_native(force: bool = False) -> None:
    sys.path.insert(0, HERE)
    try:
        from idc_models_b200 import build as native_build
        native_build.build_all(force=force)
    finally:
        sys.path.pop(0)


class BuildExt(Command):
    description = "compile csrc/*.cu for sm_100a and csrc/host/*.cpp into idc_models_b200/_native/"
    user_options = [("inplace", "i", "accepted for familiarity; the build is always in-tree"),
                    ("force", "f", "rebuild every object")]
    boolean_options = ["inplace", "force"]

    def initialize_options(self):
        self.inplace = False
        self.force = False

    def finalize_options(self):
        pass

    def run(self):
        _build_native(bool(self.force))


class BuildPy(build_py):
    def run(self):
        _build_native()
        super().run()


setup(
    name="idc_models_b200",
    version="0.2.0",
    description="B200-native (sm_100a) training framework for the IDC transfer-learning / federated workloads",
    packages=find_packages(include=["idc_models_b200", "idc_models_b200.*"]),
    package_data={"idc_models_b200": ["_native/*.so"]},
    python_requires=">=3.10",
    install_requires=[],          # torch and numpy are taken from the environment (no index access on the target boxes)
    cmdclass={"build_ext": BuildExt, "build_py": BuildPy},
    zip_safe=False,
)
