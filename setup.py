"""Packaging: builds the native pieces in-tree (build.py) and ships both import names.

The reference drives three Bazel targets from its setup.py (controller.so, fhe.so, generated pb2
files; ``setup.py:21-45``); here `python build.py` does the whole native build with nvcc / g++."""
import os
import subprocess
import sys

from setuptools import find_packages, setup
from setuptools.command.build_py import build_py


class BuildNative(build_py):
    def run(self):
        subprocess.check_call([sys.executable, os.path.join(os.path.dirname(os.path.abspath(__file__)), "build.py")])
        super().run()


setup(
    name="metisfl_b200",
    version="0.1.0",
    description="Blackwell-native federated learning framework with MetisFL's API and model format",
    packages=find_packages(include=["metisfl_b200", "metisfl_b200.*", "metisfl"]),
    package_data={"metisfl_b200": ["lib/*.so", "_core*.so", "bin/*", "resources/**/*"]},
    python_requires=">=3.10",
    install_requires=["torch>=2.4", "numpy", "protobuf>=4", "grpcio", "PyYAML", "cloudpickle", "cryptography"],
    extras_require={"ssh": ["fabric"], "metrics": ["prometheus_client"]},
    cmdclass={"build_py": BuildNative},
)
