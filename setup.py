"""Build hook: compile the native core in-tree (plain nvcc, sm_100a only) before the files are collected.
The source directory keeps the long repository name; it is installed as the package ``tadnn``."""
import os
import subprocess
import sys

from setuptools import setup
from setuptools.command.build_py import build_py

ROOT = os.path.dirname(os.path.abspath(__file__))
SRC = "torch-automatic-distributed-neural-network_b200"


class BuildWithNative(build_py):
    def run(self):
        if os.environ.get("TADNN_SKIP_NATIVE", "0") in ("0", "", "false"):
            subprocess.check_call([sys.executable, os.path.join(ROOT, "build_native.py")], cwd=ROOT)
        super().run()


SUBPACKAGES = ["models", "nn", "ops", "parallel", "utils"]

setup(
    packages=["tadnn"] + [f"tadnn.{s}" for s in SUBPACKAGES],
    package_dir={"tadnn": SRC, **{f"tadnn.{s}": f"{SRC}/{s}" for s in SUBPACKAGES}},
    package_data={"tadnn": ["lib/libtadnn.so"]},
    cmdclass={"build_py": BuildWithNative},
)
