"""pip install .  ->  CMake build of libpcp_b200.so + _pcp (the CMakeExtension / CMakeBuild pattern of the
reference's PCLKeypoints/setup.py:12-58, using the installed pybind11 instead of a vendored copy)."""
import os
import subprocess
import sys

from setuptools import Extension, setup
from setuptools.command.build_ext import build_ext


class CMakeExtension(Extension):
    def __init__(self, name, sourcedir=""):
        super().__init__(name, sources=[])
        self.sourcedir = os.path.abspath(sourcedir)


class CMakeBuild(build_ext):
    def build_extension(self, ext):
        extdir = os.path.abspath(os.path.dirname(self.get_ext_fullpath(ext.name)))
        cfg = "Debug" if self.debug else "Release"
        os.makedirs(self.build_temp, exist_ok=True)
        subprocess.check_call(["cmake", ext.sourcedir, f"-DCMAKE_LIBRARY_OUTPUT_DIRECTORY={extdir}",
                               f"-DPython3_EXECUTABLE={sys.executable}", f"-DCMAKE_BUILD_TYPE={cfg}"], cwd=self.build_temp)
        subprocess.check_call(["cmake", "--build", ".", "--", "-j4"], cwd=self.build_temp)


setup(
    name="point_cloud_process_b200",
    version="0.1.0",
    description="B200-native exact kNN / radius search and point-to-point ICP (drop-in for ruanyyyyyyy/point-cloud-process)",
    packages=["point_cloud_process_b200"],
    ext_modules=[CMakeExtension("point_cloud_process_b200._pcp")],
    cmdclass={"build_ext": CMakeBuild},
    zip_safe=False,
)
