"""In-tree build: `pip install -e .` (or `python setup.py build_ext`) compiles
libob2.so with nvcc for sm_100a and registers the algorithm entry point the
reference's factory looks for (group `OptimizationAlgorithm`, reference
setup.py:63-70)."""
import os
import subprocess

from setuptools import find_packages, setup
from setuptools.command.build_py import build_py

HERE = os.path.dirname(os.path.abspath(__file__))


class BuildWithKernels(build_py):
    def run(self):
        subprocess.check_call(['make', '-C', os.path.join(HERE, 'orion_b200', 'csrc'), '-j4'])
        super().run()


setup(
    name='orion-b200',
    version='0.1.0',
    description="B200-native candidate generation for Orion (sample / transform / contains / dedup)",
    packages=find_packages(include=['orion_b200', 'orion_b200.*']),
    package_data={'orion_b200': ['libob2.so']},
    cmdclass={'build_py': BuildWithKernels},
    entry_points={'OptimizationAlgorithm': ['random = orion_b200.algo.random:Random']},
    install_requires=['numpy', 'scipy', 'torch'],
)
