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
    # Loaded by the reference's factory (src/orion/core/utils/__init__.py:76-81) AND by the mirror's.  The
    # name must not collide with the reference's own `random` (its setup.py:67-69); loading it defines a
    # subclass of whichever `orion.algo.base.BaseAlgorithm` is importable (orion_b200/plugin.py).
    entry_points={'OptimizationAlgorithm': ['gpurandom = orion_b200.plugin:GPURandom']},
    # torch (device memory, streams, torch.distributed) is needed at run time but not listed as a hard
    # requirement: the reference's factory loads entry points through pkg_resources, which re-resolves the
    # whole requirement tree of the advertising distribution and refuses on any pin conflict inside it.
    install_requires=['numpy', 'scipy'],
    extras_require={'gpu': ['torch']},
)
