"""orion_b200 plugged into the real reference package (``oracle/_ref`` staged by
``oracle/make_ref.py``, or the checkout in the build container).  Registration
and entry-point discovery need no GPU; driving kernels from the reference's own
objects is in tests/test_gpu_reference_objects.py."""
import os
import subprocess
import sys

import pytest

from oracle import ref_harness

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.skipif(not ref_harness.available(), reason="no reference package staged (oracle/make_ref.py)")


def test_install_registers_with_the_reference_factory():
    ref_harness.load()
    try:
        from orion.algo.base import BaseAlgorithm, OptimizationAlgorithm
        from orion.algo.space import Real, Space
        from orion_b200 import compat
        klass = compat.install()
        assert issubclass(klass, BaseAlgorithm) and klass.__bases__ == (BaseAlgorithm,)
        assert 'gpurandom' in OptimizationAlgorithm.typenames
        space = Space()
        space.register(Real('x', 'uniform', 0, 1))
        algo = OptimizationAlgorithm('gpurandom', space)
        assert isinstance(algo, klass) and algo.space is space
        assert compat.install() is klass       # idempotent
        assert OptimizationAlgorithm.typenames.count('gpurandom') == 1
        import importlib
        t = importlib.import_module('orion.core.worker.transformer')      # aliased module
        assert hasattr(t, 'build_required_space')
    finally:
        ref_harness.unload()
        sys.modules.pop('orion_b200.plugin', None)      # its class is bound to the modules just unloaded


CHECK = r"""
import sys
from oracle import ref_harness
ref_harness.load()
from orion.algo.base import BaseAlgorithm, OptimizationAlgorithm
from orion.algo.space import Real, Space
assert 'gpurandom' in OptimizationAlgorithm.typenames, OptimizationAlgorithm.typenames
assert 'random' in OptimizationAlgorithm.typenames
space = Space()
space.register(Real('x', 'uniform', 0, 1))
algo = OptimizationAlgorithm('gpurandom', space)          # no compat.install(), no manual registration
assert type(algo).__module__ == 'orion_b200.plugin' and type(algo).__bases__ == (BaseAlgorithm,)
from orion.core.worker.primary_algo import PrimaryAlgo
wrapped = PrimaryAlgo(space, 'gpurandom')                  # the route Experiment takes (experiment.py:442)
assert type(wrapped.algorithm).__name__ == 'GPURandom'
# the reference's own `random` is still the reference's
assert type(OptimizationAlgorithm('random', space)).__module__ == 'orion.algo.random'
print('resolved')
"""


def test_entry_point_resolves_in_the_reference_factory_without_a_manual_call(tmp_path):
    """With this distribution's metadata visible (what `pip install -e .` leaves
    behind: an egg-info directory with entry_points.txt), the reference's
    Factory finds `gpurandom` on its own while it builds OptimizationAlgorithm
    (src/orion/core/utils/__init__.py:76-86)."""
    egg_base = str(tmp_path)
    subprocess.check_call([sys.executable, 'setup.py', '-q', 'egg_info', '--egg-base', egg_base], cwd=ROOT,
                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    listing = os.listdir(egg_base)
    assert any(name.endswith('.egg-info') for name in listing), listing
    env = dict(os.environ)
    env['PYTHONPATH'] = os.pathsep.join([egg_base, ROOT] + [p for p in env.get('PYTHONPATH', '').split(os.pathsep) if p])
    env['PYTHONWARNINGS'] = 'ignore'
    proc = subprocess.run([sys.executable, '-c', CHECK], cwd=str(tmp_path), env=env, stdout=subprocess.PIPE,
                          stderr=subprocess.STDOUT, timeout=120)
    assert proc.returncode == 0 and b'resolved' in proc.stdout, proc.stdout.decode('utf-8', 'replace')[-3000:]


def test_mirror_factory_is_not_disturbed_by_the_reference_plugin():
    """The mirror's own factory loads the same entry-point group; a plugin that
    needs the reference package must not break or pollute it."""
    from orion_b200.algo.base import OptimizationAlgorithm
    from orion_b200.space import Real, Space
    space = Space()
    space.register(Real('x', 'uniform', 0, 1))
    OptimizationAlgorithm._discover()
    assert 'random' in OptimizationAlgorithm.typenames
    assert type(OptimizationAlgorithm('random', space)).__module__ == 'orion_b200.algo.random'
