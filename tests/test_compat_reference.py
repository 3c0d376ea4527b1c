"""orion_b200.compat against the real reference package, where it is available
(the build container; the GPU box has no /root/reference, so this skips there).
Registration needs no GPU; sampling through it is exercised on the GPU with the
mirror classes elsewhere."""
import os
import sys

import pytest

REF = '/root/reference/src'
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout not present")


def test_install_registers_with_the_reference_factory():
    import numpy
    numpy.object = object                      # the reference predates numpy 1.24
    sys.path.insert(0, REF)
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
        import importlib
        t = importlib.import_module('orion.core.worker.transformer')      # aliased module
        assert hasattr(t, 'build_required_space')
    finally:
        sys.path.remove(REF)
        for name in [m for m in sys.modules if m == 'orion' or m.startswith('orion.')]:
            del sys.modules[name]
