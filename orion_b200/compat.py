"""Hook orion_b200 into an installed reference (`orion`) package.

The reference discovers algorithms as immediate subclasses of ITS
``orion.algo.base.BaseAlgorithm`` (``src/orion/core/utils/__init__.py:84-86``)
looked up by lower-cased class name, and has no ``orion.core.worker.transformer``
module (``orion.core`` is a regular package, so one cannot be contributed from
outside).  ``install()`` therefore

* defines ``GPURandom(orion.algo.base.BaseAlgorithm)`` whose ``suggest`` draws
  on the device from the reference's own ``Space`` object, and appends it to
  ``OptimizationAlgorithm.types`` / ``typenames`` (the same registration the
  reference's tests use for their stub, ``tests/conftest.py:69-70``);
* aliases ``orion.core.worker.transformer`` to :mod:`orion_b200.transformer`
  in ``sys.modules``.

Nothing here is needed when orion_b200's own mirror classes are used.
"""
import sys


def install():
    """Register with the reference; returns the GPURandom class.  Raises
    ImportError when `orion` is not importable."""
    from orion.algo.base import BaseAlgorithm, OptimizationAlgorithm   # the reference's, not the mirror
    from .descriptor import compile_space
    from .space import Space as MirrorSpace, _raise_if_improbable, philox_key

    for klass in OptimizationAlgorithm.types:
        if klass.__name__ == 'GPURandom':
            return klass

    class GPURandom(BaseAlgorithm):
        """Random search on the device over a reference `Space`."""

        def __init__(self, space):
            super(GPURandom, self).__init__(space)

        def suggest_columns(self, num=1):
            batch = compile_space(self.space).sample(num, philox_key(None))
            _raise_if_improbable(batch, list(self.space.values()))
            return batch

        def suggest(self, num=1):
            batch = self.suggest_columns(num)
            return MirrorSpace.rows_from_columns(self.space, batch.to_numpy(), num)

        def observe(self, points, results):
            pass

    OptimizationAlgorithm.types.append(GPURandom)
    OptimizationAlgorithm.typenames.append(GPURandom.__name__.lower())
    from . import transformer
    sys.modules.setdefault('orion.core.worker.transformer', transformer)
    return GPURandom
