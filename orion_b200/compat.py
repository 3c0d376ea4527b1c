"""Hook orion_b200 into an installed reference (`orion`) package.

The reference discovers algorithms as immediate subclasses of ITS
``orion.algo.base.BaseAlgorithm`` (``src/orion/core/utils/__init__.py:84-86``)
looked up by lower-cased class name, and has no ``orion.core.worker.transformer``
module (``orion.core`` is a regular package, so one cannot be contributed from
outside).  ``install()`` therefore

* takes ``GPURandom(orion.algo.base.BaseAlgorithm)`` from :mod:`orion_b200.plugin`
  (its ``suggest`` draws on the device from the reference's own ``Space``
  object) and appends it to ``OptimizationAlgorithm.types`` / ``typenames`` if
  the factory has not found it through the entry point already (the same
  registration the reference's tests use for their stub, ``tests/conftest.py:69-70``);
* aliases ``orion.core.worker.transformer`` to :mod:`orion_b200.transformer`
  in ``sys.modules``.

Nothing here is needed when orion_b200's own mirror classes are used.
"""
import sys


def install():
    """Register with an already-imported reference; returns the GPURandom
    class.  Raises ImportError when `orion` is not importable.  (With this
    distribution installed, the reference's factory finds the same class by
    itself through the ``gpurandom`` entry point, :mod:`orion_b200.plugin`.)"""
    from orion.algo.base import OptimizationAlgorithm   # the reference's, not the mirror
    from .plugin import GPURandom

    if GPURandom not in OptimizationAlgorithm.types:
        OptimizationAlgorithm.types.append(GPURandom)
        OptimizationAlgorithm.typenames.append(GPURandom.__name__.lower())
    from . import transformer
    sys.modules.setdefault('orion.core.worker.transformer', transformer)
    return GPURandom
