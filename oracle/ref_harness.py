"""Import the UNMODIFIED reference (DatCorno/orion) for tests and the CPU
baseline.  TEST INFRASTRUCTURE: only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s CPU legs may use this; the product package never does.

Looks for the package in ``oracle/_ref`` (staged by ``oracle/make_ref.py``;
travels to the GPU box) and, failing that, in ``/root/reference/src`` (the build
container).  Accommodations, none of which edits a reference file:

* ``numpy.object = object`` -- ``src/orion/algo/space.py:399,422`` use the alias
  numpy removed in 1.24;
* a stub ``pymongo`` in ``sys.modules`` -- ``orion.core.worker`` imports the
  database layer at package import (``worker/__init__.py:14`` ->
  ``io/database/mongodb.py:13``); no database call is ever made on this path;
* ``Random`` appended to ``OptimizationAlgorithm.types`` / ``typenames`` by hand,
  as the reference's own ``tests/conftest.py:69-70`` does for its stub
  algorithm: entry points are only visible for an installed distribution
  (``src/orion/core/utils/__init__.py:76-86``).
"""
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
CANDIDATES = [os.path.join(HERE, '_ref'), '/root/reference/src']


def reference_path():
    for path in CANDIDATES:
        if os.path.isdir(os.path.join(path, 'orion', 'algo')):
            return path
    return None


def available():
    return reference_path() is not None


def load():
    """Make ``import orion...`` resolve to the reference; returns its path.
    Raises ImportError when no copy is present."""
    path = reference_path()
    if path is None:
        raise ImportError("no reference package: run `python oracle/make_ref.py` where /root/reference exists")
    import numpy
    if 'object' not in vars(numpy):
        numpy.object = object
    if path not in sys.path:
        sys.path.insert(0, path)
    if 'pymongo' not in sys.modules:
        try:
            import pymongo  # noqa: F401
        except ImportError:
            stub = types.ModuleType('pymongo')
            stub.MongoClient = object
            errors = types.ModuleType('pymongo.errors')
            for name in ('ConnectionFailure', 'OperationFailure', 'DuplicateKeyError', 'BulkWriteError',
                         'PyMongoError'):
                setattr(errors, name, type(name, (Exception,), {}))
            stub.errors = errors
            sys.modules['pymongo'] = stub
            sys.modules['pymongo.errors'] = errors
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        from orion.algo.base import OptimizationAlgorithm
        from orion.algo.random import Random
    if 'random' not in OptimizationAlgorithm.typenames:
        OptimizationAlgorithm.types.append(Random)
        OptimizationAlgorithm.typenames.append('random')
    return path


def unload():
    """Forget the reference modules (tests that must not leak them)."""
    for name in [m for m in sys.modules if m == 'orion' or m.startswith('orion.')]:
        del sys.modules[name]
    for path in CANDIDATES:
        while path in sys.path:
            sys.path.remove(path)


MIXED64_ARGS = (['-u%d~uniform(-1, 1)' % i for i in range(16)] +
                ['-l%d~loguniform(1e-5, 1)' % i for i in range(8)] +
                ['-g%d~normal(0, 1)' % i for i in range(8)] +
                ['-i%d~uniform(-100, 100, discrete=True)' % i for i in range(16)] +
                ['-c%d~choices([0, 1, 2, 3, 4, 5, 6, 7])' % i for i in range(16)])
REAL16_ARGS = (['-u%d~uniform(-1, 1)' % i for i in range(8)] +
               ['-l%d~loguniform(1e-5, 1)' % i for i in range(8)])


def build_space(cmd_args):
    """A reference ``Space`` built by the reference's own DSL parser
    (``SpaceBuilder().build_from``, src/orion/core/io/space_builder.py:236-261)."""
    load()
    from orion.core.io.space_builder import SpaceBuilder
    return SpaceBuilder().build_from(list(cmd_args))


def primary_algo(space, config='random'):
    """The reference's ``PrimaryAlgo(space, config)``
    (src/orion/core/worker/primary_algo.py:25-42)."""
    load()
    from orion.core.worker.primary_algo import PrimaryAlgo
    return PrimaryAlgo(space, config)
