"""Entry-point target for the REFERENCE's algorithm factory.

``setup.py`` advertises ``gpurandom = orion_b200.plugin:GPURandom`` in the
``OptimizationAlgorithm`` group.  The reference's ``Factory`` loads that group
while it is creating ``OptimizationAlgorithm`` and then takes the immediate
subclasses of ITS ``orion.algo.base.BaseAlgorithm``
(``src/orion/core/utils/__init__.py:76-86``), looked up by lower-cased class
name (``:108-110``).  So loading the entry point has to DEFINE a subclass of
the reference's base class -- the mirror classes of :mod:`orion_b200.algo`
would be invisible to it -- under a name that does not collide with the
reference's own ``random`` (``setup.py:67-69`` of the reference).

The class is built on first attribute access (PEP 562), because at that moment
``orion.algo.base`` is only partially imported: ``BaseAlgorithm`` exists,
``OptimizationAlgorithm`` does not yet.  Nothing here touches CUDA until
``suggest`` is called.
"""


def _define():
    from orion.algo.base import BaseAlgorithm          # the reference's, not the mirror

    class GPURandom(BaseAlgorithm):
        """Random search drawn on the device from a reference ``Space``
        (drop-in for ``orion.algo.random.Random``, src/orion/algo/random.py:14-41)."""

        def __init__(self, space):
            super(GPURandom, self).__init__(space)
            self._compiled = None

        def _kernels(self):
            from .descriptor import compile_space
            if self._compiled is None or self._compiled_for is not self.space:
                self._compiled = compile_space(self.space)
                self._compiled_for = self.space
            return self._compiled

        def suggest_columns(self, num=1):
            """Tensor fast path: the columnar batch in HBM."""
            from .space import _raise_if_improbable, philox_key
            batch = self._kernels().sample(num, philox_key(None))
            _raise_if_improbable(batch, list(self.space.values()))
            return batch

        def suggest(self, num=1):
            from .space import Space as MirrorSpace
            # Random.suggest is space.sample(num) on numpy's global state (random.py:24-33): the same
            # numbers go to the kernel, so after numpy.random.seed(s) this returns what `random` returns
            cols = MirrorSpace.reference_stream_columns(self.space, num, None)
            if cols is None:                        # a prior scipy does not draw from uniforms / normals
                cols = self.suggest_columns(num).to_numpy()
            return MirrorSpace.rows_from_columns(self.space, cols, num)

        def observe(self, points, results):
            pass

    return GPURandom


def __getattr__(name):
    if name == 'GPURandom':
        klass = _define()
        globals()['GPURandom'] = klass
        return klass
    raise AttributeError("module %r has no attribute %r" % (__name__, name))
