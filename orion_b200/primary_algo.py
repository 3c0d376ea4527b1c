"""Sanitising wrapper around the configured algorithm.

Mirror of ``src/orion/core/worker/primary_algo.py:16-123``: every point going
in or out must lie in the space.  The reference checks point by point in Python
(``assert point in self._space`` for each, :53-54, :64-65); here a whole list is
checked by one ``ob2_contains`` launch, and ``suggest_columns`` uses the mask
the sampler already fused.
"""
from .algo.base import BaseAlgorithm


class PrimaryAlgo(BaseAlgorithm):
    """Wrap the primary algorithm; verify points against the space."""

    def __init__(self, space, algorithm_config):
        self.algorithm = None
        super(PrimaryAlgo, self).__init__(space, algorithm=algorithm_config)

    def _assert_in_space(self, points):
        assert all(self._space.contains_points(list(points)))

    def suggest(self, num=1):
        points = self.algorithm.suggest(num)
        self._assert_in_space(points)
        return points

    def suggest_columns(self, num=1):
        """Tensor fast path: the wrapped algorithm's ColumnBatch, asserted
        in-space through its fused mask (one device reduction)."""
        batch = self.algorithm.suggest_columns(num)
        assert bool(batch.mask[:batch.n].all().item())
        return batch

    def observe(self, points, results):
        self._assert_in_space(points)
        assert len(points) == len(results)
        self.algorithm.observe(points, results)

    @property
    def is_done(self):
        return self.algorithm.is_done

    def score(self, point):
        self._assert_in_space([point])
        return self.algorithm.score(point)

    def judge(self, point, measurements):
        self._assert_in_space([point])
        return self.algorithm.judge(point, measurements)

    @property
    def should_suspend(self):
        return self.algorithm.should_suspend

    @property
    def configuration(self):
        return self.algorithm.configuration

    @property
    def space(self):
        """Read-only here: redefining the property without a setter blocks the base setter."""
        return self._space
