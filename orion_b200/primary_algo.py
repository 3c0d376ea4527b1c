"""Sanitising wrapper around the configured algorithm.

Mirror of ``src/orion/core/worker/primary_algo.py:16-123``: every point going
in or out must lie in the space.  The reference checks point by point in Python
(``assert point in self._space`` for each, :53-54, :64-65); here a whole list is
checked by one ``ob2_contains`` launch, and ``suggest_columns`` uses the mask
the sampler already fused.

The reference leaves the space transformation as TODOs (``primary_algo.py:13-14,
42, 55, 66``: "create appropriate transformed space; needs to know the Type
of Algorithm", "transform back to the original space").  ``requirement=...``
switches that design on: the wrapped algorithm then works in
``build_required_space(requirement, space)``, ``suggest`` hands back the
reversed points, ``observe`` / ``score`` / ``judge`` transform what they are
given.  Left at ``None`` the wrapper behaves exactly like the snapshot.
"""
from .algo.base import BaseAlgorithm


class PrimaryAlgo(BaseAlgorithm):
    """Wrap the primary algorithm; verify points against the space."""

    def __init__(self, space, algorithm_config, requirement=None, quantize='round'):
        self.algorithm = None
        self.transformed_space = None
        if requirement is not None:
            from .transformer import build_required_space
            self.transformed_space = build_required_space(requirement, space, quantize=quantize)
        # the wrapped algorithm lives in the transformed space when there is one
        super(PrimaryAlgo, self).__init__(self.transformed_space or space, algorithm=algorithm_config)
        self._space = space

    def _assert_in_space(self, points):
        assert all(self._space.contains_points(list(points)))

    def _to_algo(self, points):
        """User points -> the wrapped algorithm's space."""
        if self.transformed_space is None:
            return points
        return self.transformed_space.transform_points(list(points))

    def suggest(self, num=1):
        points = self.algorithm.suggest(num)
        if self.transformed_space is not None:
            assert all(self.transformed_space.contains_points(list(points)))
            points = self.transformed_space.reverse_points(list(points))
        self._assert_in_space(points)
        return points

    def suggest_columns(self, num=1):
        """Tensor fast path: the wrapped algorithm's ColumnBatch, asserted
        in-space through its fused mask (one device reduction).  With a
        transformed space: the rows as handed out (reverse of the algorithm's
        transformed batch, which stays reachable as ``.transformed_batch``)."""
        batch = self.algorithm.suggest_columns(num)
        assert bool(batch.mask[:batch.n].all().item())
        if self.transformed_space is not None:
            back = getattr(batch, 'original', None)
            if back is None:                               # an algorithm without the fused path
                back = self.transformed_space.reverse_columns(batch)
            back.mask, back.keep, back.transformed_batch = batch.mask, batch.keep, batch
            return back
        return batch

    def observe(self, points, results):
        self._assert_in_space(points)
        assert len(points) == len(results)
        self.algorithm.observe(self._to_algo(points), results)

    @property
    def is_done(self):
        return self.algorithm.is_done

    def score(self, point):
        self._assert_in_space([point])
        return self.algorithm.score(self._to_algo([point])[0])

    def judge(self, point, measurements):
        self._assert_in_space([point])
        return self.algorithm.judge(self._to_algo([point])[0], measurements)

    def score_points(self, points):
        """``score`` for a whole list: what ``Experiment.reserve_trial`` computes as
        ``map(score_handle, map(trial_to_tuple, new_trials))``
        (src/orion/core/worker/experiment.py:173-178 over primary_algo.py:75-84), with the
        ``point in space`` assertions of all points done by ONE ``ob2_contains`` launch
        (and one transform launch under a requirement) instead of one per point."""
        points = list(points)
        if not points:
            return []
        self._assert_in_space(points)
        inner = self._to_algo(points)
        batched = getattr(self.algorithm, 'score_points', None)
        if batched is not None:
            return list(batched(inner))
        return [self.algorithm.score(p) for p in inner]

    def judge_points(self, points, measurements):
        """``judge`` for a whole list (primary_algo.py:86-95); same single membership launch."""
        points = list(points)
        assert len(points) == len(measurements)
        if not points:
            return []
        self._assert_in_space(points)
        return [self.algorithm.judge(p, m) for p, m in zip(self._to_algo(points), measurements)]

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
